#!/bin/bash
# One SASS excerpt per tcgen05 / TMA / cluster kernel of the in-tree library (no GPU needed): per kernel, the count of each
# Blackwell-native mnemonic (UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG/UTMASTG/UTMAREDG = TMA tensor load / store /
# reduce, UBLKCP = cp.async.bulk, SYNCS = mbarrier, FFMA2/FMUL2/FADD2 = packed fp32) and the first line of each.
LIB=${1:-mm-vae_b200/libmmvae_b200.so}
cuobjdump -sass "$LIB" | awk '
/Function : /{ fn=$3; next }
{ for (i = 1; i <= NF; i++) if ($i ~ /^(UTCHMMA|UTCBAR|LDTM|UTMALDG|UTMASTG|UTMAREDG|UBLKCP|FFMA2|FMUL2|FADD2|UTCATOMSWS|ATOMS)/) { split($i, a, /[.;]/); k = fn "\t" a[1]; n[k]++; if (!(k in first)) { line=$0; gsub(/^[ \t]+/, "", line); sub(/[ \t]*\/\* 0x[0-9a-f]+ \*\/.*$/, "", line); first[k]=line } } }
END { for (k in n) print k "\t" n[k] "\t" first[k] }' | sort | c++filt | awk -F'\t' '{ f=$1; gsub(/^void /, "", f); gsub(/mmv::/, "", f); gsub(/\(anonymous namespace\)::/, "", f); gsub(/\(CUtensorMap.*$|\(.*$/, "", f); printf "%-64s %-10s %5d   %s\n", substr(f,1,64), $2, $3, $4 }'
