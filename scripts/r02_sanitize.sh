#!/bin/bash
# compute-sanitizer over the hand-rolled synchronisation (SURVEY section 5 / VERDICT r01 item 8): the tcgen05 / TMEM / TMA
# GEMMs (full/empty mbarrier hand-offs), the cluster kernels K4 / K5 (st.async + remote mbarrier complete_tx), the K1 panel
# kernels and the vMF-latent warp kernels.  Small cases only: the tools serialise and instrument every access.
# usage: r02_sanitize.sh <tag> [per-tool timeout seconds]
TAG=${1:-san}
TMO=${2:-240}
O=gpurun_out
mkdir -p $O
CASES_GEMM='tests/test_gpu_gemm.py -k "(test_gemm_tc and (128-128-256 or 200-136-72 or 1000-264-520)) or (persistent and 1000-20008-200 and 0.0)"'
CASES_STEP='tests/test_gpu_parity.py -k "(test_large_feature_axis_vs_oracle and bf16 and (nb-20000 or vmf-20000)) or (test_step_matches_reference and (nb_small or vmf_small))"'
CASES_LAT='tests/test_gpu_vmf_latent.py -k "test_forward_backward_vs_oracle"'
for tool in memcheck racecheck synccheck; do
  i=0
  for cases in "$CASES_GEMM" "$CASES_STEP" "$CASES_LAT"; do
    i=$((i+1))
    log=$O/${TAG}_${tool}_$i.log
    ( time eval "timeout $TMO compute-sanitizer --tool $tool --error-exitcode 9 --print-limit 20 \
        python -m pytest -q -x -p no:cacheprovider $cases" ) > $log 2>&1
    echo "== $tool set $i: exit $? ==" | tee -a $log
    grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|Error|error:" $log | head -8
  done
done
