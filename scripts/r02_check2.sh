#!/bin/bash
# round-2 one-box check: r02_check1 (GPU tests, smoke, default bench, the other BASELINE configs, fp32 line, reference arm),
# then the ncu launch list + one --set full capture of K4 / K1 / panel kernels, then compute-sanitizer (bounded)
TAG=${1:-r02b}
O=gpurun_out
mkdir -p $O
bash scripts/r02_check1.sh $TAG
( time timeout 300 python bench.py --impl reference --steps 3 --warmup 1 ) > $O/${TAG}_bench_reference.log 2>&1; tail -c 400 $O/${TAG}_bench_reference.log
timeout 300 python bench.py --cells 16384 --steps 3 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/${TAG}_ncu_launches.csv \
    python bench.py --cells 16384 --steps 3 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_launches.log 2>&1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"nb_likelihood_fast|senc_tc_kernel|panel_build|panel_aux|senc_row_pass|nb_row_meta|colsum_bf16" -s 14 -c 9 -f -o $O/${TAG}_full \
    python bench.py --cells 16384 --steps 3 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_full.log 2>&1
tail -3 $O/${TAG}_ncu_full.log
bash scripts/r02_sanitize.sh $TAG ${2:-120}
ls -la $O | tail -40
