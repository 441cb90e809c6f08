#!/bin/bash
# K4 iteration check: parity/edge/gemm tests, smoke, short default bench, optional ncu --set full of K4 + the persistent GEMMs
# (usage: r02_k4.sh <tag> [ncu])
TAG=${1:-k4}
O=gpurun_out
mkdir -p $O
( time timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edge.py tests/test_gpu_gemm.py -q -rs ) > $O/${TAG}_pytest.log 2>&1; tail -25 $O/${TAG}_pytest.log
( time timeout 300 python __graft_entry__.py --smoke ) > $O/${TAG}_smoke.log 2>&1; tail -4 $O/${TAG}_smoke.log
( time timeout 400 python bench.py --steps 50 --warmup 5 --no-cpu-baseline ) > $O/${TAG}_bench.log 2>&1; tail -c 2500 $O/${TAG}_bench.log
if [ "$2" == "ncu" ]; then
timeout 300 python bench.py --cells 16384 --steps 3 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_plain.log 2>&1 &&
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"nb_likelihood_fast|gemm_tc_persist" -s 6 -c 4 -f -o $O/${TAG}_full \
    python bench.py --cells 16384 --steps 3 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu.log 2>&1
tail -3 $O/${TAG}_ncu.log
fi
