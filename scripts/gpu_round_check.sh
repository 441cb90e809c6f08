#!/bin/bash
# one-box check of the committed state: GPU tests, smoke, default bench, vMF bench, ncu launch list,
# --set full of the two likelihood kernels (usage: gpu_round_check.sh <tag>)
TAG=${1:-chk}
O=gpurun_out
mkdir -p $O
( time timeout 900 python -m pytest tests -m gpu -x -q ) > $O/${TAG}_pytest.log 2>&1; tail -4 $O/${TAG}_pytest.log
( time timeout 300 python __graft_entry__.py --smoke ) > $O/${TAG}_smoke.log 2>&1; tail -4 $O/${TAG}_smoke.log
( time timeout 600 python bench.py ) > $O/${TAG}_bench_default.log 2>&1; tail -c 600 $O/${TAG}_bench_default.log
( time timeout 300 python bench.py --model vmf --steps 50 --warmup 5 --no-cpu-baseline ) > $O/${TAG}_bench_vmf.log 2>&1; tail -c 300 $O/${TAG}_bench_vmf.log
timeout 300 python bench.py --cells 16384 --steps 3 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/${TAG}_ncu_launches.csv \
    python bench.py --cells 16384 --steps 3 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_launches.log 2>&1
timeout 300 python bench.py --model vmf --cells 16384 --steps 3 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_vmf_plain.log 2>&1 &&
timeout 600 ncu --set full --import-source on --clock-control none -k regex:"vmf_likelihood_fast|nb_likelihood_fast" -s 4 -c 1 -f -o $O/${TAG}_vmf_full \
    python bench.py --model vmf --cells 16384 --steps 3 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_vmf.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:"nb_likelihood_fast" -s 4 -c 1 -f -o $O/${TAG}_k4_full \
    python bench.py --cells 16384 --steps 3 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_k4.log 2>&1
( time timeout 300 python bench.py --impl reference --steps 3 --warmup 1 ) > $O/${TAG}_bench_reference.log 2>&1; tail -c 400 $O/${TAG}_bench_reference.log
ls -la $O | tail -20
