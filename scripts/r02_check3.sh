#!/bin/bash
# one-box check after the second round of launch merges: GPU tests, smoke, default bench, then the K4 wait-hint A/B
TAG=${1:-r02c}
O=gpurun_out; mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -q -rs -x ) > $O/${TAG}_pytest.log 2>&1; tail -8 $O/${TAG}_pytest.log
( time timeout 400 python __graft_entry__.py --smoke ) > $O/${TAG}_smoke.log 2>&1; tail -5 $O/${TAG}_smoke.log
( time timeout 600 python bench.py --steps 50 --warmup 5 --no-cpu-baseline ) > $O/${TAG}_bench_default.log 2>&1; tail -c 1800 $O/${TAG}_bench_default.log
( time timeout 400 python bench.py --model vmf --steps 50 --warmup 5 --no-cpu-baseline ) > $O/${TAG}_bench_vmf.log 2>&1; tail -c 600 $O/${TAG}_bench_vmf.log
shift
if [ $# -gt 0 ]; then bash scripts/r02_ab.sh ${TAG}_ab "$@" 2>&1 | tee $O/${TAG}_ab.txt; fi
