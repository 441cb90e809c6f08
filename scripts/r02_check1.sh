#!/bin/bash
# round-2 one-box check: GPU tests (no -x: list every failure), smoke, default bench, the other BASELINE configs
TAG=${1:-r02a}
O=gpurun_out
mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -q -rs ) > $O/${TAG}_pytest.log 2>&1; tail -15 $O/${TAG}_pytest.log
( time timeout 400 python __graft_entry__.py --smoke ) > $O/${TAG}_smoke.log 2>&1; tail -6 $O/${TAG}_smoke.log
( time timeout 600 python bench.py --steps 50 --warmup 5 ) > $O/${TAG}_bench_default.log 2>&1; tail -c 1500 $O/${TAG}_bench_default.log
( time timeout 400 python bench.py --config 5 --steps 30 --warmup 5 --cells 262144 --no-cpu-baseline ) > $O/${TAG}_bench_cfg5.log 2>&1; tail -c 600 $O/${TAG}_bench_cfg5.log
( time timeout 400 python bench.py --config 3 --steps 30 --warmup 5 --no-cpu-baseline ) > $O/${TAG}_bench_cfg3.log 2>&1; tail -c 400 $O/${TAG}_bench_cfg3.log
( time timeout 400 python bench.py --config 1 --steps 30 --warmup 5 --no-cpu-baseline ) > $O/${TAG}_bench_cfg1.log 2>&1; tail -c 400 $O/${TAG}_bench_cfg1.log
( time timeout 400 python bench.py --config 2 --steps 30 --warmup 5 --no-cpu-baseline ) > $O/${TAG}_bench_cfg2.log 2>&1; tail -c 400 $O/${TAG}_bench_cfg2.log
( time timeout 400 python bench.py --dtype f32 --cells 65536 --steps 10 --warmup 3 --no-cpu-baseline ) > $O/${TAG}_bench_f32.log 2>&1; tail -c 400 $O/${TAG}_bench_f32.log
ls -la $O | tail -20
