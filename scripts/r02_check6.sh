#!/bin/bash
# A/B of library builds (gpurun_libs/*.so) on one box, then the GPU suite + f32x3 bench with the in-tree library
TAG=${1:-r02g}; shift
O=gpurun_out; mkdir -p $O
bash scripts/r02_ab.sh ${TAG}_ab "$@" 2>&1 | tee $O/${TAG}_ab.txt
( time timeout 1500 python -m pytest tests -m gpu -q -x ) > $O/${TAG}_pytest.log 2>&1; tail -5 $O/${TAG}_pytest.log
( timeout 300 python bench.py --dtype f32x3 --cells 65536 --steps 10 --warmup 3 --no-cpu-baseline ) > $O/${TAG}_bench_f32x3.log 2>&1; grep -o '"ms_per_step": [0-9.]*' $O/${TAG}_bench_f32x3.log | head -1
