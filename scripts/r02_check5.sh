#!/bin/bash
# fp32-tolerance mode on the tensor cores ("dtype" 2 / --dtype f32x3): parity cases, full GPU suite, bench lines
TAG=${1:-r02e}
O=gpurun_out; mkdir -p $O
( time timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edge.py -q -rs -k "f32x3" ) > $O/${TAG}_pytest_f32x3.log 2>&1; tail -15 $O/${TAG}_pytest_f32x3.log
( timeout 300 python bench.py --dtype f32x3 --cells 65536 --steps 10 --warmup 3 --no-cpu-baseline ) > $O/${TAG}_bench_f32x3.log 2>&1; tail -c 1500 $O/${TAG}_bench_f32x3.log
( time timeout 1500 python -m pytest tests -m gpu -q -rs ) > $O/${TAG}_pytest.log 2>&1; tail -8 $O/${TAG}_pytest.log
( timeout 300 python bench.py --steps 50 --warmup 5 --no-cpu-baseline ) > $O/${TAG}_bench_default.log 2>&1; tail -c 600 $O/${TAG}_bench_default.log
