#!/bin/bash
# 2-GPU box, third run: the multi-GPU tests on the final code (incl. the fp32-tolerance data-parallel case)
TAG=${1:-r02n2c}
O=gpurun_out; mkdir -p $O
( time timeout 1200 python -m pytest tests/test_gpu_multi.py tests/test_gpu_host_main.py -q -rs -k "data_parallel or two_gpus" ) > $O/${TAG}_pytest.log 2>&1; tail -6 $O/${TAG}_pytest.log
