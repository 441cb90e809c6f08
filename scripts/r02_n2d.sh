#!/bin/bash
# 2-GPU box: the data-parallel tests on the side-stream build (bf16 small + cfg-4 widths, fp32 wide, C++ main on two GPUs)
TAG=${1:-r02n2d}
O=gpurun_out; mkdir -p $O
( time timeout 150 python -m pytest tests/test_gpu_multi.py tests/test_gpu_host_main.py -q -rs -x -k "1-small-1 or 1-wide-1 or 0-wide-1 or two_gpus" ) > $O/${TAG}_pytest.log 2>&1; tail -6 $O/${TAG}_pytest.log
