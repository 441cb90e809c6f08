#!/bin/bash
# committed-state check on one box: GPU tests, smoke, default bench (with CPU baseline), reference arm, vMF and cfg-5 lines,
# then the ncu launch list and one --set full capture of the kernels bench.py reports rooflines for
TAG=${1:-r02z}
O=gpurun_out; mkdir -p $O
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > $O/${TAG}_pytest.log 2>&1; tail -5 $O/${TAG}_pytest.log
( time timeout 400 python -c "import __graft_entry__ as g; g.smoke()" ) > $O/${TAG}_smoke.log 2>&1; tail -6 $O/${TAG}_smoke.log
( time timeout 900 python bench.py ) > $O/${TAG}_bench_default.log 2>&1; tail -c 700 $O/${TAG}_bench_default.log
( time timeout 400 python bench.py --impl reference --steps 3 --warmup 1 ) > $O/${TAG}_bench_reference.log 2>&1; tail -c 400 $O/${TAG}_bench_reference.log
( timeout 400 python bench.py --model vmf --steps 50 --warmup 5 --no-cpu-baseline ) > $O/${TAG}_bench_vmf.log 2>&1; tail -c 300 $O/${TAG}_bench_vmf.log
( timeout 400 python bench.py --config 5 --steps 30 --warmup 5 --cells 262144 --no-cpu-baseline ) > $O/${TAG}_bench_cfg5.log 2>&1; tail -c 300 $O/${TAG}_bench_cfg5.log
timeout 300 python bench.py --cells 16384 --steps 3 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/${TAG}_ncu_launches.csv \
    python bench.py --cells 16384 --steps 3 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_launches.log 2>&1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"nb_likelihood_fast|senc_tc_kernel|panel_build|panel_aux|senc_row_pass|gemm_tc_persist|adam_kernel" -s 16 -c 10 -f -o $O/${TAG}_full \
    python bench.py --cells 16384 --steps 3 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_full.log 2>&1
tail -2 $O/${TAG}_ncu_full.log
ls -la $O | grep ${TAG}
