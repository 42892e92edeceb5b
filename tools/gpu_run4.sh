#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
grep -q "pytest exit 0" gpurun_out/pytest_gpu.log || exit 1
DH_TRACE=1 timeout 600 python tools/pipe_probe.py 50e6 > gpurun_out/probe.log 2>&1; grep -v dh_trace gpurun_out/probe.log | tail -7
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
tail -3 gpurun_out/bench.err
