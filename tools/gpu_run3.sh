#!/bin/bash
# tests + read_bam bench (inflate v2) + reference arm
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/bench_v2.json 2> gpurun_out/bench_v2.err; echo "bench exit $?"
cat gpurun_out/bench_v2.json; tail -5 gpurun_out/bench_v2.err
