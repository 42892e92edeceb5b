#!/bin/bash
# one GPU session: parity tests, both bench arms, ncu launch list + full capture of the hot kernels
mkdir -p gpurun_out
(nproc; free -g; lscpu | grep -E "Model name|Socket|NUMA|Thread|Core"; nvidia-smi -L) > gpurun_out/box.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json
bash tools/gpu_ncu1.sh
