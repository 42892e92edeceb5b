#!/bin/bash
# first GPU session of the round: tests, bench, ncu launch list + full capture of the inflate kernel
mkdir -p gpurun_out
(nproc; free -g; lscpu | grep -E "Model name|Socket|NUMA|Thread|Core"; nvidia-smi -L; nvidia-smi topo -m 2>/dev/null | head -20) > gpurun_out/box.txt 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_inflate.json 2> gpurun_out/bench_inflate.err; echo "bench exit $?"
cat gpurun_out/bench_inflate.json; tail -3 gpurun_out/bench_inflate.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json
python bench.py --steps 2 --warmup 3 --target-gb 1 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --target-gb 1 > gpurun_out/ncu1.log 2>&1
python bench.py --steps 2 --warmup 3 --target-gb 1 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:bgzf_inflate -s 3 -c 1 -o gpurun_out/prof_inflate python bench.py --steps 2 --warmup 3 --target-gb 1 > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out
