#!/bin/bash
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 > gpurun_out/bench_bam.json 2> gpurun_out/bench_bam.err; echo "bench exit $?"
cat gpurun_out/bench_bam.json; tail -5 gpurun_out/bench_bam.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_bam_ref.json 2> gpurun_out/bench_bam_ref.err; cat gpurun_out/bench_bam_ref.json
