#!/bin/bash
# bench (all sub-benchmarks) + ncu launch list and full capture of the BCF kernels
mkdir -p gpurun_out
timeout 1200 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
CMD="python bench.py --steps 1 --warmup 3 --records 4e6 --bcf-sites 4000"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
echo "launch list exit $?"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'bcf_decode_kernel|bcf_walk_kernel' -s 4 -c 2 -o gpurun_out/prof_bcf $CMD > gpurun_out/ncu2.log 2>&1
echo "full capture exit $?"; tail -3 gpurun_out/ncu2.log
ls -la gpurun_out
