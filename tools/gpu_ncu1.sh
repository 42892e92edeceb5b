#!/bin/bash
# ncu launch list of one small bench run + full capture of the two inflate kernels
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --records 4e6"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
echo "launch list exit $?"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'bgzf_decode3_kernel|bgzf_resolve3_kernel|bam_decode_kernel' -s 9 -c 3 -o gpurun_out/prof_v2 $CMD > gpurun_out/ncu2.log 2>&1
echo "full capture exit $?"
tail -3 gpurun_out/ncu2.log
ls -la gpurun_out
