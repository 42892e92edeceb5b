#!/bin/bash
# ncu launch list of one small bench run + full capture of the hot kernels (only after the plain run has exited 0)
mkdir -p gpurun_out
T=${TAG:-r2}
CMD="python bench.py --steps 1 --warmup 3 --records 4e6 --no-ext --bcf-sites 0"
$CMD > gpurun_out/${T}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_launches.csv $CMD > gpurun_out/${T}_ncu1.log 2>&1
echo "launch list exit $?"
$CMD > gpurun_out/${T}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'bgzf_decode3_kernel|bgzf_resolve|bam_decode_compact_kernel' -s 9 -c 3 -o gpurun_out/${T}_prof $CMD > gpurun_out/${T}_ncu2.log 2>&1
echo "full capture exit $?"
tail -3 gpurun_out/${T}_ncu2.log
ncu -i gpurun_out/${T}_prof.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_full_raw.csv 2>/dev/null
ls -la gpurun_out
