#!/bin/bash
# ncu: launch list of one small bench run, then full captures of the column kernels and of both phase-2 kernels (plain runs first)
mkdir -p gpurun_out
T=${TAG:-r2}
CMD="python bench.py --steps 1 --warmup 3 --records 4e6 --no-ext --bcf-sites 0"
$CMD > gpurun_out/${T}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_launches.csv $CMD > gpurun_out/${T}_ncu1.log 2>&1
echo "launch list exit $?"
ncu --set full --clock-control none --import-source on -k regex:'bam_decode_compact_kernel|bam_walk_kernel' -s 8 -c 4 -o gpurun_out/${T}_prof_cols $CMD > gpurun_out/${T}_ncu2.log 2>&1
echo "column kernels capture exit $?"
ncu --set full --clock-control none --import-source on -k regex:'bgzf_decode3_kernel|bgzf_resolve3' -s 9 -c 2 -o gpurun_out/${T}_prof_inf $CMD > gpurun_out/${T}_ncu3.log 2>&1
echo "inflate capture exit $?"
DH_RESOLVE5=1 $CMD > gpurun_out/${T}_plain5.log 2>&1 &&
DH_RESOLVE5=1 ncu --set full --clock-control none --import-source on -k regex:'bgzf_resolve5' -s 5 -c 1 -o gpurun_out/${T}_prof_r5 $CMD > gpurun_out/${T}_ncu4.log 2>&1
echo "resolve5 capture exit $?"
for f in cols inf r5; do ncu -i gpurun_out/${T}_prof_$f.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_${f}_raw.csv 2>/dev/null; done
ls -la gpurun_out | tail -20
