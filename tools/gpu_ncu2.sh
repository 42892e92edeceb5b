#!/bin/bash
# ncu: launch list of one small bench run, then a full capture of the inflate pair (plain run first, never a number from under ncu);
# COLS=1 adds the column kernels, OLD=1 the previous phase-2 kernel (DH_RESOLVE5=0)
mkdir -p gpurun_out
T=${TAG:-r2}
CMD="python bench.py --steps 1 --warmup 3 --records 4e6 --no-ext --bcf-sites 0"
$CMD > gpurun_out/${T}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${T}_launches.csv $CMD > gpurun_out/${T}_ncu1.log 2>&1
echo "launch list exit $?"
ncu --set full --clock-control none --import-source on -k regex:'bgzf_decode3_kernel|bgzf_resolve5' -s 9 -c 2 -o gpurun_out/${T}_prof_inf $CMD > gpurun_out/${T}_ncu3.log 2>&1
echo "inflate capture exit $?"
ncu -i gpurun_out/${T}_prof_inf.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_inf_raw.csv 2>/dev/null
if [ -n "$COLS" ]; then
  ncu --set full --clock-control none --import-source on -k regex:'bam_decode_compact_kernel|bam_walk_kernel' -s 8 -c 4 -o gpurun_out/${T}_prof_cols $CMD > gpurun_out/${T}_ncu2.log 2>&1
  echo "column kernels capture exit $?"
  ncu -i gpurun_out/${T}_prof_cols.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_cols_raw.csv 2>/dev/null
fi
if [ -n "$OLD" ]; then
  DH_RESOLVE5=0 ncu --set full --clock-control none --import-source on -k regex:'bgzf_resolve3' -s 5 -c 1 -o gpurun_out/${T}_prof_r3 $CMD > gpurun_out/${T}_ncu4.log 2>&1
  echo "resolve3 capture exit $?"
  ncu -i gpurun_out/${T}_prof_r3.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_r3_raw.csv 2>/dev/null
fi
ls -la gpurun_out | tail -12
