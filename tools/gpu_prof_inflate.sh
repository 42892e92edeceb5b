#!/bin/bash
# ncu --set full capture (with source) of one bgzf_decode3_kernel + bgzf_resolve3_kernel launch of the reduced bench
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --records 4e6 --bcf-sites 4000"
$CMD > gpurun_out/plain3.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain3.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:'bgzf_decode3_kernel|bgzf_resolve3_kernel' -s 6 -c 2 -o gpurun_out/prof_inflate_r1h $CMD > gpurun_out/ncu3.log 2>&1
echo "capture exit $?"; tail -3 gpurun_out/ncu3.log; ls -la gpurun_out/*.ncu-rep
