#!/bin/bash
# device-resident scan rate against the number of SMs the inflate leaves to the column kernels (DH_INFLATE_SM_RESERVE)
mkdir -p gpurun_out
for r in 0 8 16 24 32 48; do
  DH_INFLATE_SM_RESERVE=$r python bench.py --steps 3 --warmup 3 --bcf-sites 0 > gpurun_out/reserve_$r.json 2> gpurun_out/reserve_$r.err
  python - $r <<'PY'
import json, sys
r = sys.argv[1]
try:
    j = json.loads(open(f"gpurun_out/reserve_{r}.json").read().strip().splitlines()[-1])
    print("reserve", r, "value", round(j["value"], 1), "e2e", round(j["e2e"]["value"], 1), {k: round(v, 1) for k, v in j["kernel_ms_per_step"].items()})
except Exception as e:
    print("reserve", r, "failed", e)
PY
done
