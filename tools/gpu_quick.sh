#!/bin/bash
# parity tests of the inflate + scan, then one bench run (short summary on stdout, JSON line in gpurun_out/bench_quick.json)
mkdir -p gpurun_out
python -m pytest tests/test_bgzf_gpu.py tests/test_bam_gpu.py -m gpu -x -q 2>&1 | tail -5
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench exit $?"
python - <<'PY'
import json
try:
    j=json.loads(open("gpurun_out/bench_quick.json").read().strip().splitlines()[-1])
    print("value", round(j["value"],1), "e2e", round(j["e2e"]["value"],1), "inflate GB/s", round(j["inflate"]["value"],1), j["inflate"]["phase_ms"], "bcf", round(j["read_bcf"]["value"]), j["kernel_ms_per_step"], "frac", j["roofline"]["frac"])
except Exception as e:
    print("failed", e); print(open("gpurun_out/bench_quick.err").read()[-2000:])
PY
