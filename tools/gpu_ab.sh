#!/bin/bash
# A/B on the GPU box: for the default library and every variants/<name>/libduckhts_cuda.so, run the inflate + scan parity
# tests and one bench (JSON line -> gpurun_out/ab_<name>.json), print one summary line each.
mkdir -p gpurun_out
cp duckhts_b200/lib/libduckhts_cuda.so /tmp/default_lib.so
run_one() {
  name=$1
  python -m pytest tests/test_bgzf_gpu.py tests/test_bam_gpu.py -m gpu -x -q 2>&1 | tail -1
  python bench.py --steps 3 --warmup 3 ${AB_BENCH_ARGS} > gpurun_out/ab_$name.json 2> gpurun_out/ab_$name.err
  python bench.py --steps 3 --warmup 3 --records 4e6 --bcf-sites 0 > gpurun_out/ab_small_$name.json 2> gpurun_out/ab_small_$name.err
  python - $name <<'PY'
import json, sys
n = sys.argv[1]
try:
    k = json.loads(open(f"gpurun_out/ab_small_{n}.json").read().strip().splitlines()[-1])
    print(n, "4M records:", "value", round(k["value"], 1), {a: round(v, 2) for a, v in k["inflate"]["phase_ms"].items()})
except Exception as e:
    print(n, "small bench failed", e)
try:
    j = json.loads(open(f"gpurun_out/ab_{n}.json").read().strip().splitlines()[-1])
    print(n, "value", round(j["value"], 1), "e2e", round(j["e2e"]["value"], 1), "inflate GB/s", round(j["inflate"]["value"], 1),
          {k: round(v, 1) for k, v in j["inflate"]["phase_ms"].items()}, "bcf", round(j.get("read_bcf", {}).get("value", 0)),
          {k: round(v, 1) for k, v in j["kernel_ms_per_step"].items()})
except Exception as e:
    print(n, "failed", e); print(open(f"gpurun_out/ab_{n}.err").read()[-1500:])
PY
}
[ -z "$AB_SKIP_DEFAULT" ] && run_one default
for d in variants/*/; do
  [ -f "$d/libduckhts_cuda.so" ] || continue
  cp "$d/libduckhts_cuda.so" duckhts_b200/lib/libduckhts_cuda.so
  run_one "$(basename $d)"
done
cp /tmp/default_lib.so duckhts_b200/lib/libduckhts_cuda.so
