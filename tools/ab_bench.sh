#!/bin/bash
# A/B bench lines on one box: each argument is "label:ENV=VAL,ENV=VAL..." run as `python bench.py --no-cpu-baseline`
#   gpurun -- 'bash tools/ab_bench.sh base: lanes4:B2PT_BPT_LANES=4'
mkdir -p gpurun_out
for spec in "$@"; do
  label=${spec%%:*}; envs=${spec#*:}
  (
    IFS=','; for kv in $envs; do [ -n "$kv" ] && export "$kv"; done; unset IFS
    timeout 600 python bench.py --steps ${AB_STEPS:-6} --warmup 3 --no-cpu-baseline $AB_ARGS > gpurun_out/ab_${label}.json 2> gpurun_out/ab_${label}.err
  )
  python - "$label" <<'PY'
import json, sys
l = sys.argv[1]
try:
    d = json.load(open("gpurun_out/ab_%s.json" % l))
    r = d["roofline"]
    print("%-14s value %.2f e2e %.2f  stage_ms %s  film_ok %s" % (l, d["value"], (d.get("e2e") or {}).get("value", 0), {k: round(v) for k, v in r["stage_ms_rank0"].items()}, d["film_ok"]))
except Exception as e:
    print(l, "FAILED", e, open("gpurun_out/ab_%s.err" % l).read()[-500:])
PY
done
