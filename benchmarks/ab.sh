#!/bin/bash
# A/B of the headline step under environment knobs: each argument is one "VAR=val VAR2=val" setting ("-" = defaults).
mkdir -p gpurun_out
i=0
for cfg in "$@"; do
  i=$((i+1))
  [ "$cfg" = "-" ] && cfg=""
  for rep in 1 2; do
    env $cfg timeout 300 python bench.py --only-headline --steps ${AB_STEPS:-40} --warmup 5 > gpurun_out/ab_$i.json 2> gpurun_out/ab_$i.err || tail -5 gpurun_out/ab_$i.err
    python - "$cfg" gpurun_out/ab_$i.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print(f"{sys.argv[1] or 'defaults':60s} {d['ms_per_step']:.4f} ms  {d['value']:.1f} clips/s  e2e {d['e2e']['value']}")
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
  done
done
