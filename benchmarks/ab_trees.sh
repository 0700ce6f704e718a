#!/bin/bash
# The headline step of several source trees (old / bisect copies under the repo root) on the same box.
mkdir -p gpurun_out
for t in "$@"; do
  for rep in 1 2; do
    (cd $t && timeout 300 python bench.py --only-headline --steps 40 --warmup 5 2> /tmp/err.txt | tail -1 > /tmp/out.json) || tail -5 /tmp/err.txt
    python - "$t" <<'PY'
import json,sys
try:
    d=json.loads(open('/tmp/out.json').read())
    print(f"{sys.argv[1]:20s} {d['ms_per_step']:.4f} ms  {d['value']:.1f} clips/s  launches {d.get('gpu_launches_per_step')}")
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
  done
done
