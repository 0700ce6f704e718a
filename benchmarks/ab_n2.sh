#!/bin/bash
# A/B of the 2-GPU headline step under environment knobs (one torchrun per setting).
mkdir -p gpurun_out
port=29520
for cfg in "$@"; do
  port=$((port+1))
  [ "$cfg" = "-" ] && cfg=""
  env $cfg timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node ${AB_GPUS:-2} --master-addr 127.0.0.1 --master-port $port bench.py --gpus ${AB_GPUS:-2} --only-headline --steps ${AB_STEPS:-50} --warmup 5 2>/dev/null | tail -1 > /tmp/n2.json
  python - "$cfg" <<'PY'
import json,sys
try:
    d=json.loads(open('/tmp/n2.json').read())
    print(f"{sys.argv[1] or 'defaults':60s} {d['ms_per_step']:.4f} ms  {d['value']:.1f} clips/s  e2e {d['e2e']['value']}")
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
done
