#!/bin/bash
# One GPU-box pass: parity tests, headline bench, side benches, ncu launch list. Outputs under gpurun_out/.
set -u
mkdir -p gpurun_out
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cut -c1-600 gpurun_out/bench.json
timeout 300 python benchmarks/stft_sweep.py > gpurun_out/stft_sweep.jsonl 2> gpurun_out/stft_sweep.err; echo "stft rc=$?"
timeout 300 python benchmarks/vit_bench.py > gpurun_out/vit_bench.json 2> gpurun_out/vit_bench.err; echo "vit rc=$?"; cut -c1-400 gpurun_out/vit_bench.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 1 --warmup 3 --no-cpu > gpurun_out/ncu_bench.log 2>&1; echo "ncu rc=$?"
