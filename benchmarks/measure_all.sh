#!/usr/bin/env bash
# One-shot measurement plan for an N-GPU box (N = number of visible GPUs): headline bench, split-backward schedules,
# fine partition, interleaved variants; results land in gpurun_out/measure_all/*.json.  Usage (on the GPU box):
#   bash benchmarks/measure_all.sh [steps] [warmup]
set -u
cd "$(dirname "$0")/.."
N=$(python -c "import torch; print(torch.cuda.device_count())")
STEPS=${1:-4}; WARM=${2:-3}
OUT=gpurun_out/measure_all; mkdir -p "$OUT"
run() {  # name, extra args...
  local name=$1; shift
  if [ "$N" -gt 1 ]; then
    timeout -s KILL 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 \
      --master-port $((29500 + RANDOM % 400)) bench.py --gpus "$N" --steps "$STEPS" --warmup "$WARM" --no-e2e "$@" 2>&1 | grep '^{' | tail -1 > "$OUT/$name.json"
  else
    timeout -s KILL 400 python bench.py --gpus 1 --steps "$STEPS" --warmup "$WARM" --no-e2e "$@" 2>&1 | grep '^{' | tail -1 > "$OUT/$name.json"
  fi
  python - "$OUT/$name.json" "$name" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    print(f"{sys.argv[2]:34s} {d['ms_per_step']:9.2f} ms  {d['value']:11.0f} tok/s  idle(crit) {d.get('bubble_fraction_critical_rank', 0):.3f}  theory {d.get('bubble_fraction_theory', 0):.3f}")
except Exception as e:  # noqa: BLE001
    print(f"{sys.argv[2]:34s} FAILED ({e})")
PY
}
run "n${N}_1f1b"
[ "$N" -gt 1 ] || exit 0
run "n${N}_zerobubble" --schedule ZeroBubble1F1B
run "n${N}_1f1b_fine" --partition fine
run "n${N}_zerobubble_fine" --schedule ZeroBubble1F1B --partition fine
run "n${N}_interleaved_v2" --schedule Interleaved1F1B --virtual-stages 2
run "n${N}_zb_interleaved_v2" --schedule ZeroBubbleInterleaved --virtual-stages 2
run "n${N}_gpipe" --schedule GPipe
