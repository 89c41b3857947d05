#!/bin/bash
# dev helper (GPU box, N GPUs): step time of reduce() per all-reduce implementation.  usage: scripts/n2_modes.sh N
N=${1:-2}
run() { tag=$1; shift; env "$@" timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $N --steps 200 --warmup 10 --skip-train --skip-e2e --skip-cpu-baseline --skip-configs --skip-baselines --skip-parity --no-clock-probe 2> gpurun_out/n${N}_$tag.err | grep '^{' > gpurun_out/n${N}_$tag.json; python - <<PY
import json
d = json.load(open("gpurun_out/n${N}_$tag.json"))
print("$tag", "N=$N", round(d["ms_per_step"] * 1e3, 1), "us |", d["allreduce"][:40], "|", {k: round(v["ms"] * 1e3, 1) for k, v in d["phases"].items()}, "| launches/step", d["gpu_launches"] / d["steps"])
PY
}
if [ -n "$EXTRA" ]; then
  run fused_pull_light PSGD_PEER_ALLREDUCE=1 PSGD_PEER_MODE=pull PSGD_PEER_LIGHT=1
  run fused_multimem_light PSGD_PEER_ALLREDUCE=1 PSGD_PEER_MODE=multimem PSGD_PEER_LIGHT=1
fi
run fused_push PSGD_PEER_ALLREDUCE=1 PSGD_PEER_MODE=push
run fused_pull PSGD_PEER_ALLREDUCE=1 PSGD_PEER_MODE=pull
run fused_multimem PSGD_PEER_ALLREDUCE=1 PSGD_PEER_MODE=multimem
if [ -z "$QUICK" ]; then
  run kernel PSGD_PEER_ALLREDUCE=1 PSGD_PEER_MODE=kernel
  run nccl PSGD_PEER_ALLREDUCE=0
fi
