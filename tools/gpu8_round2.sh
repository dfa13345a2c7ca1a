#!/bin/bash
# Second 8-GPU session of a round: headline bench at 8 GPUs (end to end after the host-side work), relaxed consistency
# models with per-worker clocks (8 clocks per launch) against the sequential headline.
tag="${1:-r2_8gpu_b}"
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
out="gpurun_out/${tag}_results.jsonl"; : > "$out"
timeout 300 $TR --nproc-per-node 8 --master-port 29508 bench.py --gpus 8 --steps 200 --warmup 5 2>> "gpurun_out/${tag}.err" | grep "^{" >> "$out"
for m in -1 10; do
  timeout 300 $TR --nproc-per-node 8 --master-port 29509 bench.py --gpus 8 --steps 200 --warmup 8 --consistency_model $m --clocks_per_launch 8 2>> "gpurun_out/${tag}.err" | grep "^{" >> "$out"
done
nproc >> "gpurun_out/${tag}.err"; nvidia-smi topo -m >> "gpurun_out/${tag}.err" 2>&1
python - "$out" <<'PY'
import json, sys
for l in open(sys.argv[1]):
    d = json.loads(l)
    print(d["n_gpus"], "c=%s" % d["config"]["consistency_model"], "cpl=%s" % d["config"]["clocks_per_launch"],
          "device %.4f ms" % d["ms_per_step"], "e2e %.4f ms" % d["e2e"]["ms_per_step"], "status", d["config"]["status"])
PY
tail -30 "gpurun_out/${tag}.err" | cut -c1-200
