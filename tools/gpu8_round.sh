#!/bin/bash
# One 8-GPU session: scaling bench (p2p and the NCCL baseline), per-worker gates on 8 real GPUs, product-path F1
# parity per consistency model, sharded sparse path vs NCCL.  Everything lands in gpurun_out/.
# Usage (on an 8-GPU box, from the repo root):  bash tools/gpu8_round.sh [tag]
tag="${1:-r2_8gpu}"
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
out="gpurun_out/${tag}_results.jsonl"; : > "$out"
for n in 8 4; do
  timeout 300 $TR --nproc-per-node $n --master-port 2950$n bench.py --gpus $n --steps 100 --warmup 5 >> "$out" 2>> "gpurun_out/${tag}.err"
done
timeout 300 $TR --nproc-per-node 8 --master-port 29511 bench.py --gpus 8 --steps 100 --warmup 5 --transport nccl >> "$out" 2>> "gpurun_out/${tag}.err"
for m in 2 -1; do
  timeout 300 $TR --nproc-per-node 8 --master-port 29521 tools/mp_worker_clocks.py --model $m 2>> "gpurun_out/${tag}.err" | grep MPWORKERCLOCKS >> "gpurun_out/${tag}_checks.log"
done
timeout 300 $TR --nproc-per-node 8 --master-port 29531 tools/mp_check.py --model 0 --impl 1 2>> "gpurun_out/${tag}.err" | grep MPCHECK >> "gpurun_out/${tag}_checks.log"
timeout 900 python evaluation/parity_engine.py --gpus 8 > "gpurun_out/${tag}_parity.md" 2>> "gpurun_out/${tag}.err"
for path in slab direct; do
  timeout 300 $TR --nproc-per-node 8 --master-port 29541 tools/bench_sparse.py --path $path --consistency_model 0 --baseline >> "gpurun_out/${tag}_sparse.jsonl" 2>> "gpurun_out/${tag}.err"
done
timeout 300 $TR --nproc-per-node 8 --master-port 29551 tools/bench_sparse.py --path direct --consistency_model 4 >> "gpurun_out/${tag}_sparse.jsonl" 2>> "gpurun_out/${tag}.err"
cut -c1-400 "$out"; cat "gpurun_out/${tag}_checks.log" "gpurun_out/${tag}_parity.md"; cut -c1-500 "gpurun_out/${tag}_sparse.jsonl"; tail -5 "gpurun_out/${tag}.err"
