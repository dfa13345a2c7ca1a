#!/bin/bash
# Starts the worker app and, a moment later, the server app (producer + parameter server), both logging
# to ./logs-worker.csv / ./logs-server.csv — the two-process deployment of the reference's run.sh
# (worker first, server with `-l -p 200`).  Ctrl-C stops both.
#
#   ./run.sh                      two processes over the local rendezvous (CPU oracle math, host event loop)
#   BACKEND=cuda ./run.sh         two processes sharing device memory (CUDA IPC): rows through the INPUT_DATA topic,
#                                 weights / gradients through the server's peer-memory block inside the sm_100a kernels
#   GPUS=8 WORKERS_PER_GPU=1 ./run.sh   fused engine: one process per GPU under torchrun
trap killgroup SIGINT
killgroup() {
  echo killing...
  kill 0
}
export PYTHONPATH="$(cd "$(dirname "$0")" && pwd):$PYTHONPATH"
BACKEND=${BACKEND:-numpy}
TRAIN=${TRAIN:-./data/train.csv}
TEST=${TEST:-./data/test.csv}

runWorker() {
  python -m psb200.apps.worker_app_runner -l -test "$TEST" --backend "$BACKEND"
}
runServer() {
  sleep 1s
  python -m psb200.apps.server_app_runner -l -p 200 -training "$TRAIN" -test "$TEST" --backend "$BACKEND"
}
runFused() {
  python -m torch.distributed.run --nnodes=1 --nproc-per-node "$GPUS" --master-addr 127.0.0.1 --master-port 29092 \
    -m psb200.apps.node -l -p 200 -training "$TRAIN" -test "$TEST" --workers_per_gpu "${WORKERS_PER_GPU:-1}"
}
if [ -n "$GPUS" ]; then runFused & wait; else runWorker & runServer & wait; fi
