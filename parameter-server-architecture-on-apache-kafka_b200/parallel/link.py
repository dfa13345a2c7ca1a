"""Two-process link between ServerAppRunner and WorkerAppRunner (CPU / gloo).

In the reference the two JVMs talk through three Kafka topics on a broker at
``localhost:29092`` / ``kafka:9092`` (``ServerAppRunner.java:63``,
``ServerApp.java:31-42``).  Here the same address is a gloo rendezvous: rank 0 is
the server app (producer + parameter server), rank 1 the worker app (all
``numWorkers`` partitions, like the reference's single worker JVM with four
stream threads, ``BaseKafkaApp.java:70``).  Three message kinds replace the
three topics: ROW (INPUT_DATA), WEIGHTS (WEIGHTS_TOPIC), GRADIENT (GRADIENTS_TOPIC),
plus STOP.

The exchange is a lock-step pair of broadcasts (server batch out, worker batch
back), which cannot deadlock the way two blocking point-to-point sends can.
On a GPU box this link is only the control plane of CLI-parity runs; the
throughput path never leaves the device (``psb200.engine.device_engine``, ``csrc/comm.cuh``).
"""
from __future__ import annotations

import datetime
from typing import List, Tuple

import torch.distributed as dist

ROW, WEIGHTS, GRADIENT, STOP = 1, 2, 3, 4
SERVER_RANK, WORKER_RANK = 0, 1

# message = (kind, worker, clock, ints | None, floats | None, aux)
Message = Tuple


def init_link(rank: int, address: str, timeout_s: float = 300.0) -> None:
    host, port = address.rsplit(":", 1)
    dist.init_process_group("gloo", init_method=f"tcp://{host}:{port}", rank=rank, world_size=2,
                            timeout=datetime.timedelta(seconds=timeout_s))


def close_link() -> None:
    if dist.is_initialized():
        dist.destroy_process_group()


class ServerLink:
    def __init__(self):
        self.outbox: List[Message] = []

    def send(self, kind: int, worker: int = 0, clock: int = 0, ints=None, floats=None, aux: int = 0) -> None:
        self.outbox.append((kind, worker, clock, ints, floats, aux))

    def pending_weights(self) -> bool:
        return any(m[0] in (WEIGHTS, STOP) for m in self.outbox)

    def exchange(self) -> List[Message]:
        """Ships the outbox, runs the worker app on it, returns its replies."""
        box = [self.outbox]
        self.outbox = []
        dist.broadcast_object_list(box, src=SERVER_RANK)
        back = [None]
        dist.broadcast_object_list(back, src=WORKER_RANK)
        return back[0]


class WorkerLink:
    def recv_batch(self) -> List[Message]:
        box = [None]
        dist.broadcast_object_list(box, src=SERVER_RANK)
        return box[0]

    def reply(self, messages: List[Message]) -> None:
        dist.broadcast_object_list([messages], src=WORKER_RANK)
