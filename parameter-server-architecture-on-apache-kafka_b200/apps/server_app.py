"""Server app: producer + parameter server (+ in-process workers when asked).

Reference: ``ServerApp.java:18-57`` creates the three topics and wires
``GRADIENTS_TOPIC -> ServerProcessor``; ``ServerAppRunner.java:86-98`` starts the
producer thread, waits, then starts the topology.  Here "creating the topics" is
allocating the mailboxes (done by the transport) and the topology is
``psb200.engine.sim.PSSimulation`` running against the wall clock with either a
remote worker app behind a ``Link`` or inline workers (``--transport loopback``).  With
``--transport p2p`` (the default of ``--backend cuda``) the app is ``psb200.apps.p2p_apps.P2PServer``:
the topics are device memory handed to the worker app over CUDA IPC and the parameter-server
protocol runs inside the worker kernels.
"""
from __future__ import annotations

from typing import Optional

from ..config import PSConfig
from ..engine.sim import NumpyBackend, PSSimulation
from ..utils.logs import LogSink
from .base_app import BaseApp, load_csv_or_synthetic


def make_backend(cfg: PSConfig, test, name: str):
    if name == "cuda":
        from ..engine.cuda_backend import CudaBackend
        return CudaBackend(cfg, test)
    return NumpyBackend(cfg, test)


class ServerApp(BaseApp):
    application_id = "parameter-server-server"        # ServerApp.java:53-55

    def __init__(self, cfg: PSConfig, in_process: bool = False, backend: str = "numpy",
                 server_log: Optional[LogSink] = None, worker_log: Optional[LogSink] = None,
                 virtual: bool = False, start_delay_ms: float = 0.0, iter_time_ms=None, ns=None):
        super().__init__(cfg)
        self.ns = ns                                  # parsed CLI namespace (p2p transport options)
        self.in_process, self.backend_name = in_process, backend
        self.server_log, self.worker_log = server_log, worker_log
        self.virtual, self.start_delay_ms, self.iter_time_ms = virtual, start_delay_ms, iter_time_ms
        self.sim: Optional[PSSimulation] = None

    def run(self) -> None:
        cfg = self.cfg
        if cfg.transport == "p2p":                    # product path: device peer memory between the two apps
            from .p2p_apps import run_server
            run_server(self, cfg, self.ns, self.server_log, self.worker_log)
            return
        train = load_csv_or_synthetic(cfg.training_data_file_path, cfg, "train")
        test = load_csv_or_synthetic(cfg.test_data_file_path, cfg, "test")
        link = None
        if not self.in_process:
            from ..parallel.link import SERVER_RANK, ServerLink, init_link
            init_link(SERVER_RANK, cfg.rendezvous)
            link = ServerLink()
        backend = make_backend(cfg, test, self.backend_name)
        it = self.iter_time_ms or ((lambda w, c: 0.0) if not self.virtual else None)
        self.sim = PSSimulation(cfg, train, test, backend, self.server_log, self.worker_log,
                                iter_time_ms=it, virtual=self.virtual,
                                server_start_ms=self.start_delay_ms, link=link)
        self.sim.run()
