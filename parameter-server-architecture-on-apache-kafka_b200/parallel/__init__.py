"""Transports.  The reference has exactly one (Kafka producers built by ``ProducerBuilder.build``,
``ProducerBuilder.java:15-23``); here the choice is explicit:

* ``loopback`` — in-process event queue (``psb200.engine.sim.PSSimulation`` with inline workers);
* ``gloo``     — two-process control plane between the Runner CLIs (``psb200.parallel.link``);
* ``p2p``      — NVLink peer-memory mailboxes, the product path (``psb200.engine.device_engine``,
  ``psb200.engine.sparse_engine``; device side ``csrc/comm.cuh``, ``csrc/ps_sparse.cu``);
* ``nccl``     — library collectives, the measured baseline (``psb200.parallel.nccl_baseline``).
"""
from typing import Any

TRANSPORTS = ("loopback", "gloo", "p2p", "nccl")


def build(name: str, *args: Any, **kwargs: Any):
    """Factory in the spirit of ``ProducerBuilder.build``: returns the engine / link class for a transport."""
    if name == "loopback":
        from ..engine.sim import PSSimulation
        return PSSimulation(*args, **kwargs)
    if name == "gloo":
        from .link import ServerLink
        return ServerLink(*args, **kwargs)
    if name == "p2p":
        from ..engine.device_engine import DeviceEngine
        return DeviceEngine(*args, **kwargs)
    if name == "nccl":
        from .nccl_baseline import NcclBaselineEngine
        return NcclBaselineEngine(*args, **kwargs)
    raise ValueError(f"unknown transport {name!r}; choose one of {TRANSPORTS}")
