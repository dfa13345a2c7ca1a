"""Transports.  The reference has exactly one (Kafka producers built by ``ProducerBuilder.build``,
``ProducerBuilder.java:15-23``); here the Runner CLIs choose with ``--transport``:

* ``p2p``      — NVLink / device peer-memory mailboxes, the product path: ``psb200.parallel.p2p`` (rendezvous, CUDA
  IPC hand-over, INPUT_DATA topic in device memory) + ``csrc/comm.cuh`` (pull / push / apply inside the kernels);
  the default of ``--backend cuda``;
* ``loopback`` — the host event loop ``psb200.engine.sim.PSSimulation`` (exact reference scheduling, virtual or wall
  clock), in one process or as two processes over the gloo control plane ``psb200.parallel.link``; the default of
  ``--backend numpy`` and the CPU oracle of the tests.

``psb200.parallel.nccl_baseline`` is the library-collective baseline that ``bench.py --transport nccl`` measures
against; it is not reachable from the Runner CLIs.
"""
RUNNER_TRANSPORTS = ("auto", "p2p", "loopback")
