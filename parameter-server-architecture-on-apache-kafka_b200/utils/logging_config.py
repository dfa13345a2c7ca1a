"""Process-wide logging defaults.

Reference: ``src/main/resources/log4j.properties:1-13`` sets the root logger to ERROR on the console and
``ServerApp.java:28`` / ``WorkerApp.java:27`` switch the ``org.*`` loggers off, so that stdout carries only
the CSV rows.  Same policy here: library loggers at ERROR (override with ``PSB_LOG_LEVEL``), NCCL quiet.
"""
import logging
import os


def configure() -> None:
    level = getattr(logging, os.environ.get("PSB_LOG_LEVEL", "ERROR").upper(), logging.ERROR)
    logging.basicConfig(level=level, format="%(asctime)s %(levelname)s %(name)s: %(message)s")
    for name in ("torch", "torch.distributed", "matplotlib"):
        logging.getLogger(name).setLevel(max(level, logging.ERROR))
    os.environ.setdefault("NCCL_DEBUG", "WARN")
