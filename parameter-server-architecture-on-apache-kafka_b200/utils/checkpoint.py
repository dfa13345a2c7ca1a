"""Checkpoint / resume of the parameter-server state.

The reference keeps server weights and the message tracker only in JVM memory
(``ServerProcessor.java:35-40``) — a restart breaks the vector-clock
asserts.  Here a checkpoint is the weights in the reference key layout
(``weights-<clock>.csv``: ``key;value`` rows, keys as in
``LogisticRegressionTaskSpark.java:98-140``) plus a small JSON with the clocks and the
ingestion cursors; ``load`` restores both.  Every rank additionally writes ``rank-<r>.pt``
(``save_rank_state``): its input rings and windows, the buffer policy state (inter-arrival
statistics), its clocks and, for the relaxed consistency models, its authoritative shard — so a
resumed run trains on the same buffers instead of refilling empty rings.
"""
from __future__ import annotations

import json
import os
from typing import Dict, Optional, Tuple

import numpy as np

from ..models.logreg import LogRegLayout


def save(directory: str, clock: int, weights: np.ndarray, layout: LogRegLayout, clocks=None,
         rows_ingested: int = 0, extra: Optional[Dict] = None) -> str:
    os.makedirs(directory, exist_ok=True)
    keys = layout.device_to_keys(weights)
    path = os.path.join(directory, f"weights-{clock}.csv")
    tmp = path + ".tmp"
    with open(tmp, "w") as f:
        f.write("key;value\n")
        for k, v in enumerate(keys):
            f.write(f"{k};{float(v)!r}\n")
    os.replace(tmp, path)
    meta = {"clock": int(clock), "clocks": [int(c) for c in (clocks or [])], "rows_ingested": int(rows_ingested),
            "num_features": layout.F, "num_class_rows": layout.R, "weights": os.path.basename(path)}
    meta.update(extra or {})
    with open(os.path.join(directory, "latest.json.tmp"), "w") as f:
        json.dump(meta, f)
    os.replace(os.path.join(directory, "latest.json.tmp"), os.path.join(directory, "latest.json"))
    return path


def load(directory: str, layout: LogRegLayout) -> Tuple[np.ndarray, Dict]:
    with open(os.path.join(directory, "latest.json")) as f:
        meta = json.load(f)
    if meta["num_features"] != layout.F or meta["num_class_rows"] != layout.R:
        raise ValueError("checkpoint was written for a different model shape")
    keys = np.loadtxt(os.path.join(directory, meta["weights"]), delimiter=";", skiprows=1, dtype=np.float64)[:, 1]
    return layout.keys_to_device(keys.astype(np.float32)), meta


def save_rank_state(directory: str, rank: int, state: Dict) -> str:
    """Atomically writes one rank's engine state (``DeviceEngine.state_dict()``)."""
    import torch
    os.makedirs(directory, exist_ok=True)
    path = os.path.join(directory, f"rank-{rank}.pt")
    torch.save(state, path + ".tmp")
    os.replace(path + ".tmp", path)
    return path


def load_rank_state(directory: str, rank: int) -> Optional[Dict]:
    """The rank's saved engine state, or None when the checkpoint holds weights only."""
    import torch
    path = os.path.join(directory, f"rank-{rank}.pt")
    if not os.path.exists(path):
        return None
    return torch.load(path, map_location="cpu", weights_only=False)
