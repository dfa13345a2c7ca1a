"""Message types of the parameter-server protocol.

Reference: ``BaseMessage.java:19-57`` {vectorClock, keyRange, values},
``WeightsMessage.java:9-15``, ``GradientMessage.java:13-26`` (+partitionKey),
``KeyRange.java:13-26``, ``LabeledData.java:14-28``, ``LabeledDataWithAge.java:17-37``.

On the device these never exist as objects: a weights message is a replica
buffer plus a version word, a gradient message is an inbox slice plus an
arrival counter.  The classes below are the host view used by the loopback
transport, checkpoints, debugging dumps and the JSON wire codec.
Values are dense ``float32`` vectors over the key range instead of boxed maps.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional

import numpy as np


@dataclass(frozen=True)
class KeyRange:
    start: int
    end: int            # exclusive (the reference worker sends an inclusive end: quirk Q1)

    def __len__(self) -> int:
        return max(0, self.end - self.start)

    def contains(self, key: int) -> bool:
        return self.start <= key < self.end


@dataclass
class BaseMessage:
    vector_clock: int
    key_range: KeyRange
    values: np.ndarray      # float32[len(key_range)]

    def get_value(self, key: int) -> Optional[float]:
        if not self.key_range.contains(key):
            return None
        return float(self.values[key - self.key_range.start])


@dataclass
class WeightsMessage(BaseMessage):
    pass


@dataclass
class GradientMessage(BaseMessage):
    partition_key: int = 0


@dataclass
class LabeledData:
    input_data: Dict[int, float]
    label: int

    @staticmethod
    def from_dense(row: np.ndarray, label: int) -> "LabeledData":
        nz = np.nonzero(row)[0]
        return LabeledData({int(i): float(row[i]) for i in nz}, int(label))


@dataclass
class LabeledDataWithAge:
    input_features: Dict[int, float]
    label: int
    insertion_id: int

    @staticmethod
    def from_labeled(d: LabeledData, insertion_id: int) -> "LabeledDataWithAge":
        return LabeledDataWithAge(d.input_data, d.label, insertion_id)
