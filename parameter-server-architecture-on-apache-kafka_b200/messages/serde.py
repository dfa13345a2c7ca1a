"""JSON wire codec, compatible with the reference's Jackson format.

Reference: ``JSONSerde.java:21-44`` + ``JSONSerdeCompatible.java:12-23``: every
object is a JSON map with a type tag ``_t`` in {``wMsg``, ``gMsg``, ``keyRange``,
``LabeledData``, ``LabeledDataWithAge``, ...}; ``values`` / ``inputData`` are maps
from decimal string keys to floats.

The hot path never serialises anything (raw fp32 in peer memory); this codec
exists so that dumps, checkpoints and traces can be exchanged with tools that
speak the reference format, and so the message sizes quoted in SURVEY §2.2 can
be reproduced.
"""
from __future__ import annotations

import json
from typing import Any, Dict, Union

import numpy as np

from .messages import (GradientMessage, KeyRange, LabeledData, LabeledDataWithAge,
                       WeightsMessage)

Message = Union[WeightsMessage, GradientMessage, LabeledData, LabeledDataWithAge, KeyRange]


def _values_to_map(kr: KeyRange, values: np.ndarray) -> Dict[str, float]:
    return {str(kr.start + i): float(v) for i, v in enumerate(values)}


def _map_to_values(kr: KeyRange, m: Dict[str, Any]) -> np.ndarray:
    out = np.zeros(len(kr), dtype=np.float32)
    for k, v in m.items():
        k = int(k)
        if kr.contains(k):
            out[k - kr.start] = v
    return out


def to_obj(msg: Message) -> Dict[str, Any]:
    if isinstance(msg, GradientMessage):
        return {"_t": "gMsg", "vectorClock": msg.vector_clock,
                "keyRange": to_obj(msg.key_range),
                "values": _values_to_map(msg.key_range, msg.values),
                "partitionKey": msg.partition_key}
    if isinstance(msg, WeightsMessage):
        return {"_t": "wMsg", "vectorClock": msg.vector_clock,
                "keyRange": to_obj(msg.key_range),
                "values": _values_to_map(msg.key_range, msg.values)}
    if isinstance(msg, KeyRange):
        return {"_t": "keyRange", "start": msg.start, "end": msg.end}
    if isinstance(msg, LabeledDataWithAge):
        return {"_t": "LabeledDataWithAge",
                "inputFeatures": {str(k): float(v) for k, v in msg.input_features.items()},
                "label": msg.label, "insertionID": msg.insertion_id}
    if isinstance(msg, LabeledData):
        return {"_t": "LabeledData",
                "inputData": {str(k): float(v) for k, v in msg.input_data.items()},
                "label": msg.label}
    raise TypeError(f"not a wire message: {type(msg).__name__}")


def from_obj(o: Dict[str, Any]) -> Message:
    t = o.get("_t")
    if t == "keyRange":
        return KeyRange(int(o["start"]), int(o["end"]))
    if t in ("wMsg", "gMsg"):
        kr = from_obj(o["keyRange"]) if "_t" in o["keyRange"] else KeyRange(int(o["keyRange"]["start"]), int(o["keyRange"]["end"]))
        vals = _map_to_values(kr, o["values"])
        if t == "wMsg":
            return WeightsMessage(int(o["vectorClock"]), kr, vals)
        return GradientMessage(int(o["vectorClock"]), kr, vals, int(o["partitionKey"]))
    if t == "LabeledData":
        return LabeledData({int(k): float(v) for k, v in o["inputData"].items()}, int(o["label"]))
    if t == "LabeledDataWithAge":
        return LabeledDataWithAge({int(k): float(v) for k, v in o["inputFeatures"].items()},
                                  int(o["label"]), int(o["insertionID"]))
    raise ValueError(f"unknown type tag _t={t!r}")


def serialize(msg: Message) -> bytes:
    return json.dumps(to_obj(msg), separators=(",", ":")).encode("utf-8")


def deserialize(data: bytes) -> Message:
    if data is None:
        return None
    return from_obj(json.loads(data.decode("utf-8")))
