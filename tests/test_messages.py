"""Wire codec round trips in the reference JSON format (JSONSerdeCompatible.java:12-23)."""
import json

import numpy as np

from psb200.messages import (GradientMessage, KeyRange, LabeledData, LabeledDataWithAge, WeightsMessage,
                             deserialize, serialize)
from psb200.models import LogRegLayout


def test_weights_and_gradient_roundtrip():
    vals = np.arange(10, dtype=np.float32) / 7
    w = WeightsMessage(3, KeyRange(5, 15), vals)
    o = json.loads(serialize(w))
    assert o["_t"] == "wMsg" and o["vectorClock"] == 3 and o["keyRange"] == {"_t": "keyRange", "start": 5, "end": 15}
    assert set(o["values"]) == {str(k) for k in range(5, 15)}
    back = deserialize(serialize(w))
    assert isinstance(back, WeightsMessage) and np.allclose(back.values, vals) and back.get_value(4) is None
    g = GradientMessage(7, KeyRange(0, 10), vals, partition_key=2)
    back = deserialize(serialize(g))
    assert isinstance(back, GradientMessage) and back.partition_key == 2 and back.vector_clock == 7


def test_full_model_message_size_matches_survey_estimate():
    lay = LogRegLayout()
    msg = WeightsMessage(0, KeyRange(0, lay.num_params), np.random.default_rng(0).standard_normal(lay.num_params).astype(np.float32))
    assert 100_000 < len(serialize(msg)) < 200_000          # ~110-125 KB per message


def test_labeled_data():
    d = LabeledData.from_dense(np.array([0, 0.5, 0, -1.0]), 4)
    assert d.input_data == {1: 0.5, 3: -1.0}
    assert deserialize(serialize(d)) == d
    a = LabeledDataWithAge.from_labeled(d, 9)
    assert deserialize(serialize(a)) == a


def test_layout_key_mapping():
    lay = LogRegLayout(1024, 5)
    assert lay.num_params == 6150 and lay.key_of(3, 2) == 20 and lay.intercept_key(5) == 6149
    flat = np.arange(6150, dtype=np.float32)
    W = lay.keys_to_device(flat)
    assert W.shape == (1025, 8) and W[3, 2] == 20 and W[1024, 5] == 6149 and W[:, 6:].sum() == 0
    assert np.array_equal(lay.device_to_keys(W), flat)
