from .messages import (BaseMessage, GradientMessage, KeyRange, LabeledData, LabeledDataWithAge,
                       WeightsMessage)
from .serde import deserialize, from_obj, serialize, to_obj

__all__ = ["BaseMessage", "GradientMessage", "KeyRange", "LabeledData", "LabeledDataWithAge",
           "WeightsMessage", "serialize", "deserialize", "to_obj", "from_obj"]
