"""b200-ps: a Blackwell-native streaming parameter server.

Same capabilities as kiminh/Parameter-Server-Architecture-On-Apache-Kafka
(ServerAppRunner / WorkerAppRunner CLIs, sequential / bounded-delay / eventual
consistency, dynamic input buffer, online multinomial logistic regression with
weighted-F1 / accuracy logs) on one 8xB200 NVSwitch box: Kafka topics become
NVLink peer-memory mailboxes with in-kernel vector clocks, the Kafka producer
becomes a pinned-ring host ingester, Spark MLlib becomes hand-written sm_100a
kernels.  Import as ``psb200``.
"""
__version__ = "0.1.0"
