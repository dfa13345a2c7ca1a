"""Global configuration: what the reference hard-codes as ``public final static``.

Reference: ``BaseKafkaApp.java:25-40`` (numWorkers, topic names, default data
paths, broker address), ``LogisticRegressionTaskSpark.java:32-35`` (numFeatures,
numClasses, numMaxIter), ``ServerProcessor.java:36`` (learning rate 1/numWorkers),
``WorkerSamplingProcessor.java:22`` (inter-arrival window 500).

Everything that was a compile-time constant there is a field here, with the
reference value as default, so the reference behaviour is the zero-flag behaviour.
"""
from __future__ import annotations

import dataclasses
from dataclasses import dataclass

# --- reference constants -----------------------------------------------------
NUM_WORKERS_DEFAULT = 4            # BaseKafkaApp.java:25
NUM_FEATURES_DEFAULT = 1024        # LogisticRegressionTaskSpark.java:32
NUM_CLASSES_DEFAULT = 5            # LogisticRegressionTaskSpark.java:33 (labels 1..5 -> 6 class rows)
NUM_MAX_ITER_DEFAULT = 2           # LogisticRegressionTaskSpark.java:35
PROCESSING_INTERVAL_SIZE = 500     # WorkerSamplingProcessor.java:22
TRAINING_DATA_FILE_PATH_DEFAULT = "./data/train.csv"   # BaseKafkaApp.java:35
TEST_DATA_FILE_PATH_DEFAULT = "./data/test.csv"        # BaseKafkaApp.java:36

# The three Kafka topics and the state store of the reference map onto named
# peer-memory regions; the names are kept for log/debug output.
INPUT_DATA_TOPIC = "INPUT_DATA"            # BaseKafkaApp.java:27 -> pinned ingest ring
GRADIENTS_TOPIC = "GRADIENTS_TOPIC"        # BaseKafkaApp.java:29 -> shard inbox mailboxes
WEIGHTS_TOPIC = "WEIGHTS_TOPIC"            # BaseKafkaApp.java:28 -> replicated weight snapshots
INPUT_DATA_BUFFER = "INPUT_DATA_BUFFER"    # BaseKafkaApp.java:30 -> device row ring

# -r keeps its meaning "use the cluster address" (ServerAppRunner.java:63); the
# address is now a TCPStore rendezvous endpoint used only to swap IPC handles.
RENDEZVOUS_LOCAL = "127.0.0.1:29092"
RENDEZVOUS_REMOTE = "kafka:9092"

SERVER_LOG_HEADER = "timestamp;partition;vectorClock;loss;fMeasure;accuracy"                 # ServerAppRunner.java:81
WORKER_LOG_HEADER = "timestamp;partition;vectorClock;loss;fMeasure;accuracy;numTuplesSeen"   # WorkerAppRunner.java:80

CONSISTENCY_SEQUENTIAL = 0
CONSISTENCY_EVENTUAL = -1          # ServerProcessor.java:44 MAX_DELAY_INFINITY

CLASS_PAD = 8                      # device class stride: 6 class rows padded to one 32-byte sector


@dataclass
class PSConfig:
    """One object carrying every knob of a run (CLI flags + former constants)."""
    num_workers: int = NUM_WORKERS_DEFAULT
    num_features: int = NUM_FEATURES_DEFAULT
    num_classes: int = NUM_CLASSES_DEFAULT
    consistency_model: int = CONSISTENCY_SEQUENTIAL
    producer_time_per_event: int = 200
    min_buffer_size: int = 128
    max_buffer_size: int = 1024
    buffer_size_coefficient: float = 0.3
    training_data_file_path: str = TRAINING_DATA_FILE_PATH_DEFAULT
    test_data_file_path: str = TEST_DATA_FILE_PATH_DEFAULT
    local_solver: str = "lbfgs2"       # "lbfgs2" (reference: 2 L-BFGS its, push delta) | "sgd"
    max_iter: int = NUM_MAX_ITER_DEFAULT
    sgd_lr: float = 0.5
    standardize: bool = True           # Spark default standardization=true
    center: bool = True                # Spark centres multinomial coefs when regParam == 0
    compat_q1: bool = False            # replicate the off-by-one on key 6149
    csv_header: bool = True            # the reference always skips one line (CsvProducer.java:41-43); --no_header
    init: str = "zero"                 # reference zero-inits despite the method name (Q3)
    seed: int = 0
    transport: str = "loopback"        # loopback (host event loop) | p2p (device peer memory)
    remote: bool = False
    verbose: bool = False
    logging: bool = False
    max_clock: int = -1                # stop after this many server clocks (-1: run until producer EOF + drain)
    eval_every: int = 1
    inject_delay: str = ""             # "rank:ms[,rank:ms]" straggler injection

    @property
    def num_class_rows(self) -> int:
        """Spark needs class ids 0..numClasses, so numClasses+1 rows (Q4)."""
        return self.num_classes + 1

    @property
    def num_params(self) -> int:
        """(numClasses+1)*numFeatures + (numClasses+1) = 6150 (LogisticRegressionTaskSpark.java:101)."""
        return self.num_class_rows * self.num_features + self.num_class_rows

    @property
    def server_learning_rate(self) -> float:
        return 1.0 / self.num_workers

    @property
    def rendezvous(self) -> str:
        return RENDEZVOUS_REMOTE if self.remote else RENDEZVOUS_LOCAL

    def replace(self, **kw) -> "PSConfig":
        return dataclasses.replace(self, **kw)
