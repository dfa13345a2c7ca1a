"""Optional Prometheus endpoint fed by the two log streams (``--metrics_port``).

The reference attaches Confluent monitoring interceptors to every producer and consumer
(``BaseKafkaApp.java:73-78``) and ships disabled Prometheus exporter stanzas for the brokers
(``kubernetes/kafka.yaml:223-312``): message flow is watched from outside the apps.  Here every
server / worker log row already passes through one ``LogSink``; an exporter attached to the sink
turns the rows into gauges, so what Control Center showed (how far each partition is, how fast rows
flow) is scraped from the app itself:

    psb_server_clock, psb_server_f1, psb_server_accuracy, psb_server_rows_total
    psb_worker_clock{partition}, psb_worker_loss{partition}, psb_worker_f1{partition},
    psb_worker_accuracy{partition}, psb_worker_tuples_seen{partition}, psb_worker_rows_total
    psb_clock_lag{partition}    (newest clock of any partition - this partition's clock)
"""
from __future__ import annotations

from typing import Dict, Optional


class MetricsExporter:
    def __init__(self, port: int, addr: str = "0.0.0.0"):
        try:
            from prometheus_client import CollectorRegistry, Counter, Gauge, start_http_server
        except ImportError as e:                               # pragma: no cover
            raise RuntimeError("--metrics_port needs the prometheus_client package") from e
        self.registry = CollectorRegistry()
        g = lambda name, doc, labels=(): Gauge(name, doc, labels, registry=self.registry)
        self.server_clock = g("psb_server_clock", "vector clock of the newest server log row")
        self.server_f1 = g("psb_server_f1", "weighted F1 of the global model on the test set")
        self.server_acc = g("psb_server_accuracy", "accuracy of the global model on the test set")
        self.server_rows = Counter("psb_server_rows", "server log rows written", registry=self.registry)
        self.worker_clock = g("psb_worker_clock", "vector clock of the partition's newest row", ("partition",))
        self.worker_loss = g("psb_worker_loss", "training loss of the partition's local fit", ("partition",))
        self.worker_f1 = g("psb_worker_f1", "weighted F1 of the partition's local model", ("partition",))
        self.worker_acc = g("psb_worker_accuracy", "accuracy of the partition's local model", ("partition",))
        self.worker_seen = g("psb_worker_tuples_seen", "rows the partition has received", ("partition",))
        self.clock_lag = g("psb_clock_lag", "newest clock of any partition minus this partition's", ("partition",))
        self.worker_rows = Counter("psb_worker_rows", "worker log rows written", registry=self.registry)
        self._clocks: Dict[str, int] = {}
        self.httpd, self.thread = start_http_server(port, addr=addr, registry=self.registry)
        self.port = self.httpd.server_port

    def observe(self, line: str) -> None:
        """One line of either log stream; anything that is not a row (headers, the EOF notice) is ignored."""
        parts = line.split(";")
        try:
            if len(parts) == 6 and parts[1] == "-1":
                self.server_clock.set(int(parts[2]))
                self.server_f1.set(float(parts[4]))
                self.server_acc.set(float(parts[5]))
                self.server_rows.inc()
            elif len(parts) == 7:
                p, vc = parts[1], int(parts[2])
                self.worker_clock.labels(p).set(vc)
                self.worker_loss.labels(p).set(float(parts[3]))
                self.worker_f1.labels(p).set(float(parts[4]))
                self.worker_acc.labels(p).set(float(parts[5]))
                self.worker_seen.labels(p).set(int(parts[6]))
                self.worker_rows.inc()
                self._clocks[p] = vc
                top = max(self._clocks.values())
                for q, c in self._clocks.items():
                    self.clock_lag.labels(q).set(top - c)
        except ValueError:
            pass

    def attach(self, *sinks) -> "MetricsExporter":
        for s in sinks:
            if s is not None:
                s.observers.append(self.observe)
        return self

    def close(self) -> None:
        self.httpd.shutdown()
        self.httpd.server_close()


def start_exporter(port: int, *sinks) -> Optional[MetricsExporter]:
    """``port`` <= 0: no exporter.  Otherwise serves /metrics on ``port`` and follows the given sinks."""
    if port is None or port <= 0:
        return None
    return MetricsExporter(port).attach(*sinks)


__all__ = ["MetricsExporter", "start_exporter"]
