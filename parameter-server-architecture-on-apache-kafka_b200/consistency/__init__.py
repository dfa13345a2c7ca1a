from .tracker import ClockTracker, ConsistencyScheduler, ProtocolViolation, WorkerStatus, max_gap_bound

__all__ = ["ClockTracker", "ConsistencyScheduler", "ProtocolViolation", "WorkerStatus", "max_gap_bound"]
