from .server_app import ServerApp
from .worker_app import WorkerApp

__all__ = ["ServerApp", "WorkerApp"]
