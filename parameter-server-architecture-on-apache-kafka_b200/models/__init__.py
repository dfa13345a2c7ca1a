from .logreg import LogRegLayout

__all__ = ["LogRegLayout"]
