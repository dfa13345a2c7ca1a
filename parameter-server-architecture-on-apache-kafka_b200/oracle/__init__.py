from .lr_task import LogisticRegressionTask, csr_matrix, feature_std, loss_and_grad
from . import lbfgs

__all__ = ["LogisticRegressionTask", "csr_matrix", "feature_std", "loss_and_grad", "lbfgs"]
