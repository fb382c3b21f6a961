"""Mirror of ``nichecompass.train`` (reference: train/__init__.py)."""
from .metrics import eval_metrics
from .optim import FlatAdam
from .trainer import Trainer
from .utils import EarlyStopping, _cycle_iterable, print_progress

__all__ = ["Trainer", "EarlyStopping", "eval_metrics", "FlatAdam", "print_progress", "_cycle_iterable"]
