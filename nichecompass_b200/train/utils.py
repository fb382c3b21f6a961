"""
Helpers of the ``train`` subpackage (reference: train/utils.py): early stopping with
learning-rate reduction on plateau, the epoch progress line, and the node-loader cycling.
Plotting helpers of the reference (matplotlib) are out of scope.
"""
import sys
from typing import Tuple

import numpy as np


class EarlyStopping:
    """Same state machine as the reference (train/utils.py:13-110).

    ``step`` is called once per epoch with the monitored metric and answers
    (continue_training, reduce_lr); ``update_state`` tells whether the metric is the best seen
    so far (the trainer then snapshots the model state)."""

    def __init__(self,
                 early_stopping_metric: str = "val_global_loss",
                 metric_improvement_threshold: float = 0.,
                 patience: int = 8,
                 reduce_lr_on_plateau: bool = True,
                 lr_patience: int = 4,
                 lr_factor: float = 0.1):
        self.early_stopping_metric = early_stopping_metric
        self.metric_improvement_threshold = metric_improvement_threshold
        self.patience = patience
        self.reduce_lr_on_plateau = reduce_lr_on_plateau
        self.lr_patience = lr_patience
        self.lr_factor = lr_factor
        self.epochs = 0
        self.epochs_not_improved = 0
        self.epochs_not_improved_lr = 0
        self.current_performance = np.inf
        self.best_performance = np.inf
        self.best_performance_state = np.inf

    def step(self, current_metric: float) -> Tuple[bool, bool]:
        self.epochs += 1
        self.current_performance = current_metric
        improvement = self.best_performance - current_metric
        if improvement > 0:
            self.best_performance = current_metric
        if improvement < self.metric_improvement_threshold:
            self.epochs_not_improved += 1
            self.epochs_not_improved_lr += 1
        else:
            self.epochs_not_improved = 0
            self.epochs_not_improved_lr = 0

        reduce_lr = False
        continue_training = True
        if self.epochs < self.patience:
            pass
        elif self.epochs_not_improved >= self.patience:
            continue_training = False
        elif self.reduce_lr_on_plateau and self.epochs_not_improved_lr >= self.lr_patience:
            reduce_lr = True
            self.epochs_not_improved_lr = 0
            print("\nReducing learning rate: metric has not improved more "
                  f"than {self.metric_improvement_threshold} in the last "
                  f"{self.lr_patience} epochs.")
        if not continue_training:
            print(f"\nStopping early: metric has not improved more than {self.metric_improvement_threshold}"
                  f" in the last {self.patience} epochs.")
            print("If the early stopping criterion is too strong, please instantiate it with different "
                  "parameters in the train method.")
        return continue_training, reduce_lr

    def update_state(self, current_metric: float) -> bool:
        improved = (self.best_performance_state - current_metric) > 0
        if improved:
            self.best_performance_state = current_metric
        return improved


def print_progress(epoch: int, logs: dict, n_epochs: int):
    """One progress line per epoch: ``Epoch i/n |bar| pct% key: value; ...``
    (reference: train/utils.py:113-184)."""
    parts = [f"{k}: {v[-1]:.4f}" for k, v in logs.items() if len(v) > 0]
    frac = (epoch + 1) / float(n_epochs)
    width = 20
    filled = int(round(width * frac))
    bar = "█" * filled + "-" * (width - filled)
    sys.stdout.write(f"Epoch {epoch + 1}/{n_epochs} |{bar}| {100 * frac:.1f}% " + "; ".join(parts) + "\n")
    sys.stdout.flush()


def _cycle_iterable(iterable):
    """Endless iteration that restarts (and therefore re-shuffles) the loader when exhausted
    (reference: train/utils.py:219-225)."""
    it = iter(iterable)
    while True:
        try:
            yield next(it)
        except StopIteration:
            it = iter(iterable)
