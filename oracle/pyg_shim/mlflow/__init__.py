"""No-op stand-in for mlflow (logging only; modules/vgpgae.py:1336-1366)."""
from . import exceptions  # noqa: F401


def log_param(*a, **k):
    pass


def log_metric(*a, **k):
    pass


def log_figure(*a, **k):
    pass
