"""Import-only stand-in (train/utils.py, train/metrics.py plot helpers are never called)."""
