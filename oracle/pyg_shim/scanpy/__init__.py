"""Import-only stand-in for scanpy (models/nichecompass.py imports it; hot path never calls it)."""
