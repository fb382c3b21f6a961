"""Mirror of ``nichecompass.models`` (reference: models/__init__.py)."""
from .nichecompass import NicheCompass

__all__ = ["NicheCompass"]
