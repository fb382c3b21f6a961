"""Generic trainer helpers (reference: train/basetrainermixin.py)."""
import inspect


class BaseTrainerMixin:
    def _get_user_attributes(self) -> list:
        members = inspect.getmembers(self, lambda m: not inspect.isroutine(m))
        return [(k, v) for k, v in members if not (k.startswith("__") and k.endswith("__"))]

    def _get_public_attributes(self) -> dict:
        """Attributes whose name ends with an underscore are public by convention."""
        return {k: v for k, v in self._get_user_attributes() if k[-1] == "_"}
