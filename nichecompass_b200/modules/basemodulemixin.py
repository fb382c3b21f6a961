"""Generic module helpers mixed into the VGPGAE module (reference: modules/basemodulemixin.py)."""
import inspect
from collections import OrderedDict

import torch


class BaseModuleMixin:
    def _get_user_attributes(self) -> list:
        members = inspect.getmembers(self, lambda m: not inspect.isroutine(m))
        return [(k, v) for k, v in members if not (k.startswith("__") and k.endswith("__"))]

    def _get_public_attributes(self) -> dict:
        """Attributes whose name ends with an underscore are public by convention."""
        return {k: v for k, v in self._get_user_attributes() if k[-1] == "_"}

    def load_and_expand_state_dict(self, model_state_dict: OrderedDict):
        """Load a state dict saved from a smaller architecture (fewer add-on gene programs or
        covariate categories): tensors whose shape grew keep the loaded block and take the
        remaining rows / columns from this (freshly initialised) module, keys unknown to the
        checkpoint keep their fresh value (reference: modules/basemodulemixin.py:50-95)."""
        incoming = model_state_dict.copy()
        fresh = self.state_dict()
        device = next(self.parameters()).device
        for key, old in list(incoming.items()):
            new = fresh[key]
            if new.size() == old.size():
                continue
            old = old.to(device)
            grow_dim = None
            index = [slice(None)] * new.dim()
            for d in range(new.dim()):
                diff = new.shape[d] - old.shape[d]
                index[d] = slice(-diff, None)
                if diff > 0:
                    grow_dim = d
                    break
            incoming[key] = torch.cat([old, new[tuple(index)]], dim=grow_dim)
        for key, new in fresh.items():
            if key not in incoming:
                incoming[key] = new
        self.load_state_dict(incoming)
