"""Generic VGAE helper mixed into the VGPGAE module (reference: modules/vgaemodulemixin.py)."""
import torch

from ..functional import ReparamFn


class VGAEModuleMixin:
    def reparameterize(self, mu: torch.Tensor, logstd: torch.Tensor) -> torch.Tensor:
        """Training: ``mu + eps * exp(logstd)`` with eps ~ N(0, 1) drawn like the reference
        (``torch.randn_like`` on the module's device); evaluation: ``mu``."""
        if self.training:
            eps = torch.randn_like(mu)
            return ReparamFn.apply(mu, logstd, eps, None)
        return mu
