"""Stand-in for torch_geometric (see ../README.md)."""
__version__ = "0.0.shim"
from . import data, loader, nn, transforms, utils  # noqa: F401
