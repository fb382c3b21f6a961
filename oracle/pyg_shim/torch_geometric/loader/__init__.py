import torch

from oracle import sampler_ref as R
from ..data import Data


def _wrap(it):
    for b in it:
        yield Data(**b.__dict__)


class NeighborLoader:
    def __init__(self, data, num_neighbors, batch_size=1, directed=True, shuffle=False,
                 input_nodes=None, **kw):
        assert not directed
        self._ref = R.NodeLoaderRef(data.x, data.edge_index, data.num_nodes, input_nodes, batch_size,
                                    num_neighbors, shuffle,
                                    getattr(data, "cat_covariates_cats", None))
        self._attr = getattr(data, "edge_attr", None)

    def __len__(self):
        return len(self._ref)

    def __iter__(self):
        for b in _wrap(iter(self._ref)):
            if self._attr is not None:
                b.edge_attr = self._attr[b.e_id]
            yield b


class LinkNeighborLoader:
    def __init__(self, data, num_neighbors, batch_size=1, edge_label=None, edge_label_index=None,
                 directed=True, shuffle=False, neg_sampling_ratio=0.0, **kw):
        assert not directed and edge_label is None
        self._ref = R.LinkLoaderRef(data.x, data.edge_index, data.num_nodes, edge_label_index,
                                    batch_size, num_neighbors, shuffle, neg_sampling_ratio,
                                    getattr(data, "cat_covariates_cats", None))
        self._attr = getattr(data, "edge_attr", None)

    def __len__(self):
        return len(self._ref)

    def __iter__(self):
        for b in _wrap(iter(self._ref)):
            if self._attr is not None:
                b.edge_attr = self._attr[b.e_id]
            yield b
