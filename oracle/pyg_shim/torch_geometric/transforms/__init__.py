import copy

import torch

from oracle import sampler_ref as R


class RandomNodeSplit:
    def __init__(self, split="train_rest", num_val=500, num_test=1000, key="y", **kw):
        assert num_test == 0
        self.num_val = num_val

    def __call__(self, data):
        train_mask, val_mask = R.node_level_split_mask(data.num_nodes, self.num_val)
        data.train_mask, data.val_mask = train_mask, val_mask
        data.test_mask = torch.zeros_like(train_mask)
        return data


class RandomLinkSplit:
    """Only the configuration the reference uses (data/dataprocessors.py:80-86)."""

    def __init__(self, num_val=0.1, num_test=0.2, is_undirected=False, key="edge_label",
                 neg_sampling_ratio=1.0, **kw):
        assert is_undirected and num_test == 0 and neg_sampling_ratio == 0
        self.num_val = num_val

    def __call__(self, data):
        ei = data.edge_index
        assert bool((ei[0] != ei[1]).all()), "self loops must be stripped first"
        cand = (ei[0] <= ei[1]).nonzero(as_tuple=False).view(-1)
        perm = cand[torch.randperm(cand.size(0))]
        num_val = int(self.num_val * perm.numel())
        num_train = perm.numel() - num_val
        splits = []
        for lab, mp in ((perm[:num_train], perm[:num_train]),
                        (perm[num_train:num_train + num_val], perm[:num_train]),
                        (perm[num_train + num_val:], perm[:num_train + num_val])):
            d = copy.copy(data)
            d.__dict__ = dict(data.__dict__)
            d.edge_index = torch.cat([ei[:, mp], ei[:, mp].flip([0])], dim=-1)
            if hasattr(data, "edge_attr"):
                d.edge_attr = torch.cat([data.edge_attr[mp], data.edge_attr[mp]], dim=0)
            d.edge_label = torch.ones(lab.numel())
            d.edge_label_index = ei[:, lab]
            splits.append(d)
        return tuple(splits)
