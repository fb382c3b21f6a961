"""
CPU restatement of the training hot loop (TEST INFRASTRUCTURE -- see oracle/__init__.py).

Follows /root/reference/src/nichecompass/train/trainer.py:
  :194-196   node batch size rule          -> sampler_ref.auto_node_batch_size
  :339-342   torch.optim.Adam over the trainable parameters
  :344-356   per-epoch flags (edge recon warm-up, all-GP epochs, contrastive warm-up)
  :364-421   zip(edge loader, cycled node loader): omics forward, graph forward, loss,
             zero_grad, backward, optional clip_grad_value_, optimizer.step
  :437-445   epoch log = mean of per-iteration losses
The forward / loss arithmetic is oracle/vgpgae_ref.py; batches come from oracle/sampler_ref.py.
Reparameterisation noise is drawn here with ``torch.randn`` from an explicit generator and
returned, so that the CUDA trainer can be driven with the identical draws.
"""
import math
import time
from types import SimpleNamespace
from typing import Dict, Iterable, List, Optional

import numpy as np
import torch

from . import sampler_ref as R
from .vgpgae_ref import VGPGAERef, trainable_keys


class SparseRows:
    """Row-indexable view of a scipy CSR count matrix: ``x[n_id]`` densifies only the sampled rows
    (the reference holds the whole matrix dense on the host, data/datasets.py:59-71; at 1M x 5k that
    is 20 GB, so the baseline loader densifies per batch instead -- same values)."""

    def __init__(self, csr, dtype=torch.float32):
        self.csr, self.dtype = csr.tocsr(), dtype

    def __getitem__(self, idx: torch.Tensor) -> torch.Tensor:
        rows = self.csr[idx.numpy()]
        return torch.from_numpy(np.asarray(rows.todense())).to(self.dtype)


def adam_step(params: Dict[str, torch.Tensor], keys: List[str], state: dict, lr: float,
              weight_decay: float = 0., betas=(0.9, 0.999), eps: float = 1e-8, clip_value: float = 0.):
    """torch.optim.Adam update (no amsgrad, not maximise) written out; parameters whose gradient is
    None are skipped exactly as torch does."""
    state["step"] = state.get("step", 0) + 1
    t = state["step"]
    b1, b2 = betas
    bc1 = 1 - b1 ** t
    bc2 = 1 - b2 ** t
    with torch.no_grad():
        for k in keys:
            p = params[k]
            if p.grad is None:
                continue
            g = p.grad
            if clip_value > 0:
                g = g.clamp(-clip_value, clip_value)
            if weight_decay != 0:
                g = g + weight_decay * p
            m = state.setdefault(("m", k), torch.zeros_like(p))
            v = state.setdefault(("v", k), torch.zeros_like(p))
            m.mul_(b1).add_(g, alpha=1 - b1)
            v.mul_(b2).addcmul_(g, g, value=1 - b2)
            denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
            p.addcdiv_(m, denom, value=-lr / bc1)


class TrainerRef:
    """One-process CPU trainer over the oracle module."""

    def __init__(self, cfg: dict, params: Dict[str, torch.Tensor], edge_loader, node_loader,
                 lr: float = 1e-3, weight_decay: float = 0., grad_clip_value: float = 0.,
                 n_epochs_all_gps: int = 25, n_epochs_no_edge_recon: int = 0,
                 noise_seed: int = 1234, loss_kwargs: Optional[dict] = None):
        self.cfg, self.p = cfg, params
        self.keys = trainable_keys(params)
        for k in self.keys:
            params[k].requires_grad_(True)
        self.model = VGPGAERef(cfg, params)
        self.model.training = True
        self.edge_loader, self.node_loader = edge_loader, node_loader
        self.lr, self.weight_decay, self.clip = lr, weight_decay, grad_clip_value
        self.n_epochs_all_gps, self.n_epochs_no_edge_recon = n_epochs_all_gps, n_epochs_no_edge_recon
        self.state: dict = {}
        self.gen = torch.Generator().manual_seed(noise_seed)
        self.loss_kwargs = loss_kwargs or {}
        self.noise_log: List[torch.Tensor] = []
        self.epoch = 0

    def _eps(self, rows: int) -> torch.Tensor:
        P = self.cfg["n_prior_gp"] + self.cfg["n_addon_gp"]
        dt = self.p["encoder.fc_l1.weight"].dtype
        e = torch.randn(rows, P, generator=self.gen, dtype=torch.float64).to(dt)
        self.noise_log.append(e)
        return e

    def iteration(self, edge_batch, node_batch) -> dict:
        only_active = not (self.epoch < self.n_epochs_all_gps)
        edge_active = not (self.epoch < self.n_epochs_no_edge_recon)
        node_out = self.model.forward(node_batch, "omics", use_only_active_gps=only_active,
                                      eps=self._eps(node_batch.batch_size))
        edge_out = self.model.forward(edge_batch, "graph", use_only_active_gps=only_active,
                                      eps=self._eps(2 * edge_batch.edge_label_index.size(1)))
        L = self.model.loss(edge_out, node_out, edge_recon_active=edge_active, **self.loss_kwargs)
        for k in self.keys:
            self.p[k].grad = None
        L["optim_loss"].backward()
        adam_step(self.p, self.keys, self.state, self.lr, self.weight_decay, clip_value=self.clip)
        return {k: float(v.detach()) for k, v in L.items()}

    @torch.no_grad()
    def eval_epoch(self, edge_val_loader, node_val_loader):
        """train/trainer.py:492-589: evaluation mode (z = mu), every (edge, cycled node) validation pair through
        both forwards and the loss with ``edge_recon_active=True``.  Returns (mean of every loss term over the
        validation iterations, edge probabilities, edge labels) -- the inputs of train/metrics.py:eval_metrics."""
        only_active = not (self.epoch < self.n_epochs_all_gps)
        was = self.model.training
        self.model.training = False
        sums, n = {}, 0
        probs, labels = [], []
        for edge_batch, node_batch in zip(edge_val_loader, R.cycle_iterable(node_val_loader)):
            node_out = self.model.forward(node_batch, "omics", use_only_active_gps=only_active)
            edge_out = self.model.forward(edge_batch, "graph", use_only_active_gps=only_active)
            L = self.model.loss(edge_out, node_out, edge_recon_active=True, **self.loss_kwargs)
            for k, v in L.items():
                sums[k] = sums.get(k, 0.0) + float(v)
            n += 1
            probs.append(torch.sigmoid(edge_out["edge_recon_logits"]))
            labels.append(edge_out["edge_recon_labels"])
        self.model.training = was
        return {k: v / max(n, 1) for k, v in sums.items()}, torch.cat(probs), torch.cat(labels)

    def run(self, max_iters: Optional[int] = None) -> List[dict]:
        """One epoch (or ``max_iters`` iterations) of zip(edge loader, cycled node loader)."""
        logs = []
        for edge_batch, node_batch in zip(self.edge_loader, R.cycle_iterable(self.node_loader)):
            logs.append(self.iteration(edge_batch, node_batch))
            if max_iters is not None and len(logs) >= max_iters:
                break
        return logs


def build_loaders(x, adj_csr, edge_batch_size: int, node_batch_size: Optional[int] = None,
                  edge_val_ratio: float = 0.1, node_val_ratio: float = 0.1, num_neighbors=(-1,),
                  cat_covariates_cats=None):
    """prepare_data + initialize_dataloaders of the reference (data/dataprocessors.py:133-221,
    data/dataloaders.py:11-102) on the CPU specification; consumes the global torch generator in
    the reference's order (edge split randperm, node split randperm)."""
    n = adj_csr.shape[0]
    ei = R.adjacency_to_edge_index(adj_csr)
    splits = R.edge_level_split(ei, n, edge_val_ratio)
    train_mask, val_mask = R.node_level_split_mask(n, node_val_ratio)
    n_edges_train = splits["train"].edge_label_index.size(1)
    if node_batch_size is None:
        node_batch_size = R.auto_node_batch_size(edge_batch_size, n_edges_train, int(train_mask.sum()))
    node_loader = R.NodeLoaderRef(x, ei, n, train_mask, node_batch_size, num_neighbors,
                                  cat_covariates_cats=cat_covariates_cats)
    edge_loader = R.LinkLoaderRef(x, splits["train"].edge_index, n, splits["train"].edge_label_index,
                                  edge_batch_size, num_neighbors, cat_covariates_cats=cat_covariates_cats)
    # validation loaders (data/dataloaders.py:71-101): the validation nodes over the full graph, the validation
    # edges over the TRAINING message-passing graph
    node_val_loader = R.NodeLoaderRef(x, ei, n, val_mask, node_batch_size, num_neighbors,
                                      cat_covariates_cats=cat_covariates_cats) if int(val_mask.sum()) > 0 else None
    edge_val_loader = None
    if "val" in splits and splits["val"].edge_label_index.size(1) > 0:
        edge_val_loader = R.LinkLoaderRef(x, splits["val"].edge_index, n, splits["val"].edge_label_index,
                                          edge_batch_size, num_neighbors, cat_covariates_cats=cat_covariates_cats)
    return SimpleNamespace(edge_val_loader=edge_val_loader, node_val_loader=node_val_loader,
                           edge_loader=edge_loader, node_loader=node_loader, node_batch_size=node_batch_size,
                           n_edges_train=n_edges_train, n_nodes_train=int(train_mask.sum()),
                           splits=splits, train_mask=train_mask, val_mask=val_mask, edge_index=ei)


def time_iterations(trainer: TrainerRef, n_iters: int, warmup: int = 0) -> float:
    """Seconds per iteration of the CPU loop (bench.py cpu_baseline leg)."""
    it = zip(trainer.edge_loader, R.cycle_iterable(trainer.node_loader))
    for _ in range(warmup):
        e, nb = next(it)
        trainer.iteration(e, nb)
    t0 = time.perf_counter()
    for _ in range(n_iters):
        e, nb = next(it)
        trainer.iteration(e, nb)
    return (time.perf_counter() - t0) / max(n_iters, 1)
