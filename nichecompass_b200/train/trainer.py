"""
Trainer of the NicheCompass model (reference: train/trainer.py:28-706).

Same constructor / ``train`` signature, hyper-parameter semantics, epoch schedule (all gene
programs for ``n_epochs_all_gps`` epochs, optional warm-up without edge reconstruction /
contrastive loss), joint iteration over the edge loader and the cycled node loader, per-epoch
validation, early stopping with learning-rate reduction and best-state reload.

What runs differently (and why it is result-preserving):
  * batches are index sets sampled on the GPU from the HBM-resident graph + count store
    (``nichecompass_b200.data``); the sampler of batch i+1 runs on a side stream while batch i
    trains, so its one host read-back (the sub-graph size) does not stall the compute stream;
  * Adam is one kernel over a flat parameter buffer (``FlatAdam``); in data-parallel runs
    (``torch.distributed`` initialised, one process per GPU) rank r trains on batches
    r, r + W, ... of the common permutation and the flat gradient buffer is all-reduced (mean)
    once per step, as is the running |mu| statistic that drives gene-program pruning;
  * losses are accumulated on the device and read back once per epoch instead of a blocking
    ``.item()`` per iteration (train/trainer.py:404-411).
"""
import copy
import math
import os
import time
import warnings
from collections import defaultdict
from types import SimpleNamespace
from typing import List, Optional

import numpy as np
import torch
import torch.nn as nn

from ..data import initialize_dataloaders, prepare_data
from .basetrainermixin import BaseTrainerMixin
from .metrics import eval_metrics, eval_metrics_device
from .optim import FlatAdam
from .utils import EarlyStopping, _cycle_iterable, print_progress


def _dist_info():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


class _Prefetcher:
    """Runs a loader's sampling kernels on a side stream.  The loader's one host read-back per batch
    (the sub-graph size written by ``nc_sample_subgraph``) then waits for the side stream only, while
    the compute stream keeps executing the previous training step; the host therefore queues step
    i + 1 while the device still runs step i.  Batches are drawn strictly on demand, so the host
    random generators are consumed in the reference's order."""

    def __init__(self, loader, stream: Optional[torch.cuda.Stream]):
        self.loader, self.stream = loader, stream

    def __len__(self):
        return len(self.loader)

    def __iter__(self):
        if self.stream is None:
            yield from self.loader
            return
        main = torch.cuda.current_stream()
        self.stream.wait_stream(main)            # one-off: uploads made on the compute stream
        it = iter(self.loader)
        while True:
            with torch.cuda.stream(self.stream):
                batch = next(it, None)
                if batch is None:
                    return
                ev = torch.cuda.Event()
                ev.record(self.stream)
            for t in batch.tensors():
                t.record_stream(main)
            main.wait_event(ev)
            yield batch


class Trainer(BaseTrainerMixin):
    def __init__(self,
                 adata,
                 model: nn.Module,
                 adata_atac=None,
                 counts_key: Optional[str] = "counts",
                 adj_key: str = "spatial_connectivities",
                 cat_covariates_keys: Optional[List[str]] = None,
                 gp_targets_mask_key: str = "nichecompass_gp_targets",
                 gp_sources_mask_key: str = "nichecompass_gp_sources",
                 edge_val_ratio: float = 0.1,
                 node_val_ratio: float = 0.1,
                 edge_batch_size: int = 512,
                 node_batch_size: Optional[int] = None,
                 n_sampled_neighbors: int = -1,
                 use_early_stopping: bool = True,
                 reload_best_model: bool = True,
                 early_stopping_kwargs: Optional[dict] = None,
                 use_cuda_if_available: bool = True,
                 seed: int = 0,
                 monitor: bool = True,
                 verbose: bool = False,
                 **kwargs):
        self.adata = adata
        self.adata_atac = adata_atac
        self.model = model
        self.counts_key = counts_key
        self.adj_key = adj_key
        self.cat_covariates_keys = cat_covariates_keys
        self.n_cat_covariates = 0 if cat_covariates_keys is None else len(cat_covariates_keys)
        self.gp_targets_mask_key = gp_targets_mask_key
        self.gp_sources_mask_key = gp_sources_mask_key
        self.edge_train_ratio_ = 1 - edge_val_ratio
        self.edge_val_ratio_ = edge_val_ratio
        self.node_train_ratio_ = 1 - node_val_ratio
        self.node_val_ratio_ = node_val_ratio
        self.edge_batch_size_ = edge_batch_size
        self.node_batch_size_ = node_batch_size
        self.n_sampled_neighbors_ = n_sampled_neighbors
        self.use_early_stopping_ = use_early_stopping
        self.reload_best_model_ = reload_best_model
        self.early_stopping_kwargs_ = early_stopping_kwargs if early_stopping_kwargs else {}
        if "early_stopping_metric" not in self.early_stopping_kwargs_:
            self.early_stopping_kwargs_["early_stopping_metric"] = (
                "val_global_loss" if (edge_val_ratio > 0 and node_val_ratio > 0) else "train_global_loss")
        self.early_stopping = EarlyStopping(**self.early_stopping_kwargs_)
        self.seed_ = seed
        self.monitor_ = monitor
        self.verbose_ = verbose
        self.loaders_n_hops_ = kwargs.pop("loaders_n_hops", 1)
        self.grad_clip_value_ = kwargs.pop("grad_clip_value", 0.)
        # extensions of this build (not in the reference signature)
        self.compute_dtype_ = kwargs.pop("compute_dtype", None)
        self.store_location_ = kwargs.pop("store_location", "device")
        self.prefetch_ = kwargs.pop("prefetch", True)
        self.joint_encoder_ = kwargs.pop("joint_encoder", True)
        self.pipeline_sampler_ = kwargs.pop("pipeline_sampler", True)
        # None: stage the next iteration's encoder input on a side stream when the count store lives in host memory
        self.stage_inputs_ = kwargs.pop("stage_inputs", None)
        # CUDA graphs: the whole iteration (encoder input gather ... Adam) is captured once per size bucket and replayed
        self.cuda_graphs_ = bool(int(kwargs.pop("cuda_graphs", os.environ.get("NC_CUDA_GRAPHS", "1"))))
        self._graph_buckets, self._graph_pool = {}, None
        self.graph_stats_ = {"replayed": 0, "captured": 0, "eager": 0}
        self.dp_exchange_ = kwargs.pop("dp_exchange", os.environ.get("NC_DP_EXCHANGE", "peer"))
        self.dp_exchange_ctas_ = int(kwargs.pop("dp_exchange_ctas", os.environ.get("NC_DP_EXCHANGE_CTAS", "0")))
        kwargs.pop("latent_dtype", None)          # forwarded by NicheCompass.train, unused here (as in the reference)
        self.epoch = -1
        self.training_time = 0
        self.optimizer = None
        self.best_epoch = None
        self.best_model_state_dict = None
        self.rank_, self.world_size_ = _dist_info()

        print("\n--- INITIALIZING TRAINER ---")
        np.random.seed(self.seed_)
        if not (use_cuda_if_available and torch.cuda.is_available()):
            raise RuntimeError("nichecompass_b200 has no CPU training path: a CUDA device is required")
        torch.cuda.manual_seed(self.seed_)
        torch.manual_seed(self.seed_)
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.model.to(self.device)
        if self.compute_dtype_ is not None:
            self.model.set_compute_dtype(self.compute_dtype_)

        data_dict = prepare_data(
            adata=adata,
            cat_covariates_label_encoders=self.model.cat_covariates_label_encoders_,
            adata_atac=adata_atac,
            counts_key=self.counts_key,
            adj_key=self.adj_key,
            cat_covariates_keys=self.cat_covariates_keys,
            edge_val_ratio=self.edge_val_ratio_,
            edge_test_ratio=0.,
            node_val_ratio=self.node_val_ratio_,
            node_test_ratio=0.,
            device=self.device,
            store_location=self.store_location_)
        self.node_masked_data = data_dict["node_masked_data"]
        self.edge_train_data = data_dict["edge_train_data"]
        self.edge_val_data = data_dict["edge_val_data"]
        self.n_nodes_train = self.node_masked_data.train_mask.sum().item()
        self.n_nodes_val = self.node_masked_data.val_mask.sum().item()
        self.n_edges_train = self.edge_train_data.edge_label_index.size(1)
        self.n_edges_val = self.edge_val_data.edge_label_index.size(1)
        print(f"Number of training nodes: {self.n_nodes_train}")
        print(f"Number of validation nodes: {self.n_nodes_val}")
        print(f"Number of training edges: {self.n_edges_train}")
        print(f"Number of validation edges: {self.n_edges_val}")

        if self.node_batch_size_ is None:
            self.node_batch_size_ = int(self.edge_batch_size_ / math.floor(
                self.n_edges_train / self.n_nodes_train))
        print(f"Edge batch size: {edge_batch_size}")
        print(f"Node batch size: {node_batch_size}")

        loader_dict = initialize_dataloaders(
            node_masked_data=self.node_masked_data,
            edge_train_data=self.edge_train_data,
            edge_val_data=self.edge_val_data,
            edge_batch_size=self.edge_batch_size_,
            node_batch_size=self.node_batch_size_,
            n_direct_neighbors=self.n_sampled_neighbors_,
            n_hops=self.loaders_n_hops_,
            edges_directed=False,
            neg_edge_sampling_ratio=1.)
        self.edge_train_loader = loader_dict["edge_train_loader"]
        self.edge_val_loader = loader_dict.pop("edge_val_loader", None)
        self.node_train_loader = loader_dict["node_train_loader"]
        self.node_val_loader = loader_dict.pop("node_val_loader", None)
        # data parallelism: training AND validation loaders are strided over the ranks (validation losses are summed
        # over the ranks, the edge probabilities gathered before the metrics: every rank sees the same numbers)
        for ld in (self.edge_train_loader, self.node_train_loader, self.edge_val_loader, self.node_val_loader):
            if ld is not None:
                ld.set_shard(self.rank_, self.world_size_)
        # sampler kernels of the next batches are launched before the sizes of the current one are read back
        for ld in (self.edge_train_loader, self.node_train_loader, self.edge_val_loader, self.node_val_loader):
            if ld is not None:
                ld.set_pipeline(self.pipeline_sampler_)
        self._side_streams = ([torch.cuda.Stream(), torch.cuda.Stream()] if self.prefetch_ else [None, None])
        # data parallelism: the module averages the |mu| EMA over the ranks right after updating it, so that the edge
        # forward of the same iteration already derives its active-GP mask from the value every rank shares
        self.model._dp_sync_ema = self.world_size_ > 1

    # ------------------------------------------------------------------ one iteration
    def _loss_kwargs(self):
        return dict(lambda_edge_recon=self.lambda_edge_recon_,
                    lambda_gene_expr_recon=self.lambda_gene_expr_recon_,
                    lambda_chrom_access_recon=self.lambda_chrom_access_recon_,
                    lambda_cat_covariates_contrastive=self.lambda_cat_covariates_contrastive_,
                    contrastive_logits_pos_ratio=self.contrastive_logits_pos_ratio_,
                    contrastive_logits_neg_ratio=self.contrastive_logits_neg_ratio_,
                    lambda_group_lasso=self.lambda_group_lasso_,
                    lambda_l1_masked=self.lambda_l1_masked_,
                    l1_targets_mask=self.l1_targets_mask,
                    l1_sources_mask=self.l1_sources_mask,
                    lambda_l1_addon=self.lambda_l1_addon_)

    def train_iteration(self, edge_batch, node_batch) -> dict:
        """train/trainer.py:368-421: two forwards, loss, zero_grad, backward, clip, Adam step.
        Returns the loss dictionary (device scalars; nothing is read back here)."""
        if self._graph_eligible(node_batch, edge_batch):
            return self._graph_iteration(node_batch, edge_batch)
        self.graph_stats_["eager"] += 1
        joint = getattr(self.model, "joint_supported", None)
        if self.joint_encoder_ and joint is not None and joint(node_batch, edge_batch):
            # one encoder pass over the block-diagonal union of the two sampled sub-graphs
            node_out, edge_out = self.model.forward_joint(node_batch, edge_batch, self.use_only_active_gps)
        else:
            node_out = self.model(data_batch=node_batch.to(self.device), decoder="omics",
                                  use_only_active_gps=self.use_only_active_gps)
            edge_out = self.model(data_batch=edge_batch.to(self.device), decoder="graph",
                                  use_only_active_gps=self.use_only_active_gps)
        loss_dict = self.model.loss(
            edge_model_output=edge_out, node_model_output=node_out, **self._loss_kwargs(),
            edge_recon_active=self.edge_recon_active,
            cat_covariates_contrastive_active=self.cat_covariates_contrastive_active)
        self.optimizer.zero_grad()
        loss_dict["optim_loss"].backward()
        if self.world_size_ > 1:
            self.optimizer.all_reduce_grads()            # (the |mu| EMA was averaged inside the node forward)
        hook = getattr(self, "after_backward", None)
        if hook is not None:                 # parity tests: the gradients live in the arena until the Adam kernel clears them
            hook(self, node_out, edge_out)
        self.optimizer.step(clip_value=self.grad_clip_value_)
        # detached: a caller that keeps the dictionary must not keep the iteration's autograd graph alive with it
        return {k: v.detach() for k, v in loss_dict.items()}

    # ------------------------------------------------------------------ CUDA-graph size buckets
    # The reference loop (train/trainer.py:364-421) is ~70 kernel launches behind ~100 small torch ops and 16
    # autograd functions per iteration: 3.2-3.8 ms of Python for 3.2 ms of device work.  Every size of an iteration is
    # known on the host before it is launched (the sampler runs one batch ahead), so the iteration is captured ONCE
    # per size bucket -- sub-graph rows rounded up to 512, live edges to 2048, edge seeds to 4 x edge batch -- and
    # replayed: the host then only merges the two sampled sub-graphs into the bucket's padded buffers, gathers the
    # parameter-free label rows, draws the reparameterisation noise and launches the graph.  Padding rows repeat a
    # real node and are referenced by no edge: they cost (<= 3 %) work and change no result
    # (tests/test_trainer_gpu.py::test_cuda_graph_iterations_equal_eager_iterations).
    def _graph_eligible(self, node_batch, edge_batch) -> bool:
        if not self.cuda_graphs_:
            return False
        from .. import _lib
        m = self.model
        joint = getattr(m, "joint_supported", None)
        return (m.training and self.joint_encoder_ and joint is not None and joint(node_batch, edge_batch)
                and m._eps_queue is None and getattr(self, "after_backward", None) is None and _lib._HOOK is None
                and (m.n_cat_covariates_ == 0 or self._covariates_capturable())
                and m.node_label_method_ != "one-hop-attention"
                and "atac" not in m.modalities_
                and int(node_batch.batch_size) == int(self.node_batch_size_)
                and edge_batch.edge_label_index.size(1) == 2 * int(self.edge_batch_size_)
                and getattr(edge_batch, "_joint_input", None) is None)

    def _covariates_capturable(self) -> bool:
        """Covariate embeddings (config 3) inside the captured iteration: the categories of the sampled cells become
        static inputs.  Not with the contrastive loss (its top-k selection reads a count back on the host)."""
        contrastive = ((self.lambda_cat_covariates_contrastive_ or 0) > 0
                       and ((self.contrastive_logits_pos_ratio_ or 0) > 0 or (self.contrastive_logits_neg_ratio_ or 0) > 0))
        return not contrastive

    def _bucket_sizes(self, node_batch, edge_batch):
        gn, ge = node_batch.graph, edge_batch.graph
        B4 = 4 * int(self.edge_batch_size_)
        # rows to 512, live edges to 2048: a handful of buckets per workload (every bucket is one capture), <= 3 %
        # padding work on average at C2 (8.5 k rows, 12 k edges per iteration)
        n_sub_pad = (gn.n_nodes + ge.n_nodes + (B4 - ge.n_dst) + 511) // 512 * 512
        e_pad = (max(gn.e_dst + ge.e_dst, 1) + 2047) // 2048 * 2048
        return B4, n_sub_pad, e_pad

    def _graph_inputs(self, node_batch, edge_batch, out, background: bool):
        """Parameter-free inputs of one iteration into ``out`` (a bucket's static buffers, or look-ahead staging
        buffers): padded merge of the two sub-graphs, label rows, aggregated neighbour labels and -- for a count
        store in host memory -- the encoder input rows."""
        from .. import ops
        from .._lib import call
        from ..ops import pad8
        m = self.model
        gn, ge = node_batch.graph, edge_batch.graph
        nb = int(self.node_batch_size_)
        store = node_batch.store
        call("nc_merge_subgraphs_padded", gn.rowptr, gn.col, node_batch.n_id, gn.n_dst, gn.n_nodes, ge.rowptr, ge.col,
             edge_batch.n_id, ge.n_dst, ge.n_nodes, out.B4, out.n_sub_pad, out.e_pad, out.m_rowptr, out.m_col, out.m_nid)
        ops.csr_gather_rows(store, node_batch.n_id, nb, False, torch.float32, out=out.x_t, background=background)
        m.rna_node_label_aggregator.forward_store(store, node_batch.n_id, gn, nb, out=out.x_n)
        if out.x_enc is not None:
            ops.csr_gather_rows(store, out.m_nid, out.n_sub_pad, m.log_variational_, m.compute_dtype,
                                pad8(m.encoder.n_input), background=background, out=out.x_enc)

    def _input_buffers(self, node_batch, edge_batch, B4, n_sub_pad, e_pad):
        from ..ops import pad8
        m, dev, store = self.model, self.device, node_batch.store
        nb = int(self.node_batch_size_)
        host = store.location == "host"
        return SimpleNamespace(
            B4=B4, n_sub_pad=n_sub_pad, e_pad=e_pad, n_dst_pad=nb + B4,
            m_rowptr=torch.zeros(nb + B4 + 1, dtype=torch.int32, device=dev),
            m_col=torch.zeros(e_pad, dtype=torch.int32, device=dev),
            m_nid=torch.zeros(n_sub_pad, dtype=torch.int32, device=dev),
            x_t=torch.zeros(nb, store.n_cols, dtype=torch.float32, device=dev),
            x_n=torch.zeros(nb, store.n_cols, dtype=torch.float32, device=dev),
            # host-resident count store: the encoder input is an INPUT of the captured iteration (its rows cross
            # the host link one iteration ahead, on a side stream), not a gather inside it
            x_enc=(torch.zeros(n_sub_pad, pad8(m.encoder.n_input), dtype=m.compute_dtype, device=dev)
                   if host else None))

    def stage_graph_inputs(self, node_batch, edge_batch, side: "torch.cuda.Stream"):
        """Look-ahead for the captured iteration: the inputs of iteration k + 1 are built on ``side`` while iteration
        k replays (``_staged_pairs``); ``_graph_iteration`` then copies them device-to-device into the bucket."""
        B4, n_sub_pad, e_pad = self._bucket_sizes(node_batch, edge_batch)
        main = torch.cuda.current_stream()
        side.wait_stream(main)
        with torch.cuda.stream(side):
            t = self._input_buffers(node_batch, edge_batch, B4, n_sub_pad, e_pad)
            self._graph_inputs(node_batch, edge_batch, t, background=True)
            ev = torch.cuda.Event()
            ev.record(side)
        cached = [x for v in node_batch.graph._norm.values() for x in v if x is not None]
        for x in [t.m_rowptr, t.m_col, t.m_nid, t.x_t, t.x_n] + ([t.x_enc] if t.x_enc is not None else []) + cached:
            x.record_stream(main)
        edge_batch._graph_staged = (t, ev)

    def _graph_iteration(self, node_batch, edge_batch):
        from .. import _lib, ops
        from .._lib import call
        from ..functional import SubgraphCSR
        m, opt = self.model, self.optimizer
        B4, n_sub_pad, e_pad = self._bucket_sizes(node_batch, edge_batch)
        store = node_batch.store
        key = (n_sub_pad, e_pad, store.location, bool(self.use_only_active_gps), bool(self.edge_recon_active),
               bool(self.cat_covariates_contrastive_active), float(self.grad_clip_value_))
        S = self._graph_buckets.get(key)
        if S is None:
            dev = self.device
            P = m.n_prior_gp_ + m.n_addon_gp_
            eli = edge_batch.edge_label_index
            nb = int(self.node_batch_size_)
            S = self._graph_buckets[key] = self._input_buffers(node_batch, edge_batch, B4, n_sub_pad, e_pad)
            S.graph, S.losses, S.n_launches, S.uses = None, None, 0, 0
            S.eli, S.elab = torch.zeros_like(eli), torch.zeros_like(edge_batch.edge_label)
            S.eps_n = torch.zeros(nb, P, dtype=torch.float32, device=dev)
            S.eps_e = torch.zeros(eli.size(1) * 2, P, dtype=torch.float32, device=dev)
            S.cats_n = S.cats_e = None
            if m.n_cat_covariates_ > 0:                    # categories of the node seeds / of every edge-batch cell
                cn = node_batch.cat_covariates_cats
                S.cats_n = torch.zeros(nb, cn.size(1), dtype=cn.dtype, device=dev)
                S.cats_e = torch.zeros(n_sub_pad, cn.size(1), dtype=cn.dtype, device=dev)
        # ---- per-iteration inputs (eager, a handful of launches)
        staged = getattr(edge_batch, "_graph_staged", None)
        if staged is not None and staged[0].n_sub_pad == n_sub_pad and staged[0].e_pad == e_pad:
            t, ev = staged
            edge_batch._graph_staged = None
            torch.cuda.current_stream().wait_event(ev)
            for name in ("m_rowptr", "m_col", "m_nid", "x_t", "x_n", "x_enc"):
                if getattr(S, name) is not None:
                    getattr(S, name).copy_(getattr(t, name))
        else:
            self._graph_inputs(node_batch, edge_batch, S, background=False)
        S.eli.copy_(edge_batch.edge_label_index)
        S.elab.copy_(edge_batch.edge_label)
        if S.cats_n is not None:
            S.cats_n.copy_(node_batch.cat_covariates_cats[:S.cats_n.size(0)])
            S.cats_e[:edge_batch.cat_covariates_cats.size(0)].copy_(edge_batch.cat_covariates_cats)
        S.eps_n.normal_()                                  # same draws, same order as the eager path's randn_like
        S.eps_e.normal_()
        opt.sync_lr()
        S.uses += 1
        warm = self.__dict__.setdefault("_graph_warm", set())
        if S.graph is None and key[2:] not in warm:
            # the first iteration of every phase (store location, active-GP / edge-loss / contrastive flags) runs
            # eagerly on the bucket's buffers: every lazily initialised piece (mask structure, kernel attributes,
            # packed-weight caches, device copies of the pruning state) must exist before anything is captured
            self.graph_stats_["eager"] += 1
            out = self._graph_body(S, store)
            warm.add(key[2:])
            return out
        if S.graph is None:
            if self._graph_pool is None:
                self._graph_pool = torch.cuda.graph_pool_handle()
            # no autograd graph of an earlier (eager, default-stream) iteration may be alive: its AccumulateGrad nodes
            # would be reused and tie the capture to the stream they were created on
            # (cudaErrorStreamCaptureIsolation); the module keeps mu / logstd, the decoder its packed z
            import gc
            from .. import functional as Fn
            m.mu = m.logstd = None
            Fn._ZA_CACHE.clear()
            gc.collect()
            g = torch.cuda.CUDAGraph()
            n0, steps0 = _lib.launch_count(), opt.step_count
            with torch.cuda.graph(g, pool=self._graph_pool):
                S.losses = self._graph_body(S, store)
            S.n_launches = _lib.launch_count() - n0
            _lib.lib().nc_launch_count_add(-S.n_launches)       # recorded, not run
            opt.step_count = steps0
            S.graph = g
            self.graph_stats_["captured"] += 1
        S.graph.replay()
        _lib.lib().nc_launch_count_add(S.n_launches)
        opt.note_replayed_step()
        self.graph_stats_["replayed"] += 1
        return S.losses

    def _graph_body(self, S, store):
        """One iteration on a bucket's static buffers: everything from the encoder-input gather to the Adam update."""
        from .. import ops
        from ..functional import SubgraphCSR
        from ..ops import pad8
        m = self.model
        nb = int(self.node_batch_size_)
        gm = SubgraphCSR(S.m_rowptr, S.m_col, S.n_sub_pad, S.n_dst_pad, self._graph_max_deg(), e_dst=S.e_pad)
        x_enc = S.x_enc if S.x_enc is not None else ops.csr_gather_rows(
            store, S.m_nid, S.n_sub_pad, m.log_variational_, m.compute_dtype, pad8(m.encoder.n_input))
        node_b = SimpleNamespace(store=store, batch_size=nb, n_id=None, graph=SimpleNamespace(n_dst=nb, n_nodes=nb),
                                 cat_covariates_cats=S.cats_n)
        edge_b = SimpleNamespace(store=store, edge_label_index=S.eli, edge_label=S.elab, n_id=None,
                                 graph=SimpleNamespace(n_dst=S.n_dst_pad - nb, n_nodes=S.n_sub_pad - nb),
                                 cat_covariates_cats=S.cats_e)
        edge_b._joint_input = (node_b, gm, x_enc, None, (S.x_t, S.x_n))
        m._eps_queue = [S.eps_n, S.eps_e]
        try:
            node_out, edge_out = m.forward_joint(node_b, edge_b, self.use_only_active_gps)
        finally:
            m._eps_queue = None
        loss_dict = m.loss(edge_model_output=edge_out, node_model_output=node_out, **self._loss_kwargs(),
                           edge_recon_active=self.edge_recon_active,
                           cat_covariates_contrastive_active=self.cat_covariates_contrastive_active)
        self.optimizer.zero_grad()
        loss_dict["optim_loss"].backward()
        if self.world_size_ > 1:
            self.optimizer.all_reduce_grads()
        self.optimizer.step(clip_value=self.grad_clip_value_)
        return {k: v.detach() for k, v in loss_dict.items()}

    def _graph_max_deg(self) -> int:
        """Upper bound of a sampled row's length: the largest in-degree of the two message-passing graphs."""
        d = getattr(self, "_graph_max_deg_", None)
        if d is None:
            d = max(self.node_masked_data.device_graph().max_deg, self.edge_train_data.device_graph().max_deg)
            self._graph_max_deg_ = d
        return d

    def _sync_running_stats(self):
        """Ranks see different node batches: average the |mu| EMA so that every rank prunes the
        same gene programs (modules/vgpgae.py:641-651 is per-process state in the reference)."""
        import torch.distributed as dist
        buf = self.model.running_mean_abs_mu
        dist.all_reduce(buf, op=dist.ReduceOp.SUM)
        buf.div_(self.world_size_)

    def _set_epoch_flags(self):
        self.edge_recon_active = not (self.epoch < self.n_epochs_no_edge_recon_)
        self.use_only_active_gps = not (self.epoch < self.n_epochs_all_gps_)
        self.cat_covariates_contrastive_active = not (self.epoch < self.n_epochs_no_cat_covariates_contrastive_)

    def _accumulate(self, acc: dict, prefix: str, loss_dict: dict):
        keys = loss_dict.keys() if self.verbose_ else ("global_loss", "optim_loss")
        for k in keys:
            v = loss_dict[k].detach()
            name = f"{prefix}_{k}"
            acc[name] = v.clone() if name not in acc else acc[name].add_(v)

    def joint_batches(self, edge_loader, node_loader):
        """zip(edge loader, cycled node loader) with both samplers running one batch ahead."""
        e = _Prefetcher(edge_loader, self._side_streams[0])
        n = _Prefetcher(node_loader, self._side_streams[1])
        pairs = zip(e, _cycle_iterable(n))
        stage = self.stage_inputs_
        if stage is None:
            store = getattr(getattr(self, "node_masked_data", None), "store", None)
            stage = store is not None and store.location == "host"
        if not (stage and self.joint_encoder_ and self.prefetch_ and hasattr(self.model, "stage_joint_input")):
            return pairs
        return self._staged_pairs(pairs)

    def _staged_pairs(self, pairs):
        """One pair of look-ahead: the joint encoder input of iteration k + 1 (sub-graph merge + gather of the count
        rows over the host link) is built on a third stream while iteration k computes."""
        if len(self._side_streams) < 3:
            self._side_streams.append(torch.cuda.Stream())
        side = self._side_streams[2]
        prev = None
        for eb, nb in pairs:
            if self._graph_eligible(nb, eb):
                self.stage_graph_inputs(nb, eb, side)
            elif self.model.training and self.model.joint_supported(nb, eb):
                self.model.stage_joint_input(nb, eb, side)
            if prev is not None:
                yield prev
            prev = (eb, nb)
        if prev is not None:
            yield prev

    # ------------------------------------------------------------------ training
    def train(self,
              n_epochs: int = 100,
              n_epochs_all_gps: int = 25,
              n_epochs_no_edge_recon: int = 0,
              n_epochs_no_cat_covariates_contrastive: int = 5,
              lr: float = 0.001,
              weight_decay: float = 0.,
              lambda_edge_recon: Optional[float] = 500000.,
              lambda_cat_covariates_contrastive: Optional[float] = 0.,
              contrastive_logits_pos_ratio: Optional[float] = 0.125,
              contrastive_logits_neg_ratio: Optional[float] = 0.125,
              lambda_gene_expr_recon: float = 100.,
              lambda_chrom_access_recon: float = 10.,
              lambda_group_lasso: float = 0.,
              lambda_l1_masked: float = 0.,
              l1_targets_mask: Optional[torch.Tensor] = None,
              l1_sources_mask: Optional[torch.Tensor] = None,
              lambda_l1_addon: float = 0.,
              mlflow_experiment_id: Optional[str] = None,
              max_iters_per_epoch: Optional[int] = None):
        self.setup_training(n_epochs, n_epochs_all_gps, n_epochs_no_edge_recon,
                            n_epochs_no_cat_covariates_contrastive, lr, weight_decay, lambda_edge_recon,
                            lambda_cat_covariates_contrastive, contrastive_logits_pos_ratio,
                            contrastive_logits_neg_ratio, lambda_gene_expr_recon, lambda_chrom_access_recon,
                            lambda_group_lasso, lambda_l1_masked, l1_targets_mask, l1_sources_mask,
                            lambda_l1_addon, mlflow_experiment_id)
        print("\n--- MODEL TRAINING ---")
        start_time = time.time()
        for self.epoch in range(n_epochs):
            self._set_epoch_flags()
            self.iter_logs = defaultdict(list)
            self.iter_logs["n_train_iter"] = 0
            self.iter_logs["n_val_iter"] = 0
            acc = {}
            for edge_batch, node_batch in self.joint_batches(self.edge_train_loader, self.node_train_loader):
                loss_dict = self.train_iteration(edge_batch, node_batch)
                self._accumulate(acc, "train", loss_dict)
                self.iter_logs["n_train_iter"] += 1
                if max_iters_per_epoch is not None and self.iter_logs["n_train_iter"] >= max_iters_per_epoch:
                    break

            if self.edge_val_loader is not None and self.node_val_loader is not None:
                self.eval_epoch(acc)
            elif self.edge_val_loader is None and self.node_val_loader is not None:
                warnings.warn("You have specified a node validation set but no edge validation set. "
                              "Skipping validation...")
            elif self.edge_val_loader is not None and self.node_val_loader is None:
                warnings.warn("You have specified an edge validation set but no node validation set. "
                              "Skipping validation...")

            self._epoch_logs_from(acc)
            if self.monitor_ and self.rank_ == 0:
                print_progress(self.epoch, self.epoch_logs, self.n_epochs_)
            if self.use_early_stopping_ and self.is_early_stopping():
                break

        self.training_time += (time.time() - start_time)
        minutes, seconds = divmod(self.training_time, 60)
        print(f"Model training finished after {int(minutes)} min {int(seconds)} sec.")
        if self.best_model_state_dict is not None and self.reload_best_model_:
            print(f"Using best model state, which was in epoch {self.best_epoch + 1}.")
            self.model.load_state_dict(self.best_model_state_dict)
        self.model.eval()
        if self.edge_val_loader is not None:
            self.eval_end()

    def setup_training(self, n_epochs=100, n_epochs_all_gps=25, n_epochs_no_edge_recon=0,
                       n_epochs_no_cat_covariates_contrastive=5, lr=0.001, weight_decay=0.,
                       lambda_edge_recon=500000., lambda_cat_covariates_contrastive=0.,
                       contrastive_logits_pos_ratio=0.125, contrastive_logits_neg_ratio=0.125,
                       lambda_gene_expr_recon=100., lambda_chrom_access_recon=10., lambda_group_lasso=0.,
                       lambda_l1_masked=0., l1_targets_mask=None, l1_sources_mask=None, lambda_l1_addon=0.,
                       mlflow_experiment_id=None):
        """Hyper-parameter bookkeeping + optimiser construction of ``train`` (train/trainer.py:310-342),
        split out so that single iterations can be driven from outside (bench, tests)."""
        self.n_epochs_ = n_epochs
        self.n_epochs_all_gps_ = n_epochs_all_gps
        self.n_epochs_no_edge_recon_ = n_epochs_no_edge_recon
        self.n_epochs_no_cat_covariates_contrastive_ = n_epochs_no_cat_covariates_contrastive
        self.lr_ = lr
        self.weight_decay_ = weight_decay
        self.lambda_edge_recon_ = lambda_edge_recon
        self.lambda_gene_expr_recon_ = lambda_gene_expr_recon
        self.lambda_chrom_access_recon_ = lambda_chrom_access_recon
        self.lambda_cat_covariates_contrastive_ = lambda_cat_covariates_contrastive
        self.contrastive_logits_pos_ratio_ = contrastive_logits_pos_ratio
        self.contrastive_logits_neg_ratio_ = contrastive_logits_neg_ratio
        self.lambda_group_lasso_ = lambda_group_lasso
        self.lambda_l1_masked_ = lambda_l1_masked
        self.l1_targets_mask = l1_targets_mask
        self.l1_sources_mask = l1_sources_mask
        self.lambda_l1_addon_ = lambda_l1_addon
        self.mlflow_experiment_id = mlflow_experiment_id     # mlflow logging is not part of this build
        self.epoch_logs = defaultdict(list)
        self.model.train()
        self.model.materialize_outputs = False              # training consumes the fused losses only
        params = [p for p in self.model.parameters() if p.requires_grad]
        enc = getattr(self.model, "encoder", None)
        groups = enc.arena_groups() if hasattr(enc, "arena_groups") else None
        shadow = getattr(self.model, "compute_dtype", torch.float32) == torch.bfloat16
        self.optimizer = self._make_optimizer(params, lr, weight_decay, groups, shadow)
        self.epoch = 0
        self._set_epoch_flags()

    def _make_optimizer(self, params, lr, weight_decay, groups, shadow):
        """N > 1 ranks: arenas in peer-mapped memory, gradient exchange fused with Adam over NVLink
        (``dp_exchange="peer"``, default; csrc/dp_exchange.cu); ``"nccl"`` keeps the all-reduce + Adam pair.  If peer
        memory cannot be set up (no cudaIpc between the ranks' devices) every rank falls back to NCCL together."""
        if self.world_size_ > 1 and self.dp_exchange_ == "peer":
            import torch.distributed as dist
            opt, ok = None, 1
            try:
                opt = FlatAdam(params, lr=lr, weight_decay=weight_decay, groups=groups, bf16_shadow=shadow,
                               peer_group=True, peer_ctas=self.dp_exchange_ctas_)
            except Exception as e:                                      # noqa: BLE001
                warnings.warn(f"rank {self.rank_}: peer-memory gradient exchange unavailable ({e}); using NCCL")
                ok = 0
            flag = torch.tensor([ok], device=self.device)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            if int(flag.item()) == 1:
                # the |mu| running mean is averaged over the ranks by the exchange kernel itself (no NCCL call in the
                # iteration at all): the module's buffer moves into peer-mapped memory
                ema = getattr(self.model, "running_mean_abs_mu", None)
                if ema is not None:
                    self.model.running_mean_abs_mu.data = opt.attach_aux(ema.data)
                    self.model._dp_sync_ema = False
                return opt
            self.dp_exchange_ = "nccl"
        return FlatAdam(params, lr=lr, weight_decay=weight_decay, groups=groups, bf16_shadow=shadow)

    def _epoch_logs_from(self, acc: dict):
        """Mean of the per-iteration losses (train/trainer.py:437-445); one read-back per epoch."""
        if not acc:
            return
        names = sorted(acc.keys())
        vals = torch.stack([acc[k].float() for k in names])
        n_train, n_val = float(self.iter_logs["n_train_iter"]), float(self.iter_logs["n_val_iter"])
        if self.world_size_ > 1:
            import torch.distributed as dist
            packed = torch.cat((vals, torch.tensor([n_train, n_val], device=vals.device)))
            dist.all_reduce(packed, op=dist.ReduceOp.SUM)         # sums and iteration counts of all ranks
            vals, n_train, n_val = packed[:-2], float(packed[-2]), float(packed[-1])
        vals = vals.cpu().tolist()
        for k, v in zip(names, vals):
            n = n_train if k.startswith("train") else n_val
            self.epoch_logs[k].append(v / max(n, 1.0))

    # ------------------------------------------------------------------ evaluation
    def _collect_edge_outputs(self, store: dict, edge_out: dict):
        store["probs"].append(torch.sigmoid(edge_out["edge_recon_logits"]).detach())
        store["labels"].append(edge_out["edge_recon_labels"].detach())
        if edge_out["edge_same_cat_covariates_cat"] is not None:
            for i, s in enumerate(edge_out["edge_same_cat_covariates_cat"]):
                store["same"][i].append(s.detach())
        if edge_out["edge_incl"] is not None:
            store["incl"].append(edge_out["edge_incl"].detach())

    def _edge_store(self):
        return {"probs": [], "labels": [], "same": [[] for _ in range(self.n_cat_covariates)], "incl": []}

    @staticmethod
    def _cat(ts):
        return torch.cat(ts) if len(ts) > 0 else None

    def _gather_ranks(self, t: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
        """Concatenation over the ranks of a 1-D device tensor whose length differs per rank (validation shards)."""
        if t is None or self.world_size_ == 1:
            return t
        import torch.distributed as dist
        n = torch.tensor([t.numel()], device=t.device)
        sizes = [torch.zeros_like(n) for _ in range(self.world_size_)]
        dist.all_gather(sizes, n)
        sizes = [int(x) for x in sizes]
        buf = torch.zeros(max(sizes), dtype=t.dtype, device=t.device)
        buf[:t.numel()] = t
        parts = [torch.empty_like(buf) for _ in range(self.world_size_)]
        dist.all_gather(parts, buf)
        return torch.cat([p_[:k] for p_, k in zip(parts, sizes)])

    def _edge_metrics(self, store: dict, omics_pred_dict: Optional[dict] = None) -> dict:
        """AUROC / AUPRC / threshold sweeps on the device (``train/metrics.py: eval_metrics_device``): the edge
        probabilities never leave HBM, one read-back of the finished scalars.  With N ranks every rank evaluated
        its shard of the validation batches: the per-edge outputs are gathered first."""
        if self.world_size_ > 1:
            store = {"probs": [self._gather_ranks(self._cat(store["probs"]))],
                     "labels": [self._gather_ranks(self._cat(store["labels"]))],
                     "incl": ([self._gather_ranks(self._cat(store["incl"]).to(torch.uint8)).bool()]
                              if store["incl"] else []),
                     "same": [[self._gather_ranks(self._cat(s_).to(torch.uint8)).bool()] if s_ else []
                              for s_ in store["same"]]}
            if omics_pred_dict is not None:
                omics_pred_dict = {k: self._gather_ranks(v) for k, v in omics_pred_dict.items()}
        incl = self._cat(store["incl"])
        same = [self._cat(s) for s in store["same"]] if incl is not None else None
        return eval_metrics_device(edge_recon_probs=self._cat(store["probs"]), edge_labels=self._cat(store["labels"]),
                                   edge_same_cat_covariates_cat=same, edge_incl=incl,
                                   omics_pred_dict=omics_pred_dict)

    @torch.no_grad()
    def eval_epoch(self, acc: Optional[dict] = None):
        """Per-epoch validation (train/trainer.py:492-589)."""
        acc = {} if acc is None else acc
        self.model.eval()
        store = self._edge_store()
        joint = getattr(self.model, "joint_supported", None)
        for edge_batch, node_batch in self.joint_batches(self.edge_val_loader, self.node_val_loader):
            if self.joint_encoder_ and joint is not None and joint(node_batch, edge_batch):
                node_out, edge_out = self.model.forward_joint(node_batch, edge_batch, self.use_only_active_gps)
            else:
                node_out = self.model(data_batch=node_batch.to(self.device), decoder="omics",
                                      use_only_active_gps=self.use_only_active_gps)
                edge_out = self.model(data_batch=edge_batch.to(self.device), decoder="graph",
                                      use_only_active_gps=self.use_only_active_gps)
            loss_dict = self.model.loss(edge_model_output=edge_out, node_model_output=node_out,
                                        **self._loss_kwargs(), edge_recon_active=True)
            self._accumulate(acc, "val", loss_dict)
            self.iter_logs["n_val_iter"] += 1
            self._collect_edge_outputs(store, edge_out)
        val_eval_dict = self._edge_metrics(store)
        if self.verbose_:
            for k in ("auroc_score", "auprc_score", "best_acc_score", "best_f1_score"):
                self.epoch_logs[f"val_{k}"].append(val_eval_dict[k])
        self.model.train()
        return val_eval_dict

    @torch.no_grad()
    def eval_end(self):
        """End-of-training validation metrics (train/trainer.py:591-683)."""
        self.model.eval()
        prev = self.model.materialize_outputs
        self.model.materialize_outputs = True
        store = self._edge_store()
        for edge_batch in _Prefetcher(self.edge_val_loader, self._side_streams[0]):
            edge_out = self.model(data_batch=edge_batch.to(self.device), decoder="graph", use_only_active_gps=True)
            self._collect_edge_outputs(store, edge_out)
        preds = defaultdict(list)
        if self.node_val_loader is not None:
            for node_batch in _Prefetcher(self.node_val_loader, self._side_streams[1]):
                node_out = self.model(data_batch=node_batch.to(self.device), decoder="omics",
                                      use_only_active_gps=True)
                for mod in self.model.modalities_:
                    for ent in ("target", "source"):
                        preds[f"{ent}_{mod}_preds"].append(node_out[f"{ent}_{mod}_nb_means"].reshape(-1))
                        preds[f"{ent}_{mod}"].append(node_out["node_labels"][f"{ent}_{mod}"].reshape(-1))
        omics = {k: torch.cat(v) for k, v in preds.items()} if preds else None
        self.model.materialize_outputs = prev
        val_eval_dict = self._edge_metrics(store, omics)
        self.val_eval_dict_ = val_eval_dict
        if self.rank_ == 0:
            print("\n--- MODEL EVALUATION ---")
            print(f"val AUROC score: {val_eval_dict['auroc_score']:.4f}")
            print(f"val AUPRC score: {val_eval_dict['auprc_score']:.4f}")
            print(f"val best accuracy score: {val_eval_dict['best_acc_score']:.4f}")
            print(f"val best F1 score: {val_eval_dict['best_f1_score']:.4f}")
            if omics is not None:
                for mod in self.model.modalities_:
                    for ent in ("target", "source"):
                        print(f"val {ent} {mod} MSE score: {val_eval_dict[f'{ent}_{mod}_mse_score']:.4f}")
            for i in range(self.n_cat_covariates):
                if f"cat_covariate{i}_mean_sim_diff" in val_eval_dict:
                    print(f"Val cat covariate{i} mean sim diff: "
                          f"{val_eval_dict[f'cat_covariate{i}_mean_sim_diff']:.4f}")
        return val_eval_dict

    def is_early_stopping(self) -> bool:
        """train/trainer.py:685-706: snapshot the best state, reduce the learning rate on plateau,
        answer whether to stop."""
        metric = self.early_stopping.early_stopping_metric
        current = self.epoch_logs[metric][-1]
        if self.world_size_ > 1:
            # validation is replicated per rank and may differ in the last ulp (float atomics): every rank acts on
            # rank 0's value, so stop / reduce-lr / best-state decisions cannot diverge (a rank that stops alone
            # would leave the others in the gradient exchange forever)
            import torch.distributed as dist
            t = torch.tensor([float(current)], dtype=torch.float64, device=self.device)
            dist.broadcast(t, src=0)
            current = float(t.item())
        if self.early_stopping.update_state(current):
            self.best_model_state_dict = copy.deepcopy(self.model.state_dict())
            self.best_epoch = self.epoch
        continue_training, reduce_lr = self.early_stopping.step(current)
        if reduce_lr:
            for param_group in self.optimizer.param_groups:
                param_group["lr"] *= self.early_stopping.lr_factor
            print(f"New learning rate is {param_group['lr']}.\n")
        return not continue_training
