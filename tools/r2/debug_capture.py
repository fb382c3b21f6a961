"""Bisects which stage of the training iteration breaks a CUDA-graph capture."""
import contextlib, io, os, sys, traceback
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from types import SimpleNamespace
from oracle.vgpgae_ref import init_params
from tests import helpers as H
from tests.test_trainer_gpu import _module, LAM
from nichecompass_b200.train import Trainer
from nichecompass_b200 import ops
from nichecompass_b200.functional import SubgraphCSR
from nichecompass_b200.ops import pad8

adata = H.tiny_problem(n_cells=600, n_genes=44, n_prior_gp=12)
cfg = H.oracle_cfg_from_adata(adata, n_addon_gp=4)
p32 = init_params(cfg, seed=3)
with contextlib.redirect_stdout(io.StringIO()):
    tr = Trainer(adata, _module(cfg, p32), edge_batch_size=32, seed=0, monitor=False, compute_dtype=torch.float32,
                 cuda_graphs=True)
    tr.setup_training(lr=1e-3, **LAM)
it = iter(tr.joint_batches(tr.edge_train_loader, tr.node_train_loader))
eb, nb = next(it)
tr.train_iteration(eb, nb)          # eager warm-up on the bucket
S = list(tr._graph_buckets.values())[0]
store = nb.store
m = tr.model
mode = sys.argv[1] if len(sys.argv) > 1 else "global"

def stage(n):
    gm = SubgraphCSR(S.m_rowptr, S.m_col, S.n_sub_pad, S.n_dst_pad, tr._graph_max_deg(), e_dst=S.e_pad)
    x_enc = ops.csr_gather_rows(store, S.m_nid, S.n_sub_pad, m.log_variational_, m.compute_dtype, pad8(m.encoder.n_input))
    if n == 1: return
    nbs = int(tr.node_batch_size_)
    node_b = SimpleNamespace(store=store, batch_size=nbs, n_id=None, graph=SimpleNamespace(n_dst=nbs, n_nodes=nbs))
    edge_b = SimpleNamespace(store=store, edge_label_index=S.eli, edge_label=S.elab, n_id=None,
                             graph=SimpleNamespace(n_dst=S.n_dst_pad - nbs, n_nodes=S.n_sub_pad - nbs))
    edge_b._joint_input = (node_b, gm, x_enc, None, (S.x_t, S.x_n))
    m._eps_queue = [S.eps_n, S.eps_e]
    if n == 2:
        ml = m.encoder.encode(x_enc, gm); m._eps_queue = None; return
    node_out, edge_out = m.forward_joint(node_b, edge_b, False)
    m._eps_queue = None
    if n == 3: return
    L = m.loss(edge_model_output=edge_out, node_model_output=node_out, **tr._loss_kwargs(), edge_recon_active=True,
               cat_covariates_contrastive_active=True)
    if n == 4: return
    tr.optimizer.zero_grad()
    L["optim_loss"].backward()
    if n == 5: return
    tr.optimizer.step()

import gc
from nichecompass_b200 import functional as Fn
for n in range(1, 7):
    m.mu = m.logstd = None; Fn._ZA_CACHE.clear(); gc.collect()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    try:
        with torch.cuda.graph(g, capture_error_mode=mode):
            stage(n)
        g.replay(); torch.cuda.synchronize()
        print("stage", n, "ok")
    except Exception as e:
        print("stage", n, "FAILED:", str(e).splitlines()[0])
        traceback.print_exc(limit=6)
        break
