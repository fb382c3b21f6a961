"""
Validation pass of the trainer (reference train/trainer.py:492-589,
train/metrics.py:14-135): ``Trainer.eval_epoch`` on the CUDA kernels -- joint encoder pass, losses accumulated on
the device, AUROC / AUPRC / threshold sweeps by sort + prefix sums on the device -- against the CPU oracle's
validation loop and scikit-learn.
"""
import contextlib
import io

import numpy as np
import pytest
import torch

from oracle import train_ref as TR
from oracle.vgpgae_ref import init_params
from tests import helpers as H
from tests.test_trainer_gpu import LAM, _module

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("joint", [True, False])
def test_eval_epoch_losses_and_metrics_match_oracle_loop(joint):
    from nichecompass_b200.train import Trainer
    from nichecompass_b200.train.metrics import eval_metrics
    adata = H.tiny_problem(n_cells=400, n_genes=44, n_prior_gp=12)
    cfg = H.oracle_cfg_from_adata(adata, n_addon_gp=4)
    p32 = init_params(cfg, seed=3)
    # oracle: splits, loaders, then the validation loop straight away (same draws from the global generator)
    torch.manual_seed(0)
    p64 = {k: v.double().clone() for k, v in p32.items()}
    ld = TR.build_loaders(H.dense_x(adata, torch.float64), adata.obsp["spatial_connectivities"], 32)
    ref = TR.TrainerRef(cfg, p64, ld.edge_loader, ld.node_loader, lr=1e-3, loss_kwargs=LAM)
    want_loss, probs, labels = ref.eval_epoch(ld.edge_val_loader, ld.node_val_loader)
    want_metrics = eval_metrics(probs.numpy(), labels.numpy())          # scikit-learn on the oracle's probabilities
    with contextlib.redirect_stdout(io.StringIO()):
        tr = Trainer(adata, _module(cfg, p32), edge_batch_size=32, seed=0, monitor=False, verbose=True,
                     compute_dtype=torch.float32, pipeline_sampler=False, joint_encoder=joint)
        tr.setup_training(lr=1e-3, **LAM)
    tr.iter_logs = {"n_val_iter": 0}
    acc = {}
    got_metrics = tr.eval_epoch(acc)
    n = tr.iter_logs["n_val_iter"]
    assert n == len(ld.edge_val_loader)
    for k, v in want_loss.items():
        got = float(acc[f"val_{k}"]) / n
        assert abs(got - v) <= 1e-5 * abs(v) + 1e-12, (k, got, v)
    for k in ("auroc_score", "auprc_score", "best_acc_score", "best_acc_threshold", "best_f1_score"):
        assert abs(got_metrics[k] - want_metrics[k]) < 1e-5, (k, got_metrics[k], want_metrics[k])
    assert tr.model.training                                            # back in training mode


def test_device_metrics_equal_sklearn_on_ties_covariates_and_exclusions():
    from nichecompass_b200.train.metrics import eval_metrics, eval_metrics_device
    rng = np.random.default_rng(0)
    n = 20_000
    labels = torch.tensor((rng.random(n) < 0.5).astype(np.float32))
    probs = torch.tensor(np.round(rng.random(n) * 0.6 + labels.numpy() * 0.3, 3).astype(np.float32))   # many ties
    same = [torch.tensor(rng.random(n) < 0.5), torch.tensor(rng.random(n) < 0.3)]
    incl = torch.tensor(rng.random(n) < 0.9)
    om = {f"{e}_rna{s}": torch.rand(500) for e in ("target", "source") for s in ("", "_preds")}
    a = eval_metrics(probs, labels, same, incl, {k: v.numpy() for k, v in om.items()})
    b = eval_metrics_device(probs.cuda(), labels.cuda(), [s.cuda() for s in same], incl.cuda(),
                            {k: v.cuda() for k, v in om.items()})
    assert set(a) == set(b)
    for k in a:
        assert abs(a[k] - b[k]) < 1e-6, (k, a[k], b[k])
