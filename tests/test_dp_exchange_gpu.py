"""
Data-parallel step over NVLink peer memory (csrc/dp_exchange.cu, train/optim.py FlatAdam(peer_group=...)): world
size 2 on two GPUs of one box (skipped on a 1-GPU box).  Every rank ends with the parameters, bf16 shadow and
cleared gradients of ONE Adam step on the mean gradient -- the same numbers as nc_adam_step on that mean (the 1-GPU
step on the concatenated batch), and the ranks' arenas are bit-identical.  Three steps, so that the flag epochs and
the shard-local moments are exercised across steps.
"""
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_params, out):
    import torch.distributed as dist
    from nichecompass_b200.train.optim import FlatAdam
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        g0 = torch.Generator().manual_seed(0)
        shapes = [(n_params // 3, 3), (1001,), (257, 5)]
        init = [torch.randn(s, generator=g0) for s in shapes]

        def params():
            return [torch.nn.Parameter(t.clone().cuda()) for t in init]
        peer = FlatAdam(params(), lr=1e-2, weight_decay=0.01, bf16_shadow=True, peer_group=True)
        assert peer.peer is not None
        plain = FlatAdam(params(), lr=1e-2, weight_decay=0.01, bf16_shadow=True)       # witness: mean gradient, nc_adam_step
        for step in range(3):
            grads = []
            for r in range(world):                                   # every rank can reproduce every rank's gradient
                gr = torch.Generator().manual_seed(100 * step + r)
                grads.append(torch.randn(peer.numel, generator=gr))
            peer.zero_grad()
            peer.flat_g.copy_(grads[rank].cuda())
            plain.zero_grad()
            plain.flat_g.copy_((sum(grads) / world).cuda())
            peer.step(clip_value=0.5)
            plain.step(clip_value=0.5)
            torch.cuda.synchronize()
            dist.barrier()
            assert float(peer.flat_g.abs().max()) == 0.0, "gradients must be cleared by the shard owners"
            err = float((peer.flat_p - plain.flat_p).abs().max())
            assert err <= 1e-6, (step, err)                          # mean in a different order: last-ulp differences
            assert torch.equal(peer.flat_bf16.float(), peer.flat_p.to(torch.bfloat16).float())
            mine = peer.flat_p.clone()
            other = [torch.empty_like(mine) for _ in range(world)]
            dist.all_gather(other, mine)
            assert all(torch.equal(o, mine) for o in other), "ranks must hold bit-identical parameters"
        out[rank] = "ok"
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs on one box")
@pytest.mark.parametrize("n_params", [30_000, 3_000_003])
def test_peer_exchange_equals_adam_on_the_mean_gradient(n_params):
    import torch.multiprocessing as mp
    world = 2
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_params, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
    for p in procs:
        if p.is_alive():
            p.kill()
            raise AssertionError("a rank hung (barrier flags?)")
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert dict(out) == {0: "ok", 1: "ok"}


def _train_worker(rank, world, port, out):
    import contextlib
    import io
    import torch.distributed as dist
    from oracle.vgpgae_ref import init_params
    from tests import helpers as H
    from tests.test_trainer_gpu import LAM, _module
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from nichecompass_b200.train import Trainer
        adata = H.tiny_problem(n_cells=600, n_genes=44, n_prior_gp=12)
        cfg = H.oracle_cfg_from_adata(adata, n_addon_gp=4)
        p32 = init_params(cfg, seed=3)
        with contextlib.redirect_stdout(io.StringIO()):
            tr = Trainer(adata, _module(cfg, p32), edge_batch_size=32, seed=0, monitor=False, verbose=True,
                         compute_dtype=torch.float32)
            tr.setup_training(lr=1e-3, **LAM)
        assert tr.optimizer.peer is not None and tr.world_size_ == 2
        losses = []
        for i, (eb, nb) in enumerate(tr.joint_batches(tr.edge_train_loader, tr.node_train_loader)):
            losses.append(float(tr.train_iteration(eb, nb)["global_loss"]))
            if i == 9:
                break
        # both ranks hold the same parameters and the same |mu| statistic after ten exchanged steps
        flat = tr.optimizer.flat_p.clone()
        others = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(others, flat)
        assert all(torch.equal(o, flat) for o in others)
        ema = tr.model.running_mean_abs_mu.clone()
        emas = [torch.empty_like(ema) for _ in range(world)]
        dist.all_gather(emas, ema)
        assert all(torch.equal(e, ema) for e in emas) and float(ema.abs().sum()) > 0
        assert tr.graph_stats_["replayed"] >= 5
        # sharded validation: every rank reports the same losses and metrics
        tr.iter_logs = {"n_train_iter": 10, "n_val_iter": 0}
        acc = {}
        metrics = tr.eval_epoch(acc)
        tr._epoch_logs_from(acc)
        got = torch.tensor([metrics["auroc_score"], metrics["auprc_score"], tr.epoch_logs["val_global_loss"][-1]],
                           dtype=torch.float64, device="cuda")
        gs = [torch.empty_like(got) for _ in range(world)]
        dist.all_gather(gs, got)
        assert all(torch.equal(g, got) for g in gs), gs
        assert 0.0 < float(got[0]) <= 1.0
        out[rank] = "ok"
        torch.cuda.synchronize()
        dist.barrier()
    finally:
        os._exit(0 if out.get(rank) == "ok" else 1)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs on one box")
def test_two_rank_training_keeps_ranks_identical_and_shards_validation():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_train_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=240)
    for p in procs:
        if p.is_alive():
            p.kill()
            raise AssertionError("a rank hung")
    assert dict(out) == {0: "ok", 1: "ok"}, (dict(out), [p.exitcode for p in procs])
