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
