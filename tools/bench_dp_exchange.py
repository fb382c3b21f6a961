"""
The fused gradient exchange + Adam kernel alone (csrc/dp_exchange.cu) at the C2 arena size (65 M parameters), all
ranks in lock step: time per step by CUDA events (max over ranks) against ncclAllReduce(AVG) + nc_adam_step_dyn.

    torchrun --nproc-per-node N tools/bench_dp_exchange.py [--numel 65000000] [--ctas 0,32,64,148,296]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from nichecompass_b200.train.optim import FlatAdam


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--numel", type=int, default=65_000_000)
    ap.add_argument("--ctas", default="0,32,64,296")
    ap.add_argument("--iters", type=int, default=30)
    a = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", torch.cuda.current_device())
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
    res = {}

    def timed(opt, exchange):
        ts = []
        for _ in range(a.iters):
            opt.flat_g.normal_()
            torch.cuda.synchronize()
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            if exchange:
                opt.all_reduce_grads()
            opt.step()
            e1.record()
            e1.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts = sorted(ts)[len(ts) // 4: -len(ts) // 4 or None]
        t = torch.tensor([sum(ts) / len(ts)], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)
    p = [torch.nn.Parameter(torch.randn(a.numel, device=dev) * 0.01)]
    plain = FlatAdam(p, lr=1e-3, bf16_shadow=True)
    res["nccl_allreduce_plus_adam_ms"] = timed(plain, True)
    del plain
    for c in [int(x) for x in a.ctas.split(",")]:
        p = [torch.nn.Parameter(torch.randn(a.numel, device=dev) * 0.01)]
        peer = FlatAdam(p, lr=1e-3, bf16_shadow=True, peer_group=True, peer_ctas=c)
        res[f"peer_fused_ms_ctas{c if c else 'default'}"] = timed(peer, False)
        peer.close()
    if rank == 0:
        n = a.numel
        res.update(n_gpus=world, numel=n,
                   link_bytes_per_rank={"read": 4 * n * (world - 1) // world, "written": 10 * n * (world - 1) // world // world * 1})
        print(json.dumps(res))
    dist.barrier()
    os._exit(0)


if __name__ == "__main__":
    main()
