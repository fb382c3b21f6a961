"""N-GPU check of the partitioned layer-wise latent inference (models/tiled_inference.py).
Run under torchrun (one rank per GPU).  Every rank computes the latent matrix with the halo exchange
(spatial tiles) and compares it with its own single-rank full-graph result (no process group involved:
computed before init on the same weights).  Prints timing of both on a larger synthetic sample.
usage: torchrun --nproc-per-node N tools/dist_inference_check.py [cells]"""
import contextlib
import io
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np                               # noqa: E402
import torch                                     # noqa: E402
import torch.distributed as dist                 # noqa: E402

from nichecompass_b200.data import make_synthetic_adata          # noqa: E402
from nichecompass_b200.models import NicheCompass                # noqa: E402


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    cells = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
    adata = make_synthetic_adata(cells, 256, 6, 40, seed=0, lib_mean=150.)
    with contextlib.redirect_stdout(io.StringIO()):
        model = NicheCompass(adata, n_addon_gp=8, active_gp_thresh_ratio=0.01)
    model.is_trained_ = True
    model.model.to(torch.device("cuda", torch.cuda.current_device()))
    kw = dict(adata=adata, only_active_gps=False, dtype=np.float32, return_mu_std=True)
    singles = {}
    for dt in (torch.float32, torch.bfloat16):          # world 1: no process group yet
        model.model.set_compute_dtype(dt)
        singles[dt] = model.get_latent_representation(layerwise=True, partition="spatial", **kw)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    for dt in (torch.float32, torch.bfloat16):
        model.model.set_compute_dtype(dt)
        for rep in range(2):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            t0 = time.time()
            multi = model.get_latent_representation(layerwise=True, partition="auto", **kw)
            torch.cuda.synchronize()
            dt_s = time.time() - t0
        err = max(float(np.abs(m - s).max()) for m, s in zip(multi, singles[dt]))
        st = model.last_inference_stats_
        print(f"rank {rank}/{world} {dt}: max |multi - single| = {err:.3e}  own {st['n_own']} halo {st['n_halo']} "
              f"({dt_s * 1e3:.1f} ms incl. host graph plumbing)", flush=True)
        assert err <= 1e-6, err
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print("dist inference check OK")


if __name__ == "__main__":
    main()
