#!/usr/bin/env python
"""
bench.py -- training throughput of the VGPGAE hot loop (BASELINE.json metric: training cells/sec
per epoch).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload c2|c1|c3|c4] [--edge-batch-size B] [--cells N_CELLS]

A "step" is one iteration of the reference hot loop (train/trainer.py:364-421): one node batch
through the omics decoders, one edge batch (B positive + B negative edges) through the graph
decoder, the loss, backward and the Adam update.  One epoch is ceil(E_train / B) steps, so

    cells/sec = n_cells / (steps_per_epoch * seconds_per_step)        (whole job, all ranks)

With N > 1 ranks (torchrun, one process per GPU) the batch stream is strided over the ranks and
gradients are all-reduced: per-GPU work per step is fixed, an epoch takes steps_per_epoch / N
global steps ("weak" scaling of the batch stream).

Legs of the default arm ("ours"):
  value         K steps with the count store resident in HBM, nothing read back per step.
  e2e           K steps through the same public Trainer API with the count matrix in pinned HOST
                memory (every step pulls the CSR rows of its sampled cells over the host link
                inside the timed region, the negatives come from the host generator) and the loss
                of EVERY step read on the host inside the timed region -- through a pinned double
                buffer, one step late (LossReader); e2e.blocking_read is the same leg with the
                reference's blocking .item() per step.
  roofline      the bf16 tcgen05 GEMM (>= 95 % of the step's FLOPs): per-launch CUDA-event times in
                a separate pass of the same K steps; the HBM-bound kernels are listed under "kernels".
  cpu_baseline  (rank 0, N = 1) the CPU oracle's training loop (oracle/train_ref.py: a restatement
                of the reference's PyTorch + PyG path) on the same data and batch sizes, timed on a
                bounded number of steps.
--impl reference times that CPU loop alone with the same metric / config; one CPU step of the C2
workload takes ~5 s, so the loop stops after --ref-wall-s seconds (default 150, at least 3 timed
steps) and reports the steps it timed ("steps") beside the K asked for ("steps_requested").
"""
import argparse
import contextlib
import io
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np
import torch

WORKLOADS = {
    # BASELINE.json configs[1]: the configuration the metric is quoted on
    "c2": dict(name="synthetic 1M-cell x 5k-gene Xenium-scale sample, kNN k=8, 1.5k prior + 100 add-on gene "
                    "programs, GATv2 encoder (4 heads), one-hop-norm labels, bf16",
               n_cells=1_000_000, n_genes=5000, k=8, n_prior_gp=1500, n_addon_gp=100, n_samples=1,
               lib_mean=300., dtype="bf16"),
    # BASELINE.json configs[0]: the reference's own CPU-runnable case (parity / quick runs)
    "c1": dict(name="synthetic 10k-cell x 2k-gene, kNN k=6, 300 prior + 100 add-on gene programs, GATv2, fp32",
               n_cells=10_000, n_genes=2000, k=6, n_prior_gp=300, n_addon_gp=100, n_samples=1,
               lib_mean=2000., dtype="f32"),
    # BASELINE.json configs[2]: multi-sample atlas with a sample covariate (needs --gpus 8 for the full size)
    "c3": dict(name="multi-sample atlas, 8 samples x 500k cells x 5k genes, kNN k=8, sample covariate embedding",
               n_cells=4_000_000, n_genes=5000, k=8, n_prior_gp=1500, n_addon_gp=100, n_samples=8,
               lib_mean=300., dtype="bf16"),
    # BASELINE.json configs[3]: multimodal RNA + ATAC (MaskedGeneExprDecoder + MaskedChromAccessDecoder)
    "c4": dict(name="multimodal RNA+ATAC, 2M cells x 5k genes + 50k peaks, kNN k=8, 1.5k prior + 100 add-on gene "
                    "programs, masked RNA and chromatin-accessibility decoders, bf16",
               n_cells=2_000_000, n_genes=5000, n_peaks=50_000, k=8, n_prior_gp=1500, n_addon_gp=100, n_samples=1,
               lib_mean=300., atac_lib_mean=600., dtype="bf16"),
}
LAMBDAS = dict(lambda_edge_recon=500000., lambda_gene_expr_recon=300., lambda_l1_addon=30.,
               lambda_group_lasso=0., lambda_l1_masked=0.)       # NicheCompass.train defaults


def config_dict(wl, edge_bs, node_bs, steps_per_epoch, world):
    """Workload description shared by both arms (the driver compares the two dicts key by key)."""
    d = {"workload": wl["name"], "n_cells": wl["n_cells"], "n_genes": wl["n_genes"], "k": wl["k"],
         "edge_batch_size": int(edge_bs), "node_batch_size": int(node_bs), "steps_per_epoch": int(steps_per_epoch),
         "global_steps_per_epoch": int(math.ceil(steps_per_epoch / world)), "parallelism": f"dp{world}",
         "l2_policy": "inputs larger than L2: every step samples a new sub-graph and reads all parameters "
                      "(260 MB fp32 at C2) plus fresh activations (> 126 MB L2); no flush needed"}
    if wl.get("n_peaks"):
        d["n_peaks"] = wl["n_peaks"]
    return d


def log(*a):
    print(*a, file=sys.stderr, flush=True)


@contextlib.contextmanager
def quiet():
    with contextlib.redirect_stdout(io.StringIO()):
        yield


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--ref-wall-s", type=float, default=150.0,
                    help="--impl reference: stop the CPU loop after this many seconds (>= 3 timed steps); a CPU step of "
                         "the C2 workload takes ~5 s, so K steps are a bounded sample of the epoch")
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--edge-batch-size", type=int, default=256, help="reference default (NicheCompass.train)")
    ap.add_argument("--cells", type=int, default=None, help="override the number of cells (quick runs)")
    ap.add_argument("--cpu-steps", type=int, default=None, help="steps of the CPU baseline (default: sized to ~20 s)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--epoch", action="store_true",
                    help="additionally run ONE REAL EPOCH (every training batch once: steps_per_epoch / N iterations "
                         "per rank) and the validation pass, and report measured (not extrapolated) cells/s")
    ap.add_argument("--seed", type=int, default=0)
    return ap.parse_args()


def build_adata(wl: dict, seed: int):
    from nichecompass_b200.data import make_synthetic_adata_large
    t0 = time.time()
    adata = make_synthetic_adata_large(wl["n_cells"], wl["n_genes"], wl["k"], wl["n_prior_gp"], seed=seed,
                                       n_samples=wl["n_samples"], lib_mean=wl["lib_mean"])
    log(f"[bench] synthetic data: {adata} in {time.time() - t0:.1f} s")
    if wl.get("n_peaks"):
        from nichecompass_b200.data import add_synthetic_atac
        adata.uns["_atac"] = add_synthetic_atac(adata, wl["n_peaks"], seed=seed, lib_mean=wl["atac_lib_mean"])
        log(f"[bench] synthetic ATAC: {adata.uns['_atac']} in {time.time() - t0:.1f} s")
    return adata


def build_adata_shared(wl: dict, seed: int, rank: int, world: int):
    """N > 1 ranks on one box: rank 0 draws the synthetic dataset once and hands it to the other ranks through
    /dev/shm (8 ranks drawing 4 M cells each would need ~3x the dataset per rank in host memory at the same time)."""
    if world == 1:
        return build_adata(wl, seed)
    import pickle
    import torch.distributed as dist
    path = f"/dev/shm/nc_bench_{os.getuid()}_{wl['n_cells']}_{wl['n_genes']}_{wl.get('n_peaks', 0)}_{seed}.pkl"
    adata = None
    if rank == 0:
        adata = build_adata(wl, seed)
        with open(path + ".tmp", "wb") as fh:
            pickle.dump(adata, fh, protocol=5)
        os.replace(path + ".tmp", path)
    dist.barrier()
    if rank != 0:
        with open(path, "rb") as fh:
            adata = pickle.load(fh)
    dist.barrier()
    if rank == 0:
        with contextlib.suppress(OSError):
            os.remove(path)
    return adata


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu_index, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.gpu_index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(smax)), "reasons": sorted(reasons),
                "power_w_max": float(max(power)), "samples": len(sm)}


# ----------------------------------------------------------------------------- per-kernel timing
class KernelTimer:
    """CUDA events around every C-ABI call on the launching stream (installed for the profiling pass
    only).  Algorithmic work per call is derived from the call's own arguments:
      GEMM      2*M*N*K flops
      GATv2     (slots + n_dst) * ld * s bytes of x_l / x_r reads + alpha + output (fwd), x2 + gradients (bwd)
      SpMM      gathered rows * n_cols * s + output; from the count store: nnz * 8 B (value + column id)
    (DESIGN.md 'algorithmic bytes' has the derivations)."""

    def __init__(self, store_nnz_per_row: float):
        self.records = []          # (name, start_event, end_event, work, unit)
        self.nnz_row = store_nnz_per_row
        self._start = None
        self._adam_untouched = {}  # exp_avg_sq pointer -> parameters in chunks the Adam kernel skips

    @staticmethod
    def _esize(code):
        return 2 if code == 1 else 4

    def work(self, name, a):
        if name == "nc_gemm_bf16":
            return 2.0 * a[9] * a[10] * a[11], "flop"
        if name == "nc_gemm_f32":
            return 2.0 * a[8] * a[9] * a[10], "flop"
        if name == "nc_gatv2_fwd":
            s, ld, n_dst, alpha, ldo = self._esize(a[2]), a[3], a[8], a[17], a[16]
            n_conv, H = a[11], a[12]
            slots = alpha.numel() // max(n_conv * H, 1)
            return (slots + n_dst) * ld * s + alpha.numel() * 4 + n_dst * ldo * 4 + slots * 4, "byte"
        if name == "nc_gatv2_bwd":
            s, ld, n_dst, n_src, alpha = self._esize(a[2]), a[3], a[7], a[8], a[20]
            n_conv, H = a[14], a[15]
            slots = alpha.numel() // max(n_conv * H, 1)
            # every x_l / x_r row read ONCE (the per-source pass works from the stored LeakyReLU sign bits, 1 bit per
            # element, not from a second read of the rows), d x_l (all sources) and d x_r (targets) written,
            # alpha read + d score written and read
            return (slots + n_dst) * ld * s + (n_src + n_dst) * ld * s + 3 * alpha.numel() * 4 \
                + 2 * slots * ld // 8, "byte"
        if name == "nc_spmm_csr_store":
            n_rows, n_cols = a[10], a[3]
            deg = 10.0                                     # k=8 symmetrised kNN: ~9.2 neighbours + self
            return n_rows * deg * self.nnz_row * 8 + n_rows * n_cols * 4, "byte"
        if name == "nc_csr_gather_rows":
            n_rows, ld, s = a[5], a[9], self._esize(a[8])
            return n_rows * self.nnz_row * 8 + n_rows * ld * s, "byte"
        if name == "nc_spmm_csr":
            s, n_cols, n_rows = self._esize(a[1]), a[3], a[10]
            return n_rows * 10.0 * n_cols * s + n_rows * n_cols * self._esize(a[12]), "byte"
        if name == "nc_adam_step":
            return a[4] * 28.0, "byte"
        if name == "nc_adam_step_dyn":              # p, g, m, v read; p, m, v, bf16 shadow written; g cleared
            # 16-byte chunks whose g, m and v are all zero (masked-out decoder entries: they never see a gradient) are
            # read (16 B per parameter) and left alone by the kernel: count them once, from the moment buffers
            n, key = a[4], a[3].data_ptr()
            if key not in self._adam_untouched:
                n4 = (n // 4) * 4
                idle = (a[3][:n4].view(-1, 4) == 0).all(1) & (a[2][:n4].view(-1, 4) == 0).all(1)
                self._adam_untouched[key] = int(idle.sum().item()) * 4 if float(a[9]) == 0.0 else 0
            u = self._adam_untouched[key]
            return (n - u) * 34.0 + u * 16.0, "byte"
        if name == "nc_dp_adam_step":               # per rank: HBM + NVLink bytes of its shard (see csrc/dp_exchange.cu)
            n, w = a[6], a[8]
            return n * (4.0 + 20.0 / w + 10.0), "byte"
        if name == "nc_nb_loss_fused":
            return a[6] * a[7] * 4.0 * 4, "byte"
        if name == "nc_cast_bf16":
            return a[4] * a[7] * 6.0, "byte"
        return 0.0, "byte"

    def before(self, name, args):
        self._start = torch.cuda.Event(enable_timing=True)
        self._start.record()

    def after(self, name, args):
        end = torch.cuda.Event(enable_timing=True)
        end.record()
        w, unit = self.work(name, args)
        self.records.append((name, self._start, end, w, unit))

    def summary(self) -> dict:
        torch.cuda.synchronize()
        agg = {}
        for name, s, e, w, unit in self.records:
            d = agg.setdefault(name, {"launches": 0, "ms": 0.0, "work": 0.0, "unit": unit})
            d["launches"] += 1
            d["ms"] += s.elapsed_time(e)
            d["work"] += w
        total = sum(d["ms"] for d in agg.values()) or 1.0
        for d in agg.values():
            d["share"] = d["ms"] / total
            rate = d["work"] / (d["ms"] * 1e-3) if d["ms"] > 0 else 0.0
            d["rate"] = rate / (1e12 if d["unit"] == "flop" else 1e9)
            d["rate_unit"] = "TFLOP/s" if d["unit"] == "flop" else "GB/s"
        return agg


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            p = json.load(fh)
        return {"hbm_gbs": p["hbm_gbs"], "bf16_burst": p["bf16_tflops"], "bf16_sustained": p["bf16_tflops_sustained"],
                "source": "MEASURED_PEAKS.json"}
    return {"hbm_gbs": 6650.0, "bf16_burst": 1590.0, "bf16_sustained": 1400.0, "source": "fallback"}


# ----------------------------------------------------------------------------- our arm
def dist_setup(n_gpus: int):
    """One process per GPU.  Under torchrun: NCCL process group; plain `python bench.py`: rank 0 of 1."""
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1 and os.environ.get("NC_PIN_CORES", "1") == "1" and hasattr(os, "sched_setaffinity"):
        # N Python processes on one box: give every rank its own slice of the cores it may run on (they all start
        # with the same affinity mask, and the slowest rank's enqueue loop sets the pace of a step)
        try:
            cores = sorted(os.sched_getaffinity(0))
            per = len(cores) // world
            if per >= 1:
                os.sched_setaffinity(0, set(cores[local * per:(local + 1) * per]))
                torch.set_num_threads(max(per, 1))
        except OSError:
            pass
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if world != n_gpus:
        log(f"[bench] WARNING --gpus {n_gpus} but WORLD_SIZE {world}: using WORLD_SIZE")
    return rank, world, local


def barrier(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(x: float, world: int) -> float:
    if world == 1:
        return x
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def make_trainer(adata, wl, args, store_location="device"):
    from nichecompass_b200.models import NicheCompass
    from nichecompass_b200.train import Trainer
    cov = ["batch"] if wl["n_samples"] > 1 else None
    atac = adata.uns.get("_atac")
    with quiet():
        model = NicheCompass(adata, adata_atac=atac, n_addon_gp=wl["n_addon_gp"], cat_covariates_keys=cov,
                             seed=args.seed)
        dtype = torch.bfloat16 if wl["dtype"] == "bf16" else torch.float32
        tr = Trainer(adata, model.model, adata_atac=atac, cat_covariates_keys=cov, edge_batch_size=args.edge_batch_size,
                     seed=args.seed, monitor=False, compute_dtype=dtype, store_location=store_location)
        tr.setup_training(lr=1e-3, **LAMBDAS)
    return model, tr


def ncu_traffic():
    """DRAM bytes per GEMM launch (read + write) from the committed ncu capture of one C2-shaped step
    (profiles/r2f_gemm_traffic.json, from the full-metric capture of tools/r2/call_final1.sh); None if absent."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r2f_gemm_traffic.json")
    try:
        with open(path) as fh:
            return float(json.load(fh)["dram_bytes_per_launch"])
    except (OSError, KeyError, ValueError):
        return None


def spmm_metric(tr):
    """BASELINE.json's second metric, "SpMM HBM GB/s": the one-hop gcn-norm node-label aggregation
    (nn/aggregators.py:176-213) of 16384 seed cells of THIS workload over their sampled neighbourhood,
    algorithmic bytes / CUDA-event time with L2 flushed -- from dense fp32 / bf16 rows (the reference's
    layout) and straight from the CSR count store (the product path).  tools/bench_spmm.py."""
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "tools"))
    import bench_spmm
    data = tr.node_masked_data
    peaks = measured_peaks()
    try:
        r = bench_spmm.measure(data.store, data.device_graph(), n_rows=16384, verbose=False)
    except Exception as e:                                  # noqa: BLE001 -- a side metric must not kill the line
        return {"error": str(e)}
    out = {"unit": "GB/s", "peak": peaks["hbm_gbs"], "shape": r["shape"]}
    for k in ("dense_f32", "dense_bf16", "csr_store"):
        out[k] = {"achieved": r[k]["gbs"], "frac": r[k]["gbs"] / peaks["hbm_gbs"], "ms": r[k]["ms"],
                  "algorithmic_bytes": r[k]["algorithmic_bytes"]}
    return out


class LossReader:
    """Every step's ``global_loss`` on the host without stalling the launch pipeline: step i's scalar is
    copied into a pinned double buffer right after its kernels are enqueued and READ while step i + 1 is
    being enqueued (one step late; the last one after the loop).  Each of the n steps is read inside the
    timed region."""

    def __init__(self):
        self.buf = torch.empty(2, dtype=torch.float32).pin_memory()
        self.ev = [None, None]
        self.i = 0
        self.values = []

    def push(self, loss: torch.Tensor):
        k = self.i & 1
        self.buf[k:k + 1].copy_(loss.detach().reshape(1), non_blocking=True)
        self.ev[k] = torch.cuda.Event()
        self.ev[k].record()
        if self.i > 0:
            self._read(k ^ 1)
        self.i += 1

    def _read(self, k):
        self.ev[k].synchronize()
        self.values.append(float(self.buf[k]))

    def flush(self):
        if self.i > 0:
            self._read((self.i - 1) & 1)
        return self.values


def run_steps(tr, it, n, read_loss=False):
    """n training iterations; returns the host losses when ``read_loss``: True = blocking ``.item()`` per
    step (the reference's train/trainer.py:404-411), "pipelined" = ``LossReader``."""
    out = []
    reader = LossReader() if read_loss == "pipelined" else None
    for _ in range(n):
        eb, nb = next(it)
        L = tr.train_iteration(eb, nb)
        if reader is not None:
            reader.push(L["global_loss"])
        elif read_loss:
            out.append(float(L["global_loss"].item()))       # D2H every step (the reference's .item())
    return reader.flush() if reader is not None else out


def timed(tr, it, steps, world, read_loss=False):
    barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    t0 = time.perf_counter()
    losses = run_steps(tr, it, steps, read_loss)
    timed.last_host_ms = (time.perf_counter() - t0) * 1e3 / max(steps, 1)      # enqueue time: no device wait inside
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    barrier(world)
    return max_over_ranks(ms, world), losses


class BytesProbe:
    """Bytes one step moves between host and device in the e2e leg, counted from the tensors involved:
    H2D = 8 B per non-zero of every count-store row a kernel reads (encoder gather of all sampled cells
    of both batches, label rows and one-hop aggregation rows of the node batch) + the negatives;
    D2H = sampler size read-backs + the loss."""

    def __init__(self, store):
        self.rowptr = store.rowptr
        self.acc = torch.zeros((), dtype=torch.int64, device=store.rowptr.device)

    def wrap(self, it):
        for eb, nb in it:
            for b, is_node in ((eb, False), (nb, True)):
                n_id = b.n_id.long()
                nnz = self.rowptr[n_id + 1] - self.rowptr[n_id]
                self.acc += nnz.sum()
                if is_node:
                    n = b.batch_size
                    g = b.graph
                    live = torch.arange(g.col.numel(), device=g.col.device) < g.rowptr[n]
                    self.acc += 2 * nnz[:n].sum() + (nnz[g.col.long()] * live).sum()
            yield eb, nb

    def per_step(self, n_steps: int, edge_batch: int):
        nnz = int(self.acc.item()) / max(n_steps, 1)
        h2d = int(nnz * 8 + 2 * edge_batch * 4 + 2 * 64)
        d2h = 2 * 32 + 4 + 4
        return h2d, d2h


def batch_stream(tr):
    from nichecompass_b200.train.utils import _cycle_iterable

    def gen():
        while True:                                           # epochs back to back
            yield from tr.joint_batches(tr.edge_train_loader, tr.node_train_loader)
    return gen()


def run_ours(args):
    from nichecompass_b200 import _lib
    from nichecompass_b200.data import CSRFeatureStore
    rank, world, local = dist_setup(args.gpus)
    wl = dict(WORKLOADS[args.workload])
    if args.cells:
        wl["n_cells"] = args.cells
    adata = build_adata_shared(wl, args.seed, rank, world)
    model, tr = make_trainer(adata, wl, args)
    steps_per_epoch = tr.edge_train_loader.n_batches_total
    global_steps_per_epoch = math.ceil(steps_per_epoch / world)
    store = tr.node_masked_data.store
    nnz_row = store.nnz / max(store.n_rows, 1)
    log(f"[bench] rank {rank}/{world}: {steps_per_epoch} steps/epoch, node batch {tr.node_batch_size_}, "
        f"edge batch {tr.edge_batch_size_}, store nnz/row {nnz_row:.1f}")

    it = batch_stream(tr)
    # warm-up: at least 40 iterations, so that the size buckets of the captured iteration (one CUDA-graph capture
    # each, ~0.1 s) exist before the timed region; `warmup` in the result line is the number actually run
    n_warm = max(args.warmup, 40) if tr.cuda_graphs_ else args.warmup
    run_steps(tr, it, n_warm)
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    n0 = _lib.launch_count()
    ms_total, _ = timed(tr, it, args.steps, world)
    host_ms = timed.last_host_ms
    launches = _lib.launch_count() - n0
    clock_info = clocks.stop() if rank == 0 else None
    ms_step = ms_total / args.steps
    value = wl["n_cells"] / (global_steps_per_epoch * ms_step * 1e-3)

    # ---- per-kernel pass (same steps, events around every launch)
    timer = KernelTimer(nnz_row)
    _lib.set_call_hook(timer)
    run_steps(tr, it, args.steps)
    _lib.set_call_hook(None)
    kern = timer.summary()

    # ---- end-to-end leg: count matrix in pinned host memory, loss read back every step
    e2e = None
    if not args.no_e2e:
        ds = tr.node_masked_data.dataset
        dev_store = ds._store
        ds._store = CSRFeatureStore.from_scipy(ds._x_rna, ds.device, ds._x_atac, location="host")
        probe = BytesProbe(ds._store)
        it2 = batch_stream(tr)
        n_probe = max(args.warmup, 40 if tr.cuda_graphs_ else 3)      # the host-store buckets are captured here
        run_steps(tr, probe.wrap(it2), n_probe, read_loss=True)      # warm-up, also counts the bytes a step moves
        h2d, d2h = probe.per_step(n_probe, tr.edge_batch_size_)
        ms_blk, losses_blk = timed(tr, it2, args.steps, world, read_loss=True)
        ms_e2e, losses = timed(tr, it2, args.steps, world, read_loss="pipelined")
        assert len(losses) == args.steps
        ds._store = dev_store
        e2e_ms_step = ms_e2e / args.steps
        blk_ms_step = ms_blk / args.steps
        e2e = {"value": wl["n_cells"] / (global_steps_per_epoch * e2e_ms_step * 1e-3), "unit": "cells/s",
               "ms_per_step": e2e_ms_step, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "how": "Trainer.train_iteration via the public Trainer API; CSR count matrix in pinned host memory, "
                      "the rows of every step pulled over the host link by the gather/SpMM kernels (those of step "
                      "i+1 on a side stream while step i computes), negatives drawn on the host and copied "
                      "H2D, sampler sizes + global_loss copied D2H each step; the loss of step i is copied to a "
                      "pinned buffer and read on the host while step i+1 is enqueued (every step's value is read "
                      "inside the timed region, LossReader); blocking_read = the same with a blocking .item() "
                      "per step as in the reference loop",
               "blocking_read": {"value": wl["n_cells"] / (global_steps_per_epoch * blk_ms_step * 1e-3),
                                 "ms_per_step": blk_ms_step},
               "last_global_loss": losses[-1] if losses else None}

    epoch = None
    if args.epoch:
        # the metric as its name says: one pass over every training edge batch (train/trainer.py:364-421 loop), then
        # the validation pass of the same epoch (train/trainer.py:492-589), timed on the device, max over ranks
        barrier(world)
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record()
        n_it = 0
        for eb, nb in tr.joint_batches(tr.edge_train_loader, tr.node_train_loader):
            tr.train_iteration(eb, nb)
            n_it += 1
        e1.record()
        tr.iter_logs = {"n_val_iter": 0}
        metrics = tr.eval_epoch({})
        e2.record()
        torch.cuda.synchronize()
        t_train = max_over_ranks(e0.elapsed_time(e1), world) * 1e-3
        t_val = max_over_ranks(e1.elapsed_time(e2), world) * 1e-3
        barrier(world)
        epoch = {"iterations_per_rank": n_it, "train_seconds": t_train, "validation_seconds": t_val,
                 "validation_iterations": tr.iter_logs["n_val_iter"],
                 "cells_per_s_train_only": wl["n_cells"] / t_train,
                 "cells_per_s_with_validation": wl["n_cells"] / (t_train + t_val),
                 "val_auroc": metrics.get("auroc_score"),
                 "note": "measured over the whole epoch; the validation batches are sharded over the ranks and the metrics "
                         "computed on the gathered logits"}
    result = None
    if rank == 0:
        peaks = measured_peaks()
        g = kern.get("nc_gemm_bf16") or kern.get("nc_gemm_f32") or {"launches": 0, "ms": 0., "work": 0., "rate": 0., "share": 0.}
        peak_tf = peaks["bf16_sustained"]
        roofline = {"kernel": "gemm_bf16_tcgen05 (nc_gemm_bf16)" if "nc_gemm_bf16" in kern else "gemm_f32_simt",
                    "bound": "tensor", "achieved": g["rate"], "peak": peak_tf, "unit": "TFLOP/s",
                    "frac": g["rate"] / peak_tf if peak_tf else None, "traffic": ncu_traffic(),
                    "traffic_note": "dram__bytes_read + dram__bytes_write per launch from the committed ncu --set full "
                                    "capture of the same 13 launches (profiles/r2f_gemm_traffic.json); ncu cannot run "
                                    "inside a timed bench, so this figure moves only when the capture is redone",
                    "peak_source": f"{peaks['source']} bf16_tflops_sustained (kernel timed inside a long step)",
                    "launches_per_step": g["launches"] / args.steps, "share_of_step_kernel_time": g["share"],
                    "avg_launch_ms": g["ms"] / max(g["launches"], 1),
                    "flop_per_launch": g["work"] / max(g["launches"], 1)}
        kernels = {k: {"launches_per_step": v["launches"] / args.steps, "ms_per_step": v["ms"] / args.steps,
                       "share": round(v["share"], 4), "achieved": round(v["rate"], 2), "unit": v["rate_unit"],
                       "frac_of_peak": round(v["rate"] / (peak_tf if v["unit"] == "flop" else peaks["hbm_gbs"]), 4)}
                   for k, v in sorted(kern.items(), key=lambda kv: -kv[1]["ms"])}
        hb = [(k, v) for k, v in kern.items() if v["unit"] == "byte" and v["work"] > 0]
        roofline_hbm = None
        if hb:
            k, v = max(hb, key=lambda kv: kv[1]["ms"])
            roofline_hbm = {"kernel": k, "bound": "hbm", "achieved": v["rate"], "peak": peaks["hbm_gbs"], "unit": "GB/s",
                            "frac": v["rate"] / peaks["hbm_gbs"], "traffic": None,
                            "avg_launch_ms": v["ms"] / max(v["launches"], 1),
                            "bytes_per_launch": v["work"] / max(v["launches"], 1),
                            "note": "the HBM-bound kernel with the largest share of the step; algorithmic bytes as "
                                    "DESIGN.md section 4 defines them (live attention slots only)"}
        result = {
            "metric": "training cells/sec per epoch", "value": value, "unit": "cells/s", "n_gpus": world,
            "steps": args.steps, "warmup": n_warm, "warmup_requested": args.warmup, "ms_per_step": ms_step,
            "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": wl["dtype"], "data": "synthetic",
            "config": config_dict(wl, tr.edge_batch_size_, tr.node_batch_size_, steps_per_epoch, world),
            "parallelism_note": (f"dp{world} over the batch stream, flat-gradient exchange over NVLink" if world > 1
                                 else "single GPU"),
            "clocks": clock_info, "e2e": e2e, "gpu_launches": int(launches),
            "host_enqueue_ms_per_step": host_ms, "cuda_graphs": dict(tr.graph_stats_),
            "roofline": roofline, "roofline_hbm": roofline_hbm, "kernels": kernels, "epoch": epoch,
        }
    return result, (adata, wl, tr, rank, world)


# ----------------------------------------------------------------------------- CPU baseline / reference arm
def cpu_reference(adata, wl, args, steps=None, warmup=1, budget_s=20.0, wall_s=None):
    """The reference's hot loop on the host cores: oracle/train_ref.py (PyTorch CPU restatement of the
    reference module + PyG semantics + loaders; the unmodified reference cannot be imported on the box --
    torch_geometric / torch_sparse / anndata / mlflow are not installed and /root/reference does not
    travel).  fp32, torch.set_num_threads(all cores).  Returns (cells/s, seconds per step, description)."""
    from oracle import train_ref as TR
    from oracle.vgpgae_ref import init_params
    from tests import helpers as H
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(args.seed)
    cfg = H.oracle_cfg_from_adata(adata, n_addon_gp=wl["n_addon_gp"], active_gp_thresh_ratio=0.01)
    params = init_params(cfg, seed=args.seed, dtype=torch.float32)
    params0 = {k: v.detach().clone() for k, v in params.items()}
    first_logs = []
    t0 = time.time()
    ld = TR.build_loaders(TR.SparseRows(adata.X), adata.obsp["spatial_connectivities"], args.edge_batch_size)
    ref = TR.TrainerRef(cfg, params, ld.edge_loader, ld.node_loader, lr=1e-3,
                        loss_kwargs={k: v for k, v in LAMBDAS.items()})
    log(f"[bench] cpu reference set-up {time.time() - t0:.1f} s, {cores} threads")
    it = zip(ref.edge_loader, TR.R.cycle_iterable(ref.node_loader))
    t_start = time.perf_counter()
    done_w = 0
    for _ in range(warmup):
        e, nb = next(it)
        t1 = time.perf_counter()
        first_logs.append(ref.iteration(e, nb))
        first = time.perf_counter() - t1
        done_w += 1
        if wall_s is not None and time.perf_counter() - t_start > 0.2 * wall_s:
            break                                      # a 5 s CPU step: the warm-up is bounded by the clock too
    warmup = done_w
    if steps is None:
        steps = int(min(max(budget_s / max(first, 1e-3), 2), 50))
    t1 = time.perf_counter()
    done = 0
    for _ in range(steps):
        e, nb = next(it)
        log_ = ref.iteration(e, nb)
        if len(first_logs) < 2:
            first_logs.append(log_)
        done += 1
        if wall_s is not None and done >= 3 and time.perf_counter() - t_start > wall_s:
            break                                      # bounded sample: the per-step time is what the metric needs
    steps = done
    sec_step = (time.perf_counter() - t1) / steps
    steps_per_epoch = math.ceil(ld.n_edges_train / args.edge_batch_size)
    value = wl["n_cells"] / (steps_per_epoch * sec_step)
    sample = (f"{steps} hot-loop steps (after {warmup} warm-up) of the same workload: edge batch "
              f"{args.edge_batch_size}, node batch {ld.node_batch_size}, {steps_per_epoch} steps/epoch; sampling + "
              f"dense row slicing + forward/backward/Adam, fp32, extrapolated to the epoch")
    cpu_reference.last = {"cfg": cfg, "params0": params0, "logs": first_logs[:2],
                          "noise": [t.clone() for t in ref.noise_log[:4]]}
    cpu_reference.last_shape = {"node_batch_size": int(ld.node_batch_size), "steps_per_epoch": int(steps_per_epoch)}
    return value, sec_step, (steps, warmup), cores, sample


def parity_check(adata, wl, args):
    """The first two training iterations of BOTH arms on identical batches (same seed -> bit-identical samplers),
    identical initial parameters (the oracle's) and identical reparameterisation noise: every loss term of the CUDA
    path (in the benchmarked dtype) beside the CPU loop's.  Runs after ``cpu_reference`` (which keeps what it drew)."""
    from nichecompass_b200.models import NicheCompass
    from nichecompass_b200.train import Trainer
    ref = cpu_reference.last
    if len(ref["logs"]) < 2 or len(ref["noise"]) < 4:
        return None
    torch.manual_seed(args.seed)
    with quiet():
        model = NicheCompass(adata, n_addon_gp=wl["n_addon_gp"], seed=args.seed)
        model.model.load_state_dict({k: v.float() for k, v in ref["params0"].items()}, strict=True)
        dtype = torch.bfloat16 if wl["dtype"] == "bf16" else torch.float32
        tr = Trainer(adata, model.model, edge_batch_size=args.edge_batch_size, seed=args.seed, monitor=False,
                     compute_dtype=dtype, pipeline_sampler=False, cuda_graphs=False)
        tr.setup_training(lr=1e-3, **LAMBDAS)
    tr.model._eps_queue = [e.clone() for e in ref["noise"]]
    gpu = []
    for i, (eb, nb) in enumerate(tr.joint_batches(tr.edge_train_loader, tr.node_train_loader)):
        gpu.append({k: float(v) for k, v in tr.train_iteration(eb, nb).items()})
        if i == 1:
            break
    worst, where = 0.0, None
    for it_, (a, b) in enumerate(zip(gpu, ref["logs"])):
        for k, v in b.items():
            d = abs(a[k] - v) / max(abs(v), 1e-30)
            if d > worst:
                worst, where = d, f"iteration {it_}: {k}"
    return {"steps": 2, "dtype": wl["dtype"], "max_rel_diff": worst, "where": where,
            "global_loss_cuda": [g["global_loss"] for g in gpu],
            "global_loss_cpu": [float(b["global_loss"]) for b in ref["logs"]],
            "how": "same seed (bit-identical sampled batches), the CPU loop's initial parameters and noise injected "
                   "into the CUDA module; all loss terms compared.  The CPU arm is the EXACT fp32 loop: the bf16 "
                   "storage of the CUDA arm costs 1e-3 .. 3e-3 here (iteration 1 runs after one Adam step of either "
                   "arm); against an oracle that rounds the same tensors to bf16 the distance is < 1e-3 "
                   "(tests/test_c1_parity_gpu.py, which also pins gradients and latents at C1)"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return None
    wl = dict(WORKLOADS[args.workload])
    if args.cells:
        wl["n_cells"] = args.cells
    adata = build_adata(wl, args.seed)
    value, sec_step, (steps, warm), cores, sample = cpu_reference(adata, wl, args, steps=args.steps,
                                                                  warmup=max(args.warmup, 1), wall_s=args.ref_wall_s)
    return {
        "impl": "reference", "metric": "training cells/sec per epoch", "value": value, "unit": "cells/s",
        "n_gpus": args.gpus, "steps": steps, "steps_requested": args.steps, "warmup": warm,
        "warmup_requested": args.warmup,
        "ms_per_step": sec_step * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(wl, args.edge_batch_size, cpu_reference.last_shape["node_batch_size"],
                              cpu_reference.last_shape["steps_per_epoch"], args.gpus),
        "parallelism_note": "rank 0 only: one process, all host threads (the reference has no data parallelism)",
        "cpu_baseline": {"value": value, "unit": "cells/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "cells/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def emit(res, fd):
    """The ONE JSON line, written to the process's original stdout."""
    os.write(fd, (json.dumps(res) + "\n").encode())


def main():
    args = parse_args()
    # Libraries write to fd 1 behind Python's back (NCCL prints "NCCL version ..." at NCCL_DEBUG=VERSION and
    # =WARN): everything but the result line goes to stderr.
    sys.stdout.flush()
    out_fd = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        res = run_reference(args)
        if res is not None:
            emit(res, out_fd)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (nichecompass_b200 has no CPU path); "
                         "use --impl reference for the CPU baseline")
    res, (adata, wl, tr, rank, world) = run_ours(args)
    if rank == 0 and world == 1:
        res["spmm"] = spmm_metric(tr)
    if rank == 0 and world == 1 and not args.no_cpu_baseline and not wl.get("n_peaks"):
        try:
            import psutil
            free_gb = psutil.virtual_memory().available / 2 ** 30
        except Exception:                                   # noqa: BLE001
            free_gb = 1e9
        if free_gb < 96 and wl["n_cells"] >= 500_000:
            res["cpu_baseline"] = {"value": None, "unit": "cells/s", "cores": os.cpu_count(), "kind": "port",
                                   "sample": f"skipped: only {free_gb:.0f} GB host memory available"}
        else:
            del tr
            torch.cuda.empty_cache()
            value, sec_step, (steps, _), cores, sample = cpu_reference(adata, wl, args, steps=args.cpu_steps)
            res["cpu_baseline"] = {"value": value, "unit": "cells/s", "cores": cores, "kind": "port",
                                   "sample": sample, "ms_per_step": sec_step * 1e3}
            try:
                res["parity"] = parity_check(adata, wl, args)
            except Exception as e:                              # noqa: BLE001 -- a side check must not kill the line
                res["parity"] = {"error": str(e)}
    if rank == 0:
        emit(res, out_fd)
    if world > 1:
        import torch.distributed as dist
        # every rank still references the peers' arenas (cudaIpc mappings): drop the CUDA graphs that hold them, then
        # unmap together before the process group goes away
        tr._graph_buckets.clear()
        torch.cuda.synchronize()
        dist.barrier()
        sys.stderr.flush()
        # leave without interpreter teardown: destructors of CUDA graphs / cudaIpc mappings / NCCL communicators of N
        # processes that reference each other's memory have hung a rank at exit (after the result line was written)
        os._exit(0)


if __name__ == "__main__":
    main()
