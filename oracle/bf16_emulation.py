"""
TEST INFRASTRUCTURE (see oracle/__init__.py) -- bf16 rounding points of the CUDA tensor-core mode, replayed
inside the CPU oracle.

The bf16 mode of nichecompass_b200 stores a handful of tensors in bf16 (operands of the tcgen05 GEMMs and the
rows the GATv2 kernels stream) and keeps everything else -- accumulators, attention scores, softmax, losses,
parameter gradients -- in fp32.  Against the exact (fp64) oracle that costs ~3e-3 on losses and up to a few
1e-2 on individual gradients, which says nothing about whether the KERNELS are right.  With
``cfg["emulate_bf16"] = True`` the oracle rounds the same tensors at the same places (forward AND backward),
so what remains between the two is accumulation order only:

  forward   x_enc = bf16(log1p x);  W1, W_l, W_r, (W * mask) of the decoders: bf16 shadows;
            hidden = bf16(relu(.));  x_l, x_r = bf16(.);  z as a decoder operand = bf16(z)
  backward  d x_l, d x_r = bf16(.) (written by the attention backward);  d hidden (after the ReLU mask) = bf16(.);
            d logits as a GEMM operand = bf16(.) -- for d z and for the weight gradient of a DENSE mask block
            (the add-on programs); the weight gradient of a SPARSE mask block (prior programs) is an SDDMM over
            the fp32 d logits (nichecompass_b200/functional.py MaskedDecoderFn.backward).
Everything here follows nichecompass_b200/functional.py (EncoderGATv2Fn, MaskedDecoderFn); reference call
sites: nn/encoders.py:144-208, nn/layercomponents.py:53-81.
"""
import torch


def to_bf16(x: torch.Tensor) -> torch.Tensor:
    """Round to the nearest bf16 (ties to even), keep the dtype."""
    return x.to(torch.bfloat16).to(x.dtype)


class _Round(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, fwd: bool, bwd: bool):
        ctx.bwd = bwd
        return to_bf16(x) if fwd else x.clone()

    @staticmethod
    def backward(ctx, g):
        return (to_bf16(g) if ctx.bwd else g), None, None


def round_fwd(x):
    """bf16 value, exact gradient (weights' shadows, x_enc, z as an operand)."""
    return _Round.apply(x, True, False)


def round_bwd(x):
    """exact value, bf16 gradient (d hidden after the ReLU mask)."""
    return _Round.apply(x, False, True)


def round_both(x):
    """bf16 value and bf16 gradient (x_l / x_r rows)."""
    return _Round.apply(x, True, True)


class MaskedLinearBF16(torch.autograd.Function):
    """logits = bf16(z) @ bf16(W * mask)^T in exact arithmetic; backward as the CUDA path runs it:
    d z = bf16(d logits) @ bf16(W * mask);  d (W * mask) = d logits^T @ bf16(z) with d logits rounded to bf16
    only when the block's weight gradient is a dense GEMM (``dense_grad``)."""

    @staticmethod
    def forward(ctx, z, w_eff, dense_grad: bool):
        zr, wr = to_bf16(z), to_bf16(w_eff)
        ctx.save_for_backward(zr, wr)
        ctx.dense_grad = dense_grad
        return zr @ wr.t()

    @staticmethod
    def backward(ctx, dlogits):
        zr, wr = ctx.saved_tensors
        dl = to_bf16(dlogits)
        dz = dl @ wr
        dw = (dl if ctx.dense_grad else dlogits).t() @ zr
        return dz, dw, None


def mask_is_sparse(mask: torch.Tensor) -> bool:
    """functional.mask_info: a mask block takes the SDDMM route when fewer than 10 % of its entries are set."""
    return int((mask != 0).sum()) < 0.1 * mask.numel()
