"""Thin, allocation-explicit Python wrappers over the C ABI (one function per native entry point).

Conventions: activations are ``torch.bfloat16`` tensors with logical shape ``[N,C,H,W]`` in
``channels_last`` memory format (physical NHWC) or plain ``[M,K]`` matrices; all wrappers launch on the
current CUDA stream and never synchronise, so everything here is CUDA-graph capturable.

What each group replaces in the reference (all library calls there): cuDNN conv / BN and cuBLAS linear of
``torchvision.models.resnet18`` and the inline ``Net`` (``notebooks/code/cifar10-distributed-smddp-gpu.py:30-51``), ATen pool / ReLU /
``CrossEntropyLoss`` / ``optim.SGD`` (``:150-156,162-168``), the torchvision transform pipeline (``:55-62``), and for the MNTD zoo
``torch.stft`` + mel matmul + ``nn.LSTM`` (``model_lib/audio_rnn_model.py:11,25-36``), the (k x 300) convs of the text CNN
(``model_lib/rtNLP_cnn_model.py:28-31,49-58``) and the trojan stamping (``model_lib/cifar10_cnn_model.py:77-84``).
"""
from __future__ import annotations

import ctypes
from ctypes import byref, c_float, c_int, c_size_t, c_void_p
from typing import Optional, Tuple

import torch

from . import _lib as L

CL = torch.channels_last


class ConvGeom(ctypes.Structure):
    _fields_ = [("N", c_int), ("H", c_int), ("W", c_int), ("C", c_int), ("K", c_int),
                ("R", c_int), ("S", c_int), ("stride", c_int), ("pad", c_int)]


# --------------------------------------------------------------------------------------------
# per-device scratch (BN partials + ticket counters, softmax ticket) — zeroed once, self-resetting
# --------------------------------------------------------------------------------------------
_SCRATCH = {}
SCRATCH_BYTES = 8 << 20


def scratch(device: torch.device) -> torch.Tensor:
    key = (device.type, device.index if device.index is not None else torch.cuda.current_device())
    t = _SCRATCH.get(key)
    if t is None:
        t = torch.zeros(SCRATCH_BYTES, dtype=torch.uint8, device=device)
        _SCRATCH[key] = t
    return t


class _SplitK:
    """``SPLIT_K[0] = False`` pins conv/linear fwd+dgrad to one CTA per tile (``B200DDP_CONV_SPLITK=0``)."""

    def __init__(self):
        self._v = True
        if __import__("os").environ.get("B200DDP_CONV_SPLITK", "1") in ("0", "false", "off"):
            self[0] = False

    def __getitem__(self, i):
        return self._v

    def __setitem__(self, i, v):
        self._v = bool(v)
        L.lib().b200_conv_set_split_k(c_int(int(self._v)))


SPLIT_K = _SplitK()


def _fn(name: str, t: torch.Tensor):
    """The entry point for ``t``'s activation dtype: ``b200_xxx`` takes NHWC bf16, ``b200_xxx_f32`` NHWC fp32."""
    if t.dtype == torch.bfloat16:
        return getattr(L.lib(), name)
    if t.dtype == torch.float32:
        return getattr(L.lib(), name + "_f32")
    raise TypeError("b200ddp activations are bfloat16 or float32, got %s" % t.dtype)


def _cl_ok(x: torch.Tensor) -> bool:
    return x.dim() == 4 and x.is_contiguous(memory_format=CL)


def to_cl(x: torch.Tensor) -> torch.Tensor:
    return x if _cl_ok(x) else x.contiguous(memory_format=CL)


def empty_act(n: int, c: int, h: int, w: int, device, dtype=torch.bfloat16) -> torch.Tensor:
    return torch.empty((n, c, h, w), dtype=dtype, device=device, memory_format=CL)


# --------------------------------------------------------------------------------------------
# tensor-core conv / linear
# --------------------------------------------------------------------------------------------
def _geom(N, H, W, C, K, R, S, stride, pad) -> ConvGeom:
    return ConvGeom(N, H, W, C, K, R, S, stride, pad)


def conv_out_hw(H, W, R, S, stride, pad) -> Tuple[int, int]:
    return (H + 2 * pad - R) // stride + 1, (W + 2 * pad - S) // stride + 1


def conv_fwd(x: torch.Tensor, w_krsc: torch.Tensor, bias: Optional[torch.Tensor], stride: int, pad: int,
             relu: bool, bn_sums: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x [N,C,H,W] cl bf16, w [K,R,S,C] bf16 contiguous -> y [N,K,Ho,Wo] cl bf16.  With ``bn_sums`` (fp32 [2,K], zero on
    entry; no bias / ReLU) the epilogue also accumulates the per-channel sum and sum of squares of the output."""
    N, C, H, W = x.shape
    K, R, S, C2 = w_krsc.shape
    assert C2 == C and _cl_ok(x) and w_krsc.is_contiguous() and w_krsc.dtype == x.dtype
    Ho, Wo = conv_out_hw(H, W, R, S, stride, pad)
    y = empty_act(N, K, Ho, Wo, x.device, x.dtype)
    g = _geom(N, H, W, C, K, R, S, stride, pad)
    if bn_sums is not None:
        assert bias is None and not relu and bn_sums.numel() >= 2 * K and bn_sums.dtype == torch.float32
        L.check(_fn("b200_conv_fwd_stats", x)(L.ptr(x), L.ptr(w_krsc), L.ptr(y), byref(g), L.ptr(bn_sums), L.stream_ptr()),
                "conv_fwd_stats")
        return y
    L.check(_fn("b200_conv_fwd", x)(L.ptr(x), L.ptr(w_krsc), L.ptr(bias), L.ptr(y), byref(g), c_int(int(relu)),
                                    L.stream_ptr()), "conv_fwd")
    return y


def conv_dgrad(dy: torch.Tensor, w_krsc: torch.Tensor, x_shape, stride: int, pad: int,
               residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    """dx = dgrad(dy, w) (+ residual, a channels-last bf16 tensor of dx's shape, added in the epilogue)."""
    N, C, H, W = x_shape
    K, R, S, _ = w_krsc.shape
    assert _cl_ok(dy) and w_krsc.dtype == dy.dtype
    dx = empty_act(N, C, H, W, dy.device, dy.dtype)
    g = _geom(N, H, W, C, K, R, S, stride, pad)
    if residual is not None:
        assert tuple(residual.shape) == (N, C, H, W) and _cl_ok(residual) and residual.dtype == dy.dtype
        L.check(_fn("b200_conv_dgrad_add", dy)(L.ptr(dy), L.ptr(w_krsc), L.ptr(dx), byref(g), L.ptr(residual), L.stream_ptr()),
                "conv_dgrad_add")
        return dx
    L.check(_fn("b200_conv_dgrad", dy)(L.ptr(dy), L.ptr(w_krsc), L.ptr(dx), byref(g), L.stream_ptr()), "conv_dgrad")
    return dx


def conv_wgrad_(dw_krsc_f32: torch.Tensor, x: torch.Tensor, dy: torch.Tensor, stride: int, pad: int,
                max_splits: int = 0) -> None:
    """dw [K,R,S,C] fp32 (contiguous storage) += wgrad(x, dy)."""
    N, C, H, W = x.shape
    K = dy.shape[1]
    R, S = dw_krsc_f32.shape[1], dw_krsc_f32.shape[2]
    assert _cl_ok(x) and _cl_ok(dy) and dw_krsc_f32.is_contiguous() and dw_krsc_f32.dtype == torch.float32 and x.dtype == dy.dtype
    g = _geom(N, H, W, C, K, R, S, stride, pad)
    L.check(_fn("b200_conv_wgrad", x)(L.ptr(x), L.ptr(dy), L.ptr(dw_krsc_f32), byref(g), c_int(max_splits),
                                    L.stream_ptr()), "conv_wgrad")


def linear_fwd(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor], relu: bool) -> torch.Tensor:
    """x [M,K] bf16, w [N,K] bf16 -> y [M,N] bf16 (a 1x1 convolution on [M,1,1,K])."""
    M, K = x.shape
    N = w.shape[0]
    assert x.is_contiguous() and w.is_contiguous() and w.dtype == x.dtype
    y = torch.empty((M, N), dtype=x.dtype, device=x.device)
    g = _geom(M, 1, 1, K, N, 1, 1, 1, 0)
    L.check(_fn("b200_conv_fwd", x)(L.ptr(x), L.ptr(w), L.ptr(bias), L.ptr(y), byref(g), c_int(int(relu)),
                                  L.stream_ptr()), "linear_fwd")
    return y


def linear_dgrad(dy: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    M, N = dy.shape
    K = w.shape[1]
    dx = torch.empty((M, K), dtype=dy.dtype, device=dy.device)
    g = _geom(M, 1, 1, K, N, 1, 1, 1, 0)
    L.check(_fn("b200_conv_dgrad", dy)(L.ptr(dy), L.ptr(w), L.ptr(dx), byref(g), L.stream_ptr()), "linear_dgrad")
    return dx


def linear_wgrad_(dw_f32: torch.Tensor, x: torch.Tensor, dy: torch.Tensor) -> None:
    M, K = x.shape
    N = dy.shape[1]
    assert dw_f32.is_contiguous() and dw_f32.dtype == torch.float32
    g = _geom(M, 1, 1, K, N, 1, 1, 1, 0)
    L.check(_fn("b200_conv_wgrad", x)(L.ptr(x), L.ptr(dy), L.ptr(dw_f32), byref(g), c_int(0), L.stream_ptr()), "linear_wgrad")


# --------------------------------------------------------------------------------------------
# batch norm (+ReLU +residual)
# --------------------------------------------------------------------------------------------
def _rows_c(x: torch.Tensor) -> Tuple[int, int]:
    if x.dim() == 4:
        n, c, h, w = x.shape
        return n * h * w, c
    return x.shape[0], x.shape[1]


def bn_stats(x: torch.Tensor, gamma, beta, eps: float, momentum: float, running_mean, running_var) -> torch.Tensor:
    rows, C = _rows_c(x)
    stat = torch.empty((4, C), dtype=torch.float32, device=x.device)
    sc = scratch(x.device)
    L.check(_fn("b200_bn_stats", x)(L.ptr(x), c_size_t(rows), c_int(C), L.ptr(gamma), L.ptr(beta), c_float(eps),
                                  c_float(momentum), L.ptr(stat), L.ptr(running_mean), L.ptr(running_var), L.ptr(sc),
                                  c_size_t(sc.numel()), L.stream_ptr()), "bn_stats")
    return stat


def bn_finalize(sums: torch.Tensor, rows: int, gamma, beta, eps: float, momentum: float, running_mean, running_var) -> torch.Tensor:
    """Training-mode statistics from the per-channel sums a conv epilogue accumulated (``conv_fwd(..., bn_sums=)``);
    leaves ``sums`` zero for the next forward."""
    C = gamma.numel()
    stat = torch.empty((4, C), dtype=torch.float32, device=gamma.device)
    L.check(L.lib().b200_bn_finalize(L.ptr(sums), c_size_t(rows), c_int(C), L.ptr(gamma), L.ptr(beta), c_float(eps),
                                     c_float(momentum), L.ptr(stat), L.ptr(running_mean), L.ptr(running_var), L.stream_ptr()),
            "bn_finalize")
    return stat


def bn_apply_sums(x, residual, sums, gamma, beta, eps: float, momentum: float, running_mean, running_var, relu: bool,
                  keep: bool = False):
    """``bn_finalize`` + ``bn_apply`` in one launch; ``sums`` holds 2C floats + 1 ticket word.  ``keep``: leave ``sums`` as
    they are (no ticket; the caller clears them before the next forward).  Returns (y, stat)."""
    rows, C = _rows_c(x)
    assert sums.numel() >= 2 * C + 1
    stat = torch.empty((4, C), dtype=torch.float32, device=x.device)
    y = torch.empty_like(x)
    L.check(_fn("b200_bn_apply_sums_keep" if keep else "b200_bn_apply_sums", x)(L.ptr(x), L.ptr(residual), L.ptr(y), c_size_t(rows), c_int(C), L.ptr(sums), L.ptr(gamma),
                                       L.ptr(beta), c_float(eps), c_float(momentum), L.ptr(stat), L.ptr(running_mean),
                                       L.ptr(running_var), c_int(int(relu)), L.stream_ptr()), "bn_apply_sums")
    return y, stat


def bn_eval_prep(gamma, beta, running_mean, running_var, eps: float) -> torch.Tensor:
    C = gamma.numel()
    stat = torch.empty((4, C), dtype=torch.float32, device=gamma.device)
    L.check(L.lib().b200_bn_eval_prep(L.ptr(gamma), L.ptr(beta), L.ptr(running_mean), L.ptr(running_var), c_float(eps),
                                      L.ptr(stat), c_int(C), L.stream_ptr()), "bn_eval_prep")
    return stat


def bn_apply(x: torch.Tensor, residual: Optional[torch.Tensor], stat: torch.Tensor, relu: bool) -> torch.Tensor:
    rows, C = _rows_c(x)
    y = torch.empty_like(x)
    L.check(_fn("b200_bn_apply", x)(L.ptr(x), L.ptr(residual), L.ptr(y), c_size_t(rows), c_int(C), L.ptr(stat),
                                  c_int(int(relu)), L.stream_ptr()), "bn_apply")
    return y


# The single-launch BN kernels accept up to b200_bn_small_rows() = 2048 rows, but one CTA per channel group only
# wins while the per-thread row loop is short: measured on B200, 1024 rows (resnet18 layer3 at batch 256) is already
# slower than the two multi-CTA launches.  B200DDP_BN_FUSED_ROWS overrides.
BN_SMALL_ROWS = int(__import__("os").environ.get("B200DDP_BN_FUSED_ROWS", "256"))
# The backward pass has a second one-launch kernel above that size: a thread-block cluster per 8 channels that keeps its
# slice of dy / x / y in registers (one pass, DSMEM reduction).  It accepts up to 16384 rows (b200_bn_bwd_fused_max_rows())
# but only wins up to ~2048: measured on resnet18, per-GPU batch 32 (rows 2048/512/128/32): 0.805 -> 0.715 ms per step;
# batch 256 with a limit of 1024 / 4096 / 16384 rows: 1.072 / 1.096 / 1.156 ms (two-kernel path everywhere: 1.101).
BN_BWD_FUSED_ROWS = int(__import__("os").environ.get("B200DDP_BN_BWD_FUSED_ROWS", "2048"))


def bn_fwd_fused(x, residual, gamma, beta, eps, momentum, running_mean, running_var, relu: bool):
    """Training-mode BN(+residual)(+ReLU) in ONE launch for small tensors. Returns (y, stat)."""
    rows, C = _rows_c(x)
    stat = torch.empty((4, C), dtype=torch.float32, device=x.device)
    y = torch.empty_like(x)
    L.check(_fn("b200_bn_fwd_fused", x)(L.ptr(x), L.ptr(residual), L.ptr(y), c_size_t(rows), c_int(C), L.ptr(gamma), L.ptr(beta),
                                      c_float(eps), c_float(momentum), L.ptr(stat), L.ptr(running_mean), L.ptr(running_var),
                                      c_int(int(relu)), L.stream_ptr()), "bn_fwd_fused")
    return y, stat


def bn_bwd_fused(dy, x, y, stat, relu: bool, want_dres: bool, dgamma=None, dbeta=None, beta_accum: float = 0.0):
    rows, C = _rows_c(x)
    dx = torch.empty_like(x)
    dres = torch.empty_like(x) if want_dres else None
    L.check(_fn("b200_bn_bwd_fused", x)(L.ptr(dy), L.ptr(x), L.ptr(y), L.ptr(dx), L.ptr(dres), c_size_t(rows), c_int(C),
                                      L.ptr(stat), c_int(int(relu)), L.ptr(dgamma), L.ptr(dbeta), c_float(beta_accum),
                                      L.stream_ptr()), "bn_bwd_fused")
    return dx, dres


def bn_bwd(dy, x, y, stat, relu: bool, want_dres: bool, dgamma=None, dbeta=None, beta_accum: float = 0.0):
    """Returns (dx, dres, sums[2,C]); dgamma/dbeta (if given) receive beta_accum*old + new."""
    rows, C = _rows_c(x)
    sums = torch.empty((2, C), dtype=torch.float32, device=x.device)
    sc = scratch(x.device)
    L.check(_fn("b200_bn_bwd_reduce", x)(L.ptr(dy), L.ptr(x), L.ptr(y), c_size_t(rows), c_int(C), L.ptr(stat),
                                       c_int(int(relu)), L.ptr(sums), L.ptr(dgamma), L.ptr(dbeta), c_float(beta_accum),
                                       L.ptr(sc), c_size_t(sc.numel()), L.stream_ptr()), "bn_bwd_reduce")
    dx = torch.empty_like(x)
    dres = torch.empty_like(x) if want_dres else None
    L.check(_fn("b200_bn_bwd_apply", x)(L.ptr(dy), L.ptr(x), L.ptr(y), L.ptr(dx), L.ptr(dres), c_size_t(rows), c_int(C),
                                      L.ptr(stat), L.ptr(sums), c_int(int(relu)), L.stream_ptr()), "bn_bwd_apply")
    return dx, dres, sums


# --------------------------------------------------------------------------------------------
# pools / relu
# --------------------------------------------------------------------------------------------
def maxpool_fwd(x: torch.Tensor, k: int, stride: int, pad: int):
    N, C, H, W = x.shape
    Ho, Wo = conv_out_hw(H, W, k, k, stride, pad)
    y = empty_act(N, C, Ho, Wo, x.device, x.dtype)
    amax = torch.empty((N, Ho, Wo, C), dtype=torch.uint8, device=x.device)
    L.check(_fn("b200_maxpool_fwd", x)(L.ptr(x), L.ptr(y), L.ptr(amax), c_int(N), c_int(H), c_int(W), c_int(C), c_int(k),
                                     c_int(stride), c_int(pad), L.stream_ptr()), "maxpool_fwd")
    return y, amax


def maxpool_bwd(dy: torch.Tensor, amax: torch.Tensor, x_shape, k: int, stride: int, pad: int) -> torch.Tensor:
    N, C, H, W = x_shape
    dx = empty_act(N, C, H, W, dy.device, dy.dtype)
    L.check(_fn("b200_maxpool_bwd", dy)(L.ptr(dy), L.ptr(amax), L.ptr(dx), c_int(N), c_int(H), c_int(W), c_int(C), c_int(k),
                                     c_int(stride), c_int(pad), L.stream_ptr()), "maxpool_bwd")
    return dx


def avgpool_fwd(x: torch.Tensor) -> torch.Tensor:
    N, C, H, W = x.shape
    y = torch.empty((N, C), dtype=x.dtype, device=x.device)
    L.check(_fn("b200_avgpool_fwd", x)(L.ptr(x), L.ptr(y), c_int(N), c_int(H * W), c_int(C), L.stream_ptr()), "avgpool_fwd")
    return y


def avgpool_bwd(dy: torch.Tensor, x_shape) -> torch.Tensor:
    N, C, H, W = x_shape
    dx = empty_act(N, C, H, W, dy.device, dy.dtype)
    L.check(_fn("b200_avgpool_bwd", dy)(L.ptr(dy), L.ptr(dx), c_int(N), c_int(H * W), c_int(C), L.stream_ptr()), "avgpool_bwd")
    return dx


def colsum_accum_(out_f32: torch.Tensor, dy: torch.Tensor, c_valid: int) -> None:
    """out[:c_valid] += column sums of dy ([rows, C] bf16 matrix or channels-last activation)."""
    rows, C = _rows_c(dy)
    L.check(_fn("b200_colsum_accum", dy)(L.ptr(dy), c_size_t(rows), c_int(C), c_int(c_valid), L.ptr(out_f32), L.stream_ptr()),
            "colsum_accum")


def relu_bwd(dy: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    dx = torch.empty_like(dy)
    L.check(_fn("b200_relu_bwd", dy)(L.ptr(dy), L.ptr(y), L.ptr(dx), c_size_t(dy.numel()), L.stream_ptr()), "relu_bwd")
    return dx


# --------------------------------------------------------------------------------------------
# losses / metrics
# --------------------------------------------------------------------------------------------
def softmax_xent(logits: torch.Tensor, targets: torch.Tensor, grad_scale: float, want_grad: bool = True):
    """Returns (loss_mean fp32 scalar tensor, dlogits or None).  logits [B,C] fp32/bf16 contiguous."""
    B, C = logits.shape
    assert logits.is_contiguous() and targets.dtype == torch.int64
    row_loss = torch.empty(B, dtype=torch.float32, device=logits.device)
    loss = torch.empty((), dtype=torch.float32, device=logits.device)
    dlogits = torch.empty_like(logits) if want_grad else None
    sc = scratch(logits.device)
    counter = sc[4096 - 64: 4096 - 60]      # a ticket word the BN counters never reach (they use [0, 4096-64))
    L.check(L.lib().b200_softmax_xent(L.ptr(logits), c_int(L.dtype_code(logits.dtype)), L.ptr(targets), c_int(B), c_int(C),
                                      L.ptr(row_loss), L.ptr(loss), L.ptr(dlogits), c_float(grad_scale), L.ptr(counter),
                                      L.stream_ptr()), "softmax_xent")
    return loss, dlogits


def bce_logits(logits: torch.Tensor, targets: torch.Tensor, grad_scale: float, want_grad: bool = True):
    n = logits.numel()
    loss = torch.empty((), dtype=torch.float32, device=logits.device)
    dlogits = torch.empty_like(logits) if want_grad else None
    L.check(L.lib().b200_bce_logits(L.ptr(logits), c_int(L.dtype_code(logits.dtype)), L.ptr(targets), c_int(n), L.ptr(loss),
                                    L.ptr(dlogits), c_float(grad_scale), L.stream_ptr()), "bce_logits")
    return loss, dlogits


def eval_count(logits: torch.Tensor, targets: torch.Tensor, correct: torch.Tensor, nll_sum: Optional[torch.Tensor]):
    B, C = logits.shape
    L.check(L.lib().b200_eval_count(L.ptr(logits), c_int(L.dtype_code(logits.dtype)), L.ptr(targets), c_int(B), c_int(C),
                                    L.ptr(correct), L.ptr(nll_sum), L.stream_ptr()), "eval_count")


# --------------------------------------------------------------------------------------------
# optimizers over flat arenas
# --------------------------------------------------------------------------------------------
def sgd_step(p, g, m, shadow, lr, momentum, dampening, weight_decay, nesterov, grad_scale=1.0, zero_grad=True):
    L.check(L.lib().b200_sgd_momentum(L.ptr(p), L.ptr(g), L.ptr(m), L.ptr(shadow), c_size_t(p.numel()), c_float(lr),
                                      c_float(momentum), c_float(dampening), c_float(weight_decay), c_int(int(nesterov)),
                                      c_float(grad_scale), c_int(int(zero_grad)), L.stream_ptr()), "sgd")


def adam_step(p, g, m, v, shadow, lr, beta1, beta2, eps, weight_decay, step_dev, grad_scale=1.0, zero_grad=True):
    L.check(L.lib().b200_step_inc(L.ptr(step_dev), L.stream_ptr()), "step_inc")
    L.check(L.lib().b200_adam(L.ptr(p), L.ptr(g), L.ptr(m), L.ptr(v), L.ptr(shadow), c_size_t(p.numel()), c_float(lr),
                              c_float(beta1), c_float(beta2), c_float(eps), c_float(weight_decay), L.ptr(step_dev),
                              c_float(grad_scale), c_int(int(zero_grad)), L.stream_ptr()), "adam")


def cast_bf16_(dst_bf16: torch.Tensor, src_f32: torch.Tensor):
    L.check(L.lib().b200_cast_f32_to_bf16(L.ptr(src_f32), L.ptr(dst_bf16), c_size_t(src_f32.numel()), L.stream_ptr()), "cast")


# --------------------------------------------------------------------------------------------
# data: augmentation, trojan stamping
# --------------------------------------------------------------------------------------------
def augment(img_u8: torch.Tensor, offsets: Optional[torch.Tensor], mean: torch.Tensor, inv_std: torch.Tensor,
            c_out: int = 8, pad: int = 4, nchw_fp32: bool = False, dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """uint8 NHWC batch -> normalised activations (NHWC with zero-padded channels in the compute dtype -- bf16 or fp32,
    default ``functional.act_dtype()`` -- or NCHW fp32)."""
    N, H, W, C = img_u8.shape
    assert img_u8.is_contiguous() and img_u8.dtype == torch.uint8
    if nchw_fp32:
        out = torch.empty((N, C, H, W), dtype=torch.float32, device=img_u8.device)
        c_out, layout = C, 1
    else:
        if dtype is None:
            from . import functional as _BF
            dtype = _BF.act_dtype()
        out = empty_act(N, c_out, H, W, img_u8.device, dtype)
        layout = 0 if dtype == torch.bfloat16 else 2
    L.check(L.lib().b200_augment(L.ptr(img_u8), L.ptr(offsets), L.ptr(out), c_int(N), c_int(H), c_int(W), c_int(C),
                                 c_int(c_out), c_int(pad), L.ptr(mean), L.ptr(inv_std), c_int(layout),
                                 L.stream_ptr()), "augment")
    return out


def trojan_stamp_(x: torch.Tensor, pattern: torch.Tensor, loc_h: int, loc_w: int, alpha: float,
                  mask: Optional[torch.Tensor] = None):
    """In-place alpha-blend of ``pattern`` [PC,ph,pw] into x [N,C,H,W] fp32 at (loc_h, loc_w)."""
    N, C, H, W = x.shape
    PC, ph, pw = pattern.shape
    assert x.is_contiguous() and x.dtype == torch.float32 and pattern.is_contiguous()
    L.check(L.lib().b200_trojan_stamp(L.ptr(x), c_int(N), c_int(C), c_int(H), c_int(W), L.ptr(pattern), c_int(PC), c_int(ph),
                                      c_int(pw), c_int(loc_h), c_int(loc_w), c_float(alpha), L.ptr(mask), L.stream_ptr()),
            "trojan_stamp")


# --------------------------------------------------------------------------------------------
# workshop Net megakernel
# --------------------------------------------------------------------------------------------
_NET_SCRATCH = {}


def net_fwd_bwd(x: torch.Tensor, targets: torch.Tensor, params, grads, krsc: bool, want_logits: bool = False,
                aug=None, loss_out: Optional[torch.Tensor] = None):
    """Forward + CE loss (+ backward when ``grads`` is a list of 10 fp32 buffers, accumulated into) of the
    workshop Net.  ``x``: normalised fp32 NCHW batch, or the raw uint8 NHWC batch with ``aug = (offsets or None,
    mean, inv_std)`` -- crop / flip / normalise then happen inside the kernel.  ``params``: the 10 tensors in
    state_dict order (storage order per ``krsc``).  Returns ``(loss_mean, logits or None)``."""
    B = x.shape[0]
    dev = x.device
    lib = L.lib()
    lib.b200_net_scratch_bytes.restype = c_size_t
    need = lib.b200_net_scratch_bytes(c_int(B))
    key = (dev.index, B)
    sc = _NET_SCRATCH.get(key)
    if sc is None:
        sc = torch.zeros(need, dtype=torch.uint8, device=dev)
        _NET_SCRATCH[key] = sc
    row_loss = torch.empty(B, dtype=torch.float32, device=dev)
    loss = loss_out if loss_out is not None else torch.empty((), dtype=torch.float32, device=dev)
    logits = torch.empty((B, 10), dtype=torch.float32, device=dev) if want_logits else None
    P = (c_void_p * 10)(*[t.data_ptr() for t in params])
    G = (c_void_p * 10)(*[t.data_ptr() for t in grads]) if grads is not None else None
    if x.dtype == torch.uint8:
        assert aug is not None and x.is_contiguous() and tuple(x.shape[1:]) == (32, 32, 3)
        offsets, mean, inv_std = aug
        L.check(lib.b200_net_fwd_bwd_u8(L.ptr(x), L.ptr(offsets), L.ptr(mean), L.ptr(inv_std), L.ptr(targets), c_int(B), P, G,
                                        c_int(int(krsc)), L.ptr(row_loss), L.ptr(loss), L.ptr(logits), L.ptr(sc),
                                        c_size_t(sc.numel()), L.stream_ptr()), "net_fwd_bwd_u8")
    else:
        assert x.dtype == torch.float32 and x.is_contiguous() and tuple(x.shape[1:]) == (3, 32, 32)
        L.check(lib.b200_net_fwd_bwd(L.ptr(x), L.ptr(targets), c_int(B), P, G, c_int(int(krsc)), L.ptr(row_loss), L.ptr(loss),
                                     L.ptr(logits), L.ptr(sc), c_size_t(sc.numel()), L.stream_ptr()), "net_fwd_bwd")
    if grads is not None:
        L.LAUNCHES[0] += 1          # two kernels per call
    return loss, logits


# ------------------------------------------------------------------------------------------------
# sequence models of the MNTD zoo (csrc/seq.cu), fp32
# ------------------------------------------------------------------------------------------------
def sgemm(A: torch.Tensor, B: torch.Tensor, bias: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
          beta: float = 0.0, relu: bool = False) -> torch.Tensor:
    """``out[M,N] = A[M,K] @ B[K,N] (+ bias) (+ beta*out)``; ``A`` and ``B`` are 2-D fp32 CUDA tensors with any strides, so
    a transposed operand is just ``.t()`` and overlapping rows (``as_strided``) need no copy."""
    from ctypes import c_longlong
    M, K = A.shape
    K2, N = B.shape
    assert K == K2 and A.dtype == torch.float32 and B.dtype == torch.float32
    if out is None:
        out = torch.empty((M, N), dtype=torch.float32, device=A.device)
    assert out.is_contiguous() and tuple(out.shape) == (M, N)
    L.check(L.lib().b200_sgemm(L.ptr(A), L.ptr(B), L.ptr(out), L.ptr(bias), c_int(M), c_int(N), c_int(K),
                               c_longlong(A.stride(0)), c_longlong(A.stride(1)), c_longlong(B.stride(0)), c_longlong(B.stride(1)),
                               c_int(N), c_float(beta), c_int(int(relu)), L.stream_ptr()), "sgemm")
    return out


def lstm_rec_fwd(pre: torch.Tensor, w_hh: torch.Tensor):
    """Recurrence of one LSTM layer.  ``pre`` [B,T,4H] = x W_ih^T + b_ih + b_hh.  Returns (h [B,T,H], gates [B,T,4H], c [B,T,H])."""
    B, T, H4 = pre.shape
    H = H4 // 4
    assert pre.is_contiguous() and w_hh.is_contiguous() and tuple(w_hh.shape) == (H4, H)
    hseq = torch.empty((B, T, H), dtype=torch.float32, device=pre.device)
    gates = torch.empty((B, T, H4), dtype=torch.float32, device=pre.device)
    cseq = torch.empty((B, T, H), dtype=torch.float32, device=pre.device)
    L.check(L.lib().b200_lstm_rec_fwd(L.ptr(pre), L.ptr(w_hh), L.ptr(hseq), L.ptr(gates), L.ptr(cseq), c_int(B), c_int(T), c_int(H),
                                      L.stream_ptr()), "lstm_rec_fwd")
    return hseq, gates, cseq


def lstm_rec_bwd(dh: torch.Tensor, gates: torch.Tensor, cseq: torch.Tensor, w_hh: torch.Tensor) -> torch.Tensor:
    """BPTT of one layer: gradient of the loss w.r.t. the pre-activation gates, [B,T,4H]."""
    B, T, H = dh.shape
    assert dh.is_contiguous() and gates.is_contiguous() and cseq.is_contiguous()
    dgates = torch.empty_like(gates)
    L.check(L.lib().b200_lstm_rec_bwd(L.ptr(dh), L.ptr(gates), L.ptr(cseq), L.ptr(w_hh), L.ptr(dgates), c_int(B), c_int(T), c_int(H),
                                      L.stream_ptr()), "lstm_rec_bwd")
    return dgates


def mel_frontend(x: torch.Tensor, window: torch.Tensor, melfb: torch.Tensor) -> torch.Tensor:
    """Waveforms [B,L] -> normalised log-mel features [B, 1 + L//512, n_mels] (2048-point Hann STFT, hop 512, reflect)."""
    B, Ls = x.shape
    assert x.is_contiguous() and x.dtype == torch.float32 and window.numel() == 2048 and melfb.shape[1] == 1025
    out = torch.empty((B, 1 + Ls // 512, melfb.shape[0]), dtype=torch.float32, device=x.device)
    L.check(L.lib().b200_mel_frontend(L.ptr(x), L.ptr(window), L.ptr(melfb.contiguous()), L.ptr(out), c_int(B), c_int(Ls),
                                      c_int(melfb.shape[0]), L.stream_ptr()), "mel_frontend")
    return out


def mel_frontend_bwd(x: torch.Tensor, window: torch.Tensor, melfb: torch.Tensor, dout: torch.Tensor) -> torch.Tensor:
    """Gradient of ``mel_frontend`` w.r.t. the waveforms: [B,L]."""
    B, Ls = x.shape
    assert x.is_contiguous() and dout.is_contiguous() and tuple(dout.shape) == (B, 1 + Ls // 512, melfb.shape[0])
    dx = torch.zeros_like(x)
    L.check(L.lib().b200_mel_frontend_bwd(L.ptr(x), L.ptr(window), L.ptr(melfb.contiguous()), L.ptr(dout), L.ptr(dx), c_int(B),
                                          c_int(Ls), c_int(melfb.shape[0]), L.stream_ptr()), "mel_frontend_bwd")
    return dx


def gather_rows(table: torch.Tensor, idx: torch.Tensor, pad_rows: int = 0) -> torch.Tensor:
    """``table[idx]`` for an int64 index tensor of any shape -> [idx.numel() + pad_rows, dim] (the pad rows are zero)."""
    from ctypes import c_longlong
    rows, dim = idx.numel(), table.shape[1]
    assert table.is_contiguous() and table.dtype == torch.float32 and idx.dtype == torch.int64
    out = torch.empty((rows + pad_rows, dim), dtype=torch.float32, device=table.device)
    if pad_rows:
        out[rows:].zero_()
    L.check(L.lib().b200_gather_rows(L.ptr(table), L.ptr(idx.contiguous()), L.ptr(out), c_longlong(rows), c_int(dim),
                                     c_longlong(table.shape[0]), L.stream_ptr()), "gather_rows")
    return out


def textcnn_pool_fwd(C: torch.Tensor, bias: torch.Tensor, feat: torch.Tensor, col: int, B: int, Ls: int, k: int) -> torch.Tensor:
    """feat[:, col:col+N] = max over valid windows of relu(C + bias); returns the arg-max positions [B,N] (int32)."""
    N = C.shape[1]
    arg = torch.empty((B, N), dtype=torch.int32, device=C.device)
    L.check(L.lib().b200_textcnn_pool_fwd(L.ptr(C), L.ptr(bias), L.ptr(feat), L.ptr(arg), c_int(B), c_int(Ls), c_int(N), c_int(k),
                                          c_int(feat.shape[1]), c_int(col), L.stream_ptr()), "textcnn_pool_fwd")
    return arg


def textcnn_pool_bwd(dfeat: torch.Tensor, feat: torch.Tensor, arg: torch.Tensor, col: int, B: int, Ls: int) -> torch.Tensor:
    N = arg.shape[1]
    dC = torch.empty((B * Ls, N), dtype=torch.float32, device=feat.device)
    L.check(L.lib().b200_textcnn_pool_bwd(L.ptr(dfeat), L.ptr(feat), L.ptr(arg), L.ptr(dC), c_int(B), c_int(Ls), c_int(N),
                                          c_int(feat.shape[1]), c_int(col), L.stream_ptr()), "textcnn_pool_bwd")
    return dC
