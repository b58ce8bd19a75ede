"""Autograd functions over the native sm_100a kernels, with the plain-PyTorch CPU fallback.

On CUDA tensors every function below runs this package's kernels (tcgen05 conv/linear, fused
BN+ReLU(+add), pools, fused softmax-CE); there is no library (cuDNN/cuBLAS) path on the GPU.  On CPU
tensors the same functions evaluate the reference semantics with torch ops in fp32 — that is the
CPU/gloo fallback of BASELINE.json config 1 and the oracle the numerics tests compare against.

Weight gradients of native layers are deposited *directly* into the flat gradient arena when the
parameter carries a ``_b200_sink`` (set by ``DistributedDataParallel``): the wgrad kernel accumulates
into the (optimizer-zeroed) slot and the reducer is notified, so there is no autograd AccumulateGrad
kernel and no separate bucket copy.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F

from . import native as nv
from . import streams

CL = torch.channels_last


def ceil8(v: int) -> int:
    return (v + 7) // 8 * 8


# Compute dtype of the native layers: activations (and the weights the tensor cores read) are stored in this type.
#   torch.float32   fp32 storage, TF32 tensor-core products (tcgen05 kind::tf32), fp32 accumulation -- the precision of the
#                   reference's own run (fp32 model, cuDNN TF32 convolutions, no autocast: …smddp-gpu.py:145,155-168)
#   torch.bfloat16  bf16 storage and products (kind::f16), fp32 accumulation, fp32 master weights / gradients
_ACT_DTYPE = [torch.bfloat16]
_DTYPE_NAMES = {"fp32": torch.float32, "float32": torch.float32, "tf32": torch.float32,
                "bf16": torch.bfloat16, "bfloat16": torch.bfloat16}


def set_compute_dtype(dtype) -> None:
    """``'fp32'`` (fp32 storage + TF32 tensor cores) or ``'bf16'``; applies to native layers run after the call."""
    if isinstance(dtype, str):
        if dtype.lower() not in _DTYPE_NAMES:
            raise ValueError("compute dtype must be 'fp32' or 'bf16', got %r" % dtype)
        dtype = _DTYPE_NAMES[dtype.lower()]
    if dtype not in (torch.float32, torch.bfloat16):
        raise ValueError("compute dtype must be torch.float32 or torch.bfloat16, got %s" % dtype)
    _ACT_DTYPE[0] = dtype


def act_dtype() -> torch.dtype:
    return _ACT_DTYPE[0]


def get_compute_dtype() -> str:
    return "fp32" if _ACT_DTYPE[0] == torch.float32 else "bf16"


def as_activation(x: torch.Tensor, channels: Optional[int] = None) -> torch.Tensor:
    """NCHW (any float dtype) -> channels-last activation in the compute dtype, channel-padded to a multiple of 8."""
    dt = _ACT_DTYPE[0]
    if x.dim() == 4:
        c = x.shape[1]
        cp = ceil8(c) if channels is None else channels
        if cp != c:
            x = F.pad(x, (0, 0, 0, 0, 0, cp - c))
        if x.dtype != dt:
            x = x.to(dt)
        return x.contiguous(memory_format=CL)
    if x.dtype != dt:
        x = x.to(dt)
    return x.contiguous()


# ------------------------------------------------------------------------------------------------
# weights: bf16 kernel-layout views ("shadow"), padded when channel counts are not multiples of 8
# ------------------------------------------------------------------------------------------------
def conv_weight_shadow(weight: torch.Tensor, cin_pad: int, cout_pad: int, dtype=torch.bfloat16) -> torch.Tensor:
    """[K,C,R,S] fp32 parameter -> [Kp,R,S,Cp] contiguous in ``dtype`` (kernel layout).  bf16: the shadow copy the fused
    optimizer maintains; fp32: the master itself (stored channels-last in the DDP arena, so the permute is a view)."""
    K, C, R, S = weight.shape
    if dtype == torch.bfloat16:
        sh = getattr(weight, "_b200_shadow", None)
        if sh is not None and cin_pad == C and cout_pad == K:
            return sh                                   # maintained by the fused optimizer, zero cost
    w = weight.detach().permute(0, 2, 3, 1)
    if cin_pad != C or cout_pad != K:
        w = F.pad(w, (0, cin_pad - C, 0, 0, 0, 0, 0, cout_pad - K))
    return w.to(dtype).contiguous()


def linear_weight_shadow(weight: torch.Tensor, in_pad: int, out_pad: int, nhwc_from=None, dtype=torch.bfloat16) -> torch.Tensor:
    N, K = weight.shape
    if dtype == torch.bfloat16:
        sh = getattr(weight, "_b200_shadow", None)
        if sh is not None and in_pad == K and out_pad == N and nhwc_from is None:
            return sh
    w = weight.detach()
    if nhwc_from is not None:           # columns are (c,h,w)-ordered in the reference; ours arrive (h,w,cp)
        c, h, ww, cp = nhwc_from
        w = w.view(N, c, h, ww).permute(0, 2, 3, 1)
        if cp != c:
            w = F.pad(w, (0, cp - c))
        w = w.reshape(N, h * ww * cp)
    if w.shape[1] != in_pad or N != out_pad:
        w = F.pad(w, (0, in_pad - w.shape[1], 0, out_pad - N))
    return w.to(dtype).contiguous()


def _pad_vec(b: Optional[torch.Tensor], n: int) -> Optional[torch.Tensor]:
    if b is None:
        return None
    b = b.detach().float()
    return b if b.numel() == n else F.pad(b, (0, n - b.numel()))


def _deposit(param: torch.Tensor, grad_native: Optional[torch.Tensor], producer):
    """Route a parameter gradient: straight into the arena slot if the parameter has a sink, else
    return it to autograd.  ``producer(dst)`` must ACCUMULATE the gradient into ``dst`` (fp32)."""
    sink = getattr(param, "_b200_sink", None)
    if sink is not None and grad_native is None:
        producer(sink.flat)
        sink.done()
        return None
    return grad_native


# ------------------------------------------------------------------------------------------------
# conv2d (+bias +ReLU)
# ------------------------------------------------------------------------------------------------
class GradMailbox:
    """Shared by the consumers of a forked activation (a ResNet block input feeds conv1 and the skip connection).  In the
    backward pass every consumer but the last deposits its gradient here and hands autograd ``None``; the last one adds the
    deposit inside its dgrad epilogue (``b200_conv_dgrad_add``), so autograd never launches an elementwise add."""
    __slots__ = ("pending", "grads")

    def __init__(self, consumers: int):
        self.pending = consumers
        self.grads = []

    def arrive(self) -> bool:
        self.pending -= 1
        return self.pending == 0

    def take(self):
        g, self.grads = self.grads, []
        if not g:
            return None
        out = g[0]
        for t in g[1:]:
            out = out + t
        return out


# Off by default: on resnet18 (batch 256 and 32, one B200) the step with the 8 elementwise adds removed measured 2 % SLOWER
# (1.122 vs 1.100 ms) although the summed kernel time went down -- the adds were hidden, the longer dgrad epilogue is not.
_FUSE_SKIP_GRAD = [__import__("os").environ.get("B200DDP_FUSE_SKIP_GRAD", "0") != "0"]


def fork(x, consumers: int = 2):
    """Declare that the CUDA activation ``x`` is about to be consumed by exactly ``consumers`` native layers (conv inputs or
    BatchNorm residuals) and nothing else; returns ``x``."""
    if _FUSE_SKIP_GRAD[0] and x.is_cuda and torch.is_grad_enabled() and x.requires_grad:
        x._b200_mailbox = GradMailbox(consumers)
    return x


class _Conv2dFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, stride, pad, relu, bn_sums=None, mailbox=None):
        K, C, R, S = weight.shape
        Cp, Kp = x.shape[1], ceil8(K)
        w = conv_weight_shadow(weight, Cp, Kp, x.dtype)
        b = _pad_vec(bias, Kp)
        y = nv.conv_fwd(x, w, b, stride, pad, relu, bn_sums)
        ctx.save_for_backward(x, w, y if relu else None)
        ctx.meta = (weight, bias, stride, pad, relu, K, C, R, S, Cp, Kp)
        ctx.mailbox = mailbox
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, y = ctx.saved_tensors
        weight, bias, stride, pad, relu, K, C, R, S, Cp, Kp = ctx.meta
        dy = nv.to_cl(dy)
        if relu:
            dy = nv.relu_bwd(dy, y)
        dx = None
        if ctx.needs_input_grad[0]:
            mb = ctx.mailbox
            if mb is None:
                dx = nv.conv_dgrad(dy, w, x.shape, stride, pad)
            elif mb.arrive():                                # last consumer of a forked input: add the other branch's gradient
                dx = nv.conv_dgrad(dy, w, x.shape, stride, pad, residual=mb.take())
            else:
                mb.grads.append(nv.conv_dgrad(dy, w, x.shape, stride, pad))
        dw = None
        if weight.requires_grad:
            # a conv whose input needs no gradient is the network's first layer: its wgrad is the tail of the backward
            # pass with nothing beside it, so it may split over (almost) the whole machine: 32 clusters of 4 = one wave
            alone = -85 if not ctx.needs_input_grad[0] else 0
            exact = (Cp == C and Kp == K)
            sink = getattr(weight, "_b200_sink", None)
            if exact and sink is not None:
                with streams.forked(x.device, x, dy):        # off the dgrad critical path
                    nv.conv_wgrad_(sink.flat.view(K, R, S, C), x, dy, stride, pad, alone)
                    sink.done()
            else:
                tmp = torch.zeros((Kp, R, S, Cp), dtype=torch.float32, device=x.device)
                nv.conv_wgrad_(tmp, x, dy, stride, pad, alone)
                dw = tmp[:K, :, :, :C].permute(0, 3, 1, 2)
                if sink is not None:
                    sink.flat.view(K, R, S, C).add_(tmp[:K, :, :, :C])
                    sink.done()
                    dw = None
        db = None
        if bias is not None and bias.requires_grad:
            bsink = getattr(bias, "_b200_sink", None)
            if bsink is not None:
                nv.colsum_accum_(bsink.flat, dy, K)
                bsink.done()
            else:
                db = torch.zeros(K, dtype=torch.float32, device=dy.device)
                nv.colsum_accum_(db, dy, K)
        return dx, dw, db, None, None, None, None, None


def conv2d(x, weight, bias=None, stride=1, padding=0, relu=False, bn_sums=None):
    """``bn_sums`` (CUDA only; fp32 [2, ceil8(Cout)], zero on entry, needs bias=None and relu=False): the conv epilogue also
    accumulates the output's per-channel sum / sum of squares there -- see ``batch_norm(..., presummed=)``."""
    if not x.is_cuda:
        y = F.conv2d(x, weight, bias, stride=stride, padding=padding)
        return F.relu(y) if relu else y
    return _Conv2dFn.apply(x, weight, bias, int(stride), int(padding), bool(relu), bn_sums, getattr(x, "_b200_mailbox", None))


# ------------------------------------------------------------------------------------------------
# linear (+bias +ReLU) — a 1x1 convolution on the same tensor-core kernel
# ------------------------------------------------------------------------------------------------
class _LinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, relu, nhwc_from):
        N, K = weight.shape
        Kp, Np = x.shape[1], ceil8(N)
        w = linear_weight_shadow(weight, Kp, Np, nhwc_from, x.dtype)
        b = _pad_vec(bias, Np)
        y = nv.linear_fwd(x, w, b, relu)
        ctx.save_for_backward(x, w, y if relu else None)
        ctx.meta = (weight, bias, relu, N, K, Kp, Np, nhwc_from)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, y = ctx.saved_tensors
        weight, bias, relu, N, K, Kp, Np, nhwc_from = ctx.meta
        dy = dy.contiguous()
        if relu:
            dy = nv.relu_bwd(dy, y)
        dx = nv.linear_dgrad(dy, w) if ctx.needs_input_grad[0] else None
        dw = None
        if weight.requires_grad:
            sink = getattr(weight, "_b200_sink", None)
            exact = (Kp == K and Np == N and nhwc_from is None)
            if exact and sink is not None:
                with streams.forked(x.device, x, dy):
                    nv.linear_wgrad_(sink.flat.view(N, K), x, dy)
                    sink.done()
            else:
                tmp = torch.zeros((Np, Kp), dtype=torch.float32, device=x.device)
                nv.linear_wgrad_(tmp, x, dy)
                g = tmp[:N]
                if nhwc_from is not None:
                    c, h, ww, cp = nhwc_from
                    g = g.view(N, h, ww, cp)[..., :c].permute(0, 3, 1, 2).reshape(N, K)
                else:
                    g = g[:, :K]
                if sink is not None:
                    sink.flat.view(N, K).add_(g)
                    sink.done()
                else:
                    dw = g
        db = None
        if bias is not None and bias.requires_grad:
            bsink = getattr(bias, "_b200_sink", None)
            if bsink is not None:
                nv.colsum_accum_(bsink.flat, dy, N)
                bsink.done()
            else:
                db = torch.zeros(N, dtype=torch.float32, device=dy.device)
                nv.colsum_accum_(db, dy, N)
        return dx, dw, db, None, None


def linear(x, weight, bias=None, relu=False, nhwc_from=None):
    if not x.is_cuda:
        y = F.linear(x, weight, bias)
        return F.relu(y) if relu else y
    return _LinearFn.apply(x, weight, bias, bool(relu), nhwc_from)


# ------------------------------------------------------------------------------------------------
# batch norm (+residual add +ReLU)
# ------------------------------------------------------------------------------------------------
def _deposit_residual_grad(ctx, dres):
    """Gradient of a BatchNorm residual input that is a forked tensor: leave it in the mailbox for the last consumer."""
    mb = ctx.res_mailbox
    if mb is None or dres is None:
        return dres
    if mb.arrive():                      # we are the last consumer after all: add what the others left (plain torch add)
        other = mb.take()
        return dres if other is None else dres + other
    mb.grads.append(dres)
    return None


class _BatchNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, residual, gamma, beta, running_mean, running_var, training, momentum, eps, relu, presummed=None,
                res_mailbox=None):
        rows = x.numel() // x.shape[1]
        if training and presummed is not None:
            # the producing conv already accumulated sum / sum of squares per channel in its epilogue
            # (nv.bn_apply_sums does both in one launch, but its per-CTA ticket -- fence + same-address atomic -- measured
            # 10% slower per resnet18 step than the tiny finalize kernel, whose launch hides behind the conv's tail)
            if getattr(presummed, "_b200_pooled", False):
                # a step engine clears every layer's sums with one memset per step: scale / shift are derived inside the
                # apply kernel, no finalize launch (nn/modules.py)
                y, stat = nv.bn_apply_sums(x, residual, presummed, gamma.detach(), beta.detach(), eps, momentum, running_mean,
                                           running_var, relu, keep=True)
            else:
                stat = nv.bn_finalize(presummed, rows, gamma.detach(), beta.detach(), eps, momentum, running_mean, running_var)
                y = nv.bn_apply(x, residual, stat, relu)
        elif training and rows <= nv.BN_SMALL_ROWS:
            y, stat = nv.bn_fwd_fused(x, residual, gamma.detach(), beta.detach(), eps, momentum, running_mean, running_var, relu)
        else:
            if training:
                stat = nv.bn_stats(x, gamma.detach(), beta.detach(), eps, momentum, running_mean, running_var)
            else:
                stat = nv.bn_eval_prep(gamma.detach(), beta.detach(), running_mean, running_var, eps)
            y = nv.bn_apply(x, residual, stat, relu)
        ctx.save_for_backward(x, y if relu else None, stat)
        ctx.meta = (gamma, beta, relu, residual is not None, training)
        ctx.res_mailbox = res_mailbox
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y, stat = ctx.saved_tensors
        gamma, beta, relu, has_res, training = ctx.meta
        if not training:
            raise RuntimeError("b200ddp BatchNorm: backward through eval-mode statistics is not supported")
        dy = nv.to_cl(dy) if dy.dim() == 4 else dy.contiguous()
        gs, bs = getattr(gamma, "_b200_sink", None), getattr(beta, "_b200_sink", None)
        C = gamma.numel()
        small = (x.numel() // x.shape[1]) <= max(nv.BN_SMALL_ROWS, nv.BN_BWD_FUSED_ROWS)      # one launch (CTA or cluster kernel)
        if gs is not None and bs is not None:
            # the arena is zeroed by the optimizer, so beta_accum=1 both overwrites and accumulates correctly
            if small:
                dx, dres = nv.bn_bwd_fused(dy, x, y, stat, relu, has_res, gs.flat[:C], bs.flat[:C], 1.0)
            else:
                dx, dres, _ = nv.bn_bwd(dy, x, y, stat, relu, has_res, gs.flat[:C], bs.flat[:C], 1.0)
            gs.done()
            bs.done()
            dres = _deposit_residual_grad(ctx, dres)
            return dx, dres, None, None, None, None, None, None, None, None, None, None
        if small:
            dg, db = torch.zeros(C, device=x.device), torch.zeros(C, device=x.device)
            dx, dres = nv.bn_bwd_fused(dy, x, y, stat, relu, has_res, dg, db, 0.0)
            dres = _deposit_residual_grad(ctx, dres)
            return dx, dres, dg, db, None, None, None, None, None, None, None, None
        dx, dres, sums = nv.bn_bwd(dy, x, y, stat, relu, has_res)
        dres = _deposit_residual_grad(ctx, dres)
        return dx, dres, sums[1].clone(), sums[0].clone(), None, None, None, None, None, None, None, None


def batch_norm(x, gamma, beta, running_mean, running_var, training, momentum=0.1, eps=1e-5, relu=False, residual=None,
               presummed=None):
    if not x.is_cuda:
        y = F.batch_norm(x, running_mean, running_var, gamma, beta, training, momentum, eps)
        if residual is not None:
            y = y + residual
        return F.relu(y) if relu else y
    return _BatchNormFn.apply(x, residual, gamma, beta, running_mean, running_var, bool(training), float(momentum),
                              float(eps), bool(relu), presummed,
                              getattr(residual, "_b200_mailbox", None) if residual is not None else None)


# ------------------------------------------------------------------------------------------------
# pools, relu, flatten
# ------------------------------------------------------------------------------------------------
class _MaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, stride, pad):
        y, amax = nv.maxpool_fwd(x, k, stride, pad)
        ctx.save_for_backward(amax)
        ctx.meta = (tuple(x.shape), k, stride, pad)
        return y

    @staticmethod
    def backward(ctx, dy):
        (amax,) = ctx.saved_tensors
        shape, k, stride, pad = ctx.meta
        return nv.maxpool_bwd(nv.to_cl(dy), amax, shape, k, stride, pad), None, None, None


def max_pool2d(x, kernel_size, stride=None, padding=0):
    stride = kernel_size if stride is None else stride
    if not x.is_cuda:
        return F.max_pool2d(x, kernel_size, stride, padding)
    return _MaxPoolFn.apply(x, int(kernel_size), int(stride), int(padding))


class _GlobalAvgPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        ctx.shape = tuple(x.shape)
        return nv.avgpool_fwd(x)

    @staticmethod
    def backward(ctx, dy):
        return nv.avgpool_bwd(dy.contiguous(), ctx.shape)


def global_avg_pool(x):
    """AdaptiveAvgPool2d(1) + flatten -> [N, C]."""
    if not x.is_cuda:
        return x.mean(dim=(2, 3))
    return _GlobalAvgPoolFn.apply(x)


class _ReluFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        y = torch.relu(x)
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        return nv.relu_bwd(dy.contiguous(memory_format=CL) if dy.dim() == 4 else dy.contiguous(), y)


def relu(x):
    return F.relu(x) if not x.is_cuda or x.dtype != _ACT_DTYPE[0] or x.numel() % 8 else _ReluFn.apply(x)


def flatten_nhwc(x):
    """[N,C,H,W] channels-last activation -> [N, H*W*C] matrix without moving data."""
    if not x.is_cuda:
        return x.flatten(1)
    n, c, h, w = x.shape
    return x.permute(0, 2, 3, 1).reshape(n, h * w * c)


# ------------------------------------------------------------------------------------------------
# losses
# ------------------------------------------------------------------------------------------------
class _CrossEntropyFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, targets, num_classes):
        lg = logits if logits.shape[1] == num_classes else logits[:, :num_classes]
        lg = lg.contiguous()
        loss, dlogits = nv.softmax_xent(lg, targets, 1.0 / lg.shape[0], want_grad=True)
        ctx.save_for_backward(dlogits)
        ctx.meta = (logits.shape[1], num_classes)
        return loss

    @staticmethod
    def backward(ctx, gout):
        (dl,) = ctx.saved_tensors
        width, nc = ctx.meta
        if not _unit_loss_grad[0]:
            dl = dl * gout.to(dl.dtype)
        if width != nc:
            dl = F.pad(dl, (0, width - nc))
        return dl, None, None


_unit_loss_grad = [False]


def assume_unit_loss_grad(flag: bool = True):
    """Trainer hint: ``loss.backward()`` is called on the raw loss (grad 1.0), skip the scale kernel."""
    _unit_loss_grad[0] = bool(flag)


def cross_entropy(logits, targets, num_classes: Optional[int] = None):
    """Mean softmax cross-entropy.  ``num_classes`` < logits width ignores padded logit columns."""
    nc = logits.shape[1] if num_classes is None else num_classes
    if not logits.is_cuda:
        return F.cross_entropy(logits[:, :nc].float(), targets)
    return _CrossEntropyFn.apply(logits, targets, nc)


class _BCEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, targets):
        lg = logits.contiguous()
        loss, dl = nv.bce_logits(lg, targets.float().contiguous(), 1.0 / lg.numel(), want_grad=True)
        ctx.save_for_backward(dl)
        ctx.shape = logits.shape
        return loss

    @staticmethod
    def backward(ctx, gout):
        (dl,) = ctx.saved_tensors
        return (dl * gout.to(dl.dtype)).view(ctx.shape), None


def bce_with_logits(logits, targets):
    if not logits.is_cuda:
        return F.binary_cross_entropy_with_logits(logits.float(), targets.float())
    return _BCEFn.apply(logits, targets)


# ------------------------------------------------------------------------------------------------
# workshop Net: forward + loss + backward as two launches
# ------------------------------------------------------------------------------------------------
class _WorkshopNetFusedFn(torch.autograd.Function):
    """loss = CE(Net(x), targets).  The megakernel produces the gradients during ``forward`` (they land in the
    flat gradient arena when the parameters carry sinks); ``backward`` only tells the reducer they are ready."""

    @staticmethod
    def forward(ctx, x, targets, krsc, aug, loss_out, *params):
        want = any(ctx.needs_input_grad[5:])
        sinks = [getattr(p, "_b200_sink", None) for p in params]
        use_sinks = want and all(s is not None for s in sinks)
        grads = None
        if want:
            grads = [s.flat for s in sinks] if use_sinks else [torch.zeros_like(p) for p in params]
        loss, _ = nv.net_fwd_bwd(x, targets, [p.detach() for p in params], grads, krsc, aug=aug, loss_out=loss_out)
        if loss_out is not None:
            ctx.mark_dirty(loss_out)
        ctx.sinks = sinks if use_sinks else None
        ctx.grads = None if use_sinks else grads
        return loss

    @staticmethod
    def backward(ctx, gout):
        if ctx.sinks is not None:
            for s_ in ctx.sinks:
                s_.done()
            return (None,) * (5 + len(ctx.sinks))
        g = ctx.grads
        if not _unit_loss_grad[0]:
            g = [t * gout for t in g]
        return (None,) * 5 + tuple(g)


def workshop_net_fused_loss(model, x, targets, aug=None):
    """Mean cross-entropy of the workshop Net with gradients for every parameter.  ``x``: a normalised fp32 NCHW batch, or
    the raw uint8 NHWC batch with ``aug = (offsets or None, mean, inv_std)`` (augmentation fused into the kernel)."""
    params = [model.conv1.weight, model.conv1.bias, model.conv2.weight, model.conv2.bias, model.fc1.weight, model.fc1.bias,
              model.fc2.weight, model.fc2.bias, model.fc3.weight, model.fc3.bias]
    w = model.conv1.weight
    krsc = (not w.is_contiguous()) and w.permute(0, 2, 3, 1).is_contiguous()
    if not (w.is_contiguous() or krsc) or model.conv2.weight.is_contiguous() == krsc:
        raise RuntimeError("workshop Net megakernel: unexpected conv weight layout")
    return _WorkshopNetFusedFn.apply(x.contiguous(), targets, bool(krsc), aug, None, *params)


# ------------------------------------------------------------------------------------------------
# fp32 sequence-model pieces of the MNTD zoo (csrc/seq.cu)
# ------------------------------------------------------------------------------------------------
class _LinearF32Fn(torch.autograd.Function):
    """y = x W^T + b in fp32 on the CUDA cores (strided sgemm) -- the heads of the audio / text models."""

    @staticmethod
    def forward(ctx, x, w, b):
        x2 = x.reshape(-1, x.shape[-1])
        y = nv.sgemm(x2, w.t(), bias=b)
        ctx.save_for_backward(x2, w)
        ctx.has_bias = b is not None
        ctx.xshape = x.shape
        return y.view(*x.shape[:-1], w.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, w = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1]).contiguous()
        dx = nv.sgemm(dy2, w).view(ctx.xshape) if ctx.needs_input_grad[0] else None
        dw = nv.sgemm(dy2.t(), x2) if ctx.needs_input_grad[1] else None
        db = None
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = nv.sgemm(torch.ones((1, dy2.shape[0]), dtype=torch.float32, device=dy2.device), dy2).view(-1)
        return dx, dw, db


def linear_f32(x, w, b=None):
    if not x.is_cuda:
        return F.linear(x, w, b)
    return _LinearF32Fn.apply(x.float(), w, b)


class _MelFn(torch.autograd.Function):
    """Waveform -> normalised log-mel features (FFT kernel), with the gradient back to the waveform for query tuning."""

    @staticmethod
    def forward(ctx, x, window, melfb):
        x = x.float().contiguous()
        ctx.save_for_backward(x, window, melfb)
        return nv.mel_frontend(x, window, melfb)

    @staticmethod
    def backward(ctx, dout):
        x, window, melfb = ctx.saved_tensors
        return nv.mel_frontend_bwd(x, window, melfb, dout.float().contiguous()), None, None


def mel_features(x, window, melfb):
    return _MelFn.apply(x, window, melfb)


class _LSTMLayerFn(torch.autograd.Function):
    """One ``nn.LSTM`` layer (batch_first, zero initial state): input projection as one GEMM over [B*T], the recurrence in
    ``lstm_rec_fwd`` (W_hh rows in registers), BPTT in ``lstm_rec_bwd`` followed by three GEMMs for dW_ih, dW_hh, dx."""

    @staticmethod
    def forward(ctx, x, w_ih, w_hh, b_ih, b_hh):
        B, T, I = x.shape
        x2 = x.reshape(B * T, I)
        pre = nv.sgemm(x2, w_ih.t(), bias=(b_ih + b_hh))
        hseq, gates, cseq = nv.lstm_rec_fwd(pre.view(B, T, -1), w_hh.contiguous())
        ctx.save_for_backward(x2, w_ih, w_hh, hseq, gates, cseq)
        return hseq

    @staticmethod
    def backward(ctx, dh):
        x2, w_ih, w_hh, hseq, gates, cseq = ctx.saved_tensors
        B, T, H = hseq.shape
        dg = nv.lstm_rec_bwd(dh.contiguous(), gates, cseq, w_hh.contiguous()).view(B * T, 4 * H)
        hprev = torch.cat([torch.zeros((B, 1, H), dtype=hseq.dtype, device=hseq.device), hseq[:, :-1]], 1).reshape(B * T, H)
        dw_hh = nv.sgemm(dg.t(), hprev)
        dw_ih = nv.sgemm(dg.t(), x2)
        db = nv.sgemm(torch.ones((1, B * T), dtype=torch.float32, device=dg.device), dg).view(-1)
        dx = nv.sgemm(dg, w_ih).view(B, T, -1) if ctx.needs_input_grad[0] else None
        return dx, dw_ih, dw_hh, db, db


def lstm(x, module: torch.nn.LSTM):
    """``nn.LSTM`` forward (all layers, batch_first, zero initial state, no dropout) on this package's kernels."""
    assert module.batch_first and not module.bidirectional and module.proj_size == 0
    h = x.float()
    for layer in range(module.num_layers):
        p = [getattr(module, "%s_l%d" % (n, layer)) for n in ("weight_ih", "weight_hh", "bias_ih", "bias_hh")]
        h = _LSTMLayerFn.apply(h, *p)
    return h


class _TextConvPoolFn(torch.autograd.Function):
    """max_t relu(conv_{k x D}(E) + b) for a [B, L, D] embedding block, conv as a GEMM over overlapping rows of E
    (row (b, t) of the im2col matrix is the k*D floats starting at E[b, t]: no copy), then fused bias+ReLU+max-over-time."""

    @staticmethod
    def forward(ctx, e_flat, B, Ls, w, bias):
        # e_flat: [B*Ls + pad, D] contiguous with >= k-1 zero pad rows so that every window read stays in bounds
        N, k, D = w.shape[0], w.shape[2], w.shape[3]
        assert e_flat.shape[0] >= B * Ls + k - 1 and e_flat.is_contiguous()
        w2 = w.reshape(N, k * D)
        A = e_flat.as_strided((B * Ls, k * D), (D, 1))
        C = nv.sgemm(A, w2.t())
        feat = torch.empty((B, N), dtype=torch.float32, device=e_flat.device)
        arg = nv.textcnn_pool_fwd(C, bias, feat, 0, B, Ls, k)
        ctx.save_for_backward(e_flat, w2, feat, arg)
        ctx.dims = (B, Ls, N, k, D)
        return feat

    @staticmethod
    def backward(ctx, dfeat):
        e_flat, w2, feat, arg = ctx.saved_tensors
        B, Ls, N, k, D = ctx.dims
        dC = nv.textcnn_pool_bwd(dfeat.contiguous(), feat, arg, 0, B, Ls)
        A = e_flat.as_strided((B * Ls, k * D), (D, 1))
        dw = nv.sgemm(dC.t(), A).view(N, 1, k, D)
        db = nv.sgemm(torch.ones((1, B * Ls), dtype=torch.float32, device=dC.device), dC).view(-1)
        de = None
        if ctx.needs_input_grad[0]:
            dA = nv.sgemm(dC, w2).view(B, Ls, k, D)          # gradient per window; overlap-add back onto the rows of E
            de = torch.zeros_like(e_flat)
            dev = de[:B * Ls].view(B, Ls, D)
            for i in range(k):
                dev[:, i:] += dA[:, :Ls - i, i]
        return de, None, None, dw, db


def text_conv_pool(e_flat, B, Ls, conv: torch.nn.Conv2d):
    return _TextConvPoolFn.apply(e_flat, B, Ls, conv.weight, conv.bias)
