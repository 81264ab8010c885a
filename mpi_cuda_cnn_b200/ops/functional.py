"""The sm_100a kernels as PyTorch ops (torch.autograd Functions over the C ABI in include/b200cnn.h).

Each op replaces a piece of the reference's CPU layer code (/root/reference/cnn.c):
  conv2d_act      Layer_feedForw_conv / feedBack_conv      :175-247
  linear_act      Layer_feedForw_full / feedBack_full      :113-173
  softmax_xent    softmax + (outputs - onehot)             :125-143, :285-286
  max_pool2d      (no counterpart; stated topology)
  sgd_update_     Layer_update                             :303-314

Activations are fused into the producing kernel, and their derivatives into the kernel that
produces the incoming gradient, exactly as the engine does; the autograd form keeps that fusion by
making "conv + activation" and "linear + activation" single Functions.
"""
import ctypes as C

import torch

from .. import _native as N

_ACT = {None: N.ACT_NONE, "none": N.ACT_NONE, "tanh": N.ACT_TANH, "relu": N.ACT_RELU}


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _st():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _shape(x, oc, oh, ow, k, pad, stride):
    n, c, h, w = x.shape
    return N.ConvShape(n, c, h, w, oc, oh, ow, k, pad, stride)


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("mpi_cuda_cnn_b200.ops.functional runs on CUDA tensors only (use ops.cpu for the CPU engine)")


def set_backend(name):
    """'auto' | 'simt' | 'tcgen05' — which implementation GEMM-shaped ops use."""
    code = {"auto": N.BACKEND_AUTO, "simt": N.BACKEND_SIMT, "tcgen05": N.BACKEND_TCGEN05}[name]
    N.check(N.cuda().b200cnn_set_backend(code), "set_backend")


class _ConvAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, stride, pad, act, out_hw):
        _need_cuda(x, w, b)
        lib = N.cuda()
        x = x.contiguous()
        w = w.contiguous().float()
        b = b.contiguous().float()
        n, c, h, wd = x.shape
        oc, _, k, _ = w.shape
        oh, ow = out_hw if out_hw is not None else ((h + 2 * pad - k) // stride + 1, (wd + 2 * pad - k) // stride + 1)
        s = _shape(x, oc, oh, ow, k, pad, stride)
        y = torch.empty((n, oc, oh, ow), device=x.device, dtype=torch.float32)
        is_u8 = 1 if x.dtype == torch.uint8 else 0
        if not is_u8:
            x = x.float()
        N.check(lib.b200cnn_conv2d_fwd(_p(x), is_u8, _p(w), _p(b), _p(y), C.byref(s), _ACT[act], _st()), "conv2d_fwd")
        ctx.save_for_backward(x, w, y)
        ctx.cfg = (s, act, is_u8)
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = N.cuda()
        x, w, y = ctx.saved_tensors
        s, act, is_u8 = ctx.cfg
        # dz = dy * act'(y): one elementwise op here; inside the engine this is fused upstream.
        if act == "relu":
            dz = (dy * (y > 0)).contiguous()
        elif act == "tanh":
            dz = (dy * (1 - y * y)).contiguous()
        else:
            dz = dy.contiguous()
        dx = None
        if ctx.needs_input_grad[0] and not is_u8:
            dx = torch.empty_like(x)
            N.check(lib.b200cnn_conv2d_bwd_data(_p(dz), _p(w), None, N.ACT_NONE, _p(dx), C.byref(s), _st()), "conv2d_bwd_data")
        dw = torch.empty_like(w)
        db = torch.empty(w.shape[0], device=w.device, dtype=torch.float32)
        scratch = torch.empty(lib.b200cnn_conv2d_bwd_weight_scratch_bytes(C.byref(s)), device=w.device, dtype=torch.uint8)
        N.check(lib.b200cnn_conv2d_bwd_weight(_p(dz), _p(x), is_u8, _p(dw), _p(db), 0, _p(scratch), C.byref(s), _st()), "conv2d_bwd_weight")
        return dx, dw, db, None, None, None, None


def conv2d_act(x, w, b, stride=1, pad=0, act="relu", out_hw=None):
    """act(conv2d(x, w) + b); x may be uint8 pixels (scaled by 1/255 in the kernel)."""
    return _ConvAct.apply(x, w, b, stride, pad, act, out_hw)


class _LinearAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, act):
        _need_cuda(x, w, b)
        lib = N.cuda()
        x = x.contiguous().float()
        w = w.contiguous().float()
        b = b.contiguous().float()
        m, k = x.shape
        n = w.shape[0]
        y = torch.empty((m, n), device=x.device, dtype=torch.float32)
        N.check(lib.b200cnn_linear_fwd(_p(x), _p(w), _p(b), _p(y), m, n, k, n, _ACT[act], _st()), "linear_fwd")
        ctx.save_for_backward(x, w, y)
        ctx.act = act
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = N.cuda()
        x, w, y = ctx.saved_tensors
        m, k = x.shape
        n = w.shape[0]
        if ctx.act == "tanh":
            dz = (dy * (1 - y * y)).contiguous()
        elif ctx.act == "relu":
            dz = (dy * (y > 0)).contiguous()
        else:
            dz = dy.contiguous()
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x)
            N.check(lib.b200cnn_linear_bwd_data(_p(dz), _p(w), None, N.ACT_NONE, _p(dx), m, n, k, n, _st()), "linear_bwd_data")
        dw = torch.empty_like(w)
        db = torch.empty(n, device=w.device, dtype=torch.float32)
        N.check(lib.b200cnn_linear_bwd_weight(_p(dz), _p(x), _p(dw), _p(db), 0, m, n, k, n, _st()), "linear_bwd_weight")
        return dx, dw, db, None


def linear_act(x, w, b, act="tanh"):
    """act(x @ w.T + b) with w laid out [out][in] like the reference (cnn.c:116-121)."""
    return _LinearAct.apply(x, w, b, act)


class _SoftmaxXent(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels):
        _need_cuda(logits, labels)
        lib = N.cuda()
        logits = logits.contiguous().float()
        m, n = logits.shape
        labels = labels.to(torch.uint8).contiguous()
        probs = torch.empty_like(logits)
        dz = torch.empty_like(logits)
        metrics = torch.zeros(4, device=logits.device, dtype=torch.float32)
        N.check(lib.b200cnn_softmax_xent(_p(logits), _p(labels), _p(probs), _p(dz), _p(metrics), m, n, n, _st()), "softmax_xent")
        ctx.save_for_backward(dz)
        ctx.mark_non_differentiable(probs, metrics)
        return metrics[0], probs, metrics

    @staticmethod
    def backward(ctx, dloss, _dprobs, _dmetrics):
        (dz,) = ctx.saved_tensors
        return dz * dloss, None


def softmax_xent(logits, labels):
    """Returns (sum cross-entropy, probabilities, metrics[4]); gradient w.r.t. logits is p - onehot."""
    return _SoftmaxXent.apply(logits, labels)


class _MaxPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, stride):
        _need_cuda(x)
        lib = N.cuda()
        x = x.contiguous().float()
        n, c, h, w = x.shape
        oh, ow = (h - k) // stride + 1, (w - k) // stride + 1
        y = torch.empty((n, c, oh, ow), device=x.device, dtype=torch.float32)
        arg = torch.empty((n, c, oh, ow), device=x.device, dtype=torch.int32)
        N.check(lib.b200cnn_maxpool_fwd(_p(x), _p(y), _p(arg), n, c, h, w, k, stride, _st()), "maxpool_fwd")
        ctx.save_for_backward(arg)
        ctx.cfg = (n, c, h, w, k, stride)
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = N.cuda()
        (arg,) = ctx.saved_tensors
        n, c, h, w, k, stride = ctx.cfg
        dx = torch.empty((n, c, h, w), device=dy.device, dtype=torch.float32)
        N.check(lib.b200cnn_maxpool_bwd(_p(dy.contiguous()), _p(arg), None, N.ACT_NONE, _p(dx), n, c, h, w, k, stride, _st()), "maxpool_bwd")
        return dx, None, None


def max_pool2d(x, k=2, stride=2):
    return _MaxPool.apply(x, k, stride)


def sgd_update_(w, g, lr, zero_grad=True):
    """In place: w -= lr * g (and g = 0), the maths of Layer_update (cnn.c:303-311)."""
    _need_cuda(w, g)
    assert w.is_contiguous() and g.is_contiguous() and w.dtype == torch.float32 and g.dtype == torch.float32
    N.check(N.cuda().b200cnn_sgd_update(_p(w), _p(g), w.numel(), float(lr), 1 if zero_grad else 0, _st()), "sgd_update")
    return w


def synthetic_batch(n, seed=0, first_index=0, h=28, w=28, nclass=10, device="cuda"):
    """Device-generated MNIST-shaped batch: (uint8 images [n, h*w], uint8 labels [n])."""
    images = torch.empty((n, h * w), device=device, dtype=torch.uint8)
    labels = torch.empty((n,), device=device, dtype=torch.uint8)
    N.check(N.cuda().b200cnn_synthetic_batch(_p(images), _p(labels), n, h, w, nclass, seed, first_index, _st()), "synthetic_batch")
    return images, labels


def gather_batch(images, labels, batch, seed, step, shard_begin=0, shard_len=None, out=None):
    """Random mini-batch with replacement out of a device-resident uint8 data set."""
    n_items = images.shape[0]
    item = images[0].numel()
    shard_len = n_items - shard_begin if shard_len is None else shard_len
    if out is None:
        out = (torch.empty((batch, item), device=images.device, dtype=torch.uint8),
               torch.empty((batch,), device=images.device, dtype=torch.uint8))
    N.check(N.cuda().b200cnn_gather_batch(_p(images), _p(labels), n_items, item, _p(out[0]), _p(out[1]), batch, seed, step,
                                          shard_begin, shard_len, _st()), "gather_batch")
    return out
