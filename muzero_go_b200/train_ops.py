"""The trainer's 3x3 convolutions on libmzgo.so's own kernels (csrc/train.cu): forward, data gradient and weight
gradient of `relu(conv3x3(x, W) + b)` as one torch.autograd.Function, so that `BatchedTrainer` keeps torch's
autograd for the bookkeeping (losses, the 1x1 heads, Adam) while more than 99.9 % of a training step's FLOPs --
/root/reference/main.py:75-77 and :90 run forward and backward -- go through hand-written CUDA instead of cuDNN.

float32 on the CUDA cores: the trainer's parity contract is float32 (tests/golden/train_*.npz, 1e-5 against the
real MuZeroAgent.train).  No CPU fallback here: CPU tensors never reach this module (trainer.py routes them to the
plain torch ops that the CPU-only parity tests of the host logic use).
"""
import torch

from . import _lib


class Conv3x3ReLU(torch.autograd.Function):
    """y = relu(conv2d(x, weight, bias, padding=1)) for dense float32 NCHW CUDA tensors."""

    @staticmethod
    def forward(ctx, x, weight, bias, relu=True):
        L, s = _lib.load(), _lib.stream_ptr()
        x = x.contiguous()
        weight = weight.contiguous()
        B, Cin, n, _ = x.shape
        Cout = weight.shape[0]
        assert x.is_cuda and x.dtype == torch.float32 and weight.dtype == torch.float32 and tuple(weight.shape) == (Cout, Cin, 3, 3)
        wf = torch.empty((Cin, 9, Cout), dtype=torch.float32, device=x.device)
        wb = torch.empty((Cout, 9, Cin), dtype=torch.float32, device=x.device)
        _lib.check(L.mzgo_train_conv3x3_prep_w(_lib.ptr(weight), _lib.ptr(wf), _lib.ptr(wb), Cin, Cout, s))
        y = torch.empty((B, Cout, n, n), dtype=torch.float32, device=x.device)
        _lib.check(L.mzgo_train_conv3x3_fwd(_lib.ptr(x), _lib.ptr(wf), _lib.ptr(bias), _lib.ptr(y), int(relu), B, Cin, Cout, n, s))
        ctx.save_for_backward(x, y, wb)
        ctx.relu, ctx.has_bias = bool(relu), bias is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y, wb = ctx.saved_tensors
        L, s = _lib.load(), _lib.stream_ptr()
        dy = dy.contiguous()
        B, Cin, n, _ = x.shape
        Cout = y.shape[1]
        gate = _lib.ptr(y) if ctx.relu else None
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x)
            _lib.check(L.mzgo_train_conv3x3_bwd_data(_lib.ptr(dy), gate, _lib.ptr(wb), _lib.ptr(dx), B, Cin, Cout, n, s))
        if ctx.needs_input_grad[1] or (ctx.has_bias and ctx.needs_input_grad[2]):
            dw = torch.empty((Cout, Cin, 3, 3), dtype=torch.float32, device=x.device)
            db = torch.empty((Cout,), dtype=torch.float32, device=x.device) if ctx.has_bias else None
            scratch = torch.empty(int(L.mzgo_train_wgrad_scratch_floats(B, Cin, Cout)), dtype=torch.float32, device=x.device)
            _lib.check(L.mzgo_train_conv3x3_bwd_weight(_lib.ptr(x), _lib.ptr(dy), gate, _lib.ptr(dw), _lib.ptr(db),
                                                       _lib.ptr(scratch), B, Cin, Cout, n, s))
        return dx, dw, db, None


def conv3x3_relu(x, weight, bias):
    """relu(conv2d(x, weight, bias, padding=1)) through libmzgo.so (CUDA tensors only)."""
    return Conv3x3ReLU.apply(x, weight, bias, True)
