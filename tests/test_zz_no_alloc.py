"""GPU: mzgo_train_conv3x3_bwd_weight_static, the weight gradient that allocates nothing (include/mzgo.h section 5;
no counterpart in the reference).

Its scratch is a device array built into libmzgo.so instead of a caller-owned buffer.  Called through the C ABI,
back to back on shapes of different sizes (each one twice in a row) so that the array still holds the previous
call's partial sums: every result must equal the existing entry point's bit for bit (same kernels, same order of
summation) and a float64 torch gradient within the trainer's float32 tolerance (1e-5 of the largest entry, as in
tests/test_train_gpu.py).  Collected last (the file name) so that a failure here hides no other test."""
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 1e-5          # float32 parity of the training kernels, relative to the largest entry (tests/test_train_gpu.py)


def _case(seed, B, Cin, Cout, n, gated):
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.randn(B, Cin, n, n, generator=g).cuda()
    dy = torch.randn(B, Cout, n, n, generator=g).cuda()
    y = torch.randn(B, Cout, n, n, generator=g).cuda() if gated else None        # ReLU mask = y > 0
    return x, dy, y


def _existing(x, dy, y, with_bias=True):
    from muzero_go_b200 import _lib
    L, s = _lib.load(), _lib.stream_ptr()
    B, Cin, n, _ = x.shape
    Cout = dy.shape[1]
    dw = torch.full((Cout, Cin, 3, 3), float("nan"), device="cuda")
    db = torch.full((Cout,), float("nan"), device="cuda") if with_bias else None
    scratch = torch.empty(int(L.mzgo_train_wgrad_scratch_floats(B, Cin, Cout)), dtype=torch.float32, device="cuda")
    _lib.check(L.mzgo_train_conv3x3_bwd_weight(_lib.ptr(x), _lib.ptr(dy), _lib.ptr(y), _lib.ptr(dw), _lib.ptr(db),
                                               _lib.ptr(scratch), B, Cin, Cout, n, s))
    return dw, db


def _static(x, dy, y, dw, db):
    """The call under test: outputs preallocated by the caller, nothing allocated in between."""
    from muzero_go_b200 import _lib
    L, s = _lib.load(), _lib.stream_ptr()
    B, Cin, n, _ = x.shape
    return L.mzgo_train_conv3x3_bwd_weight_static(_lib.ptr(x), _lib.ptr(dy), _lib.ptr(y), _lib.ptr(dw), _lib.ptr(db),
                                                  B, Cin, dy.shape[1], n, s)


def _float64(x, dy, y):
    g = (dy * (y > 0)).double() if y is not None else dy.double()
    w = torch.zeros(dy.shape[1], x.shape[1], 3, 3, dtype=torch.float64, device="cuda", requires_grad=True)
    out = torch.nn.functional.conv2d(x.double(), w, padding=1)
    (dw,) = torch.autograd.grad(out, w, g)
    return dw, g.sum(dim=(0, 2, 3))


def test_static_scratch_weight_gradient_matches_the_existing_entry_point():
    from muzero_go_b200 import _lib
    _lib.require_device()
    L = _lib.load()
    assert L.mzgo_train_wgrad_static_max_floats() == 10 * 1024 * 1024 // 4
    # large, then small (most of the array keeps the first call's sums), then a third shape with the reference's
    # widest layer (129 -> 128 channels, main.py:90) and no bias gradient
    shapes = [(1, 24, 32, 32, 9, True, True), (2, 5, 6, 16, 5, False, True), (3, 16, 129, 128, 6, True, False),
              (1, 24, 32, 32, 9, True, True)]
    torch.cuda.synchronize()
    for seed, B, Cin, Cout, n, gated, with_bias in shapes:
        x, dy, y = _case(seed, B, Cin, Cout, n, gated)
        want_dw, want_db = _existing(x, dy, y, with_bias)
        dw = torch.full((Cout, Cin, 3, 3), float("nan"), device="cuda")
        db = torch.full((Cout,), float("nan"), device="cuda") if with_bias else None
        torch.cuda.synchronize()
        held = torch.cuda.memory_allocated()
        _lib.check(_static(x, dy, y, dw, db))
        _lib.check(_static(x, dy, y, dw, db))                     # and once more, straight behind it on the stream
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == held              # nothing came from torch's allocator ...
        assert torch.equal(dw, want_dw), (B, Cin, Cout, n)
        ref_dw, ref_db = _float64(x, dy, y)
        assert (dw.double() - ref_dw).abs().max().item() <= TOL * ref_dw.abs().max().item()
        if with_bias:
            assert torch.equal(db, want_db)
            assert (db.double() - ref_db).abs().max().item() <= TOL * ref_db.abs().max().item()
    # ... nor from the driver: with the module loaded and the kernels' attributes set, a call leaves free memory alone
    x, dy, y = _case(9, 8, 16, 16, 9, True)
    dw, db = torch.empty(16, 16, 3, 3, device="cuda"), torch.empty(16, device="cuda")
    torch.cuda.synchronize()
    free = torch.cuda.mem_get_info()[0]
    _lib.check(_static(x, dy, y, dw, db))
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free


def test_static_scratch_refuses_shapes_beyond_the_array():
    from muzero_go_b200 import _lib
    _lib.require_device()
    L = _lib.load()
    B, Cin, Cout, n = 16, 256, 256, 5
    assert L.mzgo_train_wgrad_scratch_floats(B, Cin, Cout) > L.mzgo_train_wgrad_static_max_floats()
    x, dy, _ = _case(4, B, Cin, Cout, n, False)
    dw = torch.zeros(Cout, Cin, 3, 3, device="cuda")
    rc = _static(x, dy, None, dw, None)
    assert rc != 0
    with pytest.raises(_lib.MzgoError, match="floats of scratch"):
        _lib.check(rc)
    torch.cuda.synchronize()
    assert not dw.any()                                           # refused before anything was launched
    # a shape that fits is still served afterwards
    x, dy, y = _case(5, 3, 8, 8, 5, True)
    want, _ = _existing(x, dy, y, False)
    dw = torch.empty(8, 8, 3, 3, device="cuda")
    _lib.check(_static(x, dy, y, dw, None))
    torch.cuda.synchronize()
    assert torch.equal(dw, want)
