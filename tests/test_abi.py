"""CPU: libmzgo.so builds for sm_100a, loads without a GPU, exports every symbol include/mzgo.h declares,
its struct layouts match the ctypes mirrors, and compute calls fail loudly (no CPU fallback)."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    text = open(os.path.join(ROOT, "include", "mzgo.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mzgo_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from muzero_go_b200 import _lib
    lib = _lib.load()
    names = header_functions()
    assert len(names) >= 25
    for name in names:
        assert hasattr(lib, name), f"{name} declared in mzgo.h but not exported"
    # and every symbol is bound with a prototype
    assert set(names) == set(_lib._PROTOS), set(names) ^ set(_lib._PROTOS)
    assert lib.mzgo_abi_version() == 1


def test_struct_mirrors_match():
    from muzero_go_b200 import _lib
    lib = _lib.load()
    for which, struct in enumerate((_lib.TreeDesc, _lib.SearchParams, _lib.NetDesc, _lib.SelfPlayDesc)):
        assert lib.mzgo_sizeof(which) == ctypes.sizeof(struct)
    assert lib.mzgo_sizeof(99) == -1


def test_argument_errors_do_not_need_a_gpu():
    from muzero_go_b200 import _lib
    lib = _lib.load()
    assert lib.mzgo_rules_reset(None, 4, 9, None) == -1
    assert b"NULL" in lib.mzgo_last_error()
    assert lib.mzgo_rules_reset(1, 4, 40, None) == -1
    assert lib.mzgo_net_slot_bytes(19, 128, 0) == 128 * 361 * 4
    assert lib.mzgo_net_slot_bytes(19, 128, 1) == 128 * 384 * 2          # 19 * 20 - 1 positions, rounded up to 128 B
    for n in (2, 5, 6, 9, 13, 19):                                        # include/mzgo.h: PSH = n (n + 1) - 1, rounded up to 8
        assert lib.mzgo_net_slot_bytes(n, 128, 1) == 128 * ((n * (n + 1) - 1 + 7) // 8 * 8) * 2
        assert lib.mzgo_net_slot_bytes(n, 128, 0) == 128 * n * n * 4
    # one search of BASELINE configs[3] (2048 games x 801 slots) fits a B200 beside its trees: DESIGN.md section 3
    assert 2048 * 801 * (lib.mzgo_net_slot_bytes(19, 128, 1) + 362 * 18) < 0.92 * 190e9


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback():
    import muzero_go_b200 as mz
    assert mz._lib.load().mzgo_device_ok() == 0
    with pytest.raises(mz.MzgoError):
        mz.GoBoards(4, 9)
    with pytest.raises(mz.MzgoError):
        mz.Tree(2, 82, 51)
    net = mz.MuZeroNet(16, 26)
    with pytest.raises(mz.MzgoError):
        net.initial_inference(torch.zeros(1, 6, 5, 5))


def test_state_dict_layout_matches_reference():
    """same keys and shapes as /root/reference/main.py:72-142 so reference .pth files load."""
    import muzero_go_b200 as mz
    from oracle.nets_ref import MuZeroNetRef
    torch.manual_seed(3)
    a = mz.MuZeroNet(32, 37)
    torch.manual_seed(3)
    b = MuZeroNetRef(32, 37)
    sa, sb = a.state_dict(), b.state_dict()
    assert list(sa.keys()) == list(sb.keys())
    for k in sa:
        assert sa[k].shape == sb[k].shape and torch.equal(sa[k], sb[k]), k
    a.load_state_dict(sb)
