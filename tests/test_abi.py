"""CPU: the C-ABI library builds, loads and exports every symbol include/b200moe.h declares (no compute calls);
the ctypes table in b200moe._lib matches the header argument for argument count; compute calls fail loudly on a box
without an sm_100 device (no CPU fallback)."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_functions():
    h = open(os.path.join(ROOT, "include", "b200moe.h")).read()
    h = re.sub(r"/\*.*?\*/", "", h, flags=re.S)
    out = {}
    for m in re.finditer(r"(?:int|int64_t|const char\*)\s+(b200moe_\w+)\s*\(([^)]*)\)\s*;", h):
        args = m.group(2).strip()
        out[m.group(1)] = 0 if args in ("void", "") else len(args.split(","))
    return out


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as g
    g.build()
    import b200moe
    return b200moe.load_library()


def test_header_symbols_exported(lib):
    import b200moe
    fns = _header_functions()
    assert len(fns) >= 25
    raw = ctypes.CDLL(b200moe.LIB_PATH)
    for name, nargs in fns.items():
        assert hasattr(raw, name), f"{name} declared in include/b200moe.h but not exported"
        assert name in b200moe.SIGNATURES and len(b200moe.SIGNATURES[name]) == nargs, name
    assert set(b200moe.SIGNATURES) == set(fns)


def test_host_only_entry_points(lib):
    assert lib.b200moe_version() >= 100
    assert lib.b200moe_rows_upper_bound(3200, 2, 8, 1000) == 7424          # 6400 kept + 8*127, rounded to 128
    assert lib.b200moe_rows_upper_bound(3200, 2, 8, 100) == 1920           # capacity-bounded
    assert lib.b200moe_rows_upper_bound(10, 2, 4, -1) % 128 == 0
    assert [lib.b200moe_router_block(t) for t in (1, 3200, 20000, 40000)] == [128, 128, 128, 128]
    assert lib.b200moe_num_router_blocks(3200) == 25
    assert lib.b200moe_router_scratch_bytes(256, 8) == 2 * 16 * 256 * 2


@pytest.mark.skipif(torch.cuda.is_available(), reason="only meaningful on a box without a GPU")
def test_no_cpu_fallback(lib):
    import b200moe
    assert lib.b200moe_check_device() != 0
    x = torch.zeros(8, 128, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError):
        b200moe.ops.cast_bf16(torch.zeros(8))                             # C ABI refuses without an sm_100 device
    m = b200moe.ContextMoE(128, 128, 4, 2)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(x)


def test_module_is_deepcopy_and_pickle_safe_on_cpu():
    import copy
    import pickle
    import b200moe
    m = b200moe.ContextMoE(128, 256, 4, 2, context_vocab_sizes=(6, 10))
    m2 = copy.deepcopy(m)
    assert all(torch.equal(a, b) for a, b in zip(m.state_dict().values(), m2.state_dict().values()))
    m3 = pickle.loads(pickle.dumps(m))
    assert set(m3.state_dict()) == set(m.state_dict())
    assert [tuple(p.shape) for p in (m.w1, m.b1, m.w2, m.b2, m.gate.weight)] == \
        [(4, 256, 128), (4, 256), (4, 128, 256), (4, 128), (4, 128)]
