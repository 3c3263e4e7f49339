"""GPU: b200moe_router_bwd_dwg_noalloc (the gate's weight gradient with its scratch in a device array built into the
library; no counterpart in the reference) against the existing
b200moe_router_bwd_dwg, through the C ABI, on consecutive calls that share the built-in array.

Collected last on purpose (file name).
Tolerance: the one tests/test_gpu_aux.py::test_router_param_grads uses for this gradient (atol = rtol = 1e-4); the two
entry points run the same kernels, so the results are in fact bit-identical, which is asserted as well.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_FLOATS = 4 * 1024 * 1024          # B200MOE_DWG_STATIC_MAX_FLOATS (include/b200moe.h)


def _inputs(T, d, E, seed):
    g = torch.Generator().manual_seed(seed)
    dl = (torch.randn(T, E, generator=g) * 0.01).cuda()
    x = torch.randn(T, d, generator=g).to(torch.bfloat16).cuda()
    return dl, x


def _noalloc(lib_mod, dl, x, out):
    T, E = dl.shape
    lib_mod.call("b200moe_router_bwd_dwg_noalloc", dl.data_ptr(), x.data_ptr(), T, x.shape[1], E, out.data_ptr(),
                 lib_mod.stream_ptr())


def test_router_bwd_dwg_noalloc_matches_caller_scratch_entry_point():
    import b200moe
    from b200moe import _lib, ops
    b200moe.load_library()
    lib = _lib.load()
    # two calls in a row (then two more): CUDA-core path, tensor-pipe path (>= 8192 tokens, ragged last chunk), then
    # smaller problems whose scratch footprint lies inside the previous call's leftovers
    cases = [(3200, 256, 8), (9001, 128, 5), (777, 1024, 64), (130, 128, 5)]
    data = [(_inputs(T, d, E, seed=7 * i + T), (T, d, E)) for i, (T, d, E) in enumerate(cases)]
    outs = [torch.full((E, d), float("nan"), dtype=torch.float32, device="cuda") for _, (T, d, E) in data]
    for _, (T, d, E) in data:
        assert lib.b200moe_dwg_scratch_floats(T, d, E) <= MAX_FLOATS
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    for ((dl, x), _), out in zip(data, outs):
        _noalloc(_lib, dl, x, out)                      # back to back on one stream, no host sync in between
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before      # inputs and outputs are the caller's; nothing else appeared
    for ((dl, x), (T, d, E)), out in zip(data, outs):
        ref = ops.router_bwd_dwg(dl, x)                 # existing entry point, caller-provided scratch
        exact = (dl.double().t() @ x.double()).float()
        torch.cuda.synchronize()
        assert torch.allclose(out, ref, atol=1e-4, rtol=1e-4), (T, d, E)
        assert torch.equal(out, ref), (T, d, E)
        assert torch.allclose(out, exact, atol=1e-4, rtol=1e-4), (T, d, E)
    # the same problem twice in a row: identical bits (fixed summation order, no dependence on the array's contents)
    (dl, x), (T, d, E) = data[1]
    again = torch.empty_like(outs[1])
    _noalloc(_lib, dl, x, again)
    torch.cuda.synchronize()
    assert torch.equal(again, outs[1])


def test_router_bwd_dwg_noalloc_rejects_sizes_beyond_the_builtin_array():
    import b200moe
    from b200moe import _lib
    b200moe.load_library()
    T, d, E = 16384, 1024, 64
    assert _lib.load().b200moe_dwg_scratch_floats(T, d, E) > MAX_FLOATS
    dl, x = _inputs(T, d, E, seed=3)
    out = torch.zeros(E, d, device="cuda")
    with pytest.raises(RuntimeError, match="B200MOE_DWG_STATIC_MAX_FLOATS"):
        _noalloc(_lib, dl, x, out)
    torch.cuda.synchronize()
    assert float(out.abs().max()) == 0.0               # rejected before any launch
