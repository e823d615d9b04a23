"""lt_linear_f32_static (include/lt_policy.h): the allocation-free variant of lt_linear_f32 whose fp16 operand planes live in a
static device array.  No counterpart in the reference.  Collected LAST on purpose
(file name).

The GPU test calls the new entry point through the C ABI and compares it with the EXISTING entry point (lt_linear_f32, splits = 1:
same kernel, so the results should in fact be bit-identical) at the tolerance tests/test_gpu_policy.py uses, on two calls in a
row with different shapes -- the second call finds the first call's planes in the shared array -- and then on the first shape
again.  It has to pass alone in a fresh process (first use looks the symbol address up) and after the other tests."""
import ctypes as C
import os
import re

import pytest
import torch

TOL = 2e-6          # max |error| / sum |a||b|, as in tests/test_gpu_policy.py::test_split_fp16_gemm_matches_float64


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def test_static_entry_point_body_calls_no_allocator():
    """Source-level guard (no GPU needed): nothing between the entry point's signature and its closing brace allocates."""
    src = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "legged_tracking_b200", "csrc", "lt_policy.cu")).read()
    m = re.search(r"\nint lt_linear_f32_static\(.*?\n}\n", src, flags=re.S)
    assert m, "lt_linear_f32_static not found"
    body = m.group(0)
    for word in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "cudaMallocAsync", "alloc_dev", "alloc_planes", "new ", "malloc("):
        assert word not in body, word
    assert "static_scratch_ptr(dev)" in body
    assert "cudaGetSymbolAddress(&ptr, g_linear_static_scratch)" in src
    assert re.search(r"static __device__ __align__\(256\) unsigned char g_linear_static_scratch\[LT_LINEAR_STATIC_SCRATCH_BYTES\];", src)


@pytest.mark.gpu
def test_static_scratch_linear_matches_allocating_entry_point():
    from legged_tracking_b200 import _abi
    lib = _abi.load()
    limit = _abi.DEFINES["LT_LINEAR_STATIC_SCRATCH_BYTES"]
    shapes = [(1024, 256, 263), (300, 200, 100), (1024, 256, 263)]          # second call: smaller, other row pitch, K % 8 != 0
    g = torch.Generator(device="cuda").manual_seed(77)
    first = None
    for i, (M, N, K) in enumerate(shapes):
        assert 4 * (M + N) * ((K + 7) // 8 * 8) <= limit
        if i < 2:
            a = torch.randn(M, K, device="cuda", generator=g)
            b = torch.randn(N, K, device="cuda", generator=g) * 0.1
            bias = torch.randn(N, device="cuda", generator=g)
            if i == 0:
                first = (a, b, bias)
        else:
            a, b, bias = first
        want = torch.full((M, N), float("nan"), device="cuda")
        got = torch.full((M, N), float("nan"), device="cuda")
        torch.cuda.synchronize()
        stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        _abi.check_policy(lib.lt_linear_f32(_p(a), _p(b), _p(bias), _p(want), M, N, K, 1, stream), None, "lt_linear_f32")
        _abi.check_policy(lib.lt_linear_f32_static(_p(a), _p(b), _p(bias), _p(got), M, N, K, stream), None, "lt_linear_f32_static")
        torch.cuda.synchronize()
        assert not torch.isnan(got).any()
        mag = a.double().abs() @ b.double().abs().T + bias.double().abs()
        rel = ((got.double() - want.double()).abs() / mag).max().item()
        assert rel <= TOL, f"call {i} {M}x{N}x{K}: static vs allocating entry point {rel:.3e}"
        ref = a.double() @ b.double().T + bias.double()
        rel64 = ((got.double() - ref).abs() / mag).max().item()
        assert rel64 <= TOL, f"call {i} {M}x{N}x{K}: static entry point vs float64 {rel64:.3e}"
        print(f"call {i} {M}x{N}x{K}: bit-identical to lt_linear_f32: {bool(torch.equal(got, want))}; vs float64 {rel64:.2e}")


@pytest.mark.gpu
def test_static_scratch_linear_rejects_shapes_beyond_the_array():
    from legged_tracking_b200 import _abi
    lib = _abi.load()
    M, N, K = 8192, 512, 512                      # 4 * 8704 * 512 = 17.8 MB > 8 MiB
    assert 4 * (M + N) * K > _abi.DEFINES["LT_LINEAR_STATIC_SCRATCH_BYTES"]
    a = torch.zeros(M, K, device="cuda")
    b = torch.zeros(N, K, device="cuda")
    out = torch.full((M, N), 7.0, device="cuda")
    rc = lib.lt_linear_f32_static(_p(a), _p(b), None, _p(out), M, N, K, None)
    assert rc == _abi.ENUMS["LT_ERR_INVALID"]
    msg = lib.lt_policy_last_error(None).decode()
    assert "static array holds 8388608" in msg and "lt_linear_f32" in msg
    torch.cuda.synchronize()
    assert float(out.min()) == 7.0 and float(out.max()) == 7.0       # nothing was launched
