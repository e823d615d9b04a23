"""-m gpu parity of rows a14 / a15 through the C ABI of include/lt_policy.h: the tcgen05 split-fp16
contraction kernel against a float64 product, the actor-critic forward against the oracle port on the reference-recorded
weights / inputs (tests/golden/ppo_*.npz), and the minibatch gradient against torch autograd of the oracle's loss.

Tolerances: north_star's 1e-5 relative for fp32 outputs, measured against the magnitude of the contraction (sum |a||b|), which
is how fp32 GEMM error scales; gradients 1e-4 relative to the largest entry of each tensor."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


@pytest.mark.parametrize("M,N,K,splits", [(128, 128, 64, 1), (128, 16, 64, 1), (300, 200, 263, 1), (1000, 512, 263, 1), (257, 16, 128, 1),
                                          (64, 12, 512, 1), (4096, 256, 512, 1), (512, 144, 4096, 3), (12, 129, 10000, 7), (1, 1, 1, 1),
                                          (65536, 512, 265, 1)])
def test_split_fp16_gemm_matches_float64(M, N, K, splits):
    """lt_linear_f32: out = a @ b.T + bias with every product as three tcgen05 fp16 MMAs (hi*hi + hi*lo + lo*hi), fp32 accumulate.
    Ragged M / N / K (TMA zero fill + epilogue masks), bn = 16 .. 256, split-K."""
    from legged_tracking_b200 import _abi
    lib = _abi.load()
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    a = torch.randn(M, K, device="cuda", generator=g)
    b = torch.randn(N, K, device="cuda", generator=g) * 0.1
    bias = torch.randn(N, device="cuda", generator=g)
    out = torch.full((M, N), float("nan"), device="cuda")
    _abi.check_policy(lib.lt_linear_f32(_p(a), _p(b), _p(bias), _p(out), M, N, K, splits, None), None, "lt_linear_f32")
    torch.cuda.synchronize()
    want = a.double() @ b.double().T + bias.double()
    mag = a.double().abs() @ b.double().abs().T + bias.double().abs()
    rel = ((out.double() - want).abs() / mag).max().item()
    assert rel < 2e-6, f"max error / sum|a||b| = {rel:.3e}"      # measured 2e-7 .. 1e-6 (torch fp32: 2e-7 .. 4e-7)
