"""The tcgen05 GEMM kernel (dp_gemm_bf16 through the C-ABI) against a plain PyTorch fp32 reference of the same op."""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _gemm(A, B, bias, out_fp32, block_n):
    from deepore_b200 import _lib
    M, K = A.shape
    N = B.shape[0]
    out = torch.full((M, N), float("nan"), dtype=torch.float32 if out_fp32 else torch.bfloat16, device="cuda")
    rc = _lib.lib().dp_gemm_bf16(A.data_ptr(), A.stride(0), B.data_ptr(), B.stride(0),
                                 bias.data_ptr() if bias is not None else None, out.data_ptr(), out.stride(0),
                                 1 if out_fp32 else 0, M, N, K, block_n, C.c_void_p(torch.cuda.current_stream().cuda_stream))
    _lib.check(rc, "dp_gemm_bf16")
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("M,N,K,block_n,out_fp32", [
    (128, 256, 64, 256, True),          # one tile, one K chunk
    (128, 256, 256, 256, True),         # K loop, ring wrap inside a tile (4 stages)
    (300, 512, 2496, 256, False),       # M tail, several tiles per CTA, folded-operator K
    (1000, 1536, 832, 256, True),       # more tiles than one wave would need on small grids
    (257, 384, 200, 128, True),         # BN=128, K tail (200 = 3 chunks + zero-filled remainder), M tail
    (20000, 1536, 2496, 256, False),    # persistent loop: 157*6 tiles over 148 CTAs, accumulator double buffering
])
def test_gemm_matches_torch(M, N, K, block_n, out_fp32):
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N)
    ld = (K + 7) // 8 * 8
    A = torch.zeros((M, ld), dtype=torch.bfloat16, device="cuda")
    B = torch.zeros((N, ld), dtype=torch.bfloat16, device="cuda")
    A[:, :K] = torch.randn((M, K), generator=g, device="cuda").to(torch.bfloat16)
    B[:, :K] = (torch.randn((N, K), generator=g, device="cuda") / 8).to(torch.bfloat16)
    bias = torch.randn((N,), generator=g, device="cuda")
    out = _gemm(A[:, :K], B[:, :K], bias, out_fp32, block_n)
    ref = A[:, :K].float() @ B[:, :K].float().t() + bias
    err = (out.float() - ref).abs().max().item()
    scale = ref.abs().max().item()
    tol = (2e-4 if out_fp32 else 1e-2) * scale
    assert np.isfinite(err) and err <= tol, "max abs err %g (scale %g)" % (err, scale)
