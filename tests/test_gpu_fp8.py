"""fp8 GEMM numerics (per-tensor-scaled e4m3 and block-scaled MXFP8) against an fp32 matmul of the
dequantised operands."""
import os

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("M,N,K", [(256, 256, 256), (512, 384, 1024), (1024, 2048, 512)])
def test_fp8_gemm_matches_fp32_product_of_the_quantised_operands(M, N, K):
    from collective_communication_mpi_b200.ops.gemm_fp8 import gemm_fp8, quantize_e4m3
    torch.manual_seed(0)
    a = torch.randn(M, K, device="cuda")
    b = torch.randn(N, K, device="cuda")
    aq, sa = quantize_e4m3(a)
    bq, sb = quantize_e4m3(b)
    out = gemm_fp8(aq, bq, float(sa * sb), out_dtype=torch.float32)
    ref = (aq.float() @ bq.float().t()) * (sa * sb)
    torch.testing.assert_close(out, ref, rtol=1e-3, atol=1e-2)
    assert (out - a @ b.t()).abs().max().item() < 0.15 * (a @ b.t()).abs().max().item()


@pytest.mark.parametrize("M,N,K", [(256, 128, 128), (256, 256, 512), (512, 384, 1024), (384, 128, 256)])
def test_block_scaled_gemm_matches_the_dequantised_product(M, N, K):
    from collective_communication_mpi_b200.ops import gemm_fp8 as Q
    torch.manual_seed(1)
    # rows of very different magnitude: a per-tensor scale would lose the small ones
    a = torch.randn(M, K, device="cuda") * torch.logspace(-2, 2, M, device="cuda").unsqueeze(1)
    b = torch.randn(N, K, device="cuda")
    aq, asf = Q.quantize_mxfp8(a)
    bq, bsf = Q.quantize_mxfp8(b)
    out = Q.gemm_mxfp8(aq, asf, bq, bsf, out_dtype=torch.float32)
    ref = Q.dequantize_mxfp8(aq, asf) @ Q.dequantize_mxfp8(bq, bsf).t()
    torch.testing.assert_close(out, ref, rtol=2e-3, atol=1e-3 * ref.abs().max().item())
