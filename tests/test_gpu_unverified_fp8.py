"""fp8 GEMM numerics. The kernel variant was written after round 1's GPU budget was spent and has not
run on hardware yet, so this file is opt-in: set CCMPI_TEST_UNVERIFIED=1 to run it. Once it has
passed on a B200 the gate goes away."""
import os

import pytest

torch = pytest.importorskip("torch")
pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(os.environ.get("CCMPI_TEST_UNVERIFIED") != "1", reason="kernel not yet verified on hardware")]


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
