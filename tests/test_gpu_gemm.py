"""tcgen05 GEMM numerics against a plain PyTorch fp32 reference of the same product."""
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _ref(A, B):
    return A.float() @ B.float()


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (256, 256, 128), (384, 328, 200), (1024, 2048, 512), (4096, 128, 512),
                                   (136, 72, 1024)])
@pytest.mark.parametrize("a_km,b_kn", [(False, False), (False, True), (True, False), (True, True)])
def test_gemm_matches_fp32_reference_in_every_operand_layout(M, N, K, a_km, b_kn):
    from collective_communication_mpi_b200.ops.gemm import gemm
    torch.manual_seed(M + N + K)
    A = (torch.randn(M, K, device="cuda") * 0.5).bfloat16()
    B = (torch.randn(K, N, device="cuda") * 0.5).bfloat16()
    a = A.t().contiguous() if a_km else A
    b = B if b_kn else B.t().contiguous()
    ref = _ref(A, B)
    out32 = gemm(a, b, a_km=a_km, b_kn=b_kn, out_dtype=torch.float32)
    torch.testing.assert_close(out32, ref, rtol=1e-4, atol=1e-3)
    out16 = gemm(a, b, a_km=a_km, b_kn=b_kn)
    torch.testing.assert_close(out16.float(), ref, rtol=2e-2, atol=2e-2 * ref.abs().max().item())


def test_named_products_and_accumulate():
    from collective_communication_mpi_b200.ops.gemm import matmul_nn, matmul_nt, matmul_tn
    torch.manual_seed(0)
    x = torch.randn(512, 256, device="cuda").bfloat16()
    w = torch.randn(256, 384, device="cuda").bfloat16()      # stored [in, out]
    dy = torch.randn(512, 384, device="cuda").bfloat16()
    torch.testing.assert_close(matmul_nn(x, w, out_dtype=torch.float32), x.float() @ w.float(), rtol=1e-4, atol=1e-3)
    torch.testing.assert_close(matmul_nt(dy, w, out_dtype=torch.float32), dy.float() @ w.float().t(), rtol=1e-4, atol=1e-3)
    acc = torch.ones(256, 384, device="cuda")
    matmul_tn(x, dy, out=acc, accumulate=True)
    torch.testing.assert_close(acc, 1 + x.float().t() @ dy.float(), rtol=1e-4, atol=1e-3)


def test_strided_operands_and_sm_cap():
    from collective_communication_mpi_b200.ops.gemm import gemm
    big = torch.randn(300, 1024, device="cuda").bfloat16()
    a = big[:256, 128:640]                                    # row stride 1024, offset columns
    b = torch.randn(192, 512, device="cuda").bfloat16()
    out = gemm(a, b, out_dtype=torch.float32, num_sms=16)
    torch.testing.assert_close(out, a.float() @ b.float().t(), rtol=1e-4, atol=1e-3)


def test_bad_operands_are_rejected_before_launch():
    from collective_communication_mpi_b200.ops.gemm import gemm
    a = torch.randn(64, 64, device="cuda")
    with pytest.raises(TypeError):
        gemm(a, a)
    h = torch.randn(64, 60, device="cuda").bfloat16()
    with pytest.raises(Exception):
        gemm(h, h)
