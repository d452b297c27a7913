"""MX (block-scaled fp8) quantiser and scale layout of ops/gemm_fp8.py on the CPU: round-trip error,
power-of-two scales, and the chunked hardware layout against its index formula."""
import pytest

torch = pytest.importorskip("torch")

from collective_communication_mpi_b200.ops import gemm_fp8 as Q  # noqa: E402


def test_round_trip_is_within_e4m3_precision_per_block():
    torch.manual_seed(0)
    x = torch.randn(64, 256) * torch.logspace(-3, 3, 64).unsqueeze(1)        # rows of very different magnitude
    q, sf = Q.quantize_mxfp8(x)
    assert q.dtype == torch.float8_e4m3fn and sf.dtype == torch.uint8 and sf.shape == (64, 8)
    back = Q.dequantize_mxfp8(q, sf)
    blocks = x.reshape(64, 8, 32)
    amax = blocks.abs().amax(-1, keepdim=True)
    err = (back.reshape(64, 8, 32) - blocks).abs()
    assert (err <= amax * 2 ** -3).all()             # e4m3: 3 mantissa bits, scale within 2x of ideal
    assert err.mean() < 0.03 * blocks.abs().mean()
    # the block maximum must not saturate: scaled values stay within +-448
    assert (q.float().abs() <= 448).all()


def test_zero_and_tiny_blocks_do_not_produce_nan():
    x = torch.zeros(4, 64)
    x[1, :32] = 1e-30
    q, sf = Q.quantize_mxfp8(x)
    assert torch.isfinite(Q.dequantize_mxfp8(q, sf)).all()
    assert (Q.dequantize_mxfp8(q, sf)[0] == 0).all()


def test_scale_layout_matches_the_index_formula():
    R, Ks = 300, 10                                   # neither a multiple of the 128 x 4 chunk
    sf = (torch.arange(R * Ks) % 251).to(torch.uint8).reshape(R, Ks)
    flat = Q.swizzle_scales(sf)
    assert flat.numel() == 3 * 3 * 512
    for r in (0, 1, 31, 32, 33, 127, 128, 255, 299):
        for k in (0, 3, 4, 9):
            assert int(flat[Q.sf_chunk_offset(r, k, Ks)]) == int(sf[r, k]), (r, k)
    # padding carries the neutral exponent
    assert int(flat[Q.sf_chunk_offset(299, 9, Ks) + 1]) == 127


def test_rejects_shapes_that_do_not_tile():
    with pytest.raises(ValueError):
        Q.quantize_mxfp8(torch.zeros(4, 40))
