"""Token-axis (sequence-parallel) fused kernels with emulated ranks on one GPU (first run on a B200 in
round 2, profiles/r2_unverified_1gpu.txt). Oracle: the unfused math in fp32 PyTorch (concatenate rows /
matmul / rank-ordered sum / row slice)."""
import os

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

BF = torch.bfloat16


@pytest.fixture(params=[2, 4])
def emu(request):
    from collective_communication_mpi_b200.runtime.symm import SymmetricHeap, make_dev_comm
    nr = request.param
    heap = SymmetricHeap.create_emulated(nr, 256 << 20)
    comms = [make_dev_comm(heap, list(range(nr)), r, 0) for r in range(nr)]
    yield nr, heap, comms
    heap.close()


def _rand(*shape, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(*shape, device="cuda", generator=g) * 0.5).to(BF)


@pytest.mark.parametrize("rows,N,K,b_kn", [(256, 256, 512, True), (512, 384, 1024, False), (256, 128, 192, True)])
def test_allgather_gemm_row_gather(emu, rows, N, K, b_kn):
    from collective_communication_mpi_b200.ops import fused_seq as S
    nr, heap, comms = emu
    M = rows * nr
    x_off = heap.alloc(rows * K * 2)
    xf = [heap.alloc(M * K * 2) for _ in range(2)]
    xs = [_rand(rows, K, seed=10 + r) for r in range(nr)]
    ws = [_rand(K, N, seed=20 + r) for r in range(nr)]
    for epoch in (1, 2, 3):
        for r in range(nr):
            heap.view(x_off, (rows, K), BF, rank=r).copy_(xs[r] * epoch)
        x_full = torch.cat([heap.view(x_off, (rows, K), BF, rank=r) for r in range(nr)], dim=0)
        outs = [torch.zeros(M, N, device="cuda", dtype=torch.float32) for _ in range(nr)]
        B = [w if b_kn else w.t().contiguous() for w in ws]
        S.ag_gemm_m_raw(comms, x_off, K, xf[epoch & 1], B, outs, M, N, K, b_kn=b_kn, flag_base=0, epoch=epoch)
        torch.cuda.synchronize()
        for r in range(nr):
            torch.testing.assert_close(outs[r], x_full.float() @ ws[r].float(), rtol=1e-3, atol=2e-2)
            # the gathered activation is left whole on every rank (the backward dW GEMM reads it)
            assert torch.equal(heap.view(xf[epoch & 1], (M, K), BF, rank=r), x_full)


@pytest.mark.parametrize("rows,N,K,b_kn,dt", [(256, 256, 512, False, BF), (256, 384, 256, False, torch.float32),
                                              (512, 512, 1024, True, BF)])
def test_gemm_reduce_scatter_row_scatter(emu, rows, N, K, b_kn, dt):
    from collective_communication_mpi_b200.ops import fused_seq as S
    nr, heap, comms = emu
    M = rows * nr
    es = 2 if dt == BF else 4
    slots = [heap.alloc(nr * rows * N * es) for _ in range(2)]
    As = [_rand(M, K, seed=30 + r) for r in range(nr)]
    Ws = [_rand(K, N, seed=40 + r) for r in range(nr)]
    total = sum(a.float() @ w.float() for a, w in zip(As, Ws))
    B = [w if b_kn else w.t().contiguous() for w in Ws]
    results = []
    for epoch in (1, 2, 3):
        outs = [torch.zeros(rows, N, device="cuda", dtype=dt) for _ in range(nr)]
        S.gemm_rs_m_raw(comms, As, B, slots[epoch & 1], outs, M, N, K, b_kn=b_kn, flag_base=4096, epoch=epoch)
        torch.cuda.synchronize()
        for r in range(nr):
            ref = total[r * rows:(r + 1) * rows]
            tol = 0.15 if dt == BF else 2e-2
            assert (outs[r].float() - ref).abs().max().item() < tol * max(1.0, ref.abs().max().item() / 10)
        results.append(outs)
    for r in range(nr):                       # rank-ordered reduction: same inputs, same bits
        assert torch.equal(results[0][r], results[2][r])


@pytest.mark.parametrize("M,N,K,b_kn,dt", [(256, 256, 512, True, BF), (640, 384, 1024, False, BF), (200, 128, 512, True, torch.float32)])
def test_gather_gemm_gather_in_one_kernel(emu, M, N, K, b_kn, dt):
    from collective_communication_mpi_b200.ops import fused_seq as S
    nr, heap, comms = emu
    chunk = K // nr
    es = 2 if dt == BF else 4
    x_off = heap.alloc(M * chunk * 2)
    xf = [heap.alloc(M * K * 2) for _ in range(2)]
    yf = [heap.alloc(M * N * nr * es) for _ in range(2)]
    xs = [_rand(M, chunk, seed=50 + r) for r in range(nr)]
    ws = [_rand(K, N, seed=60 + r) for r in range(nr)]
    B = [w if b_kn else w.t().contiguous() for w in ws]
    for epoch in (1, 2, 3):
        for r in range(nr):
            heap.view(x_off, (M, chunk), BF, rank=r).copy_(xs[r] * epoch)
        x_full = torch.cat([heap.view(x_off, (M, chunk), BF, rank=r) for r in range(nr)], dim=1).float()
        ref = torch.cat([x_full @ w.float() for w in ws], dim=1)
        S.ag_gemm_ag_raw(comms, x_off, chunk, xf[epoch & 1], B, yf[epoch & 1], M, N, K, b_kn=b_kn, out_dtype=dt,
                         flag_base=8192, epoch=epoch)
        torch.cuda.synchronize()
        for r in range(nr):
            got = heap.view(yf[epoch & 1], (M, N * nr), dt, rank=r).float()
            tol = 0.15 if dt == BF else 2e-2
            assert (got - ref).abs().max().item() < tol * max(1.0, ref.abs().max().item() / 10)
            assert torch.equal(heap.view(xf[epoch & 1], (M, K), BF, rank=r).float(), x_full)


def test_row_constraints_are_reported(emu):
    from collective_communication_mpi_b200 import _native as Nat
    from collective_communication_mpi_b200.ops import fused_seq as S
    nr, heap, comms = emu
    a = _rand(128 * nr, 128, seed=1)
    outs = [torch.zeros(128, 128, device="cuda", dtype=BF) for _ in range(nr)]
    with pytest.raises(Nat.NativeError, match="multiple of 256"):
        S.gemm_rs_m_raw(comms, [a] * nr, [a] * nr, 0, outs, 128 * nr, 128, 128, b_kn=False, flag_base=0, epoch=1)


def test_sequence_parallel_ops_across_processes():
    from conftest import gpu_count, run_ranks
    n = min(gpu_count(), 8)
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    n = 1 << (n.bit_length() - 1)
    proc = run_ranks(n, ["tests/workers/w_seq.py"], transport="device")
    assert proc.returncode == 0, f"rc={proc.returncode}\n{proc.stdout[-3000:]}\n{proc.stderr[-3000:]}"
    assert f"SEQ OK ranks={n}" in proc.stdout


@pytest.mark.parametrize("fused", ["0", "1"])
def test_column_parallel_layer_with_gathered_output(fused):
    from conftest import gpu_count, run_ranks
    n = min(gpu_count(), 4)
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    n = 1 << (n.bit_length() - 1)
    proc = run_ranks(n, ["tests/workers/w_gather_output.py"], transport="device", env={"CCMPI_FUSED_GATHER_OUTPUT": fused})
    assert proc.returncode == 0, f"rc={proc.returncode}\n{proc.stdout[-3000:]}\n{proc.stderr[-3000:]}"
    assert f"GATHER OUTPUT OK ranks={n} fused={fused}" in proc.stdout


@pytest.mark.parametrize("mp,dp", [(2, 1), (2, 2), (4, 2)])
def test_sequence_parallel_block_matches_dense_model(mp, dp):
    from conftest import gpu_count, run_ranks
    if gpu_count() < mp * dp:
        pytest.skip(f"needs {mp * dp} GPUs")
    proc = run_ranks(mp * dp, ["tests/workers/w_layers.py", str(mp), str(dp), "--sp"], transport="device")
    assert proc.returncode == 0, f"rc={proc.returncode}\n{proc.stdout[-3000:]}\n{proc.stderr[-3000:]}"
    assert f"LAYERS OK mp={mp} dp={dp} sp=1" in proc.stdout


@pytest.mark.parametrize("mp,dp", [(2, 1), (2, 2), (8, 1)])
def test_megatron_layout_block_matches_dense_model(mp, dp):
    """Classic Megatron layout (replicated activations, GEMM -> all-reduce) against the dense model."""
    from conftest import gpu_count, run_ranks
    if gpu_count() < mp * dp:
        pytest.skip(f"needs {mp * dp} GPUs")
    proc = run_ranks(mp * dp, ["tests/workers/w_layers.py", str(mp), str(dp), "--megatron"], transport="device")
    assert proc.returncode == 0, f"rc={proc.returncode}\n{proc.stdout[-3000:]}\n{proc.stderr[-3000:]}"
    assert f"LAYERS OK mp={mp} dp={dp} sp=0 megatron=1" in proc.stdout
