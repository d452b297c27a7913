"""One launch of every kernel family of the library, between cudaProfilerStart/Stop, for `ncu --profile-from-start off`
(1 GPU; the multi-rank kernels run with two emulated ranks in one cooperative launch). Everything is launched once
before the profiled range as warm-up. scripts/ncu_all.sh drives it."""
import sys

sys.path.insert(0, ".")
import torch
from collective_communication_mpi_b200 import _native as N
from collective_communication_mpi_b200.ops import collectives as CO
from collective_communication_mpi_b200.ops import elementwise as EW
from collective_communication_mpi_b200.ops import fused as F
from collective_communication_mpi_b200.ops import gemm_fp8 as Q
from collective_communication_mpi_b200.ops.gemm import gemm, gemm_sgd_
from collective_communication_mpi_b200.runtime.symm import SymmetricHeap, make_dev_comm

BF = torch.bfloat16
g = torch.Generator(device="cuda").manual_seed(0)
rnd = lambda *s, sc=0.5: (torch.randn(*s, device="cuda", generator=g) * sc).to(BF)   # noqa: E731

M, H = 8192, 4096
x, w1 = rnd(M, H), rnd(H, 4 * H, sc=0.05)
u, act = torch.empty(M, 4 * H, device="cuda", dtype=BF), torch.empty(M, 4 * H, device="cuda", dtype=BF)
dy, w2 = rnd(M, H), rnd(4 * H, H, sc=0.05)
du = torch.empty(M, 4 * H, device="cuda", dtype=BF)
wq = rnd(H, 4 * H, sc=0.05)
qkv, da = rnd(M, 3 * H), rnd(M, H)
y, t = rnd(M, H), rnd(M, H)
wflat, gflat = rnd(1 << 24, sc=0.02), rnd(1 << 24, sc=0.01)

nr = 2
heap = SymmetricHeap.create_emulated(nr, 1536 << 20)
comms = [make_dev_comm(heap, list(range(nr)), r, 0) for r in range(nr)]
Me, Ke, Ne = 4096, 4096, 3072
chunk = Ke // nr
x_off = heap.alloc(Me * chunk * 2)
xf = [heap.alloc(Me * Ke * 2) for _ in range(2)]
for r in range(nr):
    heap.view(x_off, (Me, chunk), BF, rank=r).copy_(rnd(Me, chunk))
Bs = [rnd(Ke, Ne, sc=0.05) for _ in range(nr)]
outs = [torch.empty(Me, Ne, device="cuda", dtype=BF) for _ in range(nr)]
As = [rnd(Me, 2048) for _ in range(nr)]
Ws = [rnd(2048, 4096, sc=0.05) for _ in range(nr)]
slots = [heap.alloc(nr * Me * 2048 * 2) for _ in range(2)]
rs_outs = [torch.empty(Me, 2048, device="cuda", dtype=BF) for _ in range(nr)]
cnt = 32 << 20
s_off, d_off, t_off = heap.alloc(cnt * 2), heap.alloc(cnt * 2), heap.alloc(cnt * 2)
xq_off, sfl_off = heap.alloc(Me * chunk), heap.alloc(Q.sf_bytes(Me, chunk))
xqf, sff = heap.alloc(Me * Ke), heap.alloc(Q.sf_bytes(Me, Ke))
for r in range(nr):
    Q.quantize_mxfp8_cuda(rnd(Me, chunk), heap.view(xq_off, (Me, chunk), torch.uint8, rank=r),
                          heap.view(sfl_off, (Q.sf_bytes(Me, chunk),), torch.uint8, rank=r))
Wq = [Q.QuantizedWeight(b) for b in Bs]
aq, asf = Q.quantize_mxfp8_cuda(x)
bq = Q.QuantizedWeight(w1)
state = {"epoch": 0}


def everything():
    state["epoch"] += 1
    e = state["epoch"]
    gemm(x, w1, b_kn=True, out=u, gelu_out=act)                         # fc1 forward + GELU in the epilogue
    gemm(dy, w2, out=du, dgelu_of=u)                                      # fc2 dx times gelu'(u)
    gemm_sgd_(wq, x, du, 1e-4, a_km=True, b_kn=True, step=e, seed=1)      # dW with the SGD step (stochastic rounding)
    F.ag_gemm_raw(comms, x_off, chunk, xf[e & 1], Bs, outs, Me, Ne, Ke, b_kn=True, flag_base=0, epoch=e, use_mc=False)
    F.gemm_rs_raw(comms, As, Ws, slots[e & 1], rs_outs, Me, 4096, 2048, b_kn=True, flag_base=4096, epoch=e)
    CO.allreduce(comms, N.BF16, N.SUM, s_off, d_off, cnt, tmp_off=t_off, algo=N.AR_TWOSHOT)
    CO.allreduce(comms, N.BF16, N.SUM, s_off, d_off, 32768, tmp_off=t_off, algo=N.AR_ONESHOT)
    CO.allgather(comms, s_off, d_off, 1, cnt, 0, 0, use_mc=False)
    CO.reduce_scatter(comms, N.BF16, N.SUM, s_off, d_off, 1, cnt // 2, 0, 0, use_mc=False)
    CO.alltoall(comms, s_off, d_off, cnt)
    CO.barrier(comms)
    a = EW.gate_forward(qkv)
    EW.gate_backward(da, qkv)
    EW.mse_loss_and_grad(y, t)
    EW.sgd_sr_(wflat, gflat, 1e-3, step=e)
    Q.quantize_mxfp8_cuda(x)
    Q.gemm_mx(aq, asf, bq.q, bq.sf, out=u)
    Q.ag_gemm_mx_raw(comms, xq_off, chunk, sfl_off, xqf, sff, [w.q for w in Wq], [w.sf for w in Wq], outs, Me, Ne, Ke,
                     flag_base=8192, epoch=e)
    return a


for _ in range(2):
    everything()
torch.cuda.synchronize()
torch.cuda.profiler.start()
everything()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ncu targets ok")
