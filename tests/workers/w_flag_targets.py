"""Single-process check (two emulated ranks in one cooperative launch) that all-gather->GEMM launches on ONE
flag region may change their CTA count from call to call -- another N, another SM budget -- when the host passes
the running sum of CTA counts as the absolute ready-flag target (ops.fused.ag_ctas). A target of ``epoch * ctas``
would be met early (stale ``x_full``: wrong numbers) or never (device watchdog) here."""
import torch

BF = torch.bfloat16


def main():
    from collective_communication_mpi_b200.ops import fused as F
    from collective_communication_mpi_b200.runtime.symm import SymmetricHeap, make_dev_comm
    nr, M, K = 2, 512, 1024
    chunk = K // nr
    heap = SymmetricHeap.create_emulated(nr, 128 << 20)
    try:
        comms = [make_dev_comm(heap, list(range(nr)), r, 0) for r in range(nr)]
        x_off = heap.alloc(M * chunk * 2)
        xf = [heap.alloc(M * K * 2) for _ in range(2)]
        g = torch.Generator(device="cuda").manual_seed(3)
        total, counts = 0, []
        for step, (N, sms) in enumerate([(256, 0), (1024, 0), (128, 0), (1024, 32), (384, 0), (256, 0)], 1):
            ctas = F.ag_ctas(0, M, N, nr, sms)
            counts.append(ctas)
            total += ctas
            for r in range(nr):
                heap.view(x_off, (M, chunk), BF, rank=r).copy_((torch.randn(M, chunk, device="cuda", generator=g) * 0.5).to(BF))
            x_full = torch.cat([heap.view(x_off, (M, chunk), BF, rank=r) for r in range(nr)], dim=1).float()
            ws = [(torch.randn(K, N, device="cuda", generator=g) * 0.5).to(BF) for _ in range(nr)]
            outs = [torch.zeros(M, N, device="cuda", dtype=torch.float32) for _ in range(nr)]
            F.ag_gemm_raw(comms, x_off, chunk, xf[step & 1], ws, outs, M, N, K, b_kn=True, flag_base=0, epoch=step,
                          flag_target=total, num_sms=sms, use_mc=False)
            torch.cuda.synchronize()
            for r in range(nr):
                torch.testing.assert_close(outs[r], x_full @ ws[r].float(), rtol=1e-3, atol=2e-2)
        assert len(set(counts)) > 1, f"the shapes were meant to give different CTA counts: {counts}"
    finally:
        heap.close()
    print("FLAG TARGETS OK", counts)


if __name__ == "__main__":
    main()
