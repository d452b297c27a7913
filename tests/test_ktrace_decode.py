"""Decoding of the in-kernel device trace (utils/ktrace.py; layout ``TraceSlot`` in csrc/gemm_tcgen05.cuh) on a
synthetic buffer: 16 x uint64 per CTA, cycles per wait, nanosecond time stamps."""
import torch

from collective_communication_mpi_b200.utils import ktrace as KT


def _row(t0, t1, tiles, **cycles):
    r = [0] * KT.WORDS
    r[0], r[8], r[9] = t0, t1, tiles
    for name, v in cycles.items():
        r[KT.SLOTS.index(name)] = v
    return r


def test_summary_reports_blocked_time_per_role_in_microseconds():
    buf = torch.zeros(KT.MAX_CTAS * KT.WORDS, dtype=torch.int64)
    rows = [_row(1_000_000, 1_400_000, 12, prod_flag=2000 * 30, mma_full=2000 * 80, epi_acc=2000 * 100),
            _row(1_002_000, 1_398_000, 11, prod_flag=2000 * 10, mma_full=2000 * 40, epi_acc=0, epi_bg=2000 * 50)]
    rows[0][10], rows[0][11], rows[0][12], rows[0][13] = 4, 2000 * 7, 1_050_000, 3         # slow producer calls / longest / first done / slices
    flat = torch.tensor(rows, dtype=torch.int64).view(-1)
    buf[:flat.numel()] = flat
    s = KT.summary(buf, sm_mhz=2000.0)                          # 2000 cycles = 1 us
    assert s["ctas"] == 2
    assert s["kernel_us"] == 400.0 and s["start_skew_us"] == 2.0 and s["tiles_per_cta"] == 11.5
    assert s["prod_flag_us"] == {"mean": 20.0, "max": 30.0}
    assert s["mma_full_us"] == {"mean": 60.0, "max": 80.0}
    assert s["epi_acc_us"] == {"mean": 100.0, "max": 100.0}    # the mean is over CTAs that waited at all
    assert s["epi_bg_us"] == {"mean": 50.0, "max": 50.0}
    assert s["prod_max_us"] == 7.0 and s["first_slow_wait_done_us"] == {"mean": 50.0, "max": 50.0}
    assert KT.summary(torch.zeros(KT.MAX_CTAS * KT.WORDS, dtype=torch.int64)) == {"ctas": 0}


def test_buffers_are_per_op_only_when_tracing_is_on_and_forgotten_on_close(monkeypatch):
    class Op:
        pass
    a = Op()
    monkeypatch.delenv("CCMPI_KERNEL_TRACE", raising=False)
    assert KT.buffer_for(a, "x") is None and KT.last(a) is None
    monkeypatch.setenv("CCMPI_KERNEL_TRACE", "1")
    monkeypatch.setattr(KT, "new_buffer", lambda device=None: torch.ones(KT.MAX_CTAS * KT.WORDS, dtype=torch.int64))
    buf = KT.buffer_for(a, "ag_gemm")
    assert buf is KT.last(a) and int(buf.sum()) == 0            # zeroed before every traced launch
    buf[0] = 5
    assert KT.buffer_for(a, "ag_gemm") is buf and int(buf[0]) == 0
    KT.forget(a)
    assert KT.last(a) is None
