"""Flag-word allocator of the fused ops (ops/flags.py) and the close() life cycle (ops/fused.py:
FusedOp), without a GPU: the device part of a release is replaced by a recording stand-in."""
import random

import pytest
import torch

from collective_communication_mpi_b200 import _native as N
from collective_communication_mpi_b200.ops import fused as F
from collective_communication_mpi_b200.ops.flags import GRANULE, FlagAllocator, FlagSpaceExhausted


def test_fresh_allocator_is_a_bump_pointer():
    a = FlagAllocator(N.PAD_USER_WORDS)
    bases = [a.alloc(w) for w in (1, 16, 17, 100, 3)]
    assert bases == [0, 16, 32, 64, 176]
    assert a.used == 16 + 16 + 32 + 112 + 16


def test_capacity_is_the_pad_user_area_and_exhaustion_names_the_numbers():
    a = FlagAllocator(N.PAD_USER_WORDS)
    a.alloc(N.PAD_USER_WORDS - GRANULE)
    a.alloc(GRANULE)
    with pytest.raises(FlagSpaceExhausted, match=r"need 16 flag words, largest free run 0 of %d" % N.PAD_USER_WORDS):
        a.alloc(1)


def test_freed_runs_coalesce_and_are_reused_first_fit():
    a = FlagAllocator(1024)
    b0, b1, b2, b3 = (a.alloc(64) for _ in range(4))
    a.free(b1)
    a.free(b2)                       # neighbours merge into one 128-word run
    assert a.alloc(128) == b1
    a.free(b0)
    assert a.alloc(32) == b0         # lowest run that fits
    assert a.alloc(32) == b0 + 32
    a.free(b3)
    assert a.alloc(1024 - 256) == b3  # tail merges with the never-used remainder


def test_double_free_and_unknown_base_are_errors():
    a = FlagAllocator(256)
    b = a.alloc(10)
    assert a.free(b) == 16
    with pytest.raises(ValueError, match="double close"):
        a.free(b)
    with pytest.raises(ValueError):
        a.free(48)


def test_random_alloc_free_never_overlaps_and_never_leaks():
    rng = random.Random(7)
    a = FlagAllocator(4096)
    live = {}
    for _ in range(3000):
        if live and (rng.random() < 0.45 or a.used > 3500):
            base = rng.choice(list(live))
            assert a.free(base) == live.pop(base)
        else:
            w = rng.randint(1, 200)
            try:
                base = a.alloc(w)
            except FlagSpaceExhausted:
                continue
            size = FlagAllocator.rounded(w)
            assert base % GRANULE == 0 and base + size <= 4096
            for b, s in live.items():
                assert base + size <= b or b + s <= base, "overlapping flag runs"
            live[base] = size
        assert a.used == sum(live.values())
    for base in list(live):
        a.free(base)
    assert a.used == 0 and a.alloc(4096) == 0     # everything merged back into one run


# ---- close() of the op objects, with the device side replaced -----------------------------------
class _Heap:
    def __init__(self):
        self.freed, self._next = [], 0

    def empty(self, shape, dtype):
        off = self._next
        self._next += 256
        return off, torch.zeros(shape, dtype=dtype)

    def free(self, off):
        self.freed.append(off)


class _Ctx:
    def __init__(self):
        self.heap, self.flags, self.released = _Heap(), FlagAllocator(N.PAD_USER_WORDS), []

    def new_slot(self, words):
        return self.flags.alloc(words)

    def release_slot(self, base):
        self.released.append((base, self.flags.size_of(base)))
        self.flags.free(base)


class _Op(F.FusedOp):
    def __init__(self, ctx, words):
        self.ctx = ctx
        self.x_off, self.x = ctx.heap.empty((4, 4), torch.bfloat16)
        self.xfull = [ctx.heap.empty((4, 8), torch.bfloat16) for _ in range(2)]
        self.single = [ctx.heap.empty((2, 2), torch.bfloat16)] * 2       # the n == 1 idiom: one buffer listed twice
        self.flag_base = ctx.new_slot(words)
        self.epoch = 0

    def __call__(self):
        self._check_open()
        self.epoch += 1
        return self.xfull[self.epoch & 1][1]


def test_close_returns_every_buffer_once_and_the_flag_run():
    ctx = _Ctx()
    a, b = _Op(ctx, 40), _Op(ctx, 8)
    assert a._heap_offsets() == [0, 256, 512, 768]
    a()
    with a:
        pass                                                  # context manager closes
    assert ctx.heap.freed == [0, 256, 512, 768] and ctx.released == [(0, 48)]
    assert a.x is None and a.xfull is None
    with pytest.raises(RuntimeError, match="used after close"):
        a()
    a.close()                                                 # idempotent
    assert len(ctx.heap.freed) == 4
    c = _Op(ctx, 30)
    assert c.flag_base == 0 and b.flag_base == 48             # the freed run is handed out again
    b.close(), c.close()
    assert ctx.flags.used == 0


def test_close_fused_ops_walks_the_module_tree_in_a_fixed_order():
    ctx = _Ctx()

    class Layer(torch.nn.Module):
        def __init__(self, w):
            super().__init__()
            self.rs, self.ag, self.none = _Op(ctx, w), _Op(ctx, w + 16), None

    model = torch.nn.Sequential(Layer(16), Layer(64))
    assert F.close_fused_ops(model) == 4
    assert [b for b, _ in ctx.released] == [16, 0, 112, 48]   # per layer: 'ag' before 'rs' (sorted names)
    assert F.close_fused_ops(model) == 0 and ctx.flags.used == 0


def test_real_ops_share_the_life_cycle():
    from collective_communication_mpi_b200.ops import fused_seq as FS, gemm_fp8 as F8
    for cls in (F.AllGatherGemm, F.GemmReduceScatter, F.GemmAllReduce, F.GemmAllGather, FS.AllGatherGemmM,
                FS.GemmReduceScatterM, FS.AllGatherGemmAllGather, F8.AllGatherGemmMX, F8.GemmReduceScatterMX):
        assert issubclass(cls, F.FusedOp), cls


# ---- the symmetric heap's allocator (runtime/symm.py): the same first-fit + coalescing policy, in bytes.
# The object is built without its CUDA half (the allocator is pure bookkeeping; offsets must come out the
# same on every rank, which is why it is deterministic and has no address-dependent behaviour).
def _bare_heap(nbytes):
    from collective_communication_mpi_b200.runtime.symm import PAD_REGION, SymmetricHeap
    h = object.__new__(SymmetricHeap)
    h.bytes = PAD_REGION + nbytes
    h._free, h._live = [(PAD_REGION, nbytes)], {}
    return h, PAD_REGION


def test_symmetric_heap_offsets_are_deterministic_aligned_and_reused_after_free():
    from collective_communication_mpi_b200.runtime.symm import ALIGN
    (a, base), (b, _) = _bare_heap(1 << 20), _bare_heap(1 << 20)
    sizes = [1, 255, 256, 257, 4096, 100_000, 7]
    offs_a = [a.alloc(s) for s in sizes]
    offs_b = [b.alloc(s) for s in sizes]
    assert offs_a == offs_b, "every rank must compute the same offsets from the same call sequence"
    assert offs_a[0] == base and all(o % ALIGN == 0 for o in offs_a)
    assert all(o2 - o1 >= s for o1, o2, s in zip(offs_a, offs_a[1:], sizes)), "no overlap"
    a.free(offs_a[1])
    a.free(offs_a[2])                                    # neighbours coalesce: 512 B hole
    assert a.alloc(512) == offs_a[1]
    a.free(offs_a[4])
    assert a.alloc(4096) == offs_a[4]                    # exact fit is taken, nothing is split off
    with pytest.raises(KeyError):
        a.free(12345)                                    # not an allocation
    with pytest.raises(MemoryError, match="symmetric heap exhausted: need .* largest free block"):
        a.alloc(2 << 20)


def test_symmetric_heap_random_alloc_free_never_overlaps_and_merges_back():
    from collective_communication_mpi_b200.runtime.symm import ALIGN
    rng = random.Random(3)
    h, base = _bare_heap(4 << 20)
    live = {}
    for _ in range(4000):
        if live and rng.random() < 0.5:
            off = rng.choice(list(live))
            h.free(off)
            del live[off]
        else:
            want = rng.randint(1, 200_000)
            try:
                off = h.alloc(want)
            except MemoryError:
                continue
            size = max(ALIGN, (want + ALIGN - 1) // ALIGN * ALIGN)
            assert base <= off and off + size <= h.bytes and off % ALIGN == 0
            for o, s in live.items():
                assert off + size <= o or o + s <= off, "overlapping symmetric allocations"
            live[off] = size
    for off in list(live):
        h.free(off)
    assert h._free == [(base, 4 << 20)] and not h._live
