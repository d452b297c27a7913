"""Control path of the NVLink point-to-point route (mpi/p2p_device.py) without GPUs: descriptors,
the per-pair ack counters of the shared-memory segment, and reclamation of staging blocks. The
symmetric heap is replaced by a small in-process allocator; the payload copy itself needs two GPUs
and is covered by the opt-in multi-GPU check."""
import numpy as np
import pytest

from collective_communication_mpi_b200.mpi import p2p_device as P
from conftest import run_ranks


def test_descriptor_round_trip_and_payload_discrimination():
    d = P.make_descriptor(4096, 1000, 7)
    raw = np.zeros(1000, dtype=np.uint8)
    raw[:P.DESC_BYTES] = d.view(np.uint8)
    assert P.parse_descriptor(raw, 1000) == (4096, 1000, 7)
    assert P.parse_descriptor(raw, 999) is None                       # claims more bytes than the buffer holds
    assert P.parse_descriptor(np.arange(64, dtype=np.uint8), 64) is None   # an ordinary payload
    assert P.parse_descriptor(np.zeros(8, dtype=np.uint8), 8) is None      # shorter than a descriptor


class _Arena:
    def __init__(self, blocks):
        self.free_blocks, self.live = list(range(blocks)), set()

    def alloc(self, nbytes):
        if not self.free_blocks:
            raise MemoryError("arena full")
        b = self.free_blocks.pop(0)
        self.live.add(b)
        return b * 1024

    def free(self, off):
        self.live.remove(off // 1024)
        self.free_blocks.append(off // 1024)


def test_ledger_releases_blocks_in_send_order_per_destination():
    arena, acks = _Arena(4), {1: 0, 2: 0}
    led = P.StagingLedger(arena.alloc, arena.free, lambda d: acks[d], timeout_s=0.2)
    a = led.stage(1, 10); b = led.stage(2, 10); c = led.stage(1, 10)
    assert [a[1], b[1], c[1]] == [0, 0, 1] and led.outstanding == 3
    acks[1] = 1                                   # destination 1 is done with its first message only
    assert led.reclaim() == 1 and led.outstanding == 2 and a[0] // 1024 not in arena.live
    d = led.stage(2, 10); e = led.stage(2, 10)    # arena is now full (b, c, d, e)
    assert led.outstanding == 4
    with pytest.raises(MemoryError):
        led.stage(1, 10)                          # nothing acknowledged: gives up after the timeout
    acks[2] = 3
    f = led.stage(1, 10)                          # blocks of destination 2 come back, the send proceeds
    assert f[1] == 2 and led.outstanding == 2


CODE = r'''
import ctypes as C
import numpy as np
from mpi4py import MPI
from collective_communication_mpi_b200.mpi import p2p_device as P
w = MPI.COMM_WORLD
r, n = w.Get_rank(), w.Get_size()
world = w._world
lib, h = world._hostlib, world.host
nxt, prv = (r + 1) % n, (r - 1) % n
assert lib.ccmpi_host_ack_count(h, nxt) == 0
# three "device" messages to the next rank: only descriptors travel; the receiver acknowledges each
for k in range(3):
    w.Send(P.make_descriptor(4096 * k, 100 + k, k).view(np.uint8), dest=nxt, tag=5)
for k in range(3):
    raw = np.zeros(256, dtype=np.uint8)
    w.Recv(raw, source=prv, tag=5)
    assert P.parse_descriptor(raw, 256) == (4096 * k, 100 + k, k)
    assert lib.ccmpi_host_ack_post(h, prv) == 0
w.Barrier()
assert lib.ccmpi_host_ack_count(h, nxt) == 3, lib.ccmpi_host_ack_count(h, nxt)
assert lib.ccmpi_host_ack_count(h, prv) == (3 if n == 2 else 0)
assert lib.ccmpi_host_ack_post(h, r) < 0 and lib.ccmpi_host_ack_count(h, n) < 0     # self / out of range
w.Barrier()
if r == 0: print("ACK OK", n)
'''


@pytest.mark.parametrize("n", [2, 4])
def test_ack_counters_across_processes(n):
    p = run_ranks(n, ["-c", CODE], timeout=120)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert f"ACK OK {n}" in p.stdout
