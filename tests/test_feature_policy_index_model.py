"""Index arithmetic of the feature-axis fused policies (csrc/fused_policies.cuh: AgKPolicy, RsNPolicy) -- the
ones that run by default and were verified on B200s -- restated in Python as a regression pin: K-chunk visiting
order, flag-word uniqueness, push coverage, owner interleaving and slot coverage."""
import itertools

import pytest

from test_seq_policy_index_model import BM, EPI_THREADS, GRANULE, pair_tiles, raster

BK = 64


# ---- AgKPolicy ---------------------------------------------------------------------------------
def agk_kblock(it, rank, n, chunk):
    """(source rank, k0) of K-block `it`: the local chunk first, then rank+1, rank+2, ..."""
    per = chunk // BK
    ci, within = divmod(it, per)
    src = (rank + ci) % n
    return src, src * chunk + within * BK


@pytest.mark.parametrize("n,K", [(2, 512), (4, 1024), (8, 8192)])
def test_agk_walks_every_k_block_once_local_chunk_first(n, K):
    chunk = K // n
    for rank in range(n):
        visits = [agk_kblock(it, rank, n, chunk) for it in range(K // BK)]
        assert sorted(k0 for _, k0 in visits) == list(range(0, K, BK))
        assert [s for s, _ in visits[:chunk // BK]] == [rank] * (chunk // BK)
        order = [s for s, _ in visits[::chunk // BK]]
        assert order == [(rank + i) % n for i in range(n)]


@pytest.mark.parametrize("M,chunk,nctas", [(256, 256, 4), (8192, 2048, 148), (200, 128, 10), (1024, 1024, 74)])
def test_agk_bulk_push_covers_the_shard_once_and_flags_count_ctas(M, chunk, nctas):
    row_bytes = chunk * 2
    PB = 256 if row_bytes % 256 == 0 else 128
    ppr = row_bytes // PB
    total, per_blk, stride = M * ppr, BM * ppr, nctas * EPI_THREADS
    m_blocks = -(-M // BM)
    pieces, contributions = [], [0] * m_blocks
    for cta in range(nctas):
        base, flagged = cta * EPI_THREADS, 0
        while flagged < m_blocks:
            pieces += [base + t for t in range(EPI_THREADS) if base + t < total]
            nxt = base + stride
            done = (min(m_blocks, nxt // per_blk) // GRANULE) * GRANULE
            if nxt >= total:
                done = m_blocks
            base = nxt
            for j in range(flagged, max(done, flagged)):
                contributions[j] += 1
            flagged = max(done, flagged)
    assert sorted(pieces) == list(range(total))
    assert contributions == [nctas] * m_blocks
    flags = {src * m_blocks + m for src in range(8) for m in range(m_blocks)}       # ready_idx
    assert len(flags) == 8 * m_blocks


# ---- RsNPolicy ---------------------------------------------------------------------------------
def rsn_tile(t, rank, n, mb, nb_per_owner):
    g, in_g = t % n, t // n
    owner = (rank + 1 + g) % n
    m, nl = raster(in_g, mb, nb_per_owner)
    return m, nl + owner * nb_per_owner


@pytest.mark.parametrize("n,M,chunk,BN", [(2, 256, 256, 256), (2, 8192, 2048, 256), (4, 384, 128, 128), (8, 8192, 1024, 256)])
def test_rsn_tiles_flags_and_slots(n, M, chunk, BN):
    m_blocks = -(-M // BM)
    m_pairs = (m_blocks + 1) // 2
    per = chunk // BN
    seen_flags, written = {}, {}
    for rank in range(n):
        tiles = pair_tiles(lambda t, r, nn, mb, nb: rsn_tile(t, r, nn, mb, per), rank, n, m_pairs, per * n)
        real = [(m, nb) for m, nb in tiles if m < m_blocks]                  # phantom m-block of an odd tail
        assert sorted(real) == sorted(itertools.product(range(m_blocks), range(per * n)))
        owners = [nb // per for _, nb in tiles[::2]]
        assert owners[:n] == [(rank + 1 + g) % n for g in range(n)], "peers' tiles first, own tile last in each round"
        for m, nb in real:
            owner, nl = nb // per, nb % per
            idx = (m * per + nl) * n + rank                                  # flag_idx, in the owner's pad
            assert seen_flags.setdefault((owner, idx), (m, nl, rank)) == (m, nl, rank)
            assert idx < m_blocks * per * n
            written[(owner, rank, m, nl)] = written.get((owner, rank, m, nl), 0) + 1
    assert set(written.values()) == {1} and len(written) == n * n * m_blocks * per


# ---- RsNPolicy, round 2: peers' tiles first, own tiles last and reduced in their own epilogue -----
def rsn2_tile(t, rank, n, mb, per):
    """csrc/fused_policies.cuh: RsNPolicy::tile"""
    remote = (n - 1) * mb * per
    if t < remote:
        g, in_g = t % (n - 1), t // (n - 1)
        owner = (rank + 1 + g) % n
    else:
        owner, in_g = rank, t - remote
    m, nl = raster(in_g, mb, per)
    return m, nl + owner * per


@pytest.mark.parametrize("n,M,chunk,BN", [(2, 256, 256, 256), (2, 8192, 2048, 256), (4, 384, 128, 128), (8, 8192, 1024, 256),
                                          (4, 8192, 512, 256)])
def test_rsn_v2_own_tiles_come_last_and_never_wait_on_a_waiting_tile(n, M, chunk, BN):
    m_blocks = -(-M // BM)
    m_pairs = (m_blocks + 1) // 2
    per = chunk // BN
    remote = (n - 1) * m_pairs * per
    when = {}                                            # (rank, pair-row, n-block) -> position in the rank's schedule
    for rank in range(n):
        sched = [rsn2_tile(t, rank, n, m_pairs, per) for t in range(m_pairs * per * n)]
        assert sorted(sched) == sorted(itertools.product(range(m_pairs), range(per * n))), "every tile exactly once"
        owners = [nb // per for _, nb in sched]
        assert all(o != rank for o in owners[:remote]) and all(o == rank for o in owners[remote:])
        # peers are served round-robin, the next rank first: NVLink traffic is spread over the whole remote phase
        assert owners[:n - 1] == [(rank + 1 + g) % n for g in range(n - 1)]
        for pos, tile in enumerate(sched):
            when[(rank, *tile)] = pos
    # An own tile waits for the partials the peers push from THEIR remote phase, and remote tiles wait for
    # nothing: no cycle of waits exists whatever the CTAs' relative progress. In tile units the partial is
    # also produced no later than it is consumed (mirrored order), so own epilogues rarely wait at all.
    for rank in range(n):
        for m in range(m_pairs):
            for nl in range(per):
                own_pos = when[(rank, m, rank * per + nl)]
                assert own_pos >= remote
                for peer in range(n):
                    if peer != rank:
                        assert when[(peer, m, rank * per + nl)] < remote
                        assert when[(peer, m, rank * per + nl)] <= own_pos


@pytest.mark.parametrize("n", [2, 3, 4, 8])
def test_rsn_v2_own_epilogue_adds_the_peers_slots_in_rank_order(n):
    """RsNPolicy::own_load: pseudo-source k -> slot of rank (k < rank ? k : k + 1), then the residual."""
    for rank in range(n):
        srcs = [k if k < rank else k + 1 for k in range(n - 1)]
        assert srcs == [s for s in range(n) if s != rank]            # rank order, own partial excluded
    # the sum a rank forms: slots below it, its own accumulator, slots above it (np.sum(axis=0) order of
    # func_impl.py:186 is 0..n-1; the own partial is rounded to the slot type first, so the result is the
    # same bits as summing n slots)


# ---- AgKPolicy, round 2: the producer asks once per (tile, K-chunk) ---------------------------------
def agk_kchunk(ci, rank, n, chunk):
    """csrc/fused_policies.cuh: AgKPolicy::kchunk -> (source rank, first K of the chunk, reads the local shard?)"""
    src = rank + ci
    if src >= n:
        src -= n
    return src, src * chunk, src == rank


@pytest.mark.parametrize("n,K", [(2, 512), (4, 1024), (8, 8192)])
def test_agk_kchunk_form_visits_the_same_blocks_as_the_per_block_form(n, K):
    chunk = K // n
    per = chunk // BK
    for rank in range(n):
        by_chunk = []
        for ci in range(n):
            src, k_base, local = agk_kchunk(ci, rank, n, chunk)
            assert local == (ci == 0), "only the first chunk is this rank's own shard: no flag wait before it"
            by_chunk += [(src, k_base + j * BK) for j in range(per)]
        assert by_chunk == [agk_kblock(it, rank, n, chunk) for it in range(K // BK)]
