"""Index arithmetic of the token-axis fused policies (csrc/fused_policies_m.cuh), restated in Python
and checked for the properties the protocols rely on: every tile is computed exactly once, every flag
word has exactly one meaning, every slot byte is written exactly once per source, every pushed piece
goes out exactly once and every ready flag collects one contribution per CTA. The kernels themselves have
not run on hardware yet; this pins the arithmetic they were written from (a transcription error in either
copy shows up as a broken invariant here or as a wrong result on the GPU)."""
import itertools

import pytest

BM, EPI_THREADS, GROUP_M, GRANULE = 128, 128, 8, 16


def raster(tile, m_blocks, n_blocks):
    """csrc/gemm_tcgen05.cuh::raster"""
    per_group = GROUP_M * n_blocks
    g = tile // per_group
    first_m = g * GROUP_M
    gsize = min(m_blocks - first_m, GROUP_M)
    in_g = tile - g * per_group
    return first_m + in_g % gsize, in_g // gsize


def pair_tiles(policy_tile, rank, n, m_pairs, n_blocks):
    """What gemm_kernel_2cta's my_tile() hands the two CTAs of every pair, over the whole grid."""
    out = []
    for t in range(m_pairs * n_blocks):
        pm, nb = policy_tile(t, rank, n, m_pairs, n_blocks)
        out += [(2 * pm + cta_rank, nb) for cta_rank in (0, 1)]
    return out


# ---- AgMPolicy ---------------------------------------------------------------------------------
def agm_tile(t, rank, n, mb, nb):
    per_owner = (mb // n) * nb
    o = t // per_owner
    m, nn = raster(t - o * per_owner, mb // n, nb)
    return m + ((rank + o) % n) * (mb // n), nn


def agm_ready_idx(flag_base, src, lm, mbpr):
    return flag_base + src * mbpr + lm


@pytest.mark.parametrize("n,rows,N,BN", [(2, 256, 256, 256), (2, 512, 384, 128), (4, 256, 1024, 256), (8, 1024, 512, 256), (4, 2304, 640, 128)])
def test_agm_tiles_and_flags(n, rows, N, BN):
    M = rows * n
    m_blocks, nb = M // BM, -(-N // BN)
    m_pairs, mbpr = m_blocks // 2, rows // BM
    assert rows % 256 == 0
    for rank in range(n):
        tiles = pair_tiles(agm_tile, rank, n, m_pairs, nb)
        assert sorted(tiles) == sorted(itertools.product(range(m_blocks), range(nb))), "every tile exactly once"
        # both CTAs of a pair read rows of the same owner, and the first 1/n of the schedule is local rows
        owners = [m // mbpr for m, _ in tiles]
        assert all(owners[i] == owners[i + 1] for i in range(0, len(owners), 2))
        per_owner = len(tiles) // n
        for o in range(n):                      # own rows first, then rank+1, rank+2, ...
            assert set(owners[o * per_owner:(o + 1) * per_owner]) == {(rank + o) % n}
    flags = {agm_ready_idx(7, src, lm, mbpr) for src in range(n) for lm in range(mbpr)}
    assert len(flags) == n * mbpr == M // BM and min(flags) == 7 and max(flags) == 7 + M // BM - 1


def agm_push_schedule(rows, K, nctas):
    """AgMPolicy::background / tail: (piece index, CTA) of every bulk copy and the per-CTA flag log."""
    row_bytes = K * 2
    PB = 256 if row_bytes % 256 == 0 else 128
    ppr = row_bytes // PB
    total, per_blk, stride = rows * ppr, BM * ppr, nctas * EPI_THREADS
    m_blocks = rows // BM
    pieces, contributions = [], [0] * m_blocks
    for cta in range(nctas):
        base, flagged = cta * EPI_THREADS, 0
        while flagged < m_blocks:
            for tid in range(EPI_THREADS):
                if base + tid < total:
                    pieces.append(base + tid)
            nxt = base + stride
            done = min(m_blocks, nxt // per_blk)
            granule = GRANULE if m_blocks >= 32 else 4
            done = (done // granule) * granule
            if nxt >= total:
                done = m_blocks
            base = nxt
            for j in range(flagged, max(done, flagged)):
                contributions[j] += 1
            flagged = max(done, flagged)
    return pieces, contributions, total


@pytest.mark.parametrize("rows,K,nctas", [(256, 192, 4), (256, 512, 74), (1024, 4096, 148), (4096, 1024, 36), (2304, 320, 10)])
def test_agm_push_covers_every_piece_once_and_flags_count_ctas(rows, K, nctas):
    pieces, contributions, total = agm_push_schedule(rows, K, nctas)
    assert sorted(pieces) == list(range(total)), "every 128/256-byte piece is pushed exactly once"
    assert contributions == [nctas] * (rows // BM), "each ready flag gets one contribution per CTA and launch"


def test_agm_flag_is_complete_only_after_every_piece_of_the_block():
    """A consumer may read m-block j once its flag holds nctas contributions: by then every CTA has issued
    (and waited for) all of its pieces below the end of that block."""
    rows, K, nctas = 4096, 512, 12
    row_bytes = K * 2
    ppr = row_bytes // 256
    per_blk, stride, total = BM * ppr, nctas * EPI_THREADS, rows * ppr
    for cta in range(nctas):
        base, flagged = cta * EPI_THREADS, 0
        issued_below = 0          # this CTA has issued every piece of its share with index < issued_below
        while flagged < rows // BM:
            nxt = base + stride
            issued_below = nxt
            done = min(rows // BM, nxt // per_blk)
            done = (done // (GRANULE if rows // BM >= 32 else 4)) * (GRANULE if rows // BM >= 32 else 4)
            if nxt >= total:
                done = rows // BM
            if done > flagged:
                assert done * per_blk <= max(issued_below, total)
                assert done * per_blk <= issued_below or nxt >= total
                flagged = done
            base = nxt


# ---- RsMPolicy ---------------------------------------------------------------------------------
def rsm_tile(t, rank, n, mb, nb):
    g, in_g = t % n, t // n
    owner = (rank + 1 + g) % n
    m, nn = raster(in_g, mb // n, nb)
    return m + owner * (mb // n), nn


def rsm_flag_idx(flag_base, lm, n_blk, src, nblocks, n):
    return flag_base + ((lm * nblocks + n_blk) * n + src)


@pytest.mark.parametrize("n,rows,N,BN", [(2, 256, 256, 256), (2, 256, 384, 128), (4, 512, 512, 256), (8, 256, 1024, 256), (4, 768, 640, 128)])
def test_rsm_tiles_flags_slots_and_reduction_items(n, rows, N, BN):
    M = rows * n
    m_blocks, nb = M // BM, N // BN
    m_pairs, mbpr = m_blocks // 2, rows // BM
    written = {}                      # (owner, src, lrow block, n_blk) -> count
    flag_meaning = {}
    for rank in range(n):
        tiles = pair_tiles(rsm_tile, rank, n, m_pairs, nb)
        assert sorted(tiles) == sorted(itertools.product(range(m_blocks), range(nb)))
        order = [m // mbpr for m, _ in tiles[::2]]
        assert order[:n] == [(rank + 1 + g) % n for g in range(n)], "owners interleaved, own tiles last of each round"
        for m, nn in tiles:
            owner, lm = m // mbpr, m % mbpr
            key = (owner, rank, lm, nn)
            written[key] = written.get(key, 0) + 1
            idx = rsm_flag_idx(4096, lm, nn, rank, nb, n)
            assert flag_meaning.setdefault((owner, idx), (lm, nn, rank)) == (lm, nn, rank), "one meaning per flag word"
            assert 4096 <= idx < 4096 + mbpr * nb * n
    assert all(c == 1 for c in written.values())
    assert len(written) == n * n * mbpr * nb, "every owner gets every source's partial of every owned tile"
    # the reducers of one rank: items handed out CTA by CTA cover the owned tiles exactly once
    for nctas in (2, 4, 36):
        items = []
        for cta in range(nctas):
            nxt = cta
            while nxt < mbpr * nb:
                items.append(raster(nxt, mbpr, nb))
                nxt += nctas
        assert sorted(items) == sorted(itertools.product(range(mbpr), range(nb)))


def test_rsm_slot_addresses_do_not_overlap():
    """bulk_row's destination: slot[src][lrow][col] of the owner; distinct (src, row, column chunk) -> distinct
    byte ranges inside [n][Mr][N]."""
    n, rows, N, es, CW = 4, 256, 384, 2, 128
    seen = set()
    for src in range(n):
        for row in range(0, rows * n, 37):               # a sample of global rows
            owner, lrow = row // rows, row % rows
            for col0 in range(0, N, CW):
                start = ((src * rows + lrow) * N + col0) * es
                rng = (owner, start, start + min(CW, N - col0) * es)
                assert rng not in seen
                seen.add(rng)
                assert rng[2] <= n * rows * N * es
