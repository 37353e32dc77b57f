"""The count-free partition path (csrc/slabs.cu): fixed-capacity slabs + persistent
cursors.  Single-GPU groupBy against the oracle, the overflow -> counted rerun, and
the chunked exchange on ONE GPU with R local buffers standing in for the R ranks'
receive buffers (the kernels only see pointers).  Keys, counts exact; sums exact
(integer-valued data)."""
import numpy as np
import pytest
import torch

import pyspark_tf_gke_b200 as pkg
import pyspark_tf_gke_b200.dist as D
import pyspark_tf_gke_b200.synth as S
from oracle import etl_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    e = pkg.Engine(0)
    yield e
    e.close()


def _rows(n, groups, seed=21, row0=0, hot=False):
    keys, _ = S.host_column(seed, 0, S.KIND_KEY64, n, row0=row0, modulus=groups)
    vals, _ = S.host_column(seed, 1, S.KIND_EXACT, n, row0=row0)
    keys = keys - groups // 3
    if hot:
        keys[np.arange(n) % 5 < 3] = 7
    return keys, vals


def _same(res, want):
    assert res["keys"].cpu().numpy().tolist() == want["keys"].tolist()
    assert res["rows"].cpu().numpy().tolist() == want["rows"].tolist()
    assert res["cols"][0]["count"].cpu().numpy().tolist() == want["rows"].tolist()
    assert res["cols"][0]["sum"].cpu().numpy().tobytes() == want["cols"][0]["sum"].tobytes()


@pytest.mark.parametrize("n,groups", [(1, 1), (77, 5), (4097, 300), (300_001, 20_000), (2_000_003, 400_000), (3_000_000, 1_500_000)])
def test_slab_groupby_matches_oracle(eng, n, groups):
    keys, vals = _rows(n, groups)
    dk, dv = torch.from_numpy(keys).to(eng.device), torch.from_numpy(vals).to(eng.device)
    pg = eng.groupby_slab_sum(dk, dv, groups + 1024, groups_hint=groups)
    assert int(pg.counters[4].item()) == 0                      # uniform keys: nothing overflows at 6 sigma
    _same(eng.groupby_result(pg, sort=True), O.groupby_agg_fast(keys, [vals], [None]))
    # the same rows as {key, value} records (strides 2), as the receiver sees them
    rec = torch.stack([dk, dv.view(torch.int64)], dim=1).contiguous().view(-1)
    pg = eng.groupby_slab_sum(rec, rec[1:].view(torch.float64), groups + 1024, groups_hint=groups, n=n, key_stride=2, val_stride=2)
    _same(eng.groupby_result(pg, sort=True), O.groupby_agg_fast(keys, [vals], [None]))


def test_slab_overflow_repeats_on_the_counted_path(eng):
    # one key carries 60 % of the rows: its bucket slab (mean + 6 sigma of a uniform table) overflows, the
    # overflow is counted, and groupby_result transparently repeats the work with exact counts
    n, groups = 1_000_000, 200_000
    keys, vals = _rows(n, groups, hot=True)
    dk, dv = torch.from_numpy(keys).to(eng.device), torch.from_numpy(vals).to(eng.device)
    pg = eng.groupby_slab_sum(dk, dv, groups + 1024, groups_hint=groups)
    assert int(pg.counters[4].item()) > 0
    _same(eng.groupby_result(pg, sort=True), O.groupby_agg_fast(keys, [vals], [None]))
    # through the public entry
    pg = D.shuffle_groupby_sum(eng, dk, dv, groups + 1024, groups_hint=groups)
    _same(eng.groupby_result(pg, sort=True), O.groupby_agg_fast(keys, [vals], [None]))


def test_slab_path_is_the_default_and_counted_path_agrees(eng):
    n, groups = 1_500_000, 300_000
    keys, vals = _rows(n, groups)
    dk, dv = torch.from_numpy(keys).to(eng.device), torch.from_numpy(vals).to(eng.device)
    a = D.shuffle_groupby_sum(eng, dk, dv, groups + 1024, groups_hint=groups)
    b = D.shuffle_groupby_sum(eng, dk, dv, groups + 1024, groups_hint=groups, counted=True)
    assert a.counters.shape[0] == 5 and b.counters.shape[0] == 4
    ra, rb = eng.groupby_result(a, sort=True), eng.groupby_result(b, sort=True)
    for x, y in ((ra["keys"], rb["keys"]), (ra["rows"], rb["rows"]), (ra["cols"][0]["sum"], rb["cols"][0]["sum"])):
        assert torch.equal(x, y)
    _same(ra, O.groupby_agg_fast(keys, [vals], [None]))


def test_host_entry_copies_in_chunks_and_matches_oracle(eng):
    n, groups = 1_200_007, 100_000
    keys, vals = _rows(n, groups)
    hk, hv = torch.from_numpy(keys).pin_memory(), torch.from_numpy(vals).pin_memory()
    out = tuple(torch.empty(groups + 1024, dtype=t, pin_memory=True) for t in (torch.int64, torch.int64, torch.float64))
    g = D.shuffle_groupby_sum_host(eng, hk, hv, groups + 1024, groups_hint=groups, out=out, chunks=5)
    want = O.groupby_agg_fast(keys, [vals], [None])
    o = np.argsort(out[0][:g].numpy(), kind="stable")
    assert out[0][:g].numpy()[o].tolist() == want["keys"].tolist()
    assert out[1][:g].numpy()[o].tolist() == want["rows"].tolist()
    assert out[2][:g].numpy()[o].tobytes() == want["cols"][0]["sum"].tobytes()


@pytest.mark.parametrize("R,chunks", [(2, 1), (4, 3), (8, 4)])
def test_simulated_chunked_slab_exchange(eng, R, chunks):
    groups = 50_000
    sizes = [80_000 + 1013 * s + (s % 2) for s in range(R)]
    rows = [_rows(sizes[s], groups, row0=s * 1_000_000 // 32 * 32) for s in range(R)]
    d_rows = [(torch.from_numpy(k).to(eng.device), torch.from_numpy(v).to(eng.device)) for k, v in rows]
    slab = eng.slab_capacity(max(sizes) / R, 2.0)
    bufs = [torch.full((R * slab, 2), -1, dtype=torch.int64, device=eng.device) for _ in range(R)]        # per "rank"
    tables = [torch.zeros((chunks, R), dtype=torch.int64, device=eng.device) for _ in range(R)]
    overflow = torch.zeros(1, dtype=torch.int64, device=eng.device)
    # receivers' first level: `buckets` slabs per rank
    m = max(sizes) + max(sizes) // 4
    buckets, sub, rpk, cap1, cap2 = eng.slab_plan(m, groups // R + groups // (8 * R), max_buckets=16)
    level1 = [torch.full((buckets * cap1, 2), -1, dtype=torch.int64, device=eng.device) for _ in range(R)]
    cur1 = [torch.zeros(buckets, dtype=torch.int64, device=eng.device) for _ in range(R)]
    xcur = [torch.zeros(R, dtype=torch.int64, device=eng.device) for _ in range(R)]
    for c in range(chunks):
        for s in range(R):                 # every source sends its chunk c and publishes its cursors
            k, v = d_rows[s]
            step = -(-(-(-sizes[s] // chunks)) // 2048) * 2048
            lo, hi = min(sizes[s], c * step), min(sizes[s], (c + 1) * step)
            eng.slab_partition_rows(k[lo:hi], v[lo:hi], hi - lo, R, 1, [b.data_ptr() for b in bufs], s * slab, 0, slab, xcur[s],
                                    overflow, max_ctas=37)
            eng.peer_publish_counts(xcur[s], [t.data_ptr() for t in tables], c * R + s)
        for r in range(R):                 # every receiver partitions what arrived with chunk c
            eng.slab_partition_records(bufs[r], R, slab, tables[r][c], tables[r][c - 1] if c else None, slab, buckets, 0,
                                       level1[r], cap1, cur1[r], overflow, max_ctas=50)
    torch.cuda.synchronize()
    assert int(overflow.item()) == 0
    all_k = np.concatenate([k for k, _ in rows])
    all_v = np.concatenate([v for _, v in rows])
    dest = O.partition_dest(all_k, R)
    for r in range(R):
        # the count table: cumulative records per source after the last chunk = what the oracle sends there
        want_counts = [int((O.partition_dest(rows[s][0], R) == r).sum()) for s in range(R)]
        assert tables[r][chunks - 1].cpu().tolist() == want_counts
        got = bufs[r].cpu().numpy()
        for s in range(R):                 # source s's slab holds exactly source s's rows for rank r
            piece = got[s * slab:s * slab + want_counts[s]]
            sel = O.partition_dest(rows[s][0], R) == r
            a = np.stack([rows[s][0][sel], rows[s][1][sel].view(np.int64)], axis=1)
            assert piece[np.lexsort((piece[:, 1], piece[:, 0]))].tobytes() == a[np.lexsort((a[:, 1], a[:, 0]))].tobytes()
        # level 1: bucket b's slab holds exactly the records whose (rank, bucket) destination is (r, b) of R x buckets
        own = dest == r
        fine_dest = O.partition_dest(all_k[own], 1, buckets)
        l1, counts = level1[r].cpu().numpy(), cur1[r].cpu().tolist()
        assert counts == np.bincount(fine_dest, minlength=buckets).tolist()
        for b in range(buckets):
            piece = l1[b * cap1:b * cap1 + counts[b]]
            a = np.stack([all_k[own][fine_dest == b], all_v[own][fine_dest == b].view(np.int64)], axis=1)
            assert piece[np.lexsort((piece[:, 1], piece[:, 0]))].tobytes() == a[np.lexsort((a[:, 1], a[:, 0]))].tobytes()
        # second level + merge = the oracle's groupBy of the keys this rank owns
        counters = torch.zeros(5, dtype=torch.int64, device=eng.device)
        pg = eng.groupby_merge_slab_buckets(level1[r], cur1[r], cap1, buckets, sub, cap2, groups, counters)
        _same(eng.groupby_result(pg, sort=True), O.groupby_agg_fast(all_k[own], [all_v[own]], [None]))


@pytest.mark.parametrize("counted", [False, True])
def test_min_max_on_the_shuffle_path(eng, counted):
    # north-star item 3 names count / sum / mean / min / max: the shared-memory merge carries min and max of
    # the value column too (Spark's ordering: NaN is the largest double), on the slab and on the counted path
    n, groups = 700_003, 60_000
    keys, _ = _rows(n, groups)
    vals, _ = S.host_column(23, 1, S.KIND_REAL, n, nan_rate=0.002)
    vals = vals - 5e4
    dk, dv = torch.from_numpy(keys).to(eng.device), torch.from_numpy(vals).to(eng.device)
    pg = D.shuffle_groupby_sum(eng, dk, dv, groups + 1024, groups_hint=groups, minmax=True, counted=counted)
    res = eng.groupby_result(pg, sort=True)
    want = O.groupby_agg(keys, None, [vals], [None])
    assert res["keys"].cpu().numpy().tolist() == want["keys"].tolist()
    assert res["rows"].cpu().numpy().tolist() == want["rows"].tolist()
    col, wcol = res["cols"][0], want["cols"][0]
    assert np.array_equal(col["min"].cpu().numpy(), wcol["min"], equal_nan=True)
    assert np.array_equal(col["max"].cpu().numpy(), wcol["max"], equal_nan=True)
    np.testing.assert_allclose(col["sum"].cpu().numpy(), wcol["sum"], rtol=1e-6, atol=0, equal_nan=True)
    np.testing.assert_allclose(col["mean"].cpu().numpy(), wcol["mean"], rtol=1e-6, atol=0, equal_nan=True)


@pytest.mark.parametrize("R,groups_of_buckets,mover,waves", [(2, 4, "sm", 1), (4, 2, "ce", 2), (8, 8, "sm", 2), (8, 1, "ce", 1),
                                                              (2, 2, "sm", 3), (4, 4, "tma", 2), (8, 2, "tma", 1)])
def test_simulated_bucketed_exchange(eng, R, groups_of_buckets, mover, waves):
    # BucketExchange's data path on one GPU: every source partitions its rows, wave by wave, into R x Bx slabs of the
    # wave's send buffer, slabs are copied bucket group by bucket group into the "ranks'" receive buffers ([source][bucket]
    # layout, one set per wave), cursors go to the count tables, and each receiver cuts every group of every wave into
    # the same pieces and merges a group once its last wave is in (append mode)
    groups = 60_000
    sizes = [90_000 + 1013 * s + (s % 2) for s in range(R)]
    rows = [_rows(sizes[s], groups, row0=s * 1_000_000 // 32 * 32) for s in range(R)]
    bx = 8
    sub = 3
    W = waves
    wave_rows = -(-max(sizes) // (W * 4096)) * 4096
    slab = eng.slab_capacity(wave_rows / (R * bx), 2.0)
    cap2 = eng.slab_capacity(max(sizes) * 1.3 / (bx * sub), 2.0)
    recv = [[torch.full((R * bx * slab, 2), -1, dtype=torch.int64, device=eng.device) for _ in range(R)] for _ in range(W)]
    tables = [[torch.zeros(R * bx, dtype=torch.int64, device=eng.device) for _ in range(R)] for _ in range(W)]
    overflow = torch.zeros(1, dtype=torch.int64, device=eng.device)
    bpg = bx // groups_of_buckets
    sends, xcurs = [[] for _ in range(W)], [[] for _ in range(W)]
    for s in range(R):
        k, v = torch.from_numpy(rows[s][0]).to(eng.device), torch.from_numpy(rows[s][1]).to(eng.device)
        for w in range(W):
            send = torch.full((R * bx * slab, 2), -1, dtype=torch.int64, device=eng.device)
            xcur = torch.zeros(R * bx, dtype=torch.int64, device=eng.device)
            w_lo, w_hi = min(sizes[s], w * wave_rows), min(sizes[s], (w + 1) * wave_rows) if w < W - 1 else sizes[s]
            half = w_lo + (w_hi - w_lo) // 2 // 2 * 2               # two chunks per wave: the cursors persist
            for lo, hi in ((w_lo, half), (half, w_hi)):
                eng.slab_partition_rows(k[lo:hi], v[lo:hi], hi - lo, R, bx, [send.data_ptr()] * R, 0, slab, slab, xcur, overflow)
            eng.peer_publish_counts(xcur, [t.data_ptr() for t in tables[w]], s * bx, per_rank=bx)
            want = np.bincount(O.partition_dest(rows[s][0][w_lo:w_hi], R, bx), minlength=R * bx)
            assert xcur.cpu().tolist() == want.tolist()
            sends[w].append(send)
            xcurs[w].append(xcur)
    all_k = np.concatenate([k for k, _ in rows])
    all_v = np.concatenate([v for _, v in rows])
    for r in range(R):
        pg = eng.new_partitioned_groups(groups, 1 << 16, bx * sub)
        cur2 = torch.zeros(bx * sub, dtype=torch.int64, device=eng.device)
        pieces = torch.full((bx * sub * cap2, 2), -1, dtype=torch.int64, device=eng.device)
        for w in range(W):
            for g in range(groups_of_buckets):
                for s in range(R):
                    if mover == "ce":      # whole slabs through a copy engine
                        eng.copy_async(recv[w][r].data_ptr() + ((s * bx + g * bpg) * slab) * 16,
                                       sends[w][s].data_ptr() + ((r * bx + g * bpg) * slab) * 16, bpg * slab * 16)
                    elif r == 0:           # the store kernel serves every destination at once: the filled parts only
                        eng.peer_copy_slabs(sends[w][s], [t.data_ptr() for t in recv[w]], s, bx, g * bpg, bpg, slab, xcurs[w][s],
                                            ctas_per_slab=3 if s % 2 else 0, bulk=mover == "tma")
                eng.slab_partition_records(recv[w][r], bpg * R, slab, tables[w][r], None, slab, bx, sub, pieces, cap2, cur2,
                                           pg.counters[4:], sources=R, bucket0=g * bpg)
                if w == W - 1:
                    eng.merge_slab_pieces(pieces, cur2, cap2, g * bpg * sub, bpg * sub, g > 0, pg)
        torch.cuda.synchronize()
        assert int(pg.counters[4].item()) == 0 and int(overflow.item()) == 0
        # count tables of rank r: [source][bucket] = what the oracle sends there, summed over the waves
        want_counts = [int((O.partition_dest(rows[s][0], R, bx) == r * bx + b).sum()) for s in range(R) for b in range(bx)]
        assert sum(tables[w][r] for w in range(W)).cpu().tolist() == want_counts
        # piece f of rank r holds exactly the owned records whose fine index is f
        own = O.partition_dest(all_k, R) == r
        h = O.mix64(all_k[own].view(np.uint64))
        fine_idx = (((h & np.uint64(0xFFFFFFFF)) * np.uint64(bx * sub)) >> np.uint64(32)).astype(np.int64)
        assert cur2.cpu().tolist() == np.bincount(fine_idx, minlength=bx * sub).tolist()
        _same(eng.groupby_result(pg, sort=True), O.groupby_agg_fast(all_k[own], [all_v[own]], [None]))
