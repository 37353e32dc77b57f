"""The multi-GPU exchange kernels on ONE GPU: R local buffers stand in for the R
ranks' receive buffers (the kernels only see pointers), every simulated source
rank runs its count pass and its fused partition + scatter in turn, and what has
landed in each "rank" is compared with the oracle: the multiset of records per
(rank, bucket), the layout arithmetic of dist.peer_layout, and the groupBy of
what each rank received.  Records are compared bit for bit; group sums are exact
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


def _source_rows(src, n, groups, hot=False):
    keys, _ = S.host_column(11, 0, S.KIND_KEY64, n, row0=src * 1_000_000 // 32 * 32, modulus=groups)
    vals, _ = S.host_column(11, 1, S.KIND_EXACT, n, row0=src * 1_000_000 // 32 * 32)
    keys = keys - groups // 3          # negative keys too
    if hot:                            # skew: 60 % of the rows carry one key
        keys[np.arange(n) % 5 < 3] = 7
    return keys, vals


def _sorted_records(keys, vals):
    rec = np.stack([keys.astype(np.int64), vals.view(np.int64)], axis=1)
    return rec[np.lexsort((rec[:, 1], rec[:, 0]))]


def _run_exchange(eng, R, B, sizes, groups, hot=False, tight=False):
    """-> per-rank receive buffers, tot[r, b], bucket_start[r, b], host rows per source."""
    rows = [_source_rows(s, sizes[s], groups, hot) for s in range(R)]
    d_rows = [(torch.from_numpy(k).to(eng.device), torch.from_numpy(v).to(eng.device)) for k, v in rows]
    matrix = torch.stack([eng.hash_partition_count(k, R, B) for k, _ in d_rows])          # [source, rank * B + bucket]
    want_counts = np.stack([np.bincount(O.partition_dest(k, R, B), minlength=R * B) for k, _ in rows])
    assert matrix.cpu().numpy().tolist() == want_counts.tolist()
    layouts = [D.peer_layout(matrix, s, R, B) for s in range(R)]
    tot = layouts[0][1]
    incoming = tot.sum(dim=1).cpu().tolist()
    # tight: every buffer holds exactly what arrives (a store past the end would land in the guard words)
    cap = [m if tight else m + 1000 for m in incoming]
    guard = 0x5A5A5A5A5A5A5A5A
    bufs = [torch.full((c + 8, 2), guard, dtype=torch.int64, device=eng.device) for c in cap]
    for s in range(R):
        offsets, tot_s, bstart_s = layouts[s]
        assert torch.equal(tot_s, tot) and torch.equal(bstart_s, layouts[0][2])
        k, v = d_rows[s]
        eng.hash_partition_scatter_peers(k, v, [b.data_ptr() for b in bufs], offsets, buckets=B)
    torch.cuda.synchronize()
    for r in range(R):
        assert bool((bufs[r][incoming[r]:] == guard).all()), "a store landed past the end of a receive buffer"
    return bufs, tot.cpu().numpy(), layouts[0][2].cpu().numpy(), rows, matrix.cpu().numpy()


@pytest.mark.parametrize("R", [2, 4, 8])
@pytest.mark.parametrize("B", [1, 32])
def test_simulated_peer_exchange_matches_oracle(eng, R, B):
    groups = 20_000
    sizes = [60_000 + 1013 * s + (s % 2) for s in range(R)]          # unaligned, different per source
    bufs, tot, bstart, rows, matrix = _run_exchange(eng, R, B, sizes, groups, tight=True)
    all_k = np.concatenate([k for k, _ in rows])
    all_v = np.concatenate([v for _, v in rows])
    dest = O.partition_dest(all_k, R, B)
    for r in range(R):
        got = bufs[r].cpu().numpy()
        for b in range(B):
            lo, cnt = int(bstart[r, b]), int(tot[r, b])
            sel = dest == r * B + b
            assert cnt == int(sel.sum())
            part = got[lo:lo + cnt]
            assert _sorted_records(part[:, 0], part[:, 1].view(np.float64)).tobytes() == \
                _sorted_records(all_k[sel], all_v[sel]).tobytes()
            # inside a bucket the sources follow one another: source s's run holds exactly source s's rows
            run = lo
            for s in range(R):
                c = int(matrix[s, r * B + b])
                ks, vs = rows[s]
                ssel = O.partition_dest(ks, R, B) == r * B + b
                piece = got[run:run + c]
                assert _sorted_records(piece[:, 0], piece[:, 1].view(np.float64)).tobytes() == \
                    _sorted_records(ks[ssel], vs[ssel]).tobytes()
                run += c
        # the receiver's final merge of what arrived = the oracle's groupBy of the keys this rank owns
        m = int(tot[r].sum())
        own = O.partition_dest(all_k, R) == r
        flat = bufs[r].view(-1)
        pg = eng.groupby_partitioned_sum(flat, flat[1:].view(torch.float64), 1 << 15, groups_hint=groups // R, n=m,
                                         key_stride=2, val_stride=2)
        res = eng.groupby_result(pg, sort=True)
        want = O.groupby_agg_fast(all_k[own], [all_v[own]], [None])
        assert res["keys"].cpu().numpy().tolist() == want["keys"].tolist()
        assert res["rows"].cpu().numpy().tolist() == want["rows"].tolist()
        assert res["cols"][0]["sum"].cpu().numpy().tobytes() == want["cols"][0]["sum"].tobytes()


def test_simulated_peer_exchange_empty_destinations_and_hot_key(eng):
    # 3 distinct keys over 8 ranks: most destinations receive nothing; then one key with 60 % of the rows
    for groups, hot in ((3, False), (5000, True)):
        R, B = 8, 1
        sizes = [30_001, 0, 17, 40_000, 1, 2048, 2049, 33]
        bufs, tot, bstart, rows, _ = _run_exchange(eng, R, B, sizes, groups, hot=hot, tight=True)
        all_k = np.concatenate([k for k, _ in rows])
        all_v = np.concatenate([v for _, v in rows])
        dest = O.partition_dest(all_k, R)
        assert int(tot.sum()) == sum(sizes)
        for r in range(R):
            sel = dest == r
            m = int(tot[r].sum())
            assert m == int(sel.sum())
            got = bufs[r].cpu().numpy()[:m]
            assert _sorted_records(got[:, 0], got[:, 1].view(np.float64)).tobytes() == \
                _sorted_records(all_k[sel], all_v[sel]).tobytes()
        if not hot:
            assert int((tot.sum(axis=1) == 0).sum()) >= 5
