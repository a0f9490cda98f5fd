"""GPU tests of the tensor-core path (K2 candidates + K5 exact rescore) through the C ABI.
K2 must return exactly what the streaming kernel (K1) returns -- same rows, same score bits --
because every returned score comes from the canonical exact scorer; and both must match the
oracle within the stated tolerance."""
import numpy as np
import pytest

import oracle
from githubrepostorag_b200 import synthetic as S
from tests.test_gpu_parity import _check, _mode_name, _shard

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    assert torch.cuda.is_available()
    return torch


def _both(sh, torch, q, k, mode, filt=None):
    qd = torch.from_numpy(q).cuda()
    r1 = [t.clone() for t in sh.search_device(qd, k, metric=_mode_name(mode), filt=filt, algo=1)]
    r2 = [t.clone() for t in sh.search_device(qd, k, metric=_mode_name(mode), filt=filt, algo=2)]
    torch.cuda.synchronize()
    return r1, r2


@pytest.mark.parametrize("n,d", [(100_000, 768), (70_001, 384), (40_000, 1024), (33_000, 264)])
def test_k2_equals_k1_bitwise_and_matches_oracle(torch_cuda, n, d):
    torch = torch_cuda
    rows16 = S.corpus(300 + d, n, d)
    sh = _shard(torch, rows16)
    for mode in (oracle.MODE_COSINE, oracle.MODE_DOT):
        for batch, k in ((1, 10), (5, 10), (130, 20), (300, 5), (64, 100), (3, 128)):
            q = S.queries(1000 + batch + k, batch, d)
            (s1, r1, v1), (s2, r2, v2) = _both(sh, torch, q, k, mode)
            assert torch.equal(r1, r2), f"rows differ n={n} d={d} B={batch} k={k} {mode}"
            assert torch.equal(v1, v2)
            assert torch.equal(s1.view(torch.int32), s2.view(torch.int32))
            nq = min(batch, 3)
            _check(q[:nq], rows16, s2.cpu().numpy()[:nq], r2.cpu().numpy()[:nq], v2.cpu().numpy()[:nq], k, mode,
                   what=f"K2 n={n} d={d} B={batch} k={k} {mode}")
    sh.close()


def test_k2_with_filters_and_mixed_batch(torch_cuda):
    torch = torch_cuda
    n, d, k = 120_000, 384, 10
    seed = S.BASE_SEEDS["C4"]
    rows16 = S.corpus(seed, n, d)
    codes = S.metadata_codes(seed, n)
    code_mat = np.stack([codes[kk] for kk in S.FILTER_KEYS]).astype(np.int32)
    sh = _shard(torch, rows16, code_mat)
    batch = 200
    q = S.queries(seed, batch, d)
    filt = np.full((batch, 4), -1, np.int32)
    filt[:120] = [0, -1, -1, -1]          # 120 queries share namespace=default
    filt[120:150] = [1, -1, 1, -1]        # 30 share namespace=acme AND branch=dev
    filt[150:153] = [-1, 7, -1, -1]       # 3 with a selective repo filter (streaming kernel)
    filt[153] = [-1, -2, -1, -1]          # impossible
    (s1, r1, v1), (s2, r2, v2) = _both(sh, torch, q, k, oracle.MODE_COSINE, filt=filt)
    assert torch.equal(r1, r2) and torch.equal(v1, v2) and torch.equal(s1.view(torch.int32), s2.view(torch.int32))
    keep = np.ones((batch, n), dtype=bool)
    for b in range(batch):
        for m in range(4):
            if filt[b, m] == -2:
                keep[b] = False
            elif filt[b, m] >= 0:
                keep[b] &= code_mat[m] == filt[b, m]
    pick = [0, 119, 120, 149, 150, 153, 154, 199]
    _check(q[pick], rows16, s2.cpu().numpy()[pick], r2.cpu().numpy()[pick], v2.cpu().numpy()[pick], k,
           oracle.MODE_COSINE, keep=keep[pick], what="K2 filtered")
    # default dispatch (algo 0) gives the same answer
    s0, r0, v0 = sh.search_host(q, k, filt=filt)
    assert np.array_equal(r0, r2.cpu().numpy()) and np.array_equal(s0, s2.cpu().numpy())
    sh.close()


def test_k2_unprovable_queries_fall_back_to_streaming(torch_cuda):
    """Ties at the k-th boundary cannot be proven by the margin test: the library must notice and
    re-run those queries exactly."""
    torch = torch_cuda
    n, d, k = 50_000, 128, 10
    rng = np.random.default_rng(17)
    rows16 = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16)
    rows16[1000:30_000] = rows16[1000]      # 29k bit-identical rows
    q = np.stack([rows16[1000].astype(np.float32), rng.standard_normal(d, dtype=np.float32)] * 3)
    sh = _shard(torch, rows16)
    (s1, r1, v1), (s2, r2, v2) = _both(sh, torch, q, k, oracle.MODE_COSINE)
    assert torch.equal(r1, r2) and torch.equal(s1.view(torch.int32), s2.view(torch.int32))
    assert r2[0].tolist() == list(range(29_999, 29_989, -1))
    sh.close()


def test_k2_device_side_fallback_many_queries_filters_and_large_k(torch_cuda):
    """A corpus of duplicated rows leaves MOST queries unprovable; the device-side fallback must serve
    one, a few and hundreds of flagged queries, under a filter (non-identity query map, survivor list)
    and at both k widths -- bit-identical to the streaming kernel every time."""
    torch = torch_cuda
    n, d = 60_000, 256
    rng = np.random.default_rng(23)
    base = rng.standard_normal((200, d), dtype=np.float32).astype(np.float16)
    rows16 = base[rng.integers(0, 200, size=n)]            # every row has ~300 bit-identical twins
    rows16[:5000] = rng.standard_normal((5000, d), dtype=np.float32).astype(np.float16)
    code_mat = np.stack([np.arange(n, dtype=np.int32) % 3] + [np.zeros(n, np.int32)] * 3)
    sh = _shard(torch, rows16, code_mat)
    for batch, k in ((1, 10), (7, 10), (300, 10), (40, 100)):
        q = base[rng.integers(0, 200, size=batch)].astype(np.float32) + 0.01 * rng.standard_normal((batch, d), dtype=np.float32)
        filt = np.full((batch, 4), -1, np.int32)
        filt[batch // 3:, 0] = 1                           # the later two thirds share one filter
        for f in (None, filt):
            (s1, r1, v1), (s2, r2, v2) = _both(sh, torch, q, k, oracle.MODE_COSINE, filt=f)
            assert torch.equal(r1, r2), f"rows differ B={batch} k={k} filtered={f is not None}"
            assert torch.equal(v1, v2) and torch.equal(s1.view(torch.int32), s2.view(torch.int32))
    # keys output of the fallback feeds the cross-shard merge
    q = base[:9].astype(np.float32)
    qd = torch.from_numpy(q).cuda()
    keys = torch.empty((9, 10), dtype=torch.int64, device="cuda")
    s2, r2, v2 = [t.clone() for t in sh.search_device(qd, 10, keys_out=keys, algo=2)]
    keys1 = torch.empty_like(keys)
    sh.search_device(qd, 10, keys_out=keys1, algo=1)
    torch.cuda.synchronize()
    assert torch.equal(keys, keys1)
    sh.close()


@pytest.mark.parametrize("d", [384, 768])
def test_default_plan_for_tiny_batches_equals_streaming_kernel(torch_cuda, d):
    """Unfiltered batches of 2-4 queries take the tensor-core kernel by default on shards >= 128k rows
    (the streaming kernel pays almost one pass per query); the answer must not depend on the plan."""
    torch = torch_cuda
    n, k = 140_000, 10
    rows16 = S.corpus(500 + d, n, d)
    sh = _shard(torch, rows16)
    launches = __import__("githubrepostorag_b200._native", fromlist=["launch_count"]).launch_count
    for batch in (1, 2, 3, 4, 9):
        q = S.queries(2000 + batch, batch, d)
        qd = torch.from_numpy(q).cuda()
        l0 = launches()
        s0, r0, v0 = [t.clone() for t in sh.search_device(qd, k)]
        n_default = launches() - l0
        s1, r1, v1 = [t.clone() for t in sh.search_device(qd, k, algo=1)]
        torch.cuda.synchronize()
        assert torch.equal(r0, r1) and torch.equal(v0, v1) and torch.equal(s0.view(torch.int32), s1.view(torch.int32))
        # one streaming launch for a single query; prep + K2 + K5 + fallback for two or more
        assert n_default == (1 if batch == 1 else 4), f"B={batch}: {n_default} launches"
    sh.close()


def test_unfiltered_search_can_be_captured_in_a_cuda_graph(torch_cuda):
    """Nothing on the unfiltered path waits for the host any more (the tensor-core path re-runs unprovable
    queries on the device), so a caller can capture crb_search in a CUDA graph and replay it on new
    queries written into the captured buffer -- streaming (B=1) and tensor-core (B=8) plans alike."""
    torch = torch_cuda
    n, d, k = 150_000, 256, 10
    rows16 = S.corpus(901, n, d)
    sh = _shard(torch, rows16)
    for b in (1, 8):
        q = torch.from_numpy(S.queries(3000 + b, b, d)).cuda()
        ws = sh.new_workspace(b, k, 0)
        for _ in range(2):
            sh.search_device(q, k, workspace=ws)         # first calls set function attributes / caches
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            with torch.cuda.graph(g, stream=side):
                res = sh.search_device(q, k, workspace=ws)
        for seed in (1, 2):
            q2 = torch.from_numpy(S.queries(4000 + 10 * b + seed, b, d)).cuda()
            q.copy_(q2)
            g.replay()
            torch.cuda.synchronize()
            got = [t.clone() for t in res]
            want = sh.search_device(q2, k)
            torch.cuda.synchronize()
            assert torch.equal(got[1], want[1]) and torch.equal(got[2], want[2])
            assert torch.equal(got[0].view(torch.int32), want[0].view(torch.int32))
        del g
    sh.close()


def test_k2_sharded_equals_unsharded(torch_cuda):
    torch = torch_cuda
    from githubrepostorag_b200.index import merge_keys
    n, d, k, b = 90_000, 384, 10, 40
    rows16 = S.corpus(88, n, d)
    q = S.queries(88, b, d)
    qd = torch.from_numpy(q).cuda()
    whole = _shard(torch, rows16)
    s0, r0, v0 = [t.clone() for t in whole.search_device(qd, k, algo=2)]
    cuts = [0, 40_000, n]
    keys = torch.empty((2, b, k), dtype=torch.int64, device="cuda")
    shards = []
    for i in range(2):
        sh = _shard(torch, rows16[cuts[i]:cuts[i + 1]], row_offset=cuts[i])
        sh.search_device(qd, k, keys_out=keys[i], algo=2)
        shards.append(sh)
    s1, r1, v1, _ = merge_keys(keys, qd, k)
    torch.cuda.synchronize()
    assert torch.equal(r0, r1) and torch.equal(s0.view(torch.int32), s1.view(torch.int32))
    for sh in shards + [whole]:
        sh.close()


def test_exchange_entry_point_world_of_one(torch_cuda):
    """crb_search_exchange with a single rank (its own buffer as the only 'peer'): exercises the
    publish / wait-and-merge path on one GPU; the 2-GPU version is tests/test_gpu_multi.py."""
    import ctypes as C

    from githubrepostorag_b200 import _native as N
    torch = torch_cuda
    lib = N.load()
    n, d = 60_000, 384
    rows16 = S.corpus(99, n, d)
    sh = _shard(torch, rows16)
    cap = 4096
    buf = torch.zeros(2 * cap + 32, dtype=torch.int64, device="cuda")
    ex = N.Exchange()
    ex.rank, ex.world, ex.cap_keys = 0, 1, cap
    ex.peer_keys[0] = buf.data_ptr()
    ex.peer_counters[0] = buf.data_ptr() + 8 * 2 * cap
    published = 0
    for call, (b, k, algo) in enumerate(((1, 10, 0), (5, 7, 0), (150, 10, 2), (2, 100, 1))):
        q = torch.from_numpy(S.queries(200 + call, b, d)).cuda()
        want = [t.clone() for t in sh.search_device(q, k, algo=algo)]
        s = torch.empty((b, k), dtype=torch.float32, device="cuda")
        r = torch.empty((b, k), dtype=torch.int64, device="cuda")
        v = torch.empty((b,), dtype=torch.int32, device="cuda")
        ws = sh.new_workspace(b, k, 0)
        published += b
        N.check(lib.crb_search_exchange(sh._h, q.data_ptr(), b, k, 0, None, s.data_ptr(), r.data_ptr(), v.data_ptr(),
                                        None, ws.data_ptr(), ws.numel(), algo, torch.cuda.current_stream().cuda_stream,
                                        C.byref(ex), call, published))
        torch.cuda.synchronize()
        assert torch.equal(r, want[1]) and torch.equal(v, want[2]) and torch.equal(s.view(torch.int32), want[0].view(torch.int32))
    sh.close()


def test_row_masks_through_the_c_abi(torch_cuda):
    torch = torch_cuda
    n, d, k = 50_000, 128, 10
    rows16 = S.corpus(17, n, d)
    rng = np.random.default_rng(4)
    code = rng.integers(0, 5, n).astype(np.int32)
    sh = _shard(torch, rows16, code[None, :])
    keep_a = rng.random(n) < 0.03
    keep_b = rng.random(n) < 0.5
    ma = sh.pack_row_mask(torch.from_numpy(keep_a).cuda())
    mb = sh.pack_row_mask(torch.from_numpy(keep_b).cuda())
    batch = 40
    q = S.queries(18, batch, d)
    masks = [ma if i % 3 == 0 else (mb if i % 3 == 1 else None) for i in range(batch)]
    filt = np.array([[2 if i % 2 else -1] for i in range(batch)], np.int32)
    for algo in (1, 2, 0):
        s, r, v = sh.search_device(torch.from_numpy(q).cuda(), k, filt=filt, row_masks=masks, algo=algo)
        torch.cuda.synchronize()
        keep = np.ones((batch, n), bool)
        for i in range(batch):
            if i % 3 == 0:
                keep[i] &= keep_a
            elif i % 3 == 1:
                keep[i] &= keep_b
            if i % 2:
                keep[i] &= code == 2
        _check(q, rows16, s.cpu().numpy(), r.cpu().numpy(), v.cpu().numpy(), k, oracle.MODE_COSINE, keep=keep,
               what=f"row masks algo={algo}")
    sh.close()


def test_k2_large_batches_k_boundary_and_threads(torch_cuda):
    import threading
    torch = torch_cuda
    n, d = 40_000, 128
    rows16 = S.corpus(123, n, d)
    sh = _shard(torch, rows16)
    # a batch larger than one K2 launch serves (2048 queries) is split internally
    q = S.queries(124, 2500, d)
    s2, r2, v2 = sh.search_host(q, 10, algo=2)
    pick = [0, 1, 2047, 2048, 2049, 2499]
    s1, r1, v1 = sh.search_host(q[pick], 10, algo=1)
    assert np.array_equal(r2[pick], r1) and np.array_equal(s2[pick], s1)
    _check(q[pick[:3]], rows16, s2[pick[:3]], r2[pick[:3]], v2[pick[:3]], 10, oracle.MODE_COSINE, what="batch 2500")
    # largest k the tensor-core path takes (116), and the first it hands to the streaming kernel (117)
    for k in (116, 117, 128):
        qa = S.queries(125 + k, 9, d)
        a = sh.search_host(qa, k, algo=2)
        b = sh.search_host(qa, k, algo=1)
        assert np.array_equal(a[1], b[1]) and np.array_equal(a[0], b[0]), k
    # concurrent tensor-core batches from the worker's thread pool
    qs = [S.queries(300 + i, 130, d) for i in range(6)]
    want = [sh.search_host(x, 10, algo=1) for x in qs]
    errors = []

    def work(i):
        try:
            for _ in range(3):
                got = sh.search_host(qs[i], 10, algo=2)
                assert np.array_equal(got[1], want[i][1]) and np.array_equal(got[0], want[i][0])
        except Exception as e:  # noqa: BLE001
            errors.append(e)

    ts = [threading.Thread(target=work, args=(i,)) for i in range(6)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errors, errors[:2]
    sh.close()


def _unit_queries_exact_in_fp16(rng, batch, d):
    """Queries with ||q|| == 1 EXACTLY whose entries are fp16 numbers (+-2^-e or +-2^-(e-1), a couple of zeros when
    the count does not work out): the device normalises them to themselves and rounds them to fp16 without error,
    so the tensor-core score and the canonical fp32 score differ by accumulation rounding ONLY."""
    e = 0
    while d > 4 ** e:
        e += 1
    used = d
    while (4 ** e - used) % 3 != 0:
        used -= 1
    big = (4 ** e - used) // 3
    assert 0 <= big <= used
    q = np.zeros((batch, d), np.float32)
    for b in range(batch):
        pos = rng.permutation(d)[:used]
        mag = np.full(used, 2.0 ** -e, np.float32)
        mag[:big] = 2.0 ** -(e - 1)
        q[b, pos] = mag * rng.choice([-1.0, 1.0], used).astype(np.float32)
        assert float(np.sum(q[b].astype(np.float64) ** 2)) == 1.0
    return q


@pytest.mark.parametrize("n,d", [(40_000, 384), (40_000, 768), (30_000, 1024), (20_000, 2048), (17_000, 4096)])
def test_k2_measured_rounding_stays_inside_the_proof_margin(torch_cuda, n, d):
    """The exactness proof allows k2_rounding_allowance(dim) for the distance between the tensor-core score and the
    canonical fp32 score.  Measure it (crb_index_set_k5_stats) where it is the only error -- unit queries that are
    exact in fp16 -- on rows built to make truncation bias add up (every product positive), on Gaussian rows, and
    under both metrics: the observed maximum must stay under a QUARTER of the allowance, Gaussian queries must be
    proven (no fallback), and K2 must equal K1 bit for bit throughout (also the first K2 == K1 test at dim >= 2048)."""
    torch = torch_cuda
    rng = np.random.default_rng(d)
    batch, k = 48, 10
    q = _unit_queries_exact_in_fp16(rng, batch, d)
    rows = rng.standard_normal((n, d), dtype=np.float32)
    # rows aligned with a query so that every product q_i * x_i is positive: 40 per query, the top hits
    for b in range(batch):
        g = np.abs(rng.standard_normal((40, d), dtype=np.float32)) * np.sign(q[b]) + 0.3 * rng.standard_normal((40, d), dtype=np.float32)
        rows[1000 + 40 * b: 1040 + 40 * b] = g
    rows16 = rows.astype(np.float16)
    sh = _shard(torch, rows16)
    allowance = (d // 16 + 1) * 2.0 ** -22 + (d // 32 + 8) * 2.0 ** -24
    stats = torch.zeros((batch, 4), dtype=torch.float32, device="cuda")
    for mode in (oracle.MODE_COSINE, oracle.MODE_DOT):
        sh.set_k5_stats(stats)
        (s1, r1, v1), (s2, r2, v2) = _both(sh, torch, q, k, mode)
        sh.set_k5_stats(None)
        assert torch.equal(r1, r2) and torch.equal(v1, v2) and torch.equal(s1.view(torch.int32), s2.view(torch.int32))
        st = stats.cpu().numpy()
        max_norm = float(np.linalg.norm(rows16.astype(np.float32), axis=1).max()) if mode == oracle.MODE_DOT else 1.0
        # the margin is the allowance alone (times the largest row norm for inner product): the query rounding term is 0
        assert np.allclose(st[:, 1], allowance * max_norm, rtol=1e-3), (st[:3], allowance * max_norm)
        assert (st[:, 2] >= k).all()
        worst = float((st[:, 0] / st[:, 1]).max())
        assert worst <= 0.25, f"d={d} {mode}: observed |approx - exact| reaches {worst:.3f} of the allowed margin"
        _check(q[:2], rows16, s2.cpu().numpy()[:2], r2.cpu().numpy()[:2], v2.cpu().numpy()[:2], k, mode, what=f"margin d={d}")
    # ordinary (Gaussian) queries: rounding of the query dominates eps; nothing may need the fallback
    qg = rng.standard_normal((batch, d), dtype=np.float32)
    sh.set_k5_stats(stats)
    (s1, r1, v1), (s2, r2, v2) = _both(sh, torch, qg, k, oracle.MODE_COSINE)
    sh.set_k5_stats(None)
    st = stats.cpu().numpy()
    assert torch.equal(r1, r2) and torch.equal(s1.view(torch.int32), s2.view(torch.int32))
    assert float((st[:, 0] / st[:, 1]).max()) <= 0.25
    assert st[:, 3].sum() >= batch - 1, f"{batch - int(st[:, 3].sum())} of {batch} Gaussian queries fell back"
    sh.close()


@pytest.mark.parametrize("d,metric", [(768, oracle.MODE_COSINE), (384, oracle.MODE_COSINE), (1024, oracle.MODE_DOT)])
def test_k2_adversarial_cluster_at_the_kth_boundary(torch_cuda, d, metric):
    """Thousands of rows within 1e-5 of the k-th score (a centroid + 1e-3 noise, which fp16 quantises into
    near-duplicates and duplicates): the margin test cannot separate them, so the ids must still be the oracle's
    (tie-aware) and identical to the streaming kernel's, bit for bit -- through the proof or through the fallback."""
    torch = torch_cuda
    rng = np.random.default_rng(900 + d)
    n, k = 60_000, 10
    rows = rng.standard_normal((n, d), dtype=np.float32)
    c = rng.standard_normal(d, dtype=np.float32)
    rows[5_000:25_000] = c + 1e-3 * rng.standard_normal((20_000, d), dtype=np.float32)
    rows16 = rows.astype(np.float16)
    q = np.stack([c, c + 1e-4 * rng.standard_normal(d, dtype=np.float32), c + 0.05 * rng.standard_normal(d, dtype=np.float32),
                  rng.standard_normal(d, dtype=np.float32)] * 3).astype(np.float32)
    sh = _shard(torch, rows16)
    stats = torch.zeros((len(q), 4), dtype=torch.float32, device="cuda")
    sh.set_k5_stats(stats)
    (s1, r1, v1), (s2, r2, v2) = _both(sh, torch, q, k, metric)
    sh.set_k5_stats(None)
    assert torch.equal(r1, r2) and torch.equal(v1, v2) and torch.equal(s1.view(torch.int32), s2.view(torch.int32))
    _check(q[:4], rows16, s2.cpu().numpy()[:4], r2.cpu().numpy()[:4], v2.cpu().numpy()[:4], k, metric, what=f"cluster d={d}")
    # the cluster really is inside the margin: how many rows score within 1e-5 of the k-th
    e32 = rows16[5_000:25_000].astype(np.float32)
    if metric == oracle.MODE_COSINE:
        sc = e32 @ q[0] / (np.linalg.norm(e32, axis=1) * np.linalg.norm(q[0]))
        kth = np.sort(sc)[-k]
        assert int((np.abs(sc - kth) < 1e-5).sum()) > 1000
    else:       # inner products spread with the row norms: count the rows within 1e-4 (relative) of the k-th
        sc = e32 @ q[0]
        kth = np.sort(sc)[-k]
        assert int((np.abs(sc - kth) < 1e-4 * abs(kth)).sum()) > 1000
    sh.close()
