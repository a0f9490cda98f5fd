"""GPU test of ``crb_search_noalloc`` -- the search entry for hosts that may not allocate device
memory (a process whose HBM is already fully committed to the corpus and its own pool, or a
plugin host that owns the memory budget).  Collected last on purpose (file name), so a run that
stops at the first failure has already exercised every other test.

It calls the entry point through the C ABI on raw device pointers and checks it against the
existing ``crb_search`` on the same inputs -- the header promises identical results, so rows, counts AND score
bits must be equal (which is inside the tolerance the other GPU tests use) -- and once against the oracle
directly, on calls in a row (a call must not be disturbed by what the previous one left in the shared scratch
array).  It passes both alone in a fresh process and after the rest of the suite.
"""
import numpy as np
import pytest

import oracle
from githubrepostorag_b200 import synthetic as S
from tests.helpers import SCORE_ATOL_UNIT, SCORE_RTOL, assert_topk_parity

pytestmark = pytest.mark.gpu


def _noalloc(lib, shard, q, k, metric, torch):
    b = q.shape[0]
    scores = torch.full((b, k), float("nan"), dtype=torch.float32, device="cuda")
    rows = torch.full((b, k), -7, dtype=torch.int64, device="cuda")
    n_valid = torch.full((b,), -7, dtype=torch.int32, device="cuda")
    rc = lib.crb_search_noalloc(shard._h, q.data_ptr(), b, k, metric, scores.data_ptr(), rows.data_ptr(),
                                n_valid.data_ptr(), torch.cuda.current_stream().cuda_stream)
    return rc, scores, rows, n_valid


def test_noalloc_matches_crb_search_twice_in_a_row():
    import torch

    from githubrepostorag_b200 import _native as N
    from githubrepostorag_b200.index import DeviceShard
    assert torch.cuda.is_available()
    lib = N.load()
    for n, d in ((50_000, 768), (20_000, 384), (3_000, 2048)):
        rows16 = S.corpus(4242, n, d)
        shard = DeviceShard.from_host(rows16, None, device=0)
        for call, (batch, k, metric) in enumerate(((1, 10, N.METRIC_COSINE), (4, 100, N.METRIC_IP), (3, 5, N.METRIC_COSINE))):
            q = torch.from_numpy(S.queries(4242 + call, batch, d)).cuda()
            want_s, want_r, want_v = shard.search_device(q, k, metric=metric, algo=1)
            rc, got_s, got_r, got_v = _noalloc(lib, shard, q, k, metric, torch)
            assert rc == 0, lib.crb_last_error()
            torch.cuda.synchronize()
            assert torch.equal(got_v, want_v) and torch.equal(got_r, want_r), (n, d, call)
            assert torch.equal(got_s.view(torch.int32), want_s.view(torch.int32)), (n, d, call)     # "identical": the bits
            ws, gs = want_s.cpu().numpy(), got_s.cpu().numpy()
            scale = 1.0 if metric == N.METRIC_COSINE else float(np.abs(ws).max())
            assert np.all(np.abs(gs - ws) <= SCORE_RTOL * np.abs(ws) + SCORE_ATOL_UNIT * scale), (n, d, call)
            # and the oracle itself, without crb_search in between
            mode = oracle.MODE_COSINE if metric == N.METRIC_COSINE else oracle.MODE_DOT
            qn = q.cpu().numpy()
            e32 = rows16.astype(np.float32)
            osc = 1.0 if mode == oracle.MODE_COSINE else float(np.linalg.norm(qn[0]) * np.linalg.norm(e32, axis=1).max())
            s_ref, r_ref = oracle.query_vectorised(qn[0], rows16, k, mode=mode)
            assert_topk_parity(got_r[0].cpu().numpy(), gs[0], r_ref, s_ref,
                               lambda row: oracle.score_rows(qn[0], e32[row:row + 1], mode)[0], scale=osc,
                               what=f"noalloc n={n} d={d} call={call}")
        shard.close()


def test_noalloc_refuses_sizes_beyond_its_fixed_scratch():
    import torch

    from githubrepostorag_b200 import _native as N
    from githubrepostorag_b200.index import DeviceShard
    lib = N.load()
    shard = DeviceShard.from_host(S.corpus(1, 1000, 64), None, device=0)
    q = torch.zeros((N.NOALLOC_MAX_BATCH + 1, 64), dtype=torch.float32, device="cuda")
    rc, *_ = _noalloc(lib, shard, q, 5, N.METRIC_COSINE, torch)
    assert rc != 0 and b"CRB_NOALLOC_MAX_BATCH" in lib.crb_last_error()
    rc, *_ = _noalloc(lib, shard, q[:1], N.MAX_TOPK + 1, N.METRIC_COSINE, torch)
    assert rc != 0 and b"k must be" in lib.crb_last_error()
    shard.close()
