"""BASELINE.json's full-size configurations.  Size-independent properties over every query (planted queries
return their row first, results are sorted and duplicate-free, the tensor-core path returns bit-for-bit what the
streaming path returns, filtered results satisfy the filter, a sharded search equals the unsharded one), and for a
handful of queries per configuration -- planted and not -- the ORACLE'S IDS: the device corpus is streamed to the
host in 2**18-row blocks through `oracle.query_blocks` (the blockwise form of the restated LlamaIndex top-k), and
compared with the tie-aware comparator of tests/helpers.py.  Corpora are generated on the device (torch.randn,
seeded)."""
import numpy as np
import pytest

import oracle
from tests.helpers import assert_topk_parity

pytestmark = pytest.mark.gpu

ORACLE_BLOCK = 1 << 18


def _oracle_topk(torch, rows, q_host, k, *, keep=None, row_offset=0, mode=oracle.MODE_COSINE):
    """Oracle top-k of each query over the device-resident corpus `rows` (fp16), upcast exactly to fp32 on the
    device and streamed to the host block by block.  With `keep` (bool [n] on the device) only the survivors are
    streamed, as the oracle's prefilter would leave them."""
    n = rows.shape[0]

    def blocks():
        if keep is not None:
            surv = torch.nonzero(keep).flatten()
            for lo in range(0, surv.numel(), ORACLE_BLOCK):
                idx = surv[lo:lo + ORACLE_BLOCK]
                yield 0, rows[idx].to(torch.float32).cpu().numpy(), None, idx.cpu().numpy()
        else:
            for lo in range(0, n, ORACLE_BLOCK):
                yield lo, rows[lo:lo + ORACLE_BLOCK].to(torch.float32).cpu().numpy(), None, None

    if keep is None:
        return oracle.query_blocks(q_host, ((lo, blk) for lo, blk, _, _ in blocks()), k, mode=mode)
    # survivors only: block-local positions are mapped back to corpus rows afterwards
    out = [(np.empty(0, np.float32), np.empty(0, np.int64)) for _ in range(len(q_host))]
    for _, blk, _, idx in blocks():
        part = oracle.query_blocks(q_host, [(0, blk)], k, mode=mode)
        for b, (s, r) in enumerate(part):
            bs = np.concatenate([out[b][0], s])
            br = np.concatenate([out[b][1], idx[r].astype(np.int64)])
            order = np.lexsort((-br, -bs))[:k]
            out[b] = (bs[order], br[order])
    return out


def _assert_oracle_ids(torch, rows, q_host, got_s, got_r, got_v, ref, k, *, row_offset=0, what=""):
    for b, (s_ref, r_ref) in enumerate(ref):
        assert int(got_v[b]) == len(r_ref) == min(k, len(r_ref)), f"{what} q{b}: n_valid {int(got_v[b])} vs {len(r_ref)}"

        def ref_score(row, b=b):
            e = rows[row:row + 1].to(torch.float32).cpu().numpy()
            return oracle.score_rows(q_host[b], e)[0]
        assert_topk_parity(got_r[b, :len(r_ref)] - row_offset, got_s[b, :len(r_ref)], r_ref, s_ref, ref_score, what=f"{what} q{b}")


def _device_rows(torch, n, d, seed):
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    out = torch.empty((n, d), dtype=torch.float16, device="cuda")
    step = 1 << 20
    for lo in range(0, n, step):
        hi = min(n, lo + step)
        out[lo:hi] = torch.randn((hi - lo, d), generator=g, device="cuda").to(torch.float16)
    return out


def _planted(torch, rows, picks, noise, seed):
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    base = rows[torch.as_tensor(picks, device="cuda")].to(torch.float32)
    return (base + noise * torch.randn(base.shape, generator=g, device="cuda")).contiguous()


def _basic_properties(torch, s, r, v, k):
    assert torch.all(v == k)
    assert torch.all(s[:, :-1] >= s[:, 1:]), "scores must be descending"
    rs = torch.sort(r, dim=1).values
    assert torch.all(rs[:, 1:] != rs[:, :-1]), "duplicate rows in a result"


def _spot_check_scores(torch, rows, q, s, r, n_spot=3):
    """The reported score of a returned row equals the oracle's per-row formula (1e-5 relative)."""
    for b in range(min(n_spot, q.shape[0])):
        e = rows[r[b, :4]].to(torch.float32).cpu().numpy()
        ref = oracle.score_rows(q[b].cpu().numpy(), e)
        assert np.allclose(s[b, :4].cpu().numpy(), ref, rtol=1e-5, atol=2e-7)


def test_c3_10m_x768_batch1024_k10():
    import torch
    from githubrepostorag_b200.index import DeviceShard, merge_keys
    n, d, k, b = 10_000_000, 768, 10, 1024
    rows = _device_rows(torch, n, d, 1003)
    sh = DeviceShard.from_device(rows)
    picks = np.linspace(0, n - 1, b).astype(np.int64)
    q = _planted(torch, rows, picks, 0.05, 7)
    s2, r2, v2 = [t.clone() for t in sh.search_device(q, k, algo=2)]          # tensor-core path
    torch.cuda.synchronize()
    _basic_properties(torch, s2, r2, v2, k)
    assert torch.equal(r2[:, 0].cpu(), torch.from_numpy(picks)) and torch.all(s2[:, 0] > 0.99)
    _spot_check_scores(torch, rows, q, s2, r2)
    sub = torch.arange(0, b, 97, device="cuda")                                 # 11 queries through the streaming path
    s1, r1, v1 = sh.search_device(q[sub].contiguous(), k, algo=1)
    torch.cuda.synchronize()
    assert torch.equal(r1, r2[sub]) and torch.equal(s1.view(torch.int32), s2[sub].view(torch.int32))
    # batch 1 (north-star line) and sharded == unsharded at full size
    s0, r0, v0 = [t.clone() for t in sh.search_device(q[:1].contiguous(), k)]
    cuts = [0, 3_333_333, 7_000_001, n]
    keys = torch.empty((3, 1, k), dtype=torch.int64, device="cuda")
    shards = [DeviceShard.from_device(rows[cuts[i]:cuts[i + 1]], row_offset=cuts[i]) for i in range(3)]
    for i, p in enumerate(shards):
        p.search_device(q[:1].contiguous(), k, keys_out=keys[i])
    sm, rm, vm, _ = merge_keys(keys, q[:1].contiguous(), k)
    torch.cuda.synchronize()
    assert torch.equal(r0, rm) and torch.equal(s0.view(torch.int32), sm.view(torch.int32)) and torch.equal(r0, r2[:1])
    # the oracle's ids: 5 planted queries out of the tensor-core batch and 5 unplanted (Gaussian) ones, the
    # latter through both plans (a batch of 5 -> tensor-core kernel; one at a time -> streaming kernel)
    g = torch.Generator(device="cuda")
    g.manual_seed(11)
    qr = torch.randn((5, d), generator=g, device="cuda")
    sel = [0, 201, 517, 800, 1023]
    q_host = np.concatenate([q[sel].cpu().numpy(), qr.cpu().numpy()])
    ref = _oracle_topk(torch, rows, q_host, k)
    _assert_oracle_ids(torch, rows, q_host[:5], s2[sel].cpu().numpy(), r2[sel].cpu().numpy(), v2[sel].cpu().numpy(), ref[:5], k,
                       what="C3 tensor-core B=1024")
    sr2, rr2, vr2 = [t.clone() for t in sh.search_device(qr, k, algo=2)]
    torch.cuda.synchronize()
    _assert_oracle_ids(torch, rows, q_host[5:], sr2.cpu().numpy(), rr2.cpu().numpy(), vr2.cpu().numpy(), ref[5:], k,
                       what="C3 tensor-core B=5 unplanted")
    for i in range(5):
        s1, r1, v1 = sh.search_device(qr[i:i + 1].contiguous(), k)                # B=1: the metric's configuration
        torch.cuda.synchronize()
        assert torch.equal(r1[0], rr2[i]) and torch.equal(s1[0].view(torch.int32), sr2[i].view(torch.int32))
    for p in shards + [sh]:
        p.close()


def test_c4_10m_x1024_filtered_k20():
    import torch
    from githubrepostorag_b200.index import DeviceShard
    n, d, k = 10_000_000, 1024, 20
    rows = _device_rows(torch, n, d, 1004)
    g = torch.Generator(device="cuda")
    g.manual_seed(5)
    codes = torch.stack([torch.randint(0, 4, (n,), generator=g, device="cuda"),
                         torch.randint(0, 25, (n,), generator=g, device="cuda"),
                         torch.randint(0, 2, (n,), generator=g, device="cuda")]).to(torch.int32).contiguous()
    want = np.array([2, 7, 1], np.int32)                                        # ~0.5 % of the rows
    keep = (codes[0] == 2) & (codes[1] == 7) & (codes[2] == 1)
    surv = torch.nonzero(keep).flatten()
    assert 30_000 < surv.numel() < 70_000
    sh = DeviceShard.from_device(rows, codes)
    picks = surv[torch.linspace(0, surv.numel() - 1, 256, device="cuda").long()]
    q = _planted(torch, rows, picks.cpu().numpy(), 0.05, 8)
    for batch, algo in ((1, 0), (4, 0), (256, 0), (256, 2)):
        f = np.tile(want, (batch, 1))
        s, r, v = sh.search_device(q[:batch].contiguous(), k, filt=f, algo=algo)
        torch.cuda.synchronize()
        _basic_properties(torch, s, r, v, k)
        assert torch.equal(r[:, 0], picks[:batch]) and torch.all(keep[r.flatten()]), (batch, algo)
    # a planted row that fails the filter must not come back; without the filter it is the top hit
    outsider = int(torch.nonzero(~keep)[12345])
    qo = _planted(torch, rows, [outsider], 0.05, 9)
    s, r, v = sh.search_device(qo, k, filt=want[None, :])
    assert outsider not in r.cpu().numpy() and torch.all(keep[r.flatten()])
    s, r, v = sh.search_device(qo, k)
    assert int(r[0, 0]) == outsider
    # the oracle's ids under the filter: 4 planted + 4 unplanted queries; B=1 (streaming gather) and B=8
    g.manual_seed(12)
    q8 = torch.cat([q[[0, 85, 170, 255]], torch.randn((4, d), generator=g, device="cuda")]).contiguous()
    q_host = q8.cpu().numpy()
    ref = _oracle_topk(torch, rows, q_host, k, keep=keep)
    s8, r8, v8 = [t.clone() for t in sh.search_device(q8, k, filt=np.tile(want, (8, 1)))]
    torch.cuda.synchronize()
    _assert_oracle_ids(torch, rows, q_host, s8.cpu().numpy(), r8.cpu().numpy(), v8.cpu().numpy(), ref, k, what="C4 B=8")
    for i in range(8):
        s1, r1, v1 = sh.search_device(q8[i:i + 1].contiguous(), k, filt=want[None, :])
        torch.cuda.synchronize()
        assert torch.equal(r1[0], r8[i]) and torch.equal(s1[0].view(torch.int32), s8[i].view(torch.int32)), i
    sh.close()


def test_c5_shard_12m5_x384_batch64_k100():
    import torch
    from githubrepostorag_b200.index import DeviceShard
    n, d, k, b = 12_500_000, 384, 100, 64
    rows = _device_rows(torch, n, d, 1005)
    sh = DeviceShard.from_device(rows, row_offset=87_500_000)                  # the last of 8 shards of 100M
    picks = np.linspace(5, n - 7, b).astype(np.int64)
    q = _planted(torch, rows, picks, 0.05, 10)
    s2, r2, v2 = [t.clone() for t in sh.search_device(q, k)]                   # default dispatch -> tensor-core path
    torch.cuda.synchronize()
    _basic_properties(torch, s2, r2, v2, k)
    assert torch.equal(r2[:, 0].cpu(), torch.from_numpy(picks + 87_500_000))
    s1, r1, v1 = sh.search_device(q[:3].contiguous(), k, algo=1)
    torch.cuda.synchronize()
    assert torch.equal(r1, r2[:3]) and torch.equal(s1.view(torch.int32), s2[:3].view(torch.int32))
    # the oracle's ids at k=100: 4 planted queries of the batch and 4 unplanted ones
    g = torch.Generator(device="cuda")
    g.manual_seed(13)
    qr = torch.randn((4, d), generator=g, device="cuda")
    sel = [0, 21, 42, 63]
    q_host = np.concatenate([q[sel].cpu().numpy(), qr.cpu().numpy()])
    ref = _oracle_topk(torch, rows, q_host, k)
    _assert_oracle_ids(torch, rows, q_host[:4], s2[sel].cpu().numpy(), r2[sel].cpu().numpy(), v2[sel].cpu().numpy(), ref[:4], k,
                       row_offset=87_500_000, what="C5 shard B=64 k=100")
    sr, rr, vr = [t.clone() for t in sh.search_device(qr, k)]
    torch.cuda.synchronize()
    _assert_oracle_ids(torch, rows, q_host[4:], sr.cpu().numpy(), rr.cpu().numpy(), vr.cpu().numpy(), ref[4:], k,
                       row_offset=87_500_000, what="C5 shard unplanted")
    sh.close()
