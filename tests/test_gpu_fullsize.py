"""BASELINE.json's full-size configurations, checked through size-independent properties (the oracle
cannot finish these in seconds): planted queries return their row first, results are sorted and
duplicate-free, the tensor-core path returns bit-for-bit what the streaming path returns, filtered
results satisfy the filter, a sharded search equals the unsharded one.  Corpora are generated on the
device (torch.randn, seeded) -- nothing here reads the oracle except for spot rows."""
import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu


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
    sh.close()
