"""Multi-GPU tests (need >= 2 B200s on the box: ``gpurun --gpus 2``; skipped otherwise): the NCCL
all-gather searcher and the fused NVLink peer-memory exchange both return exactly the unsharded
result (same rows, same score bits) and match the oracle."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, out):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        import oracle
        from githubrepostorag_b200 import synthetic as S
        from githubrepostorag_b200.index import DeviceShard
        from githubrepostorag_b200.sharded import FusedExchangeSearcher, cuda_searcher, shard_range
        n, d = 120_001, 384
        rows16 = S.corpus(4321, n, d)
        lo, hi = shard_range(n, world, rank)
        shard = DeviceShard.from_host(rows16[lo:hi], None, device=rank, row_offset=lo)
        whole = DeviceShard.from_host(rows16, None, device=rank) if rank == 0 else None
        nccl = cuda_searcher(shard)
        fused = FusedExchangeSearcher(shard, cap_keys=1 << 15)
        for call, (b, k, algo) in enumerate(((1, 10, 0), (3, 5, 0), (40, 10, 2), (1, 100, 0), (200, 10, 0), (2, 10, 0))):
            q = torch.from_numpy(S.queries(50 + call, b, d)).to(dev)
            if rank != 0:
                q.zero_()                      # only rank 0 holds the request
            fused.algo = algo
            s1, r1, v1, _ = nccl.search(q.clone(), k, broadcast=True)
            s2, r2, v2, _ = fused.search(q.clone(), k, broadcast=True)
            torch.cuda.synchronize()
            assert torch.equal(r1, r2) and torch.equal(v1, v2), (rank, call)
            assert torch.equal(s1.view(torch.int32), s2.view(torch.int32)), (rank, call)
            if rank == 0:
                dist.broadcast(q, src=0)
                s0, r0, v0 = whole.search_device(q, k)
                torch.cuda.synchronize()
                assert torch.equal(r0, r2) and torch.equal(s0.view(torch.int32), s2.view(torch.int32)), call
                qn = q.cpu().numpy()
                for i in range(min(b, 2)):
                    s_ref, r_ref = oracle.query_vectorised(qn[i], rows16, k)
                    assert r2[i].cpu().numpy().tolist() == r_ref.tolist()
                    assert np.allclose(s2[i].cpu().numpy(), s_ref, rtol=1e-5, atol=2e-7)
            else:
                dist.broadcast(q, src=0)
        out.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback
        out.put((rank, traceback.format_exc()[-1500:]))
    finally:
        dist.destroy_process_group()


def test_two_gpu_sharded_nccl_and_fused_exchange():
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    world = 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, out)) for r in range(world)]
    [p.start() for p in procs]
    res = [out.get(timeout=600) for _ in procs]
    [p.join(timeout=60) for p in procs]
    assert sorted(res) == [(r, "ok") for r in range(world)], res
