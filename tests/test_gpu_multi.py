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


def _store_worker(rank, world, port, out):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    st = None
    try:
        import oracle
        from githubrepostorag_b200 import synthetic as S
        from githubrepostorag_b200.index import DeviceShard, MetadataDictionary
        from githubrepostorag_b200.sharded import shard_range
        from githubrepostorag_b200.sharded_store import CudaRankEngine, ShardedB200VectorStore
        n, d, k = 150_001, 384, 8
        rows16 = S.corpus(777, n, d)
        rng = np.random.default_rng(5)
        repo = (rng.zipf(1.3, n) % 40).astype(np.int32)
        module = ((repo * 7 + rng.integers(0, 6, n)) % 300).astype(np.int32)
        owner = (np.arange(n) % 3).astype(np.int32)
        topics = [["kafka"] if i % 5 == 0 else (["sql", "auth"] if i % 5 == 1 else []) for i in range(n)]
        dic = MetadataDictionary(["repo", "module", "owner"])
        dic.register_vocab("repo", [f"r{i}" for i in range(40)])
        dic.register_vocab("module", [f"m{i}" for i in range(300)])
        dic.register_vocab("owner", [f"o{i}" for i in range(3)])
        codes = np.stack([repo, module, owner])

        def meta(r):
            md = {"repo": f"r{repo[r]}", "module": f"m{module[r]}", "owner": f"o{owner[r]}"}
            if topics[r]:
                md["topics"] = topics[r]
            return md

        lo, hi = shard_range(n, world, rank)
        shard = DeviceShard.from_host(rows16[lo:hi], codes[:, lo:hi], device=rank, row_offset=lo, capacity=hi - lo + 64)
        st = ShardedB200VectorStore(CudaRankEngine(shard, cap_keys=1 << 15), dictionary=dic, id_fn=S.row_id, meta_fn=meta,
                                    text_fn=lambda r: f"text {r}", max_batch=256)
        if rank != 0:
            st.serve_forever()
            out.put((rank, "ok"))
            return
        for i in range(n):
            if topics[i]:
                st._post_row(i, {"topics": topics[i]})
        metas = [meta(r) for r in range(n)]
        ids = [S.row_id(i) for i in range(n)]
        q = S.queries(91, 6, d)

        def expect(qv, flt, kk=k, keep_extra=None):
            keep = oracle.prefilter_mask(metas, flt)
            if keep_extra is not None:
                keep &= keep_extra
            s, r = oracle.query_vectorised(qv, rows16, kk, keep=keep)
            return [ids[int(i)] for i in r], s

        # every face, filtered and not, against the oracle on the UNSHARDED corpus
        for flt in (None, {"repo": "r1"}, {"repo": "r3", "owner": "o2"}, {"module": "m17"}, {"topics": "sql"},
                    {"topics": "kafka", "repo": "r0"}, {"__any_of__": [[{"repo": "r5"}, {"topics": "auth"}]]},
                    {"repo": "no-such-repo"}):
            got = st.similarity_search_with_score_id_by_vector(q[0].tolist(), k=k, filter=flt)
            want_ids, want_s = expect(q[0], flt)
            assert [i for _, _, i in got] == want_ids, flt
            assert np.allclose([2 * s - 1 for _, s, _ in got], want_s, rtol=1e-5, atol=2e-6), flt
        # a batch with a different filter per query, and a big unfiltered batch (tensor-core plan on each shard)
        flts = [None, {"repo": "r2"}, {"topics": "sql"}, {"owner": "o1"}, {"repo": "r2"}, None]
        s_, r_, v_ = st.search_rows(q, k, flts)
        for b, flt in enumerate(flts):
            assert [st.row_id(r) for r in r_[b, :v_[b]]] == expect(q[b], flt)[0], (b, flt)
        qq = S.queries(92, 200, d)
        s_, r_, v_ = st.search_rows(qq, 10)
        for b in (0, 77, 199):
            assert [st.row_id(r) for r in r_[b]] == expect(qq[b], None, 10)[0]
        # adjacent(): one query vector, one filter per edge, ONE sharded search; vectors come from the owning rank
        adj = st.adjacent({("repo", "r1"), ("module", "m17"), ("topics", "kafka")}, q[1].tolist(), k=k, filter={"owner": "o0"})
        want = expect(q[1], {"owner": "o0", "__any_of__": [[{"repo": "r1"}, {"module": "m17"}, {"topics": "kafka"}]]})[0]
        assert [c.id for c in adj] == want
        for c in adj:
            assert np.array_equal(np.asarray(c.embedding, np.float32), rows16[int(c.id[4:])].astype(np.float32))
        # edges that are coded equalities go out as ONE union search over every shard (crb_search_exchange_host_any_of):
        # the same rows as the per-edge batch, and as the oracle under the OR of the edges
        edges = [("repo", "r1"), ("module", "m17"), ("repo", "r4"), ("owner", "o0")]
        adj = st.adjacent(edges, q[1].tolist(), k=k, filter={"owner": "o0"})
        assert [c.id for c in adj] == expect(q[1], {"owner": "o0"})[0]                 # the owner edge is the caller's filter
        adj = st.adjacent(edges[:3], q[1].tolist(), k=k, filter={"owner": "o0"})
        want = expect(q[1], {"owner": "o0", "__any_of__": [[{"repo": "r1"}, {"module": "m17"}, {"repo": "r4"}]]})[0]
        assert [c.id for c in adj] == want
        for c in adj:
            assert np.array_equal(np.asarray(c.embedding, np.float32), rows16[int(c.id[4:])].astype(np.float32))
        type(st).MAX_ALTERNATIVES, keep_max = 0, type(st).MAX_ALTERNATIVES            # the per-edge batch
        try:
            assert [c.id for c in st.adjacent(edges[:3], q[1].tolist(), k=k, filter={"owner": "o0"})] == want
        finally:
            type(st).MAX_ALTERNATIVES = keep_max
        # the graph retriever over the sharded store (the reference's retriever.invoke(q, filter=...))
        from githubrepostorag_b200.retrievers import Eager, EagerGraphRetriever

        class Emb:
            def embed_query(self, text):
                return q[2].tolist()
        st.embeddings = Emb()
        docs = EagerGraphRetriever(st, [("repo", "repo"), ("module", "module")], Eager(k=6, start_k=2, adjacent_k=4, max_depth=2)) \
            .invoke("anything", filter={"owner": "o1"})
        assert len(docs) == 6 and all(dd.metadata["owner"] == "o1" for dd in docs)
        assert docs[0].id == expect(q[2], {"owner": "o1"}, 1)[0][0]
        # writes at N > 1: appended rows are searchable, upserts and deletes hold
        new = rng.standard_normal((3, d), dtype=np.float32).astype(np.float16).astype(np.float32)
        st.add_embeddings(["n0", "n1", "n2"], new, ["a", "b", "c"], [{"repo": "r1", "owner": "zed", "topics": ["fresh"]}] * 3)
        got = st.similarity_search_with_score_id_by_vector(new[1].tolist(), k=2, filter={"owner": "zed"})
        assert got[0][2] == "n1" and abs(got[0][1] - 1.0) < 1e-3
        assert [i for _, _, i in st.similarity_search_with_score_id_by_vector(new[2].tolist(), k=1, filter={"topics": "fresh"})] == ["n2"]
        top = expect(q[3], None, 1)[0][0]
        st.delete(top)
        st.delete("n1")
        live = np.ones(n, bool)
        live[int(top[4:])] = False
        assert [i for _, _, i in st.similarity_search_with_score_id_by_vector(q[3].tolist(), k=k)] == expect(q[3], None, keep_extra=live)[0]
        assert "n1" not in [i for _, _, i in st.similarity_search_with_score_id_by_vector(new[1].tolist(), k=3, filter={"owner": "zed"})]
        st.close()
        out.put((rank, "ok"))
    except Exception:  # noqa: BLE001
        import traceback
        out.put((rank, traceback.format_exc()[-2500:]))
        try:
            if st is not None and rank == 0:
                st.close()
        except Exception:  # noqa: BLE001
            pass
    finally:
        dist.destroy_process_group()


def test_two_gpu_sharded_store_faces_filters_adjacent_and_writes():
    """ONE store object over the GPUs of the box -- rank 0 owns the faces, rank 1 runs the receive
    loop.  Filtered searches, row masks, adjacent(), the graph retriever and the write path at N = 2 return what the
    oracle returns on the unsharded corpus."""
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
    procs = [ctx.Process(target=_store_worker, args=(r, world, port, out)) for r in range(world)]
    [p.start() for p in procs]
    res = []
    try:
        for _ in procs:
            res.append(out.get(timeout=400))
    except Exception:  # noqa: BLE001
        pass
    [p.join(timeout=60) for p in procs]
    [p.kill() for p in procs if p.is_alive()]
    assert sorted(res) == [(r, "ok") for r in range(world)], res


def test_sharded_store_world_of_one():
    """The same store with a world of one (what a single-GPU box can run): the command channel, the exchange with
    itself as the only peer, masks by id, fetches, the graph retriever and writes -- against the oracle."""
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    p = ctx.Process(target=_store_worker, args=(0, 1, port, out))
    p.start()
    res = []
    try:
        res.append(out.get(timeout=400))
    except Exception:  # noqa: BLE001
        pass
    p.join(timeout=60)
    if p.is_alive():
        p.kill()
    assert res == [(0, "ok")], res


def test_serve_entry_point_answers_json_lines(tmp_path):
    """`python -m githubrepostorag_b200.serve --load <saved table>` with a world of one: the table is
    saved in the streaming directory format, loaded as a sharded store, and requests on stdin come back as JSON lines
    with the oracle's ids -- filtered, list-valued and unfiltered."""
    import json
    import subprocess
    import sys

    import oracle
    from githubrepostorag_b200 import synthetic as S
    from githubrepostorag_b200.vector_store import B200VectorStore
    rng = np.random.default_rng(12)
    n, d, k = 3000, 64, 5
    emb = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16).astype(np.float32)
    metas = [{"namespace": "default", "repo": f"r{i % 4}", **({"topics": ["kafka"]} if i % 3 == 0 else {})} for i in range(n)]
    ids = [S.row_id(i) for i in range(n)]
    st = B200VectorStore(dim=d, scope="chunk")
    st.add_embeddings(ids, emb, [f"text {i}" for i in range(n)], metas)
    st.save(str(tmp_path / "table"))
    st.close()
    q = rng.standard_normal((3, d), dtype=np.float32)
    reqs = [{"embedding": q[0].tolist(), "k": k}, {"embedding": q[1].tolist(), "k": k, "filter": {"repo": "r2"}},
            {"embedding": q[2].tolist(), "k": k, "filter": {"topics": "kafka", "namespace": "default"}}, {"k": 1}]
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    env = dict(os.environ, RANK="0", LOCAL_RANK="0", WORLD_SIZE="1", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "-m", "githubrepostorag_b200.serve", "--load", str(tmp_path / "table")],
                         input="\n".join(json.dumps(r) for r in reqs) + "\n", capture_output=True, text=True, env=env,
                         cwd=root, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [json.loads(ln) for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 4 and "error" in lines[3]                       # a bad request is answered, not fatal
    for rep, req in zip(lines, reqs[:3]):
        keep = oracle.prefilter_mask(metas, req.get("filter"))
        s_ref, r_ref = oracle.query_vectorised(np.asarray(req["embedding"], np.float32), emb.astype(np.float16), k, keep=keep)
        assert [h["id"] for h in rep["hits"]] == [ids[int(r)] for r in r_ref]
        assert np.allclose([h["score"] for h in rep["hits"]], s_ref, rtol=1e-5, atol=2e-7)
        assert rep["hits"][0]["text"] == f"text {int(r_ref[0])}"
