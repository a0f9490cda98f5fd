"""CPU test of the sharded store's host logic (world_size-2 ``gloo``): the shared-memory command channel, the
receive loop, filter codes and compiled row masks travelling to every rank, global-row bookkeeping on rank 0,
vector fetches, appends routed to a rank, deletes -- with a numpy engine (the oracle per shard + a gloo gather)
standing in for the CUDA engine.  What rank 0's faces return must equal the oracle on the unsharded corpus."""
import os
import socket

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from githubrepostorag_b200 import synthetic as S
from githubrepostorag_b200.index import MetadataDictionary
from githubrepostorag_b200.sharded import shard_range
from tests.test_sharded_gloo import pack_keys, unpack_keys


class NumpyEngine:
    """Stands in for CudaRankEngine: same interface, oracle arithmetic, gloo instead of NVLink."""

    def __init__(self, rows16, codes, row_offset, limit):
        self.rows16, self.codes, self.off, self.limit = rows16.copy(), codes.copy(), row_offset, limit
        self.dead = np.zeros(len(rows16), bool)
        self.masks = {}
        self.rank, self.world = dist.get_rank(), dist.get_world_size()

    def state(self):
        return {"n_rows": len(self.rows16), "row_offset": self.off, "dim": self.rows16.shape[1],
                "n_keys": self.codes.shape[0], "limit": self.limit}

    def search(self, q, k, metric, filt, mask_ids, algo, dest):
        b = q.shape[0]
        keys = np.zeros((b, k), np.int64)
        for i in range(b):
            keep = ~self.dead
            if filt is not None:
                for m, c in enumerate(filt[i]):
                    if c == -2:
                        keep = np.zeros_like(keep)
                    elif c >= 0:
                        keep = keep & (self.codes[m] == c)
            if mask_ids is not None and mask_ids[i] >= 0:
                keep = keep & self.masks[int(mask_ids[i])]
            s, r = oracle.query_vectorised(q[i], self.rows16, k, keep=keep, row_offset=self.off,
                                           mode=oracle.MODE_COSINE if metric == 0 else oracle.MODE_DOT)
            keys[i, :len(r)] = pack_keys(s, r)
        import torch
        gathered = [torch.zeros((b, k), dtype=torch.int64) for _ in range(self.world)] if self.rank == dest else None
        dist.gather(torch.from_numpy(keys), gathered, dst=dest)
        if self.rank != dest:
            return None
        allk = np.stack([g.numpy() for g in gathered]).view(np.uint64)        # [world, b, k]
        scores = np.full((b, k), -np.inf, np.float32)
        rows = np.full((b, k), -1, np.int64)
        nv = np.zeros(b, np.int32)
        for i in range(b):
            cand = np.sort(allk[:, i, :].reshape(-1))[::-1]
            cand = cand[cand != 0][:k]
            s, r = unpack_keys(cand.view(np.int64))
            scores[i, :len(r)], rows[i, :len(r)], nv[i] = s, r, len(r)
        return scores, rows, nv

    def search_any_of(self, q, alts, k, metric, dest):
        keep = np.zeros(len(self.rows16), bool)
        for alt in alts:
            one = ~self.dead
            for m, c in enumerate(alt):
                if c == -2:
                    one = np.zeros_like(one)
                elif c >= 0:
                    one = one & (self.codes[m] == c)
            keep |= one
        s, r = oracle.query_vectorised(np.asarray(q, np.float32), self.rows16, k, keep=keep, row_offset=self.off,
                                       mode=oracle.MODE_COSINE if metric == 0 else oracle.MODE_DOT)
        keys = np.zeros((1, k), np.int64)
        keys[0, :len(r)] = pack_keys(s, r)
        import torch
        gathered = [torch.zeros((1, k), dtype=torch.int64) for _ in range(self.world)] if self.rank == dest else None
        dist.gather(torch.from_numpy(keys), gathered, dst=dest)
        if self.rank != dest:
            return None
        cand = np.sort(np.stack([g.numpy() for g in gathered]).view(np.uint64).reshape(-1))[::-1]
        cand = cand[cand != 0][:k]
        s, r = unpack_keys(cand.view(np.int64))
        scores = np.full((1, k), -np.inf, np.float32)
        rows = np.full((1, k), -1, np.int64)
        scores[0, :len(r)], rows[0, :len(r)] = s, r
        return scores, rows, np.asarray([len(r)], np.int32)

    def _eval(self, tree):
        n = len(self.rows16)
        kind = tree[0]
        if kind == "all":
            return np.ones(n, bool)
        if kind == "none":
            return np.zeros(n, bool)
        if kind == "codes":
            return np.isin(self.codes[tree[1]], list(tree[2]))
        if kind == "rows":
            hit = np.zeros(n, bool)
            loc = np.asarray(tree[1], np.int64) - self.off
            hit[loc[(loc >= 0) & (loc < n)]] = True
            return hit
        parts = [self._eval(t) for t in tree[1]]
        out = parts[0]
        for p in parts[1:]:
            out = (out & p) if kind == "and" else (out | p)
        return out

    def set_mask(self, mid, tree):
        self.masks[mid] = self._eval(tree)

    def drop_masks(self):
        self.masks.clear()

    def fetch(self, rows):
        return self.rows16[[int(g) - self.off for g in rows]].astype(np.float32)

    def append(self, rows16, codes):
        first = self.off + len(self.rows16)
        self.rows16 = np.concatenate([self.rows16, rows16])
        self.codes = np.concatenate([self.codes, codes], axis=1)
        self.dead = np.concatenate([self.dead, np.zeros(len(rows16), bool)])
        return first

    def tombstone(self, rows):
        self.dead[[int(g) - self.off for g in rows]] = True

    def reset(self):
        dist.barrier()

    def close(self):
        pass


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from githubrepostorag_b200.sharded_store import ShardedB200VectorStore
        n, d, k = 3000, 24, 6
        rng = np.random.default_rng(3)
        emb16 = rng.standard_normal((n, d), dtype=np.float32).astype(np.float16)
        repos = [f"r{i % 5}" for i in range(n)]
        owners = [f"o{i % 3}" for i in range(n)]
        topics = [["kafka"] if i % 4 == 0 else (["sql", "auth"] if i % 4 == 1 else []) for i in range(n)]
        metas = [{"repo": repos[i], "owner": owners[i], **({"topics": topics[i]} if topics[i] else {})} for i in range(n)]
        dic = MetadataDictionary(["repo", "owner"])
        codes = dic.encode_rows(metas)
        lo, hi = shard_range(n, world, rank)
        eng = NumpyEngine(emb16[lo:hi], codes[:, lo:hi], lo, (1 << 32) - 1)
        st = ShardedB200VectorStore(eng, dictionary=dic, id_fn=S.row_id, meta_fn=lambda r: metas[r],
                                    text_fn=lambda r: f"text {r}", max_batch=8, payload_mb=4, timeout_s=60)
        if rank != 0:
            st.serve_forever()           # returns when rank 0 releases the peers ...
            out.put((rank, "released"))
            st.serve_forever()           # ... and can be entered again; ends with rank 0's close()
            out.put((rank, "ok"))
            return
        # postings of the bulk-loaded rows (a store that loads Cassandra rows fills them the same way)
        for i, md in enumerate(metas):
            st._post_row(i, md)
        ids = [S.row_id(i) for i in range(n)]
        q = rng.standard_normal((4, d), dtype=np.float32)

        def expect(qv, flt, kk=k, live=None, cond="and"):
            keep = oracle.prefilter_mask(metas, flt, condition=cond)
            if live is not None:
                keep &= live
            s, r = oracle.query_vectorised(qv, emb16, kk, keep=keep)
            return [ids[int(i)] for i in r], s

        assert len(st) == n
        # unfiltered, coded filter, unknown value, list-valued membership, OR, a batch with different filters
        for flt in (None, {"repo": "r2"}, {"repo": "r2", "owner": "o1"}, {"repo": "nope"}, {"topics": "sql"},
                    {"topics": "kafka", "owner": "o0"}, {"__any_of__": [[{"repo": "r1"}, {"topics": "auth"}]]},
                    {"repo": ("in", ["r0", "r4"])}):
            got = st.similarity_search_with_score_id_by_vector(q[0].tolist(), k=k, filter=flt)
            want_ids, want_s = expect(q[0], flt)
            assert [i for _, _, i in got] == want_ids, flt
            assert np.allclose([2 * s - 1 for _, s, _ in got], want_s, atol=1e-6)
        s_, r_, v_ = st.search_rows(q, k, [None, {"repo": "r3"}, {"topics": "sql"}, {"owner": "o2"}])
        for b, flt in enumerate((None, {"repo": "r3"}, {"topics": "sql"}, {"owner": "o2"})):
            assert [st.row_id(r) for r in r_[b, :v_[b]]] == expect(q[b], flt)[0]
        # a batch larger than the control block's capacity is served in slices
        qq = rng.standard_normal((19, d), dtype=np.float32)
        s_, r_, v_ = st.search_rows(qq, 3)
        assert all([st.row_id(r) for r in r_[b]] == expect(qq[b], None, 3)[0] for b in range(19))
        # stored vectors come from the rank that holds the row
        trip = st.similarity_search_with_embedding_id_by_vector(q[1].tolist(), k=k)
        for doc, vec, rid in trip:
            r = int(rid[4:])
            assert np.array_equal(np.asarray(vec, np.float32), emb16[r].astype(np.float32)) and doc.page_content == f"text {r}"
            assert doc.metadata["repo"] == repos[r]
        assert {int(i[4:]) >= hi for _, _, i in trip} == {True, False} or world == 1     # both shards contribute
        adj = st.adjacent({("repo", "r1"), ("topics", "kafka")}, q[2].tolist(), k=k)
        assert [c.id for c in adj] == expect(q[2], {"__any_of__": [[{"repo": "r1"}, {"topics": "kafka"}]]})[0]
        # edges that are coded equalities travel as ONE union search (CMD_SEARCH_ANY), vectors fetched from their ranks
        adj = st.adjacent([("repo", "r1"), ("owner", "o2"), ("repo", "r3")], q[3].tolist(), k=k, filter={"owner": "o2"})
        want = expect(q[3], {"owner": "o2"})[0]              # (the owner edge repeats the caller's filter: the union is it)
        assert [c.id for c in adj] == want
        assert all(np.array_equal(np.asarray(c.embedding, np.float32), emb16[int(c.id[4:])].astype(np.float32)) for c in adj)
        adj = st.adjacent([("repo", "r1"), ("repo", "r3")], q[3].tolist(), k=k)
        assert [c.id for c in adj] == expect(q[3], {"__any_of__": [[{"repo": "r1"}, {"repo": "r3"}]]})[0]
        assert st.get([S.row_id(7), S.row_id(n - 1), "nope"])[1].content == f"text {n - 1}"
        st.release_peers()
        # writes: new rows go to the rank with room (the last one: the first is bounded by its neighbour's range)
        new = rng.standard_normal((5, d), dtype=np.float32).astype(np.float16).astype(np.float32)
        st.add_embeddings([f"new-{i}" for i in range(5)], new, [f"nt{i}" for i in range(5)],
                          [{"repo": "r2", "owner": "zed", "topics": ["fresh"]} for _ in range(5)])
        assert len(st) == n + 5
        got = st.similarity_search_with_score_id_by_vector(new[3].tolist(), k=1, filter={"owner": "zed"})
        assert got[0][2] == "new-3" and got[0][0].page_content == "nt3" and abs(got[0][1] - 1.0) < 1e-3
        assert [i for _, _, i in st.similarity_search_with_score_id_by_vector(new[2].tolist(), k=2, filter={"topics": "fresh"})][0] == "new-2"
        st.add_embeddings(["new-3"], new[:1], ["again"], [{"repo": "r2", "owner": "zed"}])       # upsert
        got = st.similarity_search_with_score_id_by_vector(new[3].tolist(), k=5, filter={"owner": "zed"})
        assert sorted(i for _, _, i in got) == [f"new-{i}" for i in range(5)] and len(st) == n + 5
        assert st.get(["new-3"])[0].content == "again"
        # deletes: a bulk-loaded row and an added one never come back
        top = expect(q[3], None, 1)[0][0]
        st.delete(top)
        st.delete("new-0")
        live = np.ones(n, bool)
        live[int(top[4:])] = False
        assert [i for _, _, i in st.similarity_search_with_score_id_by_vector(q[3].tolist(), k=k)] == expect(q[3], None, live=live)[0]
        assert st.get([top, "new-0"]) == [] and len(st) == n + 3
        st.close()
        out.put((rank, "ok"))
    except Exception:  # noqa: BLE001
        import traceback
        out.put((rank, traceback.format_exc()[-2500:]))
        try:
            st.close()               # let the other ranks out of their receive loops
        except Exception:  # noqa: BLE001
            pass
    finally:
        dist.destroy_process_group()


def test_sharded_store_protocol_and_bookkeeping_two_ranks():
    world = 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, out)) for r in range(world)]
    [p.start() for p in procs]
    res = []
    try:
        for _ in range(world + 1):
            res.append(out.get(timeout=60))
    except Exception:  # noqa: BLE001 - report what did arrive
        pass
    [p.join(timeout=20) for p in procs]
    [p.kill() for p in procs if p.is_alive()]
    assert sorted(res) == [(0, "ok"), (1, "ok"), (1, "released")], res
