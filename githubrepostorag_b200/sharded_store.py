"""``ShardedB200VectorStore`` -- ONE store object over the GPUs of a box.

The reference calls one retriever object from one worker process (``retriever.invoke(q, filter=filters)``,
``rag_worker/src/worker/services/agent_graph.py:249,268``; ``replicas: 1``,
``helm/templates/rag-worker-deployment.yaml:8``).  Here the corpus rows are sharded over ``G`` processes, one per
GPU (``torchrun``); **rank 0 owns the Python boundary** -- the three faces of :class:`B200VectorStore`, the metadata
dictionary and the ids / text / metadata of ALL rows -- and ranks 1..G-1 run :meth:`serve_forever`, a receive loop.

Per request:

1. rank 0 writes the command (queries, filter codes, mask ids, k ...) into a small block of POSIX shared memory
   and bumps its sequence number; the other ranks spin on that word (host memory on the same box: the hand-off
   costs a cache-line transfer, not a collective launch -- the NCCL broadcast this replaces was the largest
   fixed cost of a batch-1 search on 8 GPUs);
2. every rank copies the queries from that block to its own GPU and searches its shard
   (``crb_search_exchange_host``); the search kernels store each query's k exact keys straight into rank 0's
   gather buffer over NVLink peer memory (``dest_mask`` = rank 0 only) and bump its arrival counters;
3. rank 0's wait-and-merge kernel produces the global top-k, one device-to-host copy brings it back, and the
   faces map the global rows to ids / text / metadata.

Filters travel as int32 codes; predicates one code per key cannot express (list-valued keys, IN, OR) are compiled on
rank 0 into per-rank row sets / code sets, evaluated into a device bit mask by each rank once, cached under a mask id
until the next write.  Writes (append / upsert / delete) are commands too: a row lives on the rank with room that holds
the fewest rows, its global row number is ``row_offset(rank) + local row``.

The engine behind each rank is injected, so the protocol, the row bookkeeping and the mask compiler run on CPU with
``gloo`` (tests/test_sharded_gloo.py); :class:`CudaRankEngine` is the CUDA wiring.
"""
from __future__ import annotations

import pickle
import threading
import time
from typing import Any, Callable, Dict, List, Mapping, Optional, Sequence, Tuple

import numpy as np

from . import _native as N
from .index import MetadataDictionary
from .vector_store import B200VectorStore, _Rows

(CMD_SEARCH, CMD_SET_MASK, CMD_DROP_MASKS, CMD_FETCH, CMD_APPEND, CMD_TOMBSTONE, CMD_RESET, CMD_SHUTDOWN, CMD_STATE,
 CMD_RELEASE, CMD_SEARCH_ANY) = range(1, 12)

_HDR_WORDS = 32            # int32 header words after the sequence number
_MAX_RANKS = N.MAX_PEERS


class ControlBlock:
    """The shared-memory block the ranks of one box talk through.

    Layout: ``seq`` u64 | header int32[32] | per-rank ``ack`` u64[8] | per-rank ``status`` int32[8] |
    queries fp32[max_batch * dim] | filt int32[max_batch * MAX_KEYS] | mask ids int32[max_batch] |
    payload bytes (masks, appended rows, row lists) | per-rank reply bytes (fetched vectors, state, errors).
    Rank 0 is the only writer of everything but ``ack`` / ``status`` / its own reply slot; a command is complete
    before ``seq`` moves (x86 stores are not reordered, and each side reads ``seq`` / ``ack`` before the data).
    """

    def __init__(self, world: int, dim: int, max_batch: int, payload_bytes: int, reply_bytes: int, *,
                 name: Optional[str] = None):
        from multiprocessing import shared_memory
        self.world, self.dim, self.max_batch = world, dim, max_batch
        off = 8 + 4 * _HDR_WORDS
        self._o_ack = off; off += 8 * _MAX_RANKS
        self._o_status = off; off += 4 * _MAX_RANKS
        off = (off + 255) // 256 * 256
        self._o_q = off; off += 4 * max_batch * dim
        self._o_filt = off; off += 4 * max_batch * N.MAX_KEYS
        self._o_mask = off; off += 4 * max_batch
        off = (off + 255) // 256 * 256
        self._o_payload = off; off += payload_bytes
        self._o_reply = off; off += reply_bytes * world
        self.payload_bytes, self.reply_bytes, self.size = payload_bytes, reply_bytes, off
        if name is None:
            self.shm = shared_memory.SharedMemory(create=True, size=off)
            self.owner = True
            self.shm.buf[:8 + 4 * _HDR_WORDS + 12 * _MAX_RANKS] = bytes(8 + 4 * _HDR_WORDS + 12 * _MAX_RANKS)
        else:
            self.shm = shared_memory.SharedMemory(name=name)
            self.owner = False
            try:    # attaching registers the segment with this process's resource tracker, which would unlink it at exit
                from multiprocessing import resource_tracker
                resource_tracker.unregister(self.shm._name, "shared_memory")
            except Exception:  # noqa: BLE001
                pass
        buf = self.shm.buf
        self.seq = np.frombuffer(buf, dtype=np.uint64, count=1, offset=0)
        self.hdr = np.frombuffer(buf, dtype=np.int32, count=_HDR_WORDS, offset=8)
        self.ack = np.frombuffer(buf, dtype=np.uint64, count=_MAX_RANKS, offset=self._o_ack)
        self.status = np.frombuffer(buf, dtype=np.int32, count=_MAX_RANKS, offset=self._o_status)
        self.queries = np.frombuffer(buf, dtype=np.float32, count=max_batch * dim, offset=self._o_q)
        self.filt = np.frombuffer(buf, dtype=np.int32, count=max_batch * N.MAX_KEYS, offset=self._o_filt)
        self.mask_ids = np.frombuffer(buf, dtype=np.int32, count=max_batch, offset=self._o_mask)
        self.payload = np.frombuffer(buf, dtype=np.uint8, count=payload_bytes, offset=self._o_payload)
        self.replies = [np.frombuffer(buf, dtype=np.uint8, count=reply_bytes, offset=self._o_reply + r * reply_bytes)
                        for r in range(world)]

    @property
    def name(self) -> str:
        return self.shm.name

    def base_address(self) -> int:
        return int(np.frombuffer(self.shm.buf, dtype=np.uint8).ctypes.data)

    def c_layout(self, rank: int) -> "N.Ctl":
        """``crb_ctl``: where the C serving path finds things in this block."""
        c = N.Ctl()
        c.seq_off, c.hdr_off, c.ack_off, c.status_off = 0, 8, self._o_ack, self._o_status
        c.queries_off, c.filt_off, c.rank, c.world = self._o_q, self._o_filt, rank, self.world
        return c

    def put_payload(self, obj: Any) -> int:
        raw = pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL)
        if len(raw) > self.payload_bytes:
            raise ValueError(f"command payload of {len(raw)} bytes exceeds the control block's {self.payload_bytes}")
        self.payload[:len(raw)] = np.frombuffer(raw, dtype=np.uint8)
        return len(raw)

    def get_payload(self, n: int) -> Any:
        return pickle.loads(self.payload[:n].tobytes())

    def put_reply(self, rank: int, obj: Any) -> None:
        raw = pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL)
        if len(raw) + 8 > self.reply_bytes:
            raise ValueError(f"reply of {len(raw)} bytes exceeds the control block's {self.reply_bytes}")
        slot = self.replies[rank]
        slot[8:8 + len(raw)] = np.frombuffer(raw, dtype=np.uint8)
        slot[:8].view(np.uint64)[0] = len(raw)

    def get_reply(self, rank: int) -> Any:
        slot = self.replies[rank]
        n = int(slot[:8].view(np.uint64)[0])
        return pickle.loads(slot[8:8 + n].tobytes())

    def close(self) -> None:
        # drop the numpy views before the mapping goes away
        for name in ("seq", "hdr", "ack", "status", "queries", "filt", "mask_ids", "payload", "replies"):
            setattr(self, name, None)
        try:
            self.shm.close()
            if self.owner:
                self.shm.unlink()
        except Exception:  # noqa: BLE001
            pass


def _spin_until(cond: Callable[[], bool], timeout_s: float, what: str) -> None:
    """Busy-wait (the hand-off is on the request's critical path), backing off to short sleeps when idle."""
    t0 = time.perf_counter()
    spins = 0
    while not cond():
        spins += 1
        if spins > 20_000:
            el = time.perf_counter() - t0
            if el > timeout_s:
                raise TimeoutError(f"sharded store: timed out after {el:.1f}s waiting for {what}")
            if el > 0.01:
                time.sleep(0.0002 if el < 1.0 else 0.002)


# ------------------------------------------------------------------------------------------------ engines
class CudaRankEngine:
    """One rank's GPU side: its :class:`~githubrepostorag_b200.index.DeviceShard` and the NVLink exchange."""

    def __init__(self, shard, *, group=None, cap_keys: int = 1 << 17, limit: Optional[int] = None):
        import torch.distributed as dist

        from .sharded import NVLinkExchange
        self.shard = shard
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        self.exchange = NVLinkExchange(shard.device, self.group, cap_keys)
        import ctypes as C
        self._ex_ref = C.byref(self.exchange.struct)
        self.limit = limit            # first global row this rank may NOT use (None = no bound but the 32-bit key)
        self._masks: Dict[int, Any] = {}
        self._registered = None

    def register_host(self, buf) -> None:
        """Page-lock the store's shared-memory command block so query batches DMA straight out of it."""
        addr = int(np.frombuffer(buf, dtype=np.uint8).ctypes.data)
        self._registered = addr if N.load().crb_host_register(addr, len(buf)) == 0 else None

    def unregister_host(self) -> None:
        if self._registered is not None:
            N.load().crb_host_unregister(self._registered)
            self._registered = None

    # what the store needs to know about a rank
    def state(self) -> Dict[str, int]:
        return {"n_rows": self.shard.n_rows, "row_offset": self.shard.row_offset, "dim": self.shard.dim,
                "n_keys": self.shard.n_keys,
                "limit": int(self.limit) if self.limit is not None else (1 << 32) - 1}

    def search(self, queries: np.ndarray, k: int, metric: int, filt: Optional[np.ndarray],
               mask_ids: Optional[np.ndarray], algo: int, dest_rank: int):
        """Collective over the ranks: returns ``(scores, rows, n_valid)`` on ``dest_rank``, None elsewhere."""
        import ctypes as C
        lib, ex = N.load(), self.exchange
        b = queries.shape[0]
        receives = self.rank == dest_rank
        out, optr = None, (None, None, None)
        if receives:
            # fresh arrays per call: the caller keeps them
            out = (np.empty((b, k), np.float32), np.empty((b, k), np.int64), np.empty((b,), np.int32))
            optr = (out[0].ctypes.data, out[1].ctypes.data, out[2].ctypes.data)
        ptrs = None
        if mask_ids is not None and (mask_ids >= 0).any():
            ptrs = (C.c_void_p * b)()
            for i, mid in enumerate(mask_ids.tolist()):
                if mid >= 0:
                    ptrs[i] = self._masks[mid].data_ptr()
        ex.struct.dest_mask = 1 << dest_rank
        call, total = ex.next_call(b)
        rc = lib.crb_search_exchange_host(
            self.shard._h, queries.ctypes.data, b, k, metric, None if filt is None else filt.ctypes.data, ptrs,
            optr[0], optr[1], optr[2], algo, self._ex_ref, call, total)
        if rc:
            N.check(rc)
        return out

    def search_any_of(self, query: np.ndarray, alts: np.ndarray, k: int, metric: int, dest_rank: int):
        """Collective: ONE query, a row passes when any row of ``alts`` (int32 ``[n_alt, n_keys]``) passes
        (``crb_search_exchange_host_any_of``).  ``(scores [1, k], rows [1, k], n_valid [1])`` on ``dest_rank``."""
        lib, ex = N.load(), self.exchange
        out, optr = None, (None, None, None)
        if self.rank == dest_rank:
            out = (np.empty((1, k), np.float32), np.empty((1, k), np.int64), np.empty((1,), np.int32))
            optr = (out[0].ctypes.data, out[1].ctypes.data, out[2].ctypes.data)
        q = np.ascontiguousarray(query, dtype=np.float32).reshape(-1)
        a = np.ascontiguousarray(alts, dtype=np.int32)
        ex.struct.dest_mask = 1 << dest_rank
        call, total = ex.next_call(1)
        rc = lib.crb_search_exchange_host_any_of(self.shard._h, q.ctypes.data, a.shape[0], a.ctypes.data if a.size else None,
                                                 k, metric, optr[0], optr[1], optr[2], self._ex_ref, call, total)
        if rc:
            N.check(rc)
        return out

    # ---- the request path in C (crb_serve_post_search / crb_serve_next): no Python between the command word and the launch
    def bind_control(self, ctl: "ControlBlock") -> None:
        import ctypes as C
        self._ctl_base = ctl.base_address()
        self._ctl_layout = ctl.c_layout(self.rank)
        self._ctl_ref = C.byref(self._ctl_layout)
        self.ex_state = self.exchange.state

    def post_search(self, queries: np.ndarray, k: int, metric: int, filt: Optional[np.ndarray], algo: int, seq_c,
                    timeout_s: float):
        """Rank 0: publish a plain search and run this rank's share; returns (scores, rows, n_valid), or the
        integer 2 when a rank reported a failure of the previous command."""
        import ctypes as C
        b = queries.shape[0]
        out = (np.empty((b, k), np.float32), np.empty((b, k), np.int64), np.empty((b,), np.int32))
        self.exchange.struct.dest_mask = 1
        rc = N.load().crb_serve_post_search(
            self.shard._h, self._ctl_base, self._ctl_ref, queries.ctypes.data, b, k, metric,
            None if filt is None else filt.ctypes.data, out[0].ctypes.data, out[1].ctypes.data, out[2].ctypes.data, algo,
            self._ex_ref, C.addressof(self.ex_state), C.addressof(seq_c), timeout_s)
        if rc == 2:
            return 2
        if rc:
            N.check(rc)
        return out

    def serve_next(self, last_c) -> int:
        """Ranks > 0: serve plain searches until a command needs the host side (0) or a search failed (-1)."""
        import ctypes as C
        self.exchange.struct.dest_mask = 1
        return int(N.load().crb_serve_next(self.shard._h, self._ctl_base, self._ctl_ref, self._ex_ref,
                                           C.addressof(self.ex_state), C.addressof(last_c)))

    def set_mask(self, mask_id: int, tree) -> None:
        self._masks[mask_id] = self.shard.pack_row_mask(self._eval(tree))

    def _eval(self, tree):
        import torch
        sh = self.shard
        kind = tree[0]
        if kind == "all":
            return torch.ones(sh.n_rows, dtype=torch.bool, device=sh.device)
        if kind == "none":
            return torch.zeros(sh.n_rows, dtype=torch.bool, device=sh.device)
        if kind == "codes":
            _, pos, codes = tree
            if sh.codes is None:
                return torch.zeros(sh.n_rows, dtype=torch.bool, device=sh.device)
            return torch.isin(sh.codes[pos], torch.as_tensor(list(codes), dtype=sh.codes[pos].dtype, device=sh.device))
        if kind == "rows":
            hit = torch.zeros(sh.n_rows, dtype=torch.bool, device=sh.device)
            rows = np.asarray(tree[1], dtype=np.int64) - sh.row_offset
            rows = rows[(rows >= 0) & (rows < sh.n_rows)]
            if len(rows):
                hit[torch.from_numpy(rows).to(sh.device)] = True
            return hit
        parts = [self._eval(t) for t in tree[1]]
        out = parts[0]
        for p in parts[1:]:
            out = (out & p) if kind == "and" else (out | p)
        return out

    def drop_masks(self) -> None:
        self._masks.clear()

    def fetch(self, global_rows: Sequence[int]) -> np.ndarray:
        local = [int(g) - self.shard.row_offset for g in global_rows]
        return self.shard.fetch_rows(local)

    def append(self, rows16: np.ndarray, codes: Optional[np.ndarray]) -> int:
        return self.shard.row_offset + self.shard.append(rows16, codes)

    def tombstone(self, global_rows: Sequence[int]) -> None:
        self.shard.tombstone([int(g) - self.shard.row_offset for g in global_rows])

    def reset(self) -> None:
        self.exchange.reset()

    def close(self) -> None:
        self.shard.close()


# ------------------------------------------------------------------------------------------------ host rows
class _ShardedRows(_Rows):
    """ids / text / metadata of all rows, kept on rank 0 and addressed by GLOBAL row number.  Bulk-loaded shards
    synthesise theirs from the row number (``id_fn`` ...); rows added through the faces are list-backed."""

    def __init__(self):
        super().__init__()
        self.host_of: Dict[int, int] = {}     # global row -> index into ids / texts / metas
        self.n_bulk = 0

    def __len__(self) -> int:
        return self.n_bulk + len(self.host_of)

    def id(self, r: int) -> str:
        h = self.host_of.get(r)
        return self.ids[h] if h is not None else self.id_fn(r)

    def text(self, r: int) -> str:
        h = self.host_of.get(r)
        if h is not None:
            return self.texts[h]
        return self.text_fn(r) if self.text_fn else ""

    def meta(self, r: int) -> Dict[str, Any]:
        h = self.host_of.get(r)
        if h is not None:
            return dict(self.metas[h])
        return dict(self.meta_fn(r)) if self.meta_fn else {}


# ------------------------------------------------------------------------------------------------ the store
class ShardedB200VectorStore(B200VectorStore):
    """Collective constructor (every rank calls it); afterwards rank 0 uses the faces of :class:`B200VectorStore`
    and every other rank calls :meth:`serve_forever`.  ``close()`` on rank 0 ends the peers' loops."""

    def _search_any_of(self, q, k, flts):
        """The union search over every rank's shard (``crb_search_exchange_host_any_of``), the hits' vectors fetched
        from the ranks that hold them.  None (-> the per-edge batch) when an edge needs a row mask, the engine has no
        union search, or the alternatives do not fit the command block."""
        compiled = [self._compile_filter(f) for f in flts]
        if any(extra or row is None for _, extra, row in compiled) or not hasattr(self.engine, "search_any_of") \
                or len(compiled) > self.ctl.max_batch:
            return None
        alts = np.concatenate([row for _, _, row in compiled])
        with self._gate.read(), self._cmd_lock:
            ctl = self.ctl
            ctl.queries[:self.dim] = np.asarray(q, dtype=np.float32).reshape(-1)
            ctl.filt[:alts.size] = alts
            self._post(CMD_SEARCH_ANY, batch=len(compiled), k=int(k), metric=self._metric_code())
            try:
                _, rows, nv = self._execute(CMD_SEARCH_ANY)
            except RuntimeError:
                # a peer that never published leaves the arrival counters out of step: realign before the next call
                self._wait_acks()
                self._post(CMD_RESET)
                self._execute(CMD_RESET)
                self._wait_acks()
                raise
        n = int(nv[0])
        vecs = self._fetch_vectors([int(r) for r in rows[0, :n]])
        return rows[0], vecs, n

    def __init__(self, engine, *, dictionary: Optional[MetadataDictionary] = None, embedding: Any = None,
                 id_fn: Optional[Callable[[int], str]] = None, meta_fn: Optional[Callable[[int], Dict[str, Any]]] = None,
                 text_fn: Optional[Callable[[int], str]] = None, metric: str = "cosine", table_name: str = "embeddings",
                 scope: Optional[str] = None, max_batch: int = 1024, payload_mb: int = 64, group=None,
                 multi_valued_keys: Optional[Sequence[str]] = None, timeout_s: float = 120.0):
        import torch.distributed as dist
        self.engine = engine
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        st = engine.state()
        keys = dictionary.keys if dictionary is not None else None
        super().__init__(embedding, dim=st["dim"], filter_keys=keys, table_name=table_name, metric=metric, scope=scope,
                         multi_valued_keys=multi_valued_keys)
        if dictionary is not None:
            self.dictionary = dictionary
        if self.dictionary.n_keys != st["n_keys"] and st["n_keys"] != 0:
            raise ValueError(f"the shard has {st['n_keys']} code columns but the dictionary {self.dictionary.n_keys} keys")
        self.timeout_s = timeout_s
        self._cmd_lock = threading.Lock()
        self._rows = _ShardedRows()
        self._rows.id_fn, self._rows.meta_fn, self._rows.text_fn = id_fn or (lambda r: f"row-{r:09d}"), meta_fn, text_fn
        self._mask_ids: Dict[Any, int] = {}
        self._dead_bulk: set = set()          # deleted rows of the bulk-loaded (id_fn) part
        self._closed = False
        reply = max(1 << 16, N.MAX_TOPK * st["dim"] * 4 + 4096)
        # control block: created by rank 0, attached by name everywhere else
        if self.rank == 0:
            self.ctl = ControlBlock(self.world, st["dim"], max_batch, payload_mb << 20, reply)
            names = [self.ctl.name]
        else:
            names = [None]
        if self.world > 1:
            dist.broadcast_object_list(names, src=0, group=group)
        if self.rank != 0:
            self.ctl = ControlBlock(self.world, st["dim"], max_batch, payload_mb << 20, reply, name=names[0])
        import ctypes as C
        self._seq_c = C.c_uint64(0)          # sequence number of the last command (shared with the C serving path)
        if hasattr(engine, "register_host"):
            engine.register_host(self.ctl.shm.buf)
        if hasattr(engine, "bind_control"):
            engine.bind_control(self.ctl)
        self._ack_peers = self.ctl.ack[1:self.world]
        self._status_peers = self.ctl.status[1:self.world]
        # every rank's row range
        states = [st]
        if self.world > 1:
            states = [None] * self.world
            dist.all_gather_object(states, st, group=group)
        self._ranges = [[s["row_offset"], s["n_rows"], s["limit"]] for s in states]     # offset, rows, limit
        for r in range(self.world - 1):       # a rank may grow up to the next rank's first row
            self._ranges[r][2] = min(self._ranges[r][2], self._ranges[r + 1][0])
        self._rows.n_bulk = sum(s["n_rows"] for s in states)
        if self.world > 1:
            dist.barrier(group)

    @property
    def _seq(self) -> int:
        return int(self._seq_c.value)

    @_seq.setter
    def _seq(self, v: int) -> None:
        self._seq_c.value = v

    # ---- things the single-GPU store does with its own shard --------------------------------------
    def _ensure_built(self):
        return None

    def __len__(self) -> int:
        return len(self._rows) - self._rows.n_dead

    def compact(self, cluster_by=None) -> None:
        raise NotImplementedError("compaction of a sharded store is not implemented (tombstoned rows are never returned)")

    def save(self, path: str) -> None:
        raise NotImplementedError("persist each rank's shard with B200VectorStore.save; the sharded store holds no file format")

    # ---- command channel ----------------------------------------------------------------------------
    def _post(self, cmd: int, **hdr: int) -> int:
        """Rank 0: wait until every rank finished the previous command, then publish this one."""
        ctl = self.ctl
        seq = self._seq
        if self.world > 1 and not (self._ack_peers == seq).all():
            _spin_until(lambda: bool((ctl.ack[1:self.world] == seq).all()), self.timeout_s, "the ranks to finish the previous command")
        if self.world > 1 and self._status_peers.any():
            bad = [r for r in range(1, self.world) if ctl.status[r] != 0]
            err = ctl.get_reply(bad[0])
            ctl.status[bad[0]] = 0
            raise RuntimeError(f"sharded store: rank {bad[0]} failed the previous command: {err}")
        g = hdr.get
        ctl.hdr[:9] = (cmd, g("batch", 0), g("k", 0), g("metric", 0), g("algo", 0), g("has_filt", 0), g("has_masks", 0),
                       g("payload", 0), g("aux", 0))
        self._seq = seq + 1
        ctl.seq[0] = self._seq
        return self._seq

    def _wait_acks(self) -> None:
        if self.world > 1:
            ctl, seq = self.ctl, self._seq
            _spin_until(lambda: bool((ctl.ack[1:self.world] == seq).all()), self.timeout_s, "the ranks to acknowledge a command")
            bad = [r for r in range(1, self.world) if ctl.status[r] != 0]
            if bad:
                err = ctl.get_reply(bad[0])
                ctl.status[bad[0]] = 0
                raise RuntimeError(f"sharded store: rank {bad[0]} failed: {err}")

    def serve_forever(self) -> None:
        """Ranks 1..G-1: execute rank 0's commands until it shuts the store down (``close()``) or hands the
        process back (``release_peers()``: this call returns and may be entered again later)."""
        if self.rank == 0:
            raise RuntimeError("rank 0 owns the store faces; serve_forever() is for the other ranks")
        ctl, last = self.ctl, self._seq
        fast = hasattr(self.engine, "serve_next")
        while True:
            failed = None
            if fast:
                # plain searches are served inside this call; it returns for every other command
                if self.engine.serve_next(self._seq_c) < 0:
                    failed = N.load().crb_last_error().decode("utf-8", "replace")
                last = self._seq
            else:
                _spin_until(lambda: int(ctl.seq[0]) != last, float("inf"), "a command")
                last = int(ctl.seq[0])
            cmd = int(ctl.hdr[0])
            try:
                if failed is not None:
                    raise RuntimeError(failed)
                self._execute(cmd)
            except Exception as e:  # noqa: BLE001 - reported to rank 0 with the next command it posts
                try:
                    ctl.put_reply(self.rank, f"{type(e).__name__}: {e}")
                except Exception:  # noqa: BLE001
                    pass
                ctl.status[self.rank] = 1
            ctl.ack[self.rank] = last
            self._seq = last
            if cmd == CMD_RELEASE:
                return
            if cmd == CMD_SHUTDOWN:
                break
        self._closed = True
        self._ack_peers = self._status_peers = None
        if hasattr(self.engine, "unregister_host"):
            self.engine.unregister_host()
        self.engine.close()
        self.ctl.close()

    def _execute(self, cmd: int):
        """One command on this rank (rank 0 runs the same code for its own share)."""
        ctl, eng = self.ctl, self.engine
        h = ctl.hdr[:9].tolist()
        if cmd == CMD_SEARCH:
            b, k = h[1], h[2]
            q = ctl.queries[:b * self.dim].reshape(b, self.dim)
            nk = self.dictionary.n_keys
            filt = ctl.filt[:b * nk].reshape(b, nk) if h[5] else None
            mids = ctl.mask_ids[:b] if h[6] else None
            return eng.search(q, k, h[3], filt, mids, h[4], 0)
        if cmd == CMD_SEARCH_ANY:
            # one query (the first of the query area), `batch` alternative filter rows: a row passes when any does
            n_alt, k = h[1], h[2]
            nk = self.dictionary.n_keys
            return eng.search_any_of(ctl.queries[:self.dim], ctl.filt[:n_alt * nk].reshape(n_alt, nk), k, h[3], 0)
        if cmd == CMD_SET_MASK:
            trees = ctl.get_payload(int(h[7]))
            eng.set_mask(int(h[8]), trees[self.rank] if isinstance(trees, dict) else trees)
        elif cmd == CMD_DROP_MASKS:
            eng.drop_masks()
        elif cmd == CMD_FETCH:
            rows = ctl.get_payload(int(h[7]))
            lo, n, _ = self._ranges[self.rank]
            mine = [(i, g) for i, g in enumerate(rows) if lo <= g < lo + eng.state()["n_rows"]]
            vec = eng.fetch([g for _, g in mine]) if mine else np.empty((0, self.dim), np.float32)
            ctl.put_reply(self.rank, ([i for i, _ in mine], vec))
        elif cmd == CMD_APPEND:
            target, rows16, codes = ctl.get_payload(int(h[7]))
            if target == self.rank:
                ctl.put_reply(self.rank, eng.append(rows16, codes))
        elif cmd == CMD_TOMBSTONE:
            rows = ctl.get_payload(int(h[7]))
            st = eng.state()
            mine = [g for g in rows if st["row_offset"] <= g < st["row_offset"] + st["n_rows"]]
            if mine:
                eng.tombstone(mine)
        elif cmd == CMD_RESET:
            eng.reset()
        elif cmd == CMD_STATE:
            ctl.put_reply(self.rank, eng.state())
        elif cmd in (CMD_SHUTDOWN, CMD_RELEASE):
            pass
        else:
            raise RuntimeError(f"unknown command {cmd}")
        return None

    def _run(self, cmd: int, payload: Any = None, aux: int = 0, wait: bool = True):
        """Rank 0: post a command, do this rank's share, wait for the others."""
        with self._cmd_lock:
            n = self.ctl.put_payload(payload) if payload is not None else 0
            self._post(cmd, payload=n, aux=aux)
            self._execute(cmd)
            if wait:
                self._wait_acks()

    def release_peers(self) -> None:
        """Rank 0: make the other ranks return from :meth:`serve_forever` (they can enter it again)."""
        self._run(CMD_RELEASE)

    def reset_exchange(self) -> None:
        """After a failed search (a rank lost or stalled): bring every rank's exchange counters back in step."""
        self._run(CMD_RESET)

    def close(self) -> None:
        if self._closed:
            return
        self._closed = True
        if self.rank == 0:
            try:
                self._run(CMD_SHUTDOWN)
            finally:
                self._ack_peers = self._status_peers = None
                if hasattr(self.engine, "unregister_host"):
                    self.engine.unregister_host()
                self.engine.close()
                self.ctl.close()

    # ---- the hot path ---------------------------------------------------------------------------------
    def _search_locked(self, queries, k, filt, split, algo, vectors=False):
        # (vectors: the base class fetches them from the owning ranks after the merge)
        q = np.ascontiguousarray(queries, dtype=np.float32)
        b = q.shape[0]
        if b > self.ctl.max_batch:
            # larger batches are served in slices of the control block's capacity
            parts = [self._search_locked(q[i:i + self.ctl.max_batch], k, None if filt is None else filt[i:i + self.ctl.max_batch],
                                         split[i:i + self.ctl.max_batch], algo) for i in range(0, b, self.ctl.max_batch)]
            return tuple(np.concatenate([p[j] for p in parts]) for j in range(3))
        mids = None
        if any(extra for _, extra in split):
            mids = np.asarray([self._mask_id(extra) if extra else -1 for _, extra in split], dtype=np.int32)
        if mids is None and hasattr(self.engine, "post_search"):
            f32 = None if filt is None else np.ascontiguousarray(filt, dtype=np.int32)
            with self._gate.read(), self._cmd_lock:
                try:
                    out = self.engine.post_search(q, k, self._metric_code(), f32, algo, self._seq_c, self.timeout_s)
                    if isinstance(out, int):         # a rank failed the previous command: surface its message
                        bad = [r for r in range(1, self.world) if self.ctl.status[r] != 0]
                        err = self.ctl.get_reply(bad[0]) if bad else "unknown failure"
                        for r in bad:
                            self.ctl.status[r] = 0
                        raise RuntimeError(f"sharded store: rank {bad[0] if bad else '?'} failed the previous command: {err}")
                    return out
                except RuntimeError:
                    self._wait_acks()
                    self._post(CMD_RESET)
                    self._execute(CMD_RESET)
                    self._wait_acks()
                    raise
        with self._gate.read(), self._cmd_lock:
            ctl = self.ctl
            ctl.queries[:b * self.dim] = q.reshape(-1)
            if filt is not None:
                ctl.filt[:filt.size] = np.ascontiguousarray(filt, dtype=np.int32).reshape(-1)
            if mids is not None:
                ctl.mask_ids[:b] = mids
            self._post(CMD_SEARCH, batch=b, k=k, metric=self._metric_code(), algo=algo, has_filt=int(filt is not None),
                       has_masks=int(mids is not None))
            try:
                return self._execute(CMD_SEARCH)
            except RuntimeError:
                # a peer that never published leaves the arrival counters out of step: realign before the next call
                self._wait_acks()
                self._post(CMD_RESET)
                self._execute(CMD_RESET)
                self._wait_acks()
                raise

    def search_rows(self, queries, k: int, filters=None):
        """Array-level search (see :meth:`B200VectorStore.search_rows`); unfiltered requests skip the filter plumbing."""
        if filters is None or all(f is None for f in filters):
            q = np.ascontiguousarray(queries, dtype=np.float32)
            if q.ndim == 1:
                q = q[None, :]
            k = int(k)
            if not 1 <= k <= N.MAX_TOPK:
                raise ValueError(f"k must be in [1, {N.MAX_TOPK}]")
            return self._search_locked(q, k, None, (), 0)
        return super().search_rows(queries, k, filters)

    def _metric_code(self) -> int:
        return N.METRIC_COSINE if self.metric in ("cosine", "default", "cos", N.METRIC_COSINE) else N.METRIC_IP

    def _fetch_vectors(self, rows: Sequence[int]) -> np.ndarray:
        rows = [int(r) for r in rows]
        out = np.empty((len(rows), self.dim), dtype=np.float32)
        if not rows:
            return out
        with self._cmd_lock:
            n = self.ctl.put_payload(rows)
            self._post(CMD_FETCH, payload=n)
            self._execute(CMD_FETCH)
            self._wait_acks()
            for r in range(self.world):
                idx, vec = self.ctl.get_reply(r)
                if idx:
                    out[idx] = vec
        return out

    # ---- masks: compiled on rank 0, evaluated by every rank on its own rows -----------------------------
    def _compile(self, items) -> tuple:
        """Canonical predicates -> a tree every rank can evaluate without the dictionary or the postings:
        ("codes", column, codes) | ("rows", global rows) | ("and" | "or", [trees]) | ("none",) | ("all",)."""
        parts = []
        for key, spec in items:
            if key == "__any_of__":
                for group in spec:
                    parts.append(("or", [self._compile(alt) for alt in group]) if group else ("none",))
                continue
            values = [spec[1]] if spec[0] == "eq" else list(spec[1])
            if key in self.multi_valued_keys:
                rows = sorted({r for v in values for r in self._postings.get((key, v), ())})
                parts.append(("rows", np.asarray(rows, dtype=np.int64)) if rows else ("none",))
                continue
            pos = self.dictionary._pos.get(key)
            codes = [] if pos is None else [self.dictionary.values[pos][v] for v in values if v in self.dictionary.values[pos]]
            parts.append(("codes", pos, tuple(codes)) if codes else ("none",))
        if not parts:
            return ("all",)
        return parts[0] if len(parts) == 1 else ("and", parts)

    def _mask_id(self, extra) -> int:
        with self._mask_lock:
            mid = self._mask_ids.get(extra)
        if mid is not None:
            return mid
        tree = self._compile(extra)
        with self._mask_lock:
            mid = len(self._mask_ids)
            self._mask_ids[extra] = mid
        self._run(CMD_SET_MASK, payload=tree, aux=mid)
        return mid

    def _row_mask(self, extra):          # the single-GPU store's hook; the sharded store ships mask ids instead
        raise RuntimeError("internal: the sharded store resolves row masks by id")

    def _invalidate_masks(self) -> None:
        with self._mask_lock:
            had = bool(self._mask_ids)
            self._mask_ids = {}
        if had:
            self._run(CMD_DROP_MASKS)

    # ---- writes -------------------------------------------------------------------------------------------
    def _append(self, ids, embeddings, texts, metadatas) -> List[str]:
        emb = np.asarray(embeddings, dtype=np.float32)
        if emb.ndim != 2 or len(ids) != emb.shape[0] or emb.shape[1] != self.dim:
            raise ValueError(f"embeddings must be [len(ids), {self.dim}]")
        if self.rank != 0:
            raise RuntimeError("writes go through rank 0")
        with self._gate.write():
            last = {str(rid): i for i, rid in enumerate(ids)}          # duplicates inside the batch: last wins
            order = sorted(last.values())
            r = self._rows
            dead = []
            for i in order:
                old = r.pos.get(str(ids[i]))
                if old is not None:
                    dead.append(old)
            # the rank with room that holds the fewest rows takes the batch
            room = [(self._ranges[k][1], k) for k in range(self.world)
                    if self._ranges[k][0] + self._ranges[k][1] + len(order) <= self._ranges[k][2]]
            if not room:
                raise RuntimeError("no rank has room for the new rows")
            target = min(room)[1]
            metas = [{str(a): v for a, v in dict(metadatas[i] or {}).items() if v is not None} for i in order]
            codes = np.asarray([self.dictionary.encode_row(md) for md in metas], dtype=np.int32).T.copy() \
                if metas else np.empty((self.dictionary.n_keys, 0), np.int32)
            if self.dictionary.n_keys == 0:
                codes = None
            self._invalidate_masks()
            with self._cmd_lock:
                n = self.ctl.put_payload((target, emb[order].astype(np.float16), codes))
                self._post(CMD_APPEND, payload=n)
                self._execute(CMD_APPEND)
                self._wait_acks()
                first = int(self.ctl.get_reply(target))
            self._ranges[target][1] += len(order)
            for j, i in enumerate(order):
                rid, g = str(ids[i]), first + j
                r.host_of[g] = len(r.ids)
                r.ids.append(rid)
                r.texts.append(texts[i] if texts is not None else "")
                r.metas.append(metas[j])
                r.pos[rid] = g
                self._post_row(g, metas[j])
            if dead:
                self._tombstone(dead)
        return [str(i) for i in ids]

    def _post_row(self, row: int, md: Mapping[str, Any]) -> None:
        for key in self.multi_valued_keys:
            if key in md:
                for val in self._members(md[key]):
                    self._postings.setdefault((key, val), []).append(row)

    def _tombstone(self, rows: Sequence[int]) -> None:
        r = self._rows
        for g in rows:
            h = r.host_of.pop(g, None)
            if h is not None:
                r.ids[h] = None
        self._invalidate_masks()
        self._run(CMD_TOMBSTONE, payload=[int(g) for g in rows])

    def add_embeddings(self, ids, embeddings, texts=None, metadatas=None) -> List[str]:
        n = len(ids)
        return self._append(ids, embeddings, texts or [""] * n, metadatas or [{}] * n)

    def delete(self, ref_doc_id: str, **kwargs: Any) -> None:
        with self._gate.write():
            g = self._rows.pos.pop(str(ref_doc_id), None)
            if g is None:
                g = self._row_of(ref_doc_id)        # a bulk-loaded row
                if g is None:
                    return
                self._rows.n_dead += 1
                self._dead_bulk.add(g)
            self._tombstone([g])

    def _row_of(self, rid: str) -> Optional[int]:
        g = self._rows.pos.get(str(rid))
        if g is not None:
            return g
        try:
            g = int(str(rid).rsplit("-", 1)[-1])
        except ValueError:
            return None
        for lo, n, _ in self._ranges:
            if lo <= g < lo + n and g not in self._rows.host_of and g not in self._dead_bulk \
                    and self._rows.id_fn(g) == str(rid):
                return g
        return None

    def get(self, ids: Sequence[str], filter: Optional[Dict[str, Any]] = None, **kw):
        rows = []
        for rid in dict.fromkeys(ids):
            r = self._row_of(rid)
            if r is None:
                continue
            if filter and any(str(self._rows.meta(r).get(a)) != str(v) for a, v in filter.items()):
                continue
            rows.append(r)
        vecs = self._fetch_vectors(rows)
        return [self._content(r, vecs[i]) for i, r in enumerate(rows)]
