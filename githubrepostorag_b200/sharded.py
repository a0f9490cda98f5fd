"""Row-sharded search over the GPUs of one box: one process per GPU, one exchange step.

The reference runs a single worker against a single Cassandra node (``replicas: 1``,
``helm/values.yaml:1,113``); the north star shards the corpus rows over 8 GPUs instead.
Rank ``r`` of ``G`` owns rows ``[r*ceil(N/G), (r+1)*ceil(N/G))``.  Queries are
replicated (broadcast from rank 0, ``B*d*4`` bytes); every rank searches its shard and emits
``k`` exact ``(score, global row)`` keys per query; ONE all-gather of the packed ``[B, k]``
uint64 keys (``B*k*8`` bytes per rank -- latency-bound on NVLink/NVSwitch) is followed by a
``G*k -> k`` merge on every rank.  Because keys are exact and globally ordered, the sharded
result equals the unsharded one bit for bit.

The local search and the merge are injected so the rendezvous / layout logic is testable on
CPU with ``gloo`` (tests/test_sharded_gloo.py); :func:`cuda_searcher` wires the CUDA path.
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import torch
import torch.distributed as dist


def shard_range(n_rows: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous row range of ``rank``: ``ceil(N/G)`` rows each, the last one short."""
    per = -(-n_rows // world)
    lo = min(n_rows, rank * per)
    return lo, min(n_rows, lo + per)


class ShardedSearcher:
    """Collective search: every rank calls :meth:`search` with the same ``batch``/``k``.

    ``local_search(queries, k) -> keys int64 [batch, k]`` searches this rank's shard and returns
    order-preserving keys (0 = empty slot); ``merge(gathered [G, batch, k], queries, k)`` returns
    ``(scores, rows, n_valid, keys)``.
    """

    def __init__(self, local_search: Callable[[torch.Tensor, int], torch.Tensor],
                 merge: Callable[[torch.Tensor, torch.Tensor, int], tuple],
                 group: Optional[dist.ProcessGroup] = None):
        self.local_search = local_search
        self.merge = merge
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0

    def search(self, queries: torch.Tensor, k: int, *, broadcast: bool = False):
        """``queries`` fp32 ``[batch, dim]`` on this rank's device.  With ``broadcast`` the
        contents of rank 0's tensor replace everyone else's first (the serving flow: only rank 0
        received the request)."""
        if self.world > 1 and broadcast:
            dist.broadcast(queries, src=0, group=self.group)
        keys = self.local_search(queries, k)
        if self.world == 1:
            gathered = keys.unsqueeze(0)
        else:
            b, kk = keys.shape
            flat = torch.empty((self.world * b, kk), dtype=keys.dtype, device=keys.device)
            dist.all_gather_into_tensor(flat, keys.contiguous(), group=self.group)
            gathered = flat.view(self.world, b, kk)
        return self.merge(gathered, queries, k)


def cuda_searcher(shard, *, metric: str = "cosine", algo: int = 0,
                  group: Optional[dist.ProcessGroup] = None) -> ShardedSearcher:
    """The CUDA wiring: ``DeviceShard.search_device`` for the local search (keys out),
    ``crb_merge_keys`` for the merge, NCCL for the all-gather."""
    from .index import merge_keys

    state = {}

    def local_search(q: torch.Tensor, k: int) -> torch.Tensor:
        b = q.shape[0]
        buf = state.get((b, k))
        if buf is None:
            buf = (torch.empty((b, k), dtype=torch.float32, device=q.device),
                   torch.empty((b, k), dtype=torch.int64, device=q.device),
                   torch.empty((b,), dtype=torch.int32, device=q.device),
                   torch.empty((b, k), dtype=torch.int64, device=q.device))
            state[(b, k)] = buf
        shard.search_device(q, k, metric=metric, algo=algo, out=buf[:3], keys_out=buf[3])
        return buf[3]

    def merge(gathered: torch.Tensor, q: torch.Tensor, k: int):
        return merge_keys(gathered, q, k, metric)

    return ShardedSearcher(local_search, merge, group)


class NVLinkExchange:
    """The peer-memory side of a row-sharded search (``crb_exchange`` of include/coderag_b200.h): one
    symmetric-memory buffer per rank (``torch.distributed._symmetric_memory``: the same allocation mapped into
    every rank of the box over NVLink) holding this rank's gather slots ``uint64 [2][world][cap_keys]`` and its
    ``uint32 [world]`` arrival counters, plus the call bookkeeping every rank must keep in step."""

    COUNTER_SLOTS = 64   # uint32 counters live in the last 256 bytes of the buffer

    def __init__(self, device, group, cap_keys: int = 1 << 17):
        import torch.distributed._symmetric_memory as symm

        from . import _native as N
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        if self.world > N.MAX_PEERS:
            raise ValueError(f"at most {N.MAX_PEERS} ranks per box")
        self.device = device
        self.cap_keys = int(cap_keys)
        n_words = 2 * self.world * self.cap_keys + self.COUNTER_SLOTS // 2
        self.buf = symm.empty(n_words, dtype=torch.int64, device=device)
        self.buf.zero_()
        torch.cuda.synchronize(device)
        self.handle = symm.rendezvous(self.buf, self.group.group_name)
        ptrs = [int(p) for p in self.handle.buffer_ptrs]
        ex = N.Exchange()
        ex.rank, ex.world, ex.cap_keys, ex.dest_mask = self.rank, self.world, self.cap_keys, 0
        for p in range(self.world):
            ex.peer_keys[p] = ptrs[p]
            ex.peer_counters[p] = ptrs[p] + 8 * 2 * self.world * self.cap_keys
        self.struct = ex
        import ctypes as C
        self.state = (C.c_uint32 * 2)()      # {calls so far, queries published so far}; the C serving path updates it too
        dist.barrier(self.group)      # every rank's counters are zero before anyone publishes

    def next_call(self, batch: int) -> Tuple[int, int]:
        """(call_index, expected_total) of the next collective search of ``batch`` queries."""
        st = self.state
        st[1] = (st[1] + batch) & 0xFFFFFFFF
        call = st[0]
        st[0] = (call + 1) & 0xFFFFFFFF
        return call, st[1]

    def reset(self) -> None:
        """Collective: after a failed call (a rank lost or stalled, ``n_valid == -2``) the arrival counters and the
        call parity of the ranks are out of step; with no search in flight anywhere, zero them and start over."""
        torch.cuda.synchronize(self.device)
        dist.barrier(self.group)
        self.buf.zero_()
        self.state[0] = self.state[1] = 0
        torch.cuda.synchronize(self.device)
        dist.barrier(self.group)


class FusedExchangeSearcher:
    """Row-sharded search with the exchange fused into the search kernels (``crb_search_exchange``).

    Each rank allocates one symmetric-memory buffer (``torch.distributed._symmetric_memory``: the
    same allocation mapped into every rank of the box over NVLink) holding its gather slots
    ``uint64 [2][world][cap_keys]`` and its ``uint32 [world]`` arrival counters.  A call launches the
    shard search, whose result sinks store keys into every peer's slots and bump the counters, and
    one wait-and-merge kernel -- no NCCL collective on the critical path (for a
    batch-1 query the ~20 us collective is a 5-9 % tax on a 0.29 ms shard scan).

    Collective: all ranks must call :meth:`search` the same number of times with the same batch/k.
    """

    def __init__(self, shard, *, cap_keys: int = 1 << 17, metric: str = "cosine", algo: int = 0,
                 group: Optional[dist.ProcessGroup] = None, dest_rank: Optional[int] = None):
        import ctypes as C

        from . import _native as N
        from .index import DeviceShard
        self._N, self._C = N, C
        self.shard = shard
        self.metric = DeviceShard._metric(metric)
        self.algo = algo
        self.exchange = NVLinkExchange(shard.device, group, cap_keys)
        # dest_rank: only that rank receives the keys and merges (the serving layout: one rank owns the store
        # face); None = every rank ends up with the merged result
        self.exchange.struct.dest_mask = 0 if dest_rank is None else (1 << dest_rank)
        self.group, self.world, self.rank = self.exchange.group, self.exchange.world, self.exchange.rank
        self.cap_keys = self.exchange.cap_keys
        self._ws = None

    def search(self, queries: torch.Tensor, k: int, *, broadcast: bool = False, filt=None):
        N, C = self._N, self._C
        if self.world > 1 and broadcast:
            dist.broadcast(queries, src=0, group=self.group)
        b, d = queries.shape
        if b * k > self.cap_keys:
            raise ValueError(f"batch*k = {b * k} exceeds cap_keys = {self.cap_keys}")
        dev = self.shard.device
        # one allocation for the three results (rows first: 8-byte aligned), so a caller that wants them on the
        # host moves ``last_packed`` with a single copy
        packed = torch.empty((b * k * 12 + b * 4,), dtype=torch.uint8, device=dev)
        rows = packed[: b * k * 8].view(torch.int64).view(b, k)
        scores = packed[b * k * 8: b * k * 12].view(torch.float32).view(b, k)
        n_valid = packed[b * k * 12:].view(torch.int32)
        self.last_packed = packed
        f = self.shard._filt(filt, b)
        nf = 0 if f is None else len({tuple(r) for r in f.tolist() if any(c != N.FILT_ANY for c in r)})
        need = self.shard.workspace_bytes(b, k, nf)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.zeros(need, dtype=torch.uint8, device=dev)
        call, total = self.exchange.next_call(b)
        lib = N.load()
        N.check(lib.crb_search_exchange(
            self.shard._h, queries.data_ptr(), b, k, self.metric, None if f is None else f.ctypes.data_as(C.c_void_p),
            scores.data_ptr(), rows.data_ptr(), n_valid.data_ptr(), None, self._ws.data_ptr(), self._ws.numel(),
            self.algo, torch.cuda.current_stream(dev).cuda_stream, C.byref(self.exchange.struct), call, total))
        return scores, rows, n_valid, None

    @staticmethod
    def check(n_valid: torch.Tensor) -> None:
        """Raise if a query of a finished call timed out waiting for a peer (``n_valid == -2``).  Synchronises; call
        it where the result is consumed -- a ``range(n_valid)`` over -2 would read as an empty, successful answer.
        Afterwards :meth:`reset` (collective) brings the ranks back in step."""
        if bool((n_valid == -2).any()):
            raise RuntimeError("sharded search: a peer did not publish before the wait ran out (lost or stalled rank); "
                               "call reset() on every rank before searching again")

    def reset(self) -> None:
        self.exchange.reset()
