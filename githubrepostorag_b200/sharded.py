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
