"""Device-resident corpus shard and the host half of the search call.

A :class:`DeviceShard` is what one ``LCCassandra(table_name=...)`` store stands for in the
reference (``rag_worker/src/worker/services/graph_rag_retrievers.py:68-80``): the rows of one
embeddings table (``helm/templates/cassandra-initdb-configmap.yaml:86-102``) laid out in HBM as

* ``rows``      fp16 ``[n_rows, dim]`` row-major (the north star's corpus dtype),
* ``inv_norm``  fp32 ``[n_rows]``,
* ``codes``     one column per filterable ``metadata_s`` key: dictionary codes ``[n_rows]``, int8 / int16 / int32 --
  as narrow as the key's dictionary allows, because the filter kernel is bound by these bytes (an int32
  ``[n_keys, n_rows]`` matrix is accepted too).

torch owns the device memory and streams; every computation goes through the C ABI
(``include/coderag_b200.h``) by raw pointer.  There is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import threading
from typing import Dict, Iterable, List, Mapping, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _native as N

UPLOAD_BLOCK_ROWS = 1 << 18


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _np_ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _addr(a: np.ndarray) -> int:
    """Address of a WRITABLE array's first byte, without building ``ndarray.ctypes`` (several microseconds)."""
    return C.addressof(C.c_char.from_buffer(a)) if a.size else 0


class MetadataDictionary:
    """Dictionary encoding of the filterable metadata keys (one int32 code per key per row).

    The reference stores every metadata value as text in ``metadata_s MAP<TEXT,TEXT>``
    (``ingest/src/app/services/vector_write_service.py:45-98``) and filters by equality on the
    allow-listed keys (``:28-34,143``).  Values are coerced with ``str()`` exactly as the writer
    does; a row lacking a key gets ``CODE_MISSING`` and can never match a filter on it.
    """

    def __init__(self, keys: Sequence[str]):
        if len(keys) > N.MAX_KEYS:
            raise ValueError(f"at most {N.MAX_KEYS} filterable keys (got {len(keys)})")
        self.keys: List[str] = list(keys)
        self._pos = {k: i for i, k in enumerate(self.keys)}
        self.values: List[Dict[str, int]] = [dict() for _ in self.keys]

    @property
    def n_keys(self) -> int:
        return len(self.keys)

    def encode_row(self, metadata: Mapping[str, object]) -> List[int]:
        out = []
        for i, k in enumerate(self.keys):
            v = metadata.get(k)
            if v is None:
                out.append(N.CODE_MISSING)
                continue
            table = self.values[i]
            out.append(table.setdefault(str(v), len(table)))
        return out

    def encode_rows(self, metadatas: Sequence[Mapping[str, object]]) -> np.ndarray:
        codes = np.empty((self.n_keys, len(metadatas)), dtype=np.int32)
        for r, md in enumerate(metadatas):
            codes[:, r] = self.encode_row(md)
        return codes

    def register_vocab(self, key: str, values: Sequence[str]) -> None:
        """Pre-register values so code ``i`` means ``values[i]`` (synthetic corpora)."""
        table = self.values[self._pos[key]]
        for v in values:
            table.setdefault(str(v), len(table))

    def encode_filter(self, flt: Optional[Mapping[str, object]]) -> np.ndarray:
        """``{key: value}`` (AND of equalities, the only shape the reference emits:
        ``agent_graph.py:198-239,357-364,532-534``) -> int32 ``[n_keys]``.  An unknown key or an
        unknown value can match no row, as with Cassandra's ``metadata_s['k']='v'``."""
        row = np.full(self.n_keys, N.FILT_ANY, dtype=np.int32)
        if not flt:
            return row
        for k, v in flt.items():
            i = self._pos.get(k)
            if i is None or v is None:
                row[:] = N.FILT_IMPOSSIBLE
                return row
            code = self.values[i].get(str(v))
            row[i] = N.FILT_IMPOSSIBLE if code is None else code
        return row


class DeviceShard:
    """One row-shard of a corpus resident on one GPU.

    The buffers may be over-allocated (``capacity >= n_rows``) so rows can be appended without
    re-uploading the shard (:meth:`append`); deleted rows are tombstoned in place
    (:meth:`tombstone`, inverse norm := NaN) and never returned.
    """

    @staticmethod
    def _width_for(max_code: int) -> torch.dtype:
        """Narrowest signed type that holds codes ``-1 .. max_code`` (-1 = key absent)."""
        return torch.int8 if max_code <= 127 else (torch.int16 if max_code <= 32767 else torch.int32)

    def __init__(self, rows: torch.Tensor, inv_norm: torch.Tensor, codes, row_offset: int = 0,
                 n_rows: Optional[int] = None):
        lib = N.load()
        if not rows.is_cuda or rows.dtype != torch.float16 or rows.dim() != 2 or not rows.is_contiguous():
            raise ValueError("rows must be a contiguous fp16 CUDA tensor [capacity, dim]")
        cap, dim = rows.shape
        n_rows = cap if n_rows is None else int(n_rows)
        if not 0 <= n_rows <= cap:
            raise ValueError("n_rows must be in [0, capacity]")
        if inv_norm.dtype != torch.float32 or inv_norm.numel() != cap or not inv_norm.is_cuda:
            raise ValueError("inv_norm must be fp32 CUDA [capacity]")
        n_keys = 0
        self._cols: Optional[List[torch.Tensor]] = None        # per-key columns (mixed widths) ...
        if isinstance(codes, (list, tuple)):
            cols = list(codes)
            for c in cols:
                if c.dim() != 1 or c.numel() != cap or not c.is_cuda or not c.is_contiguous() or \
                        c.dtype not in (torch.int8, torch.int16, torch.int32):
                    raise ValueError("code columns must be contiguous int8/int16/int32 CUDA tensors [capacity]")
            n_keys = len(cols)
            self._cols = cols if cols else None
            codes = None
        elif codes is not None:                                 # ... or one int32 matrix
            if codes.dtype != torch.int32 or codes.dim() != 2 or codes.shape[1] != cap or not codes.is_contiguous():
                raise ValueError("codes must be a contiguous int32 CUDA tensor [n_keys, capacity]")
            n_keys = codes.shape[0]
        self._rows_buf, self._inv_buf, self._codes_buf = rows, inv_norm, codes
        self.n_rows, self.dim, self.n_keys = n_rows, int(dim), int(n_keys)
        self.row_offset = int(row_offset)
        self.device = rows.device
        handle = C.c_void_p()
        if self._cols is not None:
            # created empty, then bound to the columns
            N.check(lib.crb_index_create(C.byref(handle), _ptr(rows), _ptr(inv_norm), None, 0, self.dim, self.n_keys,
                                         self.row_offset, self.device.index or 0))
        else:
            N.check(lib.crb_index_create(C.byref(handle), _ptr(rows), _ptr(inv_norm), _ptr(codes),
                                         self.n_rows, self.dim, self.n_keys, self.row_offset,
                                         self.device.index or 0))
        self._h = handle
        self._lib = lib
        if self._cols is not None or cap != n_rows:
            self._bind()
        # scratch for search_device calls that bring no workspace: one per calling thread, because the worker's
        # threads search concurrently (rag_worker/src/worker/worker.py:136-139,185) and a workspace serves one
        # call at a time -- counters, survivor lists and candidate lists all live in it
        self._ws_local = threading.local()

    @property
    def capacity(self) -> int:
        return int(self._rows_buf.shape[0])

    @property
    def rows(self) -> torch.Tensor:
        return self._rows_buf[:self.n_rows]

    @property
    def inv_norm(self) -> torch.Tensor:
        return self._inv_buf[:self.n_rows]

    def _bind(self) -> None:
        """Point the index at the current buffers and row count."""
        if self._cols is not None:
            ptrs = (C.c_void_p * self.n_keys)(*[c.data_ptr() for c in self._cols])
            widths = (C.c_int32 * self.n_keys)(*[c.element_size() for c in self._cols])
            N.check(self._lib.crb_index_rebind_columns(self._h, _ptr(self._rows_buf), _ptr(self._inv_buf), ptrs, widths,
                                                       self.n_rows))
        else:
            N.check(self._lib.crb_index_rebind(self._h, _ptr(self._rows_buf), _ptr(self._inv_buf), _ptr(self._codes_buf),
                                               self.n_rows, self.capacity))

    @property
    def codes(self):
        """Key m's codes are ``codes[m]`` (``[n_rows]``): a list of columns, or rows of the int32 matrix."""
        if self._cols is not None:
            return [c[:self.n_rows] for c in self._cols]
        return None if self._codes_buf is None else self._codes_buf[:, :self.n_rows]

    @property
    def code_widths(self) -> List[int]:
        """Bytes per row of each key's column."""
        if self._cols is not None:
            return [c.element_size() for c in self._cols]
        return [4] * self.n_keys

    # ---- construction ------------------------------------------------------------------
    @classmethod
    def from_host(cls, embeddings: np.ndarray, codes: Optional[np.ndarray] = None, *,
                  device: int | torch.device = 0, row_offset: int = 0, capacity: Optional[int] = None,
                  narrow: bool = True) -> "DeviceShard":
        """Upload host rows.  fp32 input is rounded to fp16 on the device (``crb_encode_rows``);
        fp16 input is copied as is and its inverse norms are computed on the device.  ``codes`` is int32
        ``[n_keys, n_rows]`` on the host; with ``narrow`` each key gets its own column of the narrowest type
        its codes fit (widened on a later append if the dictionary outgrows it)."""
        dev = torch.device("cuda", device) if isinstance(device, int) else device
        emb = np.ascontiguousarray(embeddings)
        if emb.ndim != 2:
            raise ValueError("embeddings must be [n_rows, dim]")
        n, d = emb.shape
        cap = max(n, int(capacity or 0))
        rows = torch.empty((cap, d), dtype=torch.float16, device=dev)
        inv = torch.empty((cap,), dtype=torch.float32, device=dev)
        codes_t = None
        if codes is not None and codes.shape[0] > 0:
            if narrow:
                codes_t = [torch.full((cap,), N.CODE_MISSING, device=dev,
                                      dtype=cls._width_for(int(codes[m].max()) if codes.shape[1] else 0))
                           for m in range(codes.shape[0])]
            else:
                codes_t = torch.full((codes.shape[0], cap), N.CODE_MISSING, dtype=torch.int32, device=dev)
        cls._upload(rows, inv, codes_t, 0, emb, codes)
        return cls(rows, inv, codes_t, row_offset, n_rows=n)

    @staticmethod
    def _upload(rows: torch.Tensor, inv: torch.Tensor, codes_t, at: int,
                emb: np.ndarray, codes: Optional[np.ndarray]) -> None:
        """Encode ``emb`` into ``rows[at:at+m]`` / ``inv[at:at+m]`` and copy its code columns."""
        lib = N.load()
        dev = rows.device
        m, d = emb.shape
        stream = torch.cuda.current_stream(dev).cuda_stream
        for lo in range(0, m, UPLOAD_BLOCK_ROWS):
            hi = min(m, lo + UPLOAD_BLOCK_ROWS)
            blk = torch.from_numpy(emb[lo:hi])
            if emb.dtype == np.float16:
                rows[at + lo:at + hi].copy_(blk, non_blocking=False)
            elif emb.dtype == np.float32:
                stage = blk.to(dev)
                N.check(lib.crb_encode_rows(stage.data_ptr(), rows[at + lo:at + hi].data_ptr(),
                                            inv[at + lo:at + hi].data_ptr(), hi - lo, d, dev.index or 0, stream))
                torch.cuda.current_stream(dev).synchronize()
            else:
                raise ValueError("embeddings must be float16 or float32")
        if emb.dtype == np.float16 and m > 0:
            N.check(lib.crb_row_inv_norms(rows[at:at + m].data_ptr(), inv[at:at + m].data_ptr(), m, d,
                                          dev.index or 0, stream))
        if codes_t is not None and m > 0:
            nk = len(codes_t)
            if codes is None or codes.shape != (nk, m):
                raise ValueError(f"codes must be [{nk}, {m}]")
            host = torch.from_numpy(np.ascontiguousarray(codes, dtype=np.int32))
            if isinstance(codes_t, list):
                for k_, col in enumerate(codes_t):
                    col[at:at + m].copy_(host[k_].to(col.dtype))
            else:
                codes_t[:, at:at + m].copy_(host)
        torch.cuda.current_stream(dev).synchronize()

    @classmethod
    def from_device(cls, rows: torch.Tensor, codes: Optional[torch.Tensor] = None, *,
                    row_offset: int = 0) -> "DeviceShard":
        """Wrap fp16 rows already in HBM; computes the inverse norms on the device.  ``codes``: an int32 matrix
        ``[n_keys, n_rows]`` or a list of per-key columns (int8 / int16 / int32)."""
        lib = N.load()
        n, d = rows.shape
        inv = torch.empty((n,), dtype=torch.float32, device=rows.device)
        if n > 0:
            N.check(lib.crb_row_inv_norms(rows.data_ptr(), inv.data_ptr(), n, d, rows.device.index or 0,
                                          torch.cuda.current_stream(rows.device).cuda_stream))
        return cls(rows, inv, codes, row_offset)

    # ---- write path -----------------------------------------------------
    def append(self, embeddings: np.ndarray, codes: Optional[np.ndarray] = None) -> int:
        """Append rows in place (growing the buffers geometrically when full) and re-bind the
        index.  Returns the local index of the first new row.  No search may be in flight."""
        emb = np.ascontiguousarray(embeddings)
        if emb.ndim != 2 or emb.shape[1] != self.dim:
            raise ValueError(f"embeddings must be [m, {self.dim}]")
        m = emb.shape[0]
        first = self.n_rows
        if m == 0:
            return first
        torch.cuda.synchronize(self.device)
        if first + m > self.capacity:
            cap = max(first + m, 2 * self.capacity, 1024)
            rows = torch.empty((cap, self.dim), dtype=torch.float16, device=self.device)
            inv = torch.empty((cap,), dtype=torch.float32, device=self.device)
            rows[:first].copy_(self._rows_buf[:first])
            inv[:first].copy_(self._inv_buf[:first])
            if self._codes_buf is not None:
                codes_t = torch.full((self.n_keys, cap), N.CODE_MISSING, dtype=torch.int32, device=self.device)
                codes_t[:, :first].copy_(self._codes_buf[:, :first])
                self._codes_buf = codes_t
            if self._cols is not None:
                grown = []
                for col in self._cols:
                    g = torch.full((cap,), N.CODE_MISSING, dtype=col.dtype, device=self.device)
                    g[:first].copy_(col[:first])
                    grown.append(g)
                self._cols = grown
            self._rows_buf, self._inv_buf = rows, inv
        if self._cols is not None and codes is not None and codes.shape[1] > 0:
            # a dictionary that outgrew its column's type: widen that column (codes keep their values)
            for k_, col in enumerate(self._cols):
                need = self._width_for(int(codes[k_].max()))
                if need.itemsize > col.element_size():
                    self._cols[k_] = col.to(need)
        self._upload(self._rows_buf, self._inv_buf, self._cols if self._cols is not None else self._codes_buf, first, emb, codes)
        self.n_rows = first + m
        self._bind()
        return first

    def tombstone(self, local_rows: Sequence[int]) -> None:
        """Delete rows in place: they are never returned again (``crb_tombstone_rows``)."""
        idx = np.ascontiguousarray(local_rows, dtype=np.int64)
        if len(idx):
            N.check(self._lib.crb_tombstone_rows(self._h, _np_ptr(idx), len(idx)))

    def reload_tuning(self) -> None:
        """Re-read the ``CODERAG_B200_*`` knobs (they are otherwise read once, when the shard is created).
        No search may be in flight."""
        N.check(self._lib.crb_index_reload_tuning(self._h))

    def set_k5_stats(self, stats: Optional[torch.Tensor]) -> None:
        """Diagnostics (``crb_index_set_k5_stats``): fp32 CUDA ``[batch, 4]`` the tensor-core plan fills with
        {max |approx - exact|, margin, re-scored candidates, proven} per query; ``None`` switches it off."""
        if stats is not None and (not stats.is_cuda or stats.dtype != torch.float32 or not stats.is_contiguous()):
            raise ValueError("stats must be a contiguous fp32 CUDA tensor")
        self._k5_stats = stats
        N.check(self._lib.crb_index_set_k5_stats(self._h, _ptr(stats)))

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.crb_index_destroy(self._h)
            self._h = None

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    # ---- search -------------------------------------------------------------------------
    @staticmethod
    def _metric(metric: str | int) -> int:
        if metric in (N.METRIC_COSINE, "cosine", "default", "cos"):
            return N.METRIC_COSINE
        if metric in (N.METRIC_IP, "ip", "dot_product", "dot"):
            return N.METRIC_IP
        raise ValueError(f"unsupported metric {metric!r} (cosine and inner product only)")

    def _filt(self, filt: Optional[np.ndarray], batch: int) -> Optional[np.ndarray]:
        if filt is None or self.n_keys == 0:
            if filt is not None and self.n_keys == 0 and np.any(np.asarray(filt) != N.FILT_ANY):
                raise ValueError("filter given but the shard has no filterable keys")
            return None
        f = np.ascontiguousarray(filt, dtype=np.int32)
        if f.shape != (batch, self.n_keys):
            raise ValueError(f"filt must be [{batch}, {self.n_keys}] (got {f.shape})")
        return f

    # Small requests (the per-``invoke`` call) go through per-thread staging arrays whose addresses were taken once:
    # ``ndarray.ctypes`` costs microseconds per array, five arrays per call, against a ~10 us kernel.
    _STAGE_ELEMS = 8192

    def _host_stage(self, b: int, k: int):
        tl = self._ws_local
        cache = getattr(tl, "stage", None)
        if cache is None:
            cache = tl.stage = {}
        ent = cache.get((b, k))
        if ent is None:
            q = np.empty((b, self.dim), dtype=np.float32)
            f = np.empty((b, max(self.n_keys, 1)), dtype=np.int32)
            scores = np.empty((b, k), dtype=np.float32)
            rows = np.empty((b, k), dtype=np.int64)
            n_valid = np.empty((b,), dtype=np.int32)
            if len(cache) >= 64:
                cache.clear()
            ent = cache[(b, k)] = (q, f, scores, rows, n_valid, q.ctypes.data, f.ctypes.data, scores.ctypes.data,
                                   rows.ctypes.data, n_valid.ctypes.data)
        return ent

    def search_host(self, queries: np.ndarray, k: int, *, metric: str | int = "cosine",
                    filt: Optional[np.ndarray] = None, algo: int = 0, vectors: bool = False):
        """Host buffers in, host buffers out (``crb_search_host``): the per-``invoke`` call.  With ``vectors``
        the hits' stored rows come back too, fp32 ``[batch, k, dim]`` (``crb_search_host_vectors``), as a
        fourth result."""
        q = queries if type(queries) is np.ndarray else np.asarray(queries, dtype=np.float32)
        if q.ndim == 1:
            q = q[None, :]
        if q.shape[1] != self.dim:
            raise ValueError(f"query dimension {q.shape[1]} != index dimension {self.dim}")
        b = q.shape[0]
        k = int(k)
        m = metric if metric.__class__ is int and metric in (N.METRIC_COSINE, N.METRIC_IP) else self._metric(metric)
        if b * self.dim <= self._STAGE_ELEMS and b * k <= self._STAGE_ELEMS:
            sq, sf, s_scores, s_rows, s_valid, pq, pf, ps, pr, pv = self._host_stage(b, k)
            sq[...] = q
            f = self._filt(filt, b) if filt is not None else None
            if f is not None:
                sf[...] = f
            else:
                pf = None
        else:
            q = np.ascontiguousarray(q, dtype=np.float32)
            f = self._filt(filt, b)
            s_scores = np.empty((b, k), dtype=np.float32)
            s_rows = np.empty((b, k), dtype=np.int64)
            s_valid = np.empty((b,), dtype=np.int32)
            pq, pf, ps, pr, pv = q.ctypes.data, (None if f is None else f.ctypes.data), _addr(s_scores), _addr(s_rows), \
                _addr(s_valid)
        if vectors:
            vecs = np.empty((b, k, self.dim), dtype=np.float32)
            N.check(self._lib.crb_search_host_vectors(self._h, pq, b, k, m, pf, ps, pr, pv, _addr(vecs), algo))
            return s_scores.copy(), s_rows.copy(), s_valid.copy(), vecs
        N.check(self._lib.crb_search_host(self._h, pq, b, k, m, pf, ps, pr, pv, algo))
        return s_scores.copy(), s_rows.copy(), s_valid.copy()

    def search_host_any_of(self, query: np.ndarray, filt: np.ndarray, k: int, *, metric: str | int = "cosine",
                           vectors: bool = False):
        """ONE query under a disjunction of filters (``crb_search_host_any_of``): ``filt`` int32 ``[n_alt, n_keys]``, a row
        passes when any of its rows passes -- the traversal's ``adjacent`` step in one search over the union of the
        edges' row sets.  Returns ``(scores [k], rows [k], n_valid)`` and, with ``vectors``, the hits' rows ``[k, dim]``."""
        q = np.ascontiguousarray(query, dtype=np.float32).reshape(-1)
        if q.shape[0] != self.dim:
            raise ValueError(f"query dimension {q.shape[0]} != index dimension {self.dim}")
        f = np.ascontiguousarray(filt, dtype=np.int32)
        if f.ndim != 2 or f.shape[0] < 1 or f.shape[1] != self.n_keys:
            raise ValueError(f"filt must be [n_alt >= 1, {self.n_keys}] (got {f.shape})")
        k = int(k)
        scores = np.empty((k,), dtype=np.float32)
        rows = np.empty((k,), dtype=np.int64)
        n_valid = np.empty((1,), dtype=np.int32)
        vecs = np.empty((k, self.dim), dtype=np.float32) if vectors else None
        N.check(self._lib.crb_search_host_any_of(self._h, q.ctypes.data, f.shape[0], f.ctypes.data if f.size else None, k,
                                                 self._metric(metric), _addr(scores), _addr(rows), _addr(n_valid),
                                                 None if vecs is None else _addr(vecs)))
        return (scores, rows, int(n_valid[0]), vecs) if vectors else (scores, rows, int(n_valid[0]))

    def workspace_bytes(self, batch: int, k: int, n_filters: int) -> int:
        return int(self._lib.crb_workspace_bytes(self._h, batch, k, n_filters))

    def new_workspace(self, batch: int, k: int, n_filters: int) -> torch.Tensor:
        """Zero-filled scratch for :meth:`search_device` (the header must start at zero)."""
        return torch.zeros(self.workspace_bytes(batch, k, n_filters), dtype=torch.uint8, device=self.device)

    def search_device(self, queries: torch.Tensor, k: int, *, metric: str | int = "cosine",
                      filt: Optional[np.ndarray] = None, algo: int = 0,
                      out: Optional[Tuple[torch.Tensor, torch.Tensor, torch.Tensor]] = None,
                      keys_out: Optional[torch.Tensor] = None,
                      workspace: Optional[torch.Tensor] = None,
                      row_masks: Optional[Sequence[Optional[torch.Tensor]]] = None):
        """Device-resident queries and results, asynchronous on torch's current stream.
        ``row_masks[b]`` (optional) is an int32 CUDA tensor of ``ceil(n_rows/32)`` words: one
        pass-bit per row, ANDed with query b's equality filter (``crb_search_masked``)."""
        if not queries.is_cuda or queries.dtype != torch.float32 or not queries.is_contiguous():
            raise ValueError("queries must be a contiguous fp32 CUDA tensor [batch, dim]")
        b, d = queries.shape
        if d != self.dim:
            raise ValueError(f"query dimension {d} != index dimension {self.dim}")
        f = self._filt(filt, b)
        if out is None:
            out = (torch.empty((b, k), dtype=torch.float32, device=self.device),
                   torch.empty((b, k), dtype=torch.int64, device=self.device),
                   torch.empty((b,), dtype=torch.int32, device=self.device))
        scores, rows, n_valid = out
        mask_ptrs = None
        if row_masks is not None and any(m is not None for m in row_masks):
            if len(row_masks) != b:
                raise ValueError("row_masks must have one entry per query")
            words = (self.n_rows + 31) // 32
            arr = (C.c_void_p * b)()
            for i, m in enumerate(row_masks):
                if m is None:
                    continue
                if not m.is_cuda or m.dtype != torch.int32 or m.numel() < words or not m.is_contiguous():
                    raise ValueError(f"row mask must be a contiguous int32 CUDA tensor of >= {words} words")
                arr[i] = m.data_ptr()
            mask_ptrs = arr
        if workspace is None:
            rows_f = [tuple(r) for r in f.tolist()] if f is not None else [()] * b
            ids_m = [0 if (row_masks is None or row_masks[i] is None) else row_masks[i].data_ptr() for i in range(b)]
            nf = len({(rows_f[i], ids_m[i]) for i in range(b)
                      if ids_m[i] or any(c != N.FILT_ANY for c in rows_f[i])})
            need = self.workspace_bytes(b, k, nf)
            stream_id = torch.cuda.current_stream(self.device).cuda_stream
            held = getattr(self._ws_local, "ws", None)
            # (a thread that moves to another stream gets a fresh workspace: work queued on the old stream may
            #  still be using the old one; torch's allocator keeps a freed block until its stream has passed it)
            if held is None or held[0] != stream_id or held[1].numel() < need:
                held = (stream_id, torch.zeros(need, dtype=torch.uint8, device=self.device))
                self._ws_local.ws = held
            workspace = held[1]
        stream = torch.cuda.current_stream(self.device).cuda_stream
        if mask_ptrs is None:
            N.check(self._lib.crb_search(self._h, queries.data_ptr(), b, k, self._metric(metric), _np_ptr(f),
                                         scores.data_ptr(), rows.data_ptr(), n_valid.data_ptr(), _ptr(keys_out),
                                         workspace.data_ptr(), workspace.numel(), algo, stream))
        else:
            N.check(self._lib.crb_search_masked(self._h, queries.data_ptr(), b, k, self._metric(metric), _np_ptr(f),
                                                mask_ptrs, scores.data_ptr(), rows.data_ptr(), n_valid.data_ptr(),
                                                _ptr(keys_out), workspace.data_ptr(), workspace.numel(), algo, stream))
        return scores, rows, n_valid

    def pack_row_mask(self, keep: torch.Tensor) -> torch.Tensor:
        """bool ``[n_rows]`` on the device -> the int32 pass-bit words ``crb_search_masked`` reads."""
        n = self.n_rows
        words = (n + 31) // 32
        bits = torch.zeros(words * 32, dtype=torch.int64, device=self.device)
        bits[:n] = keep.to(torch.int64)
        w = (bits.view(words, 32) << torch.arange(32, device=self.device, dtype=torch.int64)).sum(dim=1)
        return (w & 0xFFFFFFFF).to(torch.int64).where(w < 2**31, w - 2**32).to(torch.int32).contiguous()

    def fetch_rows(self, local_rows: Sequence[int]) -> np.ndarray:
        idx = np.ascontiguousarray(local_rows, dtype=np.int64)
        out = np.empty((len(idx), self.dim), dtype=np.float32)
        if len(idx):
            N.check(self._lib.crb_fetch_rows_host(self._h, _np_ptr(idx), len(idx), _np_ptr(out)))
        return out


def merge_keys(keys_in: torch.Tensor, queries: torch.Tensor, k: int, metric: str | int = "cosine"):
    """Cross-shard merge of all-gathered keys ``[n_lists, batch, k]`` (``crb_merge_keys``)."""
    lib = N.load()
    n_lists, b, kk = keys_in.shape
    assert kk == k and keys_in.dtype == torch.int64 or keys_in.dtype == torch.uint64
    dev = keys_in.device
    scores = torch.empty((b, k), dtype=torch.float32, device=dev)
    rows = torch.empty((b, k), dtype=torch.int64, device=dev)
    n_valid = torch.empty((b,), dtype=torch.int32, device=dev)
    keys_out = torch.empty((b, k), dtype=torch.int64, device=dev)
    N.check(lib.crb_merge_keys(keys_in.data_ptr(), n_lists, b, k, DeviceShard._metric(metric),
                               queries.data_ptr(), queries.shape[1], scores.data_ptr(), rows.data_ptr(),
                               n_valid.data_ptr(), keys_out.data_ptr(), dev.index or 0,
                               torch.cuda.current_stream(dev).cuda_stream))
    return scores, rows, n_valid, keys_out
