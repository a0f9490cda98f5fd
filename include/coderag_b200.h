/*
 * coderag_b200.h -- C ABI of the B200-native retrieval hot path for CodeRAG.
 *
 * The reference (jasonbuchanan145/GithubReposToRag) has no FFI of its own: its boundary is a
 * duck-typed Python vector-store object.  These entry points are what a
 * ctypes/cffi binding behind that object binds; every one cites the reference interface it
 * stands in for.  Plain pointers and sizes only -- no torch, no CUDA types (streams travel as
 * `void*` holding a `cudaStream_t`).
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on failure; `crb_last_error()` returns the
 *     calling thread's last message (the reference propagates Python exceptions,
 *     rag_worker/src/worker/services/agent_graph.py:276-278 -- the Python host raises
 *     RuntimeError with this text);
 *   - fewer than k survivors is NOT an error: the short list is returned, padded with
 *     score = -inf, row = -1, and `n_valid[b]` tells how many slots are real (Cassandra LIMIT
 *     semantics; agent_graph.py:249,253 tolerate short/empty lists);
 *   - `d_` pointers are device memory on the index's device, `h_` pointers are host memory;
 *   - an index borrows the corpus buffers for its lifetime (the store object owns them,
 *     mirroring GraphRetrieverFactory.close(), graph_rag_retrievers.py:136-140);
 *   - all entry points are re-entrant: up to `max_jobs=10` worker threads share one store
 *     (rag_worker/src/worker/worker.py:136-139,185).
 */
#ifndef CODERAG_B200_H
#define CODERAG_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CRB_ABI_VERSION 2

/* similarity modes: VectorStoreQueryMode.DEFAULT (cosine; the SAI index is created with
 * similarity_function cosine, helm/templates/cassandra-initdb-configmap.yaml:99-102) and
 * inner product. */
#define CRB_METRIC_COSINE 0
#define CRB_METRIC_IP 1

/* filter codes in a `filt[B][n_keys]` table */
#define CRB_FILT_ANY (-1)        /* key not constrained */
#define CRB_FILT_IMPOSSIBLE (-2) /* unknown key or value: no row can match (a Cassandra row
                                    lacking the map entry never matches metadata_s['k']='v') */
#define CRB_CODE_MISSING (-1)    /* row has no value for this key */

#define CRB_MAX_KEYS 16   /* filterable metadata keys per index: covers the writer's allow-lists plus its
                             always-kept keys (ingest/src/app/services/vector_write_service.py:28-34,58-64) */
#define CRB_MAX_TOPK 128  /* largest k (reference uses k <= 10, agent_graph.py:171-176) */
#define CRB_MAX_DIM 4096

typedef struct crb_index crb_index;

int crb_abi_version(void);
const char *crb_last_error(void);

/* ---- corpus side: what LCCassandra(table) stands for (graph_rag_retrievers.py:68-80) and the
 *      row layout of helm/templates/cassandra-initdb-configmap.yaml:86-102 ------------------ */

/* Wrap a resident shard: rows fp16 [n_rows, dim] row-major (dim % 8 == 0, 16-byte aligned),
 * inv_norm fp32 [n_rows] (1/||row||, 0 for a zero row), codes int32 [n_keys, n_rows]
 * (dictionary codes of the filterable metadata_s entries, CRB_CODE_MISSING where absent; may
 * be NULL when n_keys == 0).  `row_offset` is the global index of local row 0 (row-sharding).
 * `device` is the CUDA ordinal that owns the buffers. */
int crb_index_create(crb_index **out, const void *d_rows_f16, const float *d_inv_norm,
                     const int32_t *d_codes, int64_t n_rows, int32_t dim, int32_t n_keys,
                     int64_t row_offset, int32_t device);
int crb_index_destroy(crb_index *index);

/* Tuning knobs (the CODERAG_B200_* environment variables of INTEGRATION.md; the reference's config is
 * environment variables read at import, rag_shared/config.py:1-47) are read once, inside
 * crb_index_create -- no search call reads the environment.  A host that changed them afterwards
 * (A/B probes, tests) applies them with this call; no search may be in flight on the index. */
int crb_index_reload_tuning(crb_index *index);

/* Write path (stands for LCCassandra.add_documents / CQL INSERT and DELETE on the embeddings
 * tables: ingest/src/app/services/vector_write_service.py:136-159).  The store appends rows to
 * buffers it grows geometrically and then re-binds the index to them: new pointers / row count,
 * `code_stride` = elements between consecutive key columns of `codes` (>= n_rows, so columns can
 * be over-allocated).  No search may be in flight on the index during the call. */
int crb_index_rebind(crb_index *index, const void *d_rows_f16, const float *d_inv_norm,
                     const int32_t *d_codes, int64_t n_rows, int64_t code_stride);
/* The same re-bind with one code column per key, each as narrow as its dictionary allows: h_code_cols[m] is a
 * DEVICE pointer to key m's column, [>= n_rows] signed codes of h_code_widths[m] = 1, 2 or 4 bytes
 * (CRB_CODE_MISSING = -1 in every width).  The filter kernel is bound by the bytes of these columns, and the
 * reference's filter keys have 2 (`component_kind`, ingest/src/app/services/transform_service.py:112-127), 5
 * (`scope`, ingest_controller.py:179-188), a few (`namespace`, `branch`) or thousands (`repo`) of values: one or
 * two bytes instead of four.  The host widens a column when its dictionary outgrows it. */
int crb_index_rebind_columns(crb_index *index, const void *d_rows_f16, const float *d_inv_norm,
                             const void *const *h_code_cols, const int32_t *h_code_widths, int64_t n_rows);
/* Delete rows (LOCAL indices): their inverse norm and the first 16 bytes of the row become NaN, so
 * no search returns them under either metric.  Returns when the device has been updated. */
int crb_tombstone_rows(crb_index *index, const int64_t *h_rows_idx, int32_t n);

/* Ingest-side conversion (stands for the embedding column write of
 * ingest/src/app/services/vector_write_service.py:158-159): fp32 rows -> fp16 rows and their
 * inverse norms (norm taken of the fp16-rounded row, which is the corpus searched). */
int crb_encode_rows(const float *d_src_f32, void *d_dst_f16, float *d_inv_norm, int64_t n_rows,
                    int32_t dim, int32_t device, void *stream);
/* Inverse norms of rows that are already fp16. */
int crb_row_inv_norms(const void *d_rows_f16, float *d_inv_norm, int64_t n_rows, int32_t dim,
                      int32_t device, void *stream);

/* ---- query side: VectorStore.query(VectorStoreQuery) -> VectorStoreQueryResult
 *      (rest_api/src/app/health.py:104-107) and
 *      LCCassandra.similarity_search_with_embedding_id_by_vector(embedding, k, filter)
 *      (call sites graph_rag_retrievers.py:104-134 via GraphRetriever.invoke,
 *      agent_graph.py:249,268) ----------------------------------------------------------- */

/* Scratch bytes `crb_search` needs for a batch of `batch` queries with `n_filters` distinct
 * non-trivial filters. */
size_t crb_workspace_bytes(const crb_index *index, int32_t batch, int32_t k, int32_t n_filters);

/* Top-k search, queries and results resident on the device; asynchronous on `stream` (one
 * exception: a FILTERED row set that qualifies for the tensor-core kernel reads its survivor
 * count back once, a small synchronising copy, to pick its plan).
 *   d_queries  fp32 [batch, dim]
 *   h_filt     int32 [batch, n_keys] on the HOST (NULL = no filter): AND of equalities, one
 *              code per key (CRB_FILT_ANY / CRB_FILT_IMPOSSIBLE as above).  Rows with the same
 *              filter share one survivor list ("many q, one filter"); distinct filters on one
 *              query vector are the graph-retriever adjacent() fan-out ("one q, many filters").
 *   d_scores   fp32 [batch, k]   similarity, descending (raw cosine or inner product)
 *   d_rows     int64 [batch, k]  global row index (row_offset + local), -1 padding
 *   d_n_valid  int32 [batch]
 *   d_keys     uint64 [batch, k] or NULL: order-preserving (score, row) keys, the payload of
 *              the cross-shard all-gather (see crb_merge_keys)
 *   d_workspace / workspace_bytes  scratch of at least crb_workspace_bytes(...).  Its first
 *              4096 bytes must be ZERO before the first call that uses it (cudaMemset /
 *              torch.zeros); the kernels leave them zero again.  One workspace serves one call
 *              at a time (concurrent calls need one workspace each).  It must be ordinary device
 *              memory of the index's own GPU (cudaMalloc / a torch tensor) -- never symmetric,
 *              peer-mapped or IPC memory: the tensor-core kernel reads its query tile from here
 *              through a TMA tensor map.
 *   algo       0 = choose by batch size, 1 = force the streaming kernel, 2 = force the
 *              tensor-core kernel */
int crb_search(crb_index *index, const float *d_queries, int32_t batch, int32_t k, int32_t metric,
               const int32_t *h_filt, float *d_scores, int64_t *d_rows, int32_t *d_n_valid,
               uint64_t *d_keys, void *d_workspace, size_t workspace_bytes, int32_t algo,
               void *stream);

/* crb_search with one optional caller-built ROW MASK per query, ANDed with the equality filter:
 * h_row_mask_ptrs is a host array of `batch` DEVICE pointers (NULL entries = no mask; the array
 * may be NULL), each to ceil(n_rows/32) uint32 words, bit r%32 of word r/32 set when local row r
 * may be returned.  This carries predicates one code per key cannot express -- membership in
 * list-valued metadata (`topics`, `imports`, `labels`, which the ingest writer shreds into
 * key->value entries, ingest/src/app/services/vector_write_service.py:28-34,152-154) and IN
 * lists.  Queries with the same filter codes AND the same mask pointer share one survivor list. */
int crb_search_masked(crb_index *index, const float *d_queries, int32_t batch, int32_t k,
                      int32_t metric, const int32_t *h_filt, const uint32_t *const *h_row_mask_ptrs,
                      float *d_scores, int64_t *d_rows, int32_t *d_n_valid, uint64_t *d_keys,
                      void *d_workspace, size_t workspace_bytes, int32_t algo, void *stream);

/* Same search through HOST buffers: copies the queries in, runs, copies the results out and
 * returns when they are in the caller's arrays.  This is the call the Python store makes per
 * retriever.invoke(); it uses an internal pool of streams/staging buffers, one per concurrent
 * caller. */
int crb_search_host(crb_index *index, const float *h_queries, int32_t batch, int32_t k,
                    int32_t metric, const int32_t *h_filt, float *h_scores, int64_t *h_rows,
                    int32_t *h_n_valid, int32_t algo);

/* crb_search_host that also returns the hits' STORED vectors, fp32 [batch, k, dim] (zeros in padding slots), in the
 * same device-to-host copy: LCCassandra.similarity_search_with_embedding_id_by_vector returns
 * (Document, embedding, id) triples and graph-retriever re-scores from `embedding`
 * (call sites rag_worker/src/worker/services/graph_rag_retrievers.py:104-134), so every traversal step needs the
 * vectors -- one round trip instead of a search followed by crb_fetch_rows_host. */
int crb_search_host_vectors(crb_index *index, const float *h_queries, int32_t batch, int32_t k,
                            int32_t metric, const int32_t *h_filt, float *h_scores, int64_t *h_rows,
                            int32_t *h_n_valid, float *h_vectors, int32_t algo);

/* ONE query under a DISJUNCTION of filters: h_filt int32 [n_alt, n_keys], a row passes when ANY of the n_alt rows (each an
 * AND of equalities, as in crb_search_host) passes; at most 32 distinct alternatives.  Replaces the per-edge searches of
 * the graph traversal's adjacent step: rag_worker/src/worker/services/graph_rag_retrievers.py:82-134 gives the
 * retriever its edges ((repo, repo), (module, module), (file_path, file_path) ...), and graph-retriever's store adapter
 * runs one filtered search per (field, value) edge with the same query vector, then keeps the best k of the merged,
 * de-duplicated hits.  The best k of a union equals the best k of the per-set best k's, so this call returns exactly
 * those rows while reading and scoring a row that lies in several sets (a file's rows are its module's and its
 * repository's) once.  Results for the one query: h_scores fp32 [k], h_rows int64 [k], h_n_valid int32 [1];
 * h_vectors fp32 [k, dim] (the hits' stored vectors) or NULL. */
int crb_search_host_any_of(crb_index *index, const float *h_query, int32_t n_alt, const int32_t *h_filt,
                           int32_t k, int32_t metric, float *h_scores, int64_t *h_rows,
                           int32_t *h_n_valid, float *h_vectors);

/* Cross-shard merge (no reference counterpart in-tree: the reference runs `replicas: 1`,
 * helm/values.yaml:113; this is the k-way merge a row-sharded store needs after the
 * all-gather of per-shard keys).  d_keys_in uint64 [n_lists, batch, k] -> the best k per
 * query.  d_queries (fp32 [batch, dim]) supplies the query norm for cosine scores. */
int crb_merge_keys(const uint64_t *d_keys_in, int32_t n_lists, int32_t batch, int32_t k,
                   int32_t metric, const float *d_queries, int32_t dim, float *d_scores,
                   int64_t *d_rows, int32_t *d_n_valid, uint64_t *d_keys_out, int32_t device,
                   void *stream);

/* Row-sharded search with the cross-shard exchange fused in (no reference counterpart: the
 * reference runs one replica).  Every rank of one box calls this with the same batch/k; each
 * rank's search kernels store the k exact keys of every finished query straight into EVERY
 * rank's gather buffer over NVLink peer memory and bump a per-source arrival counter there
 * (system-scope release); a final kernel on each rank waits for `expected_total` arrivals from
 * every source and merges -- no collective launch between search and merge.  A query some source never
 * published (bounded wait) comes back with n_valid = -2: the caller must treat that as a failure.
 *   peer_keys[p]     device pointer, valid on this rank, to rank p's gather buffer:
 *                    uint64 [2 (call parity)][world][cap_keys]   (e.g. torch symmetric memory)
 *   peer_counters[p] device pointer to rank p's uint32 [world] arrival counters (zeroed once)
 *   call_index       0, 1, 2, ... (its parity picks the gather half; same on all ranks)
 *   expected_total   queries every source has published once this call is done: the running sum
 *                    of `batch` over all calls so far, this one included
 * Outputs are the merged (global) top-k on every rank.  Other arguments as crb_search. */
#define CRB_MAX_PEERS 8
typedef struct crb_exchange {
    int32_t rank;
    int32_t world;
    uint64_t cap_keys;
    uint64_t *peer_keys[CRB_MAX_PEERS];
    uint32_t *peer_counters[CRB_MAX_PEERS];
    uint32_t dest_mask; /* bit p set: rank p receives every rank's keys and merges; 0 = all ranks.  The
                           serving store sets only the bit of the rank that owns the Python face (the
                           reference runs ONE worker process, helm/templates/rag-worker-deployment.yaml:8):
                           the other ranks send, wait for nothing and write no result. */
} crb_exchange;
int crb_search_exchange(crb_index *index, const float *d_queries, int32_t batch, int32_t k,
                        int32_t metric, const int32_t *h_filt, float *d_scores, int64_t *d_rows,
                        int32_t *d_n_valid, uint64_t *d_keys, void *d_workspace,
                        size_t workspace_bytes, int32_t algo, void *stream,
                        const crb_exchange *exchange, uint32_t call_index, uint32_t expected_total);

/* crb_search_exchange through HOST buffers, the call a serving rank makes per request (the sharded counterpart
 * of crb_search_host; stands for the same retriever.invoke() -> store call, agent_graph.py:249,268): copies the
 * queries in from `h_queries` (every rank is handed the same request by the host-side dispatcher), searches its
 * shard, sends its keys to the ranks of `dest_mask`, and -- on those ranks -- waits for all sources, merges, copies
 * the merged result out and checks it: a source that never published makes the call FAIL (it does not return a
 * short result).  Ranks outside `dest_mask` may pass NULL outputs.  `h_row_mask_ptrs` as in crb_search_masked
 * (device pointers to THIS rank's slice of each mask; may be NULL).  Returns when this rank's work is done. */
int crb_search_exchange_host(crb_index *index, const float *h_queries, int32_t batch, int32_t k,
                             int32_t metric, const int32_t *h_filt, const uint32_t *const *h_row_mask_ptrs,
                             float *h_scores, int64_t *h_rows, int32_t *h_n_valid, int32_t algo,
                             const crb_exchange *exchange, uint32_t call_index, uint32_t expected_total);

/* crb_search_host_any_of over a row-sharded store: ONE query, a row passes when any of the n_alt filter rows passes;
 * every rank searches the union inside its shard and the keys are exchanged and merged as in crb_search_exchange_host
 * (one query per call: expected_total advances by 1). */
int crb_search_exchange_host_any_of(crb_index *index, const float *h_query, int32_t n_alt, const int32_t *h_filt,
                                    int32_t k, int32_t metric, float *h_scores, int64_t *h_rows,
                                    int32_t *h_n_valid, const crb_exchange *exchange, uint32_t call_index,
                                    uint32_t expected_total);

/* The serving store's request path without the host language in it (no reference counterpart: the reference has one
 * worker process and no second GPU).  The ranks of a box share one block of host memory (POSIX shared memory, laid out
 * by the host side: githubrepostorag_b200/sharded_store.py::ControlBlock); `crb_ctl` says where things live in it.
 *   crb_serve_post_search  rank 0: wait until every rank acknowledged the previous command, publish a plain SEARCH
 *                          (queries, filter codes, k ...; bumps the sequence word), run this rank's share through
 *                          crb_search_exchange_host and return the merged result.  Returns 2 when a rank reported a
 *                          failure of the previous command (the host reads its reply).
 *   crb_serve_next         ranks > 0: wait for the next command; run it and acknowledge it if it is a plain SEARCH, and
 *                          go on waiting; return 0 when the command needs the host side (masks, writes, fetches ...,
 *                          left unacknowledged, *last_seq = its sequence number), -1 on a failed search (unacknowledged).
 * `call_state` is {calls so far, queries published so far} of this rank's exchange, shared with the host side. */
typedef struct crb_ctl {
    uint64_t seq_off, hdr_off, ack_off, status_off, queries_off, filt_off;
    int32_t rank, world;
} crb_ctl;
int crb_serve_post_search(crb_index *index, void *ctl_base, const crb_ctl *ctl, const float *h_queries, int32_t batch,
                          int32_t k, int32_t metric, const int32_t *h_filt, float *h_scores, int64_t *h_rows,
                          int32_t *h_n_valid, int32_t algo, const crb_exchange *exchange, uint32_t *call_state,
                          uint64_t *seq_state, double timeout_s);
int crb_serve_next(crb_index *index, void *ctl_base, const crb_ctl *ctl, const crb_exchange *exchange,
                   uint32_t *call_state, uint64_t *last_seq);

/* Page-lock / release a host range the caller owns (no reference counterpart; plumbing of the serving store): the
 * shared-memory block its ranks exchange requests through is registered so that crb_search_exchange_host copies
 * large query batches to the GPU straight out of it.  Failure leaves the CUDA runtime's error state clean. */
int crb_host_register(void *h_ptr, size_t bytes);
int crb_host_unregister(void *h_ptr);

/* Fetch stored rows as fp32 (the `embedding` member of the (Document, embedding, id) triples
 * LCCassandra returns; graph-retriever re-scores from it).  h_rows_idx are LOCAL row indices. */
int crb_fetch_rows_host(crb_index *index, const int64_t *h_rows_idx, int32_t n, float *h_out);

/* Search that allocates nothing -- no counterpart in the reference (the reference never manages
 * device memory; this exists for hosts of this library that do).
 *
 * Who needs it: a serving process that has already given the whole card to the corpus and to its
 * own memory pool has nothing left for cudaMalloc to hand out, and a plugin host that owns the
 * device-memory budget may forbid its plugins from allocating at all; in a request path an
 * allocation is also a device-wide synchronisation point that stalls every other stream.  Those
 * callers cannot provide or grow a workspace for crb_search and cannot use crb_search_host (its
 * lanes cudaMalloc their staging buffers).  For them this entry runs the same streaming kernel
 * with its scratch (per-CTA candidate lists + arrival counters) in a fixed array that is part
 * of the library image: CRB_NOALLOC_SCRATCH_BYTES of HBM on every GPU that loads the library,
 * held until the process exits.  The caller passes only inputs and outputs, all on the index's
 * device:
 *   d_queries fp32 [batch, dim];  d_scores fp32 [batch, k];  d_rows int64 [batch, k];
 *   d_n_valid int32 [batch].  Unfiltered search only (survivor lists do not fit the array).
 *
 * Limits: batch <= CRB_NOALLOC_MAX_BATCH and k <= CRB_MAX_TOPK; anything larger is refused with
 * an error, never truncated.  Results are identical to crb_search on the same inputs.
 *
 * Set-up: the first call on each device looks up the array's device address (once per device,
 * with that device current).  Make that first call before starting any stream capture.
 * Sharing: every call on a device uses the same array, which still holds the previous call's
 * lists; the entry clears the counters it needs zeroed at the start of each call.  Two calls
 * whose kernels overlap on the GPU (different streams, or different threads) race on the array
 * and return garbage -- serialise them on one stream.
 * It calls no cudaMalloc / cudaFree / cudaMallocHost or any other allocating routine. */
#define CRB_NOALLOC_MAX_BATCH 4
#define CRB_NOALLOC_SCRATCH_BYTES (4096 + (1 << 20))
int crb_search_noalloc(crb_index *index, const float *d_queries, int32_t batch, int32_t k,
                       int32_t metric, float *d_scores, int64_t *d_rows, int32_t *d_n_valid,
                       void *stream);

/* Diagnostics of the tensor-core plan (no counterpart in the reference; used by this repository's tests to
 * MEASURE the margin its exactness proof relies on).  While `d_stats` (fp32 [4 * batch] on the index's device) is
 * set, every tensor-core search writes for query b:  [4b+0] max |approximate - exact| score over the candidates it
 * re-scored (normalised-query units), [4b+1] the margin the proof allowed, [4b+2] how many it re-scored,
 * [4b+3] 1 if the result was proven exact, 0 if the query was handed to the streaming fallback.  NULL switches it
 * off (the default).  Results are not affected.  Not thread-safe: set it with no search in flight. */
int crb_index_set_k5_stats(crb_index *index, float *d_stats);

/* Number of kernels this library has launched in the calling process (bench.py's
 * `gpu_launches`). */
int64_t crb_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* CODERAG_B200_H */
