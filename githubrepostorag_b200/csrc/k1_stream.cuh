// K1 -- HBM-streaming exact scorer with fused per-warp top-k and last-block merge.
//
// Replaces the scoring that the reference delegates to Cassandra's SAI ANN
// (`ORDER BY vector ANN OF ? LIMIT k`, reached from
// rag_worker/src/worker/services/graph_rag_retrievers.py:104-134) with an exact scan of the
// fp16 shard resident in HBM.  One warp scores one row at a time with 16-byte coalesced loads
// (R rows in flight per warp), fp32 accumulation in the library's canonical order
// (common.cuh), so a K1 score *is* the final score: no rescore pass follows.
//
// A launch covers up to kMaxGroups "groups" (blockIdx.y).  A group is up to NQ queries that
// share one row set: either the whole shard or the survivor list K3 built for their filter.
// Bound: HBM.  Algorithmic bytes per group = items * dim * 2 (+ 4 B/row inverse norm,
// + 4 B/survivor list entry).
#pragma once
#include "topk.cuh"

namespace crb {

constexpr int kK1Threads = 256;
constexpr int kK1Warps = kK1Threads / 32;
constexpr int kK1MaxQ = 4;
constexpr int kK1MergeLPT = 4;  // last-block merge: lists per thread -> grid.x <= 1024
constexpr int kK1MaxGridX = kK1Threads * kK1MergeLPT;

struct K1Group {
    int32_t q_idx[kK1MaxQ];  // query index in the batch, -1 = unused slot
    int32_t filter;          // -1 = whole shard, else survivor list slot
};

struct K1Params {
    const __half* rows;
    const float* inv_norm;
    const float* queries;          // [batch, dim] fp32
    const uint32_t* lists;         // [n_filters, list_stride] local row indices
    const uint32_t* list_counts;   // [n_filters]
    uint64_t* cand;                // [groups, NQ, grid.x, k] per-CTA sorted lists (scratch)
    unsigned* counters;            // [groups] arrival counters, zero between launches
    float* out_scores;             // [batch, k]
    int64_t* out_rows;             // [batch, k]
    int32_t* out_nvalid;           // [batch]
    uint64_t* out_keys;            // [batch, k] or nullptr
    int64_t n_rows;
    int64_t list_stride;
    uint32_t row_offset;
    int32_t dim;
    int32_t k;
    int32_t metric;
    ExchangeDev ex;                // ex.world == 0: local search only
    K1Group groups[kMaxGroups];
};

// Last-CTA work for one finished query: merge the per-CTA lists, emit (and publish to peers).  Kept out
// of line so its registers and code stay out of the streaming loop's allocation (inlined, it pushed
// the d=768 kernel against its 128-register cap and cost 7 % of the bandwidth).
__device__ __noinline__ void k1_merge_query(const K1Params& p, int qi, int64_t list_slot, const float* q_ptr,
                                            uint64_t (*xchg)[32], float* s_invq) {
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        const float iq = (p.metric == 0) ? warp_inv_qnorm(q_ptr, p.dim, lane) : 1.0f;
        if (lane == 0) *s_invq = iq;
    }
    __syncthreads();
    TopKSink sink;
    sink.scores = p.out_scores + static_cast<int64_t>(qi) * p.k;
    sink.rows = p.out_rows + static_cast<int64_t>(qi) * p.k;
    sink.n_valid = p.out_nvalid + qi;
    sink.keys = p.out_keys ? p.out_keys + static_cast<int64_t>(qi) * p.k : nullptr;
    sink.scale = *s_invq;
    sink.ex = p.ex.world > 0 ? &p.ex : nullptr;
    sink.qi = qi;
    sink.k = p.k;
    const uint64_t* lists = p.cand + list_slot * gridDim.x * p.k;
    block_merge_sorted_lists<kK1MergeLPT>(lists, gridDim.x, p.k, p.k, sink, xchg, lane, warp, kK1Warps);
}

// Row-sharded search with the receiving half folded in (ExchangeDev::fold): the group's last CTA has sent the
// shard's keys of ALL its queries to the destination ranks; if this rank is one of them it now waits for every
// source and merges their lists (its own included) into the final answers -- the step costs one launch per rank,
// not two.  Called by the whole CTA after the group's local merges.
__device__ __noinline__ void k1_fold_exchange(const K1Params& p, const int32_t* q_idx, int nq, uint64_t (*xchg)[32],
                                              float* s_invq) {
    if (!(p.ex.world > 0 && p.ex.fold && ((p.ex.dest_mask >> p.ex.rank) & 1u))) return;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const bool arrived = exchange_wait_all(p.ex);
    for (int q = 0; q < nq; ++q) {
        const int qi = q_idx[q];
        if (qi < 0) continue;
        if (!arrived) {
            if (threadIdx.x == 0) p.out_nvalid[qi] = -2;
            continue;
        }
        if (warp == 0) {
            const float iq = (p.metric == 0) ? warp_inv_qnorm(p.queries + static_cast<int64_t>(qi) * p.dim, p.dim, lane) : 1.0f;
            if (lane == 0) *s_invq = iq;
        }
        __syncthreads();
        TopKSink sink;
        sink.scores = p.out_scores + static_cast<int64_t>(qi) * p.k;
        sink.rows = p.out_rows + static_cast<int64_t>(qi) * p.k;
        sink.n_valid = p.out_nvalid + qi;
        sink.keys = p.out_keys ? p.out_keys + static_cast<int64_t>(qi) * p.k : nullptr;
        sink.scale = *s_invq;
        sink.qi = qi;
        sink.k = p.k;
        const uint64_t* gather = p.ex.keys[p.ex.rank] + static_cast<uint64_t>(p.ex.parity) * p.ex.world * p.ex.cap +
                                 static_cast<uint64_t>(qi) * p.k;
        block_merge_sorted_lists<1>(gather, p.ex.world, p.k, static_cast<int64_t>(p.ex.cap), sink, xchg, lane, warp, kK1Warps);
    }
}

// EX: the row has exactly 32 * CH chunks (dim = 256 * CH, e.g. 768 and 1024): no per-chunk range tests in the loop.
template <int NQ, int CH, int R, int KPL, int MINB = 1, bool EX = false>
__global__ void __launch_bounds__(kK1Threads, MINB)
k1_stream_kernel(const __grid_constant__ K1Params p) {
    __shared__ uint64_t xchg[2][32];
    __shared__ float s_invq;
    __shared__ int s_last;

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const K1Group& G = p.groups[blockIdx.y];
    const int chunks = p.dim >> 3;
    const bool cosine = (p.metric == 0);

    // ---- query chunks this lane needs, in registers ----
    float qr[NQ][CH][8];
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
#pragma unroll
        for (int c = 0; c < CH; ++c) {
            const int ch = lane + 32 * c;
            if (G.q_idx[q] >= 0 && ch < chunks) {
                const float4* src = reinterpret_cast<const float4*>(
                    p.queries + static_cast<int64_t>(G.q_idx[q]) * p.dim + ch * 8);
                const float4 a = __ldg(src), b = __ldg(src + 1);
                qr[q][c][0] = a.x; qr[q][c][1] = a.y; qr[q][c][2] = a.z; qr[q][c][3] = a.w;
                qr[q][c][4] = b.x; qr[q][c][5] = b.y; qr[q][c][6] = b.z; qr[q][c][7] = b.w;
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) qr[q][c][e] = 0.f;
            }
        }
    }

    WarpTopK<KPL> top[NQ];
#pragma unroll
    for (int q = 0; q < NQ; ++q) top[q].init(p.k, lane);

    // ---- row set of this group ----
    const bool use_list = G.filter >= 0;
    int64_t n_items = p.n_rows;
    const uint32_t* list = nullptr;
    if (use_list) {
        const int64_t cnt = static_cast<int64_t>(p.list_counts[G.filter]);
        n_items = cnt < p.n_rows ? cnt : p.n_rows;
        list = p.lists + static_cast<int64_t>(G.filter) * p.list_stride;
    }
    const int64_t stride = static_cast<int64_t>(gridDim.x) * kK1Warps * R;
    int64_t item0 = (static_cast<int64_t>(blockIdx.x) * kK1Warps + warp) * R;

    uint32_t idx_next = 0u;
    if (use_list && lane < R && item0 + lane < n_items) idx_next = list[item0 + lane];

    for (; item0 < n_items; item0 += stride) {
        const uint32_t idx_cur = idx_next;
        idx_next = 0u;
        if (use_list && lane < R && item0 + stride + lane < n_items)
            idx_next = list[item0 + stride + lane];

        int64_t row[R];
        bool ok[R];
#pragma unroll
        for (int j = 0; j < R; ++j) {
            ok[j] = (item0 + j) < n_items;
            const uint32_t li = __shfl_sync(CRB_FULL_MASK, idx_cur, j);
            row[j] = use_list ? static_cast<int64_t>(li) : (item0 + j);
            if (!ok[j]) row[j] = 0;
        }

        uint4 v[R][CH];
        float invn[R];
#pragma unroll
        for (int j = 0; j < R; ++j) {
            const __half* rp = p.rows + row[j] * p.dim;
#pragma unroll
            for (int c = 0; c < CH; ++c) {
                const int ch = lane + 32 * c;
                if (EX || ch < chunks) v[j][c] = ld_stream_16(rp + ch * 8);
            }
        }
#pragma unroll
        // (keep this load predicated on the metric: the d=768 instantiation sits at 125 of the 128 registers
        //  that two resident CTAs allow; an unconditional load + select tipped ptxas into re-materialising
        //  row pointers between the 12 LDG.128s and cost 7 % of the bandwidth -- profiles/_ab/k1_ab.py)
        for (int j = 0; j < R; ++j) invn[j] = cosine ? __ldg(p.inv_norm + row[j]) : 1.0f;

        float acc[NQ][R];
#pragma unroll
        for (int q = 0; q < NQ; ++q)
#pragma unroll
            for (int j = 0; j < R; ++j) acc[q][j] = 0.f;
#pragma unroll
        for (int c = 0; c < CH; ++c) {
            if (EX || lane + 32 * c < chunks) {
#pragma unroll
                for (int j = 0; j < R; ++j)
#pragma unroll
                    for (int q = 0; q < NQ; ++q) acc[q][j] = chunk_fma(v[j][c], qr[q][c], acc[q][j]);
            }
        }
#pragma unroll
        for (int j = 0; j < R; ++j) {
            const bool row_ok = ok[j] && !(cosine && invn[j] == 0.f);
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                const float s = warp_butterfly_sum(acc[q][j]) * invn[j];
                if (row_ok && G.q_idx[q] >= 0 && s == s) {
                    const uint64_t key = make_key(s, p.row_offset + static_cast<uint32_t>(row[j]));
                    if (key > top[q].thr) top[q].insert(key, lane);
                }
            }
        }
    }

    // ---- per-CTA lists -> scratch ----
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
        if (G.q_idx[q] < 0) continue;
        uint64_t* out = p.cand +
            ((static_cast<int64_t>(blockIdx.y) * NQ + q) * gridDim.x + blockIdx.x) * p.k;
        block_extract_topk<KPL>(top[q], p.k, out, xchg, lane, warp, kK1Warps);
    }

    // ---- the last CTA of the group merges all per-CTA lists and writes the results ----
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned t = atomicAdd(&p.counters[blockIdx.y], 1u);
        s_last = (t == gridDim.x - 1) ? 1 : 0;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();

#pragma unroll
    for (int q = 0; q < NQ; ++q) {
        const int qi = G.q_idx[q];
        if (qi < 0) continue;
        k1_merge_query(p, qi, static_cast<int64_t>(blockIdx.y) * NQ + q,
                       p.queries + static_cast<int64_t>(qi) * p.dim, xchg, &s_invq);
    }
    k1_fold_exchange(p, G.q_idx, NQ, xchg, &s_invq);
    if (threadIdx.x == 0) p.counters[blockIdx.y] = 0u;
}

// Half-warp variant for short rows (dim <= 512, i.e. <= 64 chunks): two rows per warp step, 16 lanes
// each, so no lane idles on a 768-byte row and the butterfly needs four shuffle levels instead of
// five.  Scores are BIT-IDENTICAL to the canonical 32-lane order: half-lane h computes the canonical
// partials of lanes h (chunks h, h+32) and h+16 (chunks h+16, h+48) in two accumulators, adds them
// (the canonical xor-16 level, locally), then runs the xor-8/4/2/1 levels inside its half.
// E3: the row has exactly 48 chunks (dim = 384, the reference's embedding width): every half-lane owns exactly
// three chunks, so the per-chunk range tests and the fourth chunk slot disappear from the loop.
template <int NQ, int R, int KPL, bool E3 = false>
__global__ void __launch_bounds__(kK1Threads, 2)
k1_stream_half_kernel(const __grid_constant__ K1Params p) {
    __shared__ uint64_t xchg[2][32];
    __shared__ float s_invq;
    __shared__ int s_last;

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int half = lane >> 4;
    const int hl = lane & 15;
    const K1Group& G = p.groups[blockIdx.y];
    const int chunks = p.dim >> 3;
    const bool cosine = (p.metric == 0);

    constexpr int NC = E3 ? 3 : 4;      // chunk slots per half-lane
    float qr[NQ][NC][8];
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            const int ch = hl + 16 * c;
            if (G.q_idx[q] >= 0 && ch < chunks) {
                const float4* src = reinterpret_cast<const float4*>(
                    p.queries + static_cast<int64_t>(G.q_idx[q]) * p.dim + ch * 8);
                const float4 a = __ldg(src), b = __ldg(src + 1);
                qr[q][c][0] = a.x; qr[q][c][1] = a.y; qr[q][c][2] = a.z; qr[q][c][3] = a.w;
                qr[q][c][4] = b.x; qr[q][c][5] = b.y; qr[q][c][6] = b.z; qr[q][c][7] = b.w;
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) qr[q][c][e] = 0.f;
            }
        }
    }

    WarpTopK<KPL> top[NQ];
#pragma unroll
    for (int q = 0; q < NQ; ++q) top[q].init(p.k, lane);

    const bool use_list = G.filter >= 0;
    int64_t n_items = p.n_rows;
    const uint32_t* list = nullptr;
    if (use_list) {
        const int64_t cnt = static_cast<int64_t>(p.list_counts[G.filter]);
        n_items = cnt < p.n_rows ? cnt : p.n_rows;
        list = p.lists + static_cast<int64_t>(G.filter) * p.list_stride;
    }
    constexpr int RW = 2 * R;   // rows per warp step
    const int64_t stride = static_cast<int64_t>(gridDim.x) * kK1Warps * RW;
    int64_t item0 = (static_cast<int64_t>(blockIdx.x) * kK1Warps + warp) * RW;

    uint32_t idx_next = 0u;
    if (use_list && lane < RW && item0 + lane < n_items) idx_next = list[item0 + lane];

    for (; item0 < n_items; item0 += stride) {
        const uint32_t idx_cur = idx_next;
        idx_next = 0u;
        if (use_list && lane < RW && item0 + stride + lane < n_items) idx_next = list[item0 + stride + lane];

        int64_t row[R];     // this half's row of pair j
        bool ok[R];
#pragma unroll
        for (int j = 0; j < R; ++j) {
            const int it = 2 * j + half;
            ok[j] = (item0 + it) < n_items;
            const uint32_t li = __shfl_sync(CRB_FULL_MASK, idx_cur, it);
            row[j] = use_list ? static_cast<int64_t>(li) : (item0 + it);
            if (!ok[j]) row[j] = 0;
        }
        uint4 v[R][NC];
        float invn[R];
#pragma unroll
        for (int j = 0; j < R; ++j) {
            const __half* rp = p.rows + row[j] * p.dim;
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                const int ch = hl + 16 * c;
                if (E3 || ch < chunks) v[j][c] = ld_stream_16(rp + ch * 8);
            }
        }
#pragma unroll
        for (int j = 0; j < R; ++j) invn[j] = cosine ? __ldg(p.inv_norm + row[j]) : 1.0f;
#pragma unroll
        for (int j = 0; j < R; ++j) {
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                // canonical lane hl: chunks hl, hl+32 ; canonical lane hl+16: chunks hl+16, hl+48
                float a0 = 0.f, a1 = 0.f;
                if constexpr (E3) {
                    a0 = chunk_fma(v[j][0], qr[q][0], a0);
                    a0 = chunk_fma(v[j][2], qr[q][2], a0);
                    a1 = chunk_fma(v[j][1], qr[q][1], a1);
                } else {
                    if (hl < chunks) a0 = chunk_fma(v[j][0], qr[q][0], a0);
                    if (hl + 32 < chunks) a0 = chunk_fma(v[j][2], qr[q][2], a0);
                    if (hl + 16 < chunks) a1 = chunk_fma(v[j][1], qr[q][1], a1);
                    if (hl + 48 < chunks) a1 = chunk_fma(v[j][NC - 1], qr[q][NC - 1], a1);
                }
                float s = a0 + a1;                                   // canonical xor-16 level
                s += __shfl_xor_sync(CRB_FULL_MASK, s, 8);
                s += __shfl_xor_sync(CRB_FULL_MASK, s, 4);
                s += __shfl_xor_sync(CRB_FULL_MASK, s, 2);
                s += __shfl_xor_sync(CRB_FULL_MASK, s, 1);
                s *= invn[j];
                const bool good = ok[j] && !(cosine && invn[j] == 0.f) && G.q_idx[q] >= 0 && s == s;
                const uint64_t key = good ? make_key(s, p.row_offset + static_cast<uint32_t>(row[j])) : 0ull;
                // one vote per row pair; only a pair that holds a candidate is broadcast and admitted,
                // one row after the other (warp-uniform keys)
                if (__any_sync(CRB_FULL_MASK, key > top[q].thr)) {
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const uint64_t kh = __shfl_sync(CRB_FULL_MASK, key, h * 16);
                        if (kh > top[q].thr) top[q].insert(kh, lane);
                    }
                }
            }
        }
    }

#pragma unroll
    for (int q = 0; q < NQ; ++q) {
        if (G.q_idx[q] < 0) continue;
        uint64_t* out = p.cand +
            ((static_cast<int64_t>(blockIdx.y) * NQ + q) * gridDim.x + blockIdx.x) * p.k;
        block_extract_topk<KPL>(top[q], p.k, out, xchg, lane, warp, kK1Warps);
    }

    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned t = atomicAdd(&p.counters[blockIdx.y], 1u);
        s_last = (t == gridDim.x - 1) ? 1 : 0;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();

#pragma unroll
    for (int q = 0; q < NQ; ++q) {
        const int qi = G.q_idx[q];
        if (qi < 0) continue;
        k1_merge_query(p, qi, static_cast<int64_t>(blockIdx.y) * NQ + q,
                       p.queries + static_cast<int64_t>(qi) * p.dim, xchg, &s_invq);
    }
    k1_fold_exchange(p, G.q_idx, NQ, xchg, &s_invq);
    if (threadIdx.x == 0) p.counters[blockIdx.y] = 0u;
}

// Generic-dimension variant (dim up to CRB_MAX_DIM, one query per group): the query lives in
// shared memory and chunks are walked with a runtime loop; same canonical summation order.
template <int R, int KPL>
__global__ void __launch_bounds__(kK1Threads)
k1_stream_generic_kernel(const __grid_constant__ K1Params p) {
    extern __shared__ float s_q[];  // [dim]
    __shared__ uint64_t xchg[2][32];
    __shared__ float s_invq;
    __shared__ int s_last;

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const K1Group& G = p.groups[blockIdx.y];
    const int chunks = p.dim >> 3;
    const bool cosine = (p.metric == 0);
    const int qi = G.q_idx[0];

    for (int i = threadIdx.x; i < p.dim; i += kK1Threads)
        s_q[i] = p.queries[static_cast<int64_t>(qi) * p.dim + i];
    __syncthreads();

    WarpTopK<KPL> top;
    top.init(p.k, lane);

    const bool use_list = G.filter >= 0;
    int64_t n_items = p.n_rows;
    const uint32_t* list = nullptr;
    if (use_list) {
        const int64_t cnt = static_cast<int64_t>(p.list_counts[G.filter]);
        n_items = cnt < p.n_rows ? cnt : p.n_rows;
        list = p.lists + static_cast<int64_t>(G.filter) * p.list_stride;
    }
    const int64_t stride = static_cast<int64_t>(gridDim.x) * kK1Warps * R;
    for (int64_t item0 = (static_cast<int64_t>(blockIdx.x) * kK1Warps + warp) * R; item0 < n_items;
         item0 += stride) {
        int64_t row[R];
        bool ok[R];
        float acc[R];
#pragma unroll
        for (int j = 0; j < R; ++j) {
            ok[j] = (item0 + j) < n_items;
            row[j] = ok[j] ? (use_list ? static_cast<int64_t>(list[item0 + j]) : item0 + j) : 0;
            acc[j] = 0.f;
        }
        for (int ch = lane; ch < chunks; ch += 32) {
            uint4 v[R];
#pragma unroll
            for (int j = 0; j < R; ++j) v[j] = ld_stream_16(p.rows + row[j] * p.dim + ch * 8);
            float q8[8];
            const float4 a = *reinterpret_cast<const float4*>(s_q + ch * 8);
            const float4 b = *reinterpret_cast<const float4*>(s_q + ch * 8 + 4);
            q8[0] = a.x; q8[1] = a.y; q8[2] = a.z; q8[3] = a.w;
            q8[4] = b.x; q8[5] = b.y; q8[6] = b.z; q8[7] = b.w;
#pragma unroll
            for (int j = 0; j < R; ++j) acc[j] = chunk_fma(v[j], q8, acc[j]);
        }
#pragma unroll
        for (int j = 0; j < R; ++j) {
            const float invn = cosine ? __ldg(p.inv_norm + row[j]) : 1.0f;
            const float s = warp_butterfly_sum(acc[j]) * invn;
            if (ok[j] && !(cosine && invn == 0.f) && s == s) {
                const uint64_t key = make_key(s, p.row_offset + static_cast<uint32_t>(row[j]));
                if (key > top.thr) top.insert(key, lane);
            }
        }
    }

    uint64_t* out = p.cand + (static_cast<int64_t>(blockIdx.y) * gridDim.x + blockIdx.x) * p.k;
    block_extract_topk<KPL>(top, p.k, out, xchg, lane, warp, kK1Warps);

    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned t = atomicAdd(&p.counters[blockIdx.y], 1u);
        s_last = (t == gridDim.x - 1) ? 1 : 0;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();

    k1_merge_query(p, qi, static_cast<int64_t>(blockIdx.y), s_q, xchg, &s_invq);
    k1_fold_exchange(p, G.q_idx, 1, xchg, &s_invq);
    if (threadIdx.x == 0) p.counters[blockIdx.y] = 0u;
}

// ---------------------------------------------------------------------------------------------
// Device-side fallback for tensor-core batches: K5 appends the queries it could not prove exact to
// `flagged`; this kernel, launched unconditionally right behind K5, re-runs exactly those through
// the streaming scan -- no host round trip, so the tensor-core path stays asynchronous on its stream.
// The grid is fixed (about one wave of the machine); the CTAs split themselves over however many
// queries were flagged: cpq = CTAs per query (all of them for one query, one each for many), query
// f is served by CTA columns f, f + n_cols, ...  Zero flagged queries: every CTA reads one word and exits.
// Same canonical summation order as every other scoring kernel, so the result is bit-identical to K1's.
struct FbParams {
    const __half* rows;
    const float* inv_norm;
    const float* queries;        // [batch, dim] fp32
    const int32_t* q_map;        // chunk slot -> query index in the caller's batch (nullptr = identity)
    const int32_t* flagged;      // [n_flagged] chunk slots
    const int32_t* n_flagged;
    const uint32_t* list;        // survivors of the set's filter (nullptr = whole shard)
    const uint32_t* list_count;
    uint64_t* cand;              // [flagged query][cpq][k] scratch
    unsigned* counters;          // [chunk] arrival counters, zero before the launch
    float* out_scores;
    int64_t* out_rows;
    int32_t* out_nvalid;
    uint64_t* out_keys;
    int64_t n_rows;
    int64_t cand_lists;          // lists of k keys that fit in `cand`
    uint32_t row_offset;
    int32_t dim;
    int32_t k;
    int32_t metric;
    ExchangeDev ex;
};

__device__ __noinline__ void fb_merge_query(const FbParams& p, int qi, const uint64_t* lists, int n_lists,
                                            const float* s_q, uint64_t (*xchg)[32], float* s_invq) {
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        const float iq = (p.metric == 0) ? warp_inv_qnorm(s_q, p.dim, lane) : 1.0f;
        if (lane == 0) *s_invq = iq;
    }
    __syncthreads();
    TopKSink sink;
    sink.scores = p.out_scores + static_cast<int64_t>(qi) * p.k;
    sink.rows = p.out_rows + static_cast<int64_t>(qi) * p.k;
    sink.n_valid = p.out_nvalid + qi;
    sink.keys = p.out_keys ? p.out_keys + static_cast<int64_t>(qi) * p.k : nullptr;
    sink.scale = *s_invq;
    sink.ex = p.ex.world > 0 ? &p.ex : nullptr;
    sink.qi = qi;
    sink.k = p.k;
    block_merge_sorted_lists<kK1MergeLPT>(lists, n_lists, p.k, p.k, sink, xchg, lane, warp, kK1Warps);
}

template <int R, int KPL>
__global__ void __launch_bounds__(kK1Threads)
k1_fallback_kernel(const __grid_constant__ FbParams p) {
    extern __shared__ float s_q[];  // [dim]
    __shared__ uint64_t xchg[2][32];
    __shared__ float s_invq;
    __shared__ int s_last;

    const int nf = *p.n_flagged;
    if (nf <= 0) return;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int chunks = p.dim >> 3;
    const bool cosine = (p.metric == 0);

    // CTAs per query: the whole grid for one query, never more lists than the scratch or the merge hold
    int64_t cpq = static_cast<int64_t>(gridDim.x) / nf;
    const int64_t fit = p.cand_lists / nf;
    if (cpq > fit) cpq = fit;
    if (cpq > kK1MaxGridX) cpq = kK1MaxGridX;
    if (cpq < 1) cpq = 1;
    const int n_cols = static_cast<int>(gridDim.x / cpq);
    const int col = static_cast<int>(blockIdx.x / cpq);
    const int x = static_cast<int>(blockIdx.x % cpq);
    if (col >= n_cols) return;

    int64_t n_items = p.n_rows;
    if (p.list) {
        const int64_t cnt = static_cast<int64_t>(*p.list_count);
        n_items = cnt < p.n_rows ? cnt : p.n_rows;
    }
    const int64_t stride = cpq * kK1Warps * R;

    for (int f = col; f < nf; f += n_cols) {
        const int slot_q = p.flagged[f];
        const int qi = p.q_map ? p.q_map[slot_q] : slot_q;
        __syncthreads();   // the previous query's merge is done with s_q
        for (int i = threadIdx.x; i < p.dim; i += kK1Threads)
            s_q[i] = p.queries[static_cast<int64_t>(qi) * p.dim + i];
        __syncthreads();

        WarpTopK<KPL> top;
        top.init(p.k, lane);
        for (int64_t item0 = (static_cast<int64_t>(x) * kK1Warps + warp) * R; item0 < n_items; item0 += stride) {
            int64_t row[R];
            bool ok[R];
            float acc[R];
#pragma unroll
            for (int j = 0; j < R; ++j) {
                ok[j] = (item0 + j) < n_items;
                row[j] = ok[j] ? (p.list ? static_cast<int64_t>(p.list[item0 + j]) : item0 + j) : 0;
                acc[j] = 0.f;
            }
            for (int ch = lane; ch < chunks; ch += 32) {
                uint4 v[R];
#pragma unroll
                for (int j = 0; j < R; ++j) v[j] = ld_stream_16(p.rows + row[j] * p.dim + ch * 8);
                float q8[8];
                const float4 a = *reinterpret_cast<const float4*>(s_q + ch * 8);
                const float4 b = *reinterpret_cast<const float4*>(s_q + ch * 8 + 4);
                q8[0] = a.x; q8[1] = a.y; q8[2] = a.z; q8[3] = a.w;
                q8[4] = b.x; q8[5] = b.y; q8[6] = b.z; q8[7] = b.w;
#pragma unroll
                for (int j = 0; j < R; ++j) acc[j] = chunk_fma(v[j], q8, acc[j]);
            }
#pragma unroll
            for (int j = 0; j < R; ++j) {
                const float invn = cosine ? __ldg(p.inv_norm + row[j]) : 1.0f;
                const float s = warp_butterfly_sum(acc[j]) * invn;
                if (ok[j] && !(cosine && invn == 0.f) && s == s) {
                    const uint64_t key = make_key(s, p.row_offset + static_cast<uint32_t>(row[j]));
                    if (key > top.thr) top.insert(key, lane);
                }
            }
        }

        uint64_t* lists = p.cand + static_cast<int64_t>(f) * cpq * p.k;
        block_extract_topk<KPL>(top, p.k, lists + static_cast<int64_t>(x) * p.k, xchg, lane, warp, kK1Warps);

        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            const unsigned t = atomicAdd(&p.counters[f], 1u);
            s_last = (t == static_cast<unsigned>(cpq) - 1u) ? 1 : 0;
        }
        __syncthreads();
        if (s_last) {
            __threadfence();
            fb_merge_query(p, qi, lists, static_cast<int>(cpq), s_q, xchg, &s_invq);
        }
    }
}


// ---------------------------------------------------------------------------------------------
// Balanced multi-group variant: "one query vector, many filters" (the graph retriever's adjacent()
// fan-out) produces single-query groups whose row sets differ by orders of magnitude -- a file's 20
// rows next to a repository's 500 000.  k1_stream_kernel gives every group the same share of the
// grid, so the big group crawls on 1/groups of the machine.  Here the groups' row sets are laid end to
// end in units of one CTA step (kK1Warps * R rows) and the 1-D grid cuts that line into equal pieces:
// CTA c owns units [c*T/C, (c+1)*T/C), which may span several groups (it then serves them one after
// the other) or be one of many pieces of a big group.  The CTAs first(g)..last(g) that touch group g
// each write one sorted list (list index c + g is collision-free) and the last to arrive merges them.
// Survivor counts are read on the device, so nothing here waits for the host.
struct BalParams {
    const __half* rows;
    const float* inv_norm;
    const float* queries;          // [batch, dim] fp32
    const uint32_t* lists;         // [n_filters, list_stride]
    const uint32_t* list_counts;   // [n_filters]
    uint64_t* cand;                // [(grid + 2 * groups), k] scratch
    unsigned* counters;            // [groups] arrival counters, zero between launches
    float* out_scores;
    int64_t* out_rows;
    int32_t* out_nvalid;
    uint64_t* out_keys;
    int64_t n_rows;
    int64_t list_stride;
    uint32_t row_offset;
    int32_t dim;
    int32_t k;
    int32_t metric;
    int32_t n_groups;
    ExchangeDev ex;
    int32_t q_idx[kMaxGroups];     // one query per group
    int32_t filter[kMaxGroups];    // -1 = whole shard, else survivor list slot
};

__device__ __noinline__ void bal_merge_query(const BalParams& p, int qi, const uint64_t* lists, int n_lists,
                                             const float* s_q, uint64_t (*xchg)[32], float* s_invq) {
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        const float iq = (p.metric == 0) ? warp_inv_qnorm(s_q, p.dim, lane) : 1.0f;
        if (lane == 0) *s_invq = iq;
    }
    __syncthreads();
    TopKSink sink;
    sink.scores = p.out_scores + static_cast<int64_t>(qi) * p.k;
    sink.rows = p.out_rows + static_cast<int64_t>(qi) * p.k;
    sink.n_valid = p.out_nvalid + qi;
    sink.keys = p.out_keys ? p.out_keys + static_cast<int64_t>(qi) * p.k : nullptr;
    sink.scale = *s_invq;
    sink.ex = p.ex.world > 0 ? &p.ex : nullptr;
    sink.qi = qi;
    sink.k = p.k;
    block_merge_sorted_lists<kK1MergeLPT>(lists, n_lists, p.k, p.k, sink, xchg, lane, warp, kK1Warps);
}

// CH = 16-byte chunks per lane (dim <= 256 * CH), all loads of a step issued before the first FMA;
// CH = 0 walks any dimension with a run-time loop.
// HC > 0 selects the half-warp layout of k1_stream_half_kernel for dim <= 512 (two rows per warp step, 16 lanes
// each, HC chunk slots per half-lane; HC = 3 means exactly 48 chunks, i.e. dim = 384, without range tests); R then
// counts row PAIRS per step and CH is ignored.  Same canonical summation order, so the same score bits.
template <int CH, int R, int KPL, int HC = 0>
__global__ void __launch_bounds__(kK1Threads, CH == 0 ? 1 : 2)
k1_balanced_kernel(const __grid_constant__ BalParams p) {
    extern __shared__ float s_q[];  // [dim]
    __shared__ uint64_t xchg[2][32];
    __shared__ int64_t s_upref[kMaxGroups + 1];   // units before group g
    __shared__ int64_t s_items[kMaxGroups];
    __shared__ float s_invq;
    __shared__ int s_last;

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int chunks = p.dim >> 3;
    const bool cosine = (p.metric == 0);
    constexpr int RW = HC > 0 ? 2 * R : R;                      // rows per warp step
    constexpr int64_t U = static_cast<int64_t>(kK1Warps) * RW;  // rows per CTA step

    if (threadIdx.x == 0) {
        int64_t acc = 0;
        for (int g = 0; g < p.n_groups; ++g) {
            int64_t n = p.n_rows;
            if (p.filter[g] >= 0) {
                const int64_t cnt = static_cast<int64_t>(p.list_counts[p.filter[g]]);
                n = cnt < p.n_rows ? cnt : p.n_rows;
            }
            s_items[g] = n;
            s_upref[g] = acc;
            acc += (n + U - 1) / U;
        }
        s_upref[p.n_groups] = acc;
    }
    __syncthreads();
    const int64_t T = s_upref[p.n_groups];
    const int64_t C = gridDim.x;
    const int64_t c = blockIdx.x;
    const int64_t my_lo = (c * T) / C, my_hi = ((c + 1) * T) / C;

    int loaded_q = -1;
    for (int g = 0; g < p.n_groups; ++g) {
        const int64_t u0 = s_upref[g], u1 = s_upref[g + 1];
        // CTAs that own a list of this group (every CTA evaluates this identically)
        int64_t first = 0, last = 0;
        if (u1 > u0) {
            first = ((u0 + 1) * C + T - 1) / T - 1;
            last = (u1 * C + T - 1) / T - 1;
        }
        const bool empty = (u1 == u0);           // no rows: CTA 0 emits "no hits" from a list of its own
        if (c < first || c > last) continue;     // uniform per CTA
        const int qi = p.q_idx[g];
        if (qi != loaded_q) {
            __syncthreads();
            for (int i = threadIdx.x; i < p.dim; i += kK1Threads)
                s_q[i] = p.queries[static_cast<int64_t>(qi) * p.dim + i];
            loaded_q = qi;
        }
        __syncthreads();

        const int64_t n_items = s_items[g];
        const uint32_t* list = p.filter[g] >= 0 ? p.lists + static_cast<int64_t>(p.filter[g]) * p.list_stride : nullptr;
        const int64_t lo = (my_lo > u0 ? my_lo : u0) - u0, hi = (my_hi < u1 ? my_hi : u1) - u0;   // units inside the group

        WarpTopK<KPL> top;
        top.init(p.k, lane);
        // survivor indices of the NEXT step are fetched (lane j holds row j's) while this step's rows stream
        uint32_t idx_next = 0u;
        if (list && lane < RW && lo < hi && lo * U + static_cast<int64_t>(warp) * RW + lane < n_items)
            idx_next = list[lo * U + static_cast<int64_t>(warp) * RW + lane];
        for (int64_t u = lo; u < hi; ++u) {
            const int64_t item0 = u * U + static_cast<int64_t>(warp) * RW;
            const uint32_t idx_cur = idx_next;
            idx_next = 0u;
            if (list && lane < RW && u + 1 < hi && item0 + U + lane < n_items) idx_next = list[item0 + U + lane];
            if constexpr (HC > 0) {
                const int half = lane >> 4, hl = lane & 15;
                int64_t hrow[R];
                bool hok[R];
#pragma unroll
                for (int j = 0; j < R; ++j) {
                    const int it = 2 * j + half;
                    hok[j] = (item0 + it) < n_items;
                    const uint32_t li = __shfl_sync(CRB_FULL_MASK, idx_cur, it);
                    hrow[j] = hok[j] ? (list ? static_cast<int64_t>(li) : item0 + it) : 0;
                }
                uint4 hv[R][HC];
#pragma unroll
                for (int j = 0; j < R; ++j) {
                    const __half* rp = p.rows + hrow[j] * p.dim;
#pragma unroll
                    for (int cc = 0; cc < HC; ++cc)
                        if (HC == 3 || hl + 16 * cc < chunks) hv[j][cc] = ld_stream_16(rp + (hl + 16 * cc) * 8);
                }
                float hinv[R];
#pragma unroll
                for (int j = 0; j < R; ++j) hinv[j] = cosine ? __ldg(p.inv_norm + hrow[j]) : 1.0f;
                float a0[R], a1[R];
#pragma unroll
                for (int j = 0; j < R; ++j) { a0[j] = 0.f; a1[j] = 0.f; }
                // canonical lane hl: chunks hl, hl+32 (slots 0, 2); canonical lane hl+16: chunks hl+16, hl+48 (slots 1, 3)
#pragma unroll
                for (int cc = 0; cc < HC; ++cc) {
                    const int slot = cc == 0 ? 0 : (cc == 1 ? 2 : (cc == 2 ? 1 : 3));   // a0's chunks first, in order
                    if (slot >= HC) continue;
                    const int ch = hl + 16 * slot;
                    if (HC == 3 || ch < chunks) {
                        float q8[8];
                        const float4 a = *reinterpret_cast<const float4*>(s_q + ch * 8);
                        const float4 b = *reinterpret_cast<const float4*>(s_q + ch * 8 + 4);
                        q8[0] = a.x; q8[1] = a.y; q8[2] = a.z; q8[3] = a.w;
                        q8[4] = b.x; q8[5] = b.y; q8[6] = b.z; q8[7] = b.w;
#pragma unroll
                        for (int j = 0; j < R; ++j) {
                            if (slot == 0 || slot == 2) a0[j] = chunk_fma(hv[j][slot], q8, a0[j]);
                            else a1[j] = chunk_fma(hv[j][slot], q8, a1[j]);
                        }
                    }
                }
#pragma unroll
                for (int j = 0; j < R; ++j) {
                    float sc = a0[j] + a1[j];                               // canonical xor-16 level
                    sc += __shfl_xor_sync(CRB_FULL_MASK, sc, 8);
                    sc += __shfl_xor_sync(CRB_FULL_MASK, sc, 4);
                    sc += __shfl_xor_sync(CRB_FULL_MASK, sc, 2);
                    sc += __shfl_xor_sync(CRB_FULL_MASK, sc, 1);
                    sc *= hinv[j];
                    const bool good = hok[j] && !(cosine && hinv[j] == 0.f) && sc == sc;
                    const uint64_t key = good ? make_key(sc, p.row_offset + static_cast<uint32_t>(hrow[j])) : 0ull;
                    if (__any_sync(CRB_FULL_MASK, key > top.thr)) {
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            const uint64_t kh = __shfl_sync(CRB_FULL_MASK, key, h * 16);
                            if (kh > top.thr) top.insert(kh, lane);
                        }
                    }
                }
                continue;
            }
            int64_t row[R];
            bool ok[R];
            float acc[R];
#pragma unroll
            for (int j = 0; j < R; ++j) {
                ok[j] = (item0 + j) < n_items;
                const uint32_t li = __shfl_sync(CRB_FULL_MASK, idx_cur, j);
                row[j] = ok[j] ? (list ? static_cast<int64_t>(li) : item0 + j) : 0;
                acc[j] = 0.f;
            }
            if constexpr (CH > 0) {
                uint4 v[R][CH];
#pragma unroll
                for (int j = 0; j < R; ++j) {
                    const __half* rp = p.rows + row[j] * p.dim;
#pragma unroll
                    for (int cc = 0; cc < CH; ++cc)
                        if (lane + 32 * cc < chunks) v[j][cc] = ld_stream_16(rp + (lane + 32 * cc) * 8);
                }
#pragma unroll
                for (int cc = 0; cc < CH; ++cc) {
                    const int ch = lane + 32 * cc;
                    if (ch < chunks) {
                        float q8[8];
                        const float4 a = *reinterpret_cast<const float4*>(s_q + ch * 8);
                        const float4 b = *reinterpret_cast<const float4*>(s_q + ch * 8 + 4);
                        q8[0] = a.x; q8[1] = a.y; q8[2] = a.z; q8[3] = a.w;
                        q8[4] = b.x; q8[5] = b.y; q8[6] = b.z; q8[7] = b.w;
#pragma unroll
                        for (int j = 0; j < R; ++j) acc[j] = chunk_fma(v[j][cc], q8, acc[j]);
                    }
                }
            } else {
                for (int ch = lane; ch < chunks; ch += 32) {
                    uint4 v[R];
#pragma unroll
                    for (int j = 0; j < R; ++j) v[j] = ld_stream_16(p.rows + row[j] * p.dim + ch * 8);
                    float q8[8];
                    const float4 a = *reinterpret_cast<const float4*>(s_q + ch * 8);
                    const float4 b = *reinterpret_cast<const float4*>(s_q + ch * 8 + 4);
                    q8[0] = a.x; q8[1] = a.y; q8[2] = a.z; q8[3] = a.w;
                    q8[4] = b.x; q8[5] = b.y; q8[6] = b.z; q8[7] = b.w;
#pragma unroll
                    for (int j = 0; j < R; ++j) acc[j] = chunk_fma(v[j], q8, acc[j]);
                }
            }
#pragma unroll
            for (int j = 0; j < R; ++j) {
                const float invn = cosine ? __ldg(p.inv_norm + row[j]) : 1.0f;
                const float s = warp_butterfly_sum(acc[j]) * invn;
                if (ok[j] && !(cosine && invn == 0.f) && s == s) {
                    const uint64_t key = make_key(s, p.row_offset + static_cast<uint32_t>(row[j]));
                    if (key > top.thr) top.insert(key, lane);
                }
            }
        }

        uint64_t* lists_g = p.cand + (empty ? C + p.n_groups + g : first + g) * p.k;
        block_extract_topk<KPL>(top, p.k, lists_g + (c - first) * p.k, xchg, lane, warp, kK1Warps);

        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            const unsigned t = atomicAdd(&p.counters[g], 1u);
            s_last = (t == static_cast<unsigned>(last - first)) ? 1 : 0;
        }
        __syncthreads();
        if (s_last) {
            __threadfence();
            bal_merge_query(p, qi, lists_g, static_cast<int>(last - first + 1), s_q, xchg, &s_invq);
            if (threadIdx.x == 0) p.counters[g] = 0u;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Fused filter + score: ONE query under ONE filter -- what every retriever.invoke(q, filter={namespace,
// repo, ...}) starts with (agent_graph.py:249) -- in a single pass and a single launch.  K3 + K1 read the
// code columns, write a survivor list, launch again, read the list and gather: two ramps, two tails, and
// neither short kernel gets near the HBM rate (ncu, 10M x 1024 at 1 %: 40 MB in 22 us + 205 MB in 52 us).
// Here a warp owns 512-row stripes: a lane reads the codes of 16 consecutive rows of each constrained key
// (one 16-byte load for an int8 column, two for int16), compares them to the wanted code four (two) at a
// time inside 32-bit words, ANDs the keys' 16-bit pass masks (and the caller's row mask), queues the
// survivors' row numbers in shared memory and scores them R at a time in the canonical summation order as
// soon as R are waiting.  The NEXT stripe's codes are in flight while the current stripe's survivors are
// gathered, so a warp's dependent round trips are the gathers only.  No list, no counters, no second
// launch: algorithmic bytes = n_rows * sum(widths of the constrained keys) + survivors * dim * 2.
constexpr int kFusedMaxKeys = 4;     // constrained keys the fused plan takes (each a column of 1 or 2 bytes)
constexpr int kFusedStripe = 512;    // rows per warp step of the scan: 16 consecutive rows per lane

struct FusedParams {
    const __half* rows;
    const float* inv_norm;
    const float* query;            // [dim] fp32
    const void* cols[kFusedMaxKeys];   // columns of the constrained keys, 16-byte aligned
    int32_t want[kFusedMaxKeys];       // wanted code of each (>= 0)
    int32_t width[kFusedMaxKeys];      // 1 or 2 bytes
    int32_t n_used;
    const uint32_t* row_mask;      // optional caller pass bits (or nullptr)
    uint64_t* cand;                // [grid, k] scratch
    unsigned* counter;             // arrival counter, zero between launches
    float* out_scores;             // [k] of this query
    int64_t* out_rows;
    int32_t* out_nvalid;
    uint64_t* out_keys;            // or nullptr
    int64_t n_rows;
    uint32_t row_offset;
    int32_t dim;
    int32_t k;
    int32_t metric;
    int32_t qi;                    // query index in the batch (exchange slot)
    ExchangeDev ex;
};

// Shared memory of the fused kernel, per warp: the survivor queue and the NEXT stripe's code words (staged with
// cp.async, so the prefetch costs no registers and the registers go to rows in flight instead).
// (Tried on top of this and dropped, profiles/README.md: prefetching the survivors' rows into L2 when they are queued and
// scoring them one stripe later -- no gain at 1 %, 30-60 % slower from 10 % of the rows passing upwards.)
// `depth` stripes of codes are in flight per warp (a ring of stages, one cp.async group each): with one stripe in flight
// the scan alone was a chain of ~8 dependent HBM round trips per warp -- a filter nothing passes took 24 us over 40 MB
// of codes (ncu), a third of the whole C4 search.
constexpr int kFusedQueue = kFusedStripe + 32;
constexpr int kFusedCodeBytes = 8192;                   // staged code words per warp: `depth` stages of slots x 512 B
constexpr size_t kFusedWarpSmem = kFusedQueue * 4 + kFusedCodeBytes;
CRB_DEVINL void cp_async_16(void* smem_dst, const void* gmem_src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst))), "l"(gmem_src) : "memory");
}
CRB_DEVINL void cp_async_wait_pending(int n) {      // wait until at most n of this thread's copy groups are pending
    switch (n) {
        case 0: asm volatile("cp.async.wait_group 0;" ::: "memory"); break;
        case 1: asm volatile("cp.async.wait_group 1;" ::: "memory"); break;
        case 2: asm volatile("cp.async.wait_group 2;" ::: "memory"); break;
        default: asm volatile("cp.async.wait_group 3;" ::: "memory"); break;
    }
}

template <int CH, int R, int KPL>
__global__ void __launch_bounds__(kK1Threads, 2)
k1_fused_filter_kernel(const __grid_constant__ FusedParams p) {
    extern __shared__ __align__(16) uint8_t fused_smem_raw[];
    __shared__ uint64_t xchg[2][32];
    __shared__ float s_invq;
    __shared__ int s_last;

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int chunks = p.dim >> 3;
    const bool cosine = (p.metric == 0);
    uint8_t* my_smem = fused_smem_raw + static_cast<size_t>(warp) * kFusedWarpSmem;
    uint32_t* queue = reinterpret_cast<uint32_t*>(my_smem);
    uint4* codes = reinterpret_cast<uint4*>(my_smem + kFusedQueue * 4);      // [depth][slots][32 lanes]
    // 16-byte slots one stripe needs per lane (an int8 column one, an int16 column two) and how many stripes fit
    int slots = 0;
#pragma unroll
    for (int u = 0; u < kFusedMaxKeys; ++u) slots += u < p.n_used ? p.width[u] : 0;
    const int depth = slots == 0 ? 1 : min(4, max(1, kFusedCodeBytes / (slots * 512)));

    float qr[CH][8];
#pragma unroll
    for (int c = 0; c < CH; ++c) {
        const int ch = lane + 32 * c;
        if (ch < chunks) {
            const float4* src = reinterpret_cast<const float4*>(p.query + ch * 8);
            const float4 a = __ldg(src), b = __ldg(src + 1);
            qr[c][0] = a.x; qr[c][1] = a.y; qr[c][2] = a.z; qr[c][3] = a.w;
            qr[c][4] = b.x; qr[c][5] = b.y; qr[c][6] = b.z; qr[c][7] = b.w;
        } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) qr[c][e] = 0.f;
        }
    }

    WarpTopK<KPL> top;
    top.init(p.k, lane);
    int queued = 0;                                   // warp-uniform

    auto score_batch = [&](int first, int n_take) {   // rows queue[first .. first + n_take), n_take <= R
        int64_t row[R];
        bool ok[R];
#pragma unroll
        for (int j = 0; j < R; ++j) {
            ok[j] = j < n_take;
            row[j] = ok[j] ? static_cast<int64_t>(queue[first + j]) : 0;
        }
        uint4 v[R][CH];
#pragma unroll
        for (int j = 0; j < R; ++j) {
            const __half* rp = p.rows + row[j] * p.dim;
#pragma unroll
            for (int c = 0; c < CH; ++c)
                if (lane + 32 * c < chunks) v[j][c] = ld_stream_16(rp + (lane + 32 * c) * 8);
        }
        float invn[R];
#pragma unroll
        for (int j = 0; j < R; ++j) invn[j] = cosine ? __ldg(p.inv_norm + row[j]) : 1.0f;
        float acc[R];
#pragma unroll
        for (int j = 0; j < R; ++j) acc[j] = 0.f;
#pragma unroll
        for (int c = 0; c < CH; ++c) {
            if (lane + 32 * c < chunks) {
#pragma unroll
                for (int j = 0; j < R; ++j) acc[j] = chunk_fma(v[j][c], qr[c], acc[j]);
            }
        }
#pragma unroll
        for (int j = 0; j < R; ++j) {
            const float s = warp_butterfly_sum(acc[j]) * invn[j];
            if (ok[j] && !(cosine && invn[j] == 0.f) && s == s) {
                const uint64_t key = make_key(s, p.row_offset + static_cast<uint32_t>(row[j]));
                if (key > top.thr) top.insert(key, lane);
            }
        }
    };

    // stage the codes of stripe `st` into ring stage `stage` (always one commit group, empty past the last stripe); a
    // lane whose 16 rows are not all inside the shard copies nothing
    auto fetch = [&](int stage, int64_t st, bool live) {
        const int64_t r0 = st * kFusedStripe + 16 * lane;
        if (live && r0 + 16 <= p.n_rows) {
            uint4* dst = codes + static_cast<size_t>(stage) * slots * 32 + lane;
            int slot = 0;
#pragma unroll
            for (int u = 0; u < kFusedMaxKeys; ++u) {
                if (u < p.n_used) {
                    if (p.width[u] == 1) {
                        cp_async_16(dst + slot * 32, static_cast<const uint8_t*>(p.cols[u]) + r0);
                        slot += 1;
                    } else {
                        const uint16_t* src = static_cast<const uint16_t*>(p.cols[u]) + r0;
                        cp_async_16(dst + slot * 32, src);
                        cp_async_16(dst + (slot + 1) * 32, src + 8);
                        slot += 2;
                    }
                }
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    // 16-bit pass mask of this lane's rows from the staged codes
    auto evaluate = [&](int stage) -> uint32_t {
        const uint4* src = codes + static_cast<size_t>(stage) * slots * 32 + lane;
        uint32_t pass = 0xffffu;
        int slot = 0;
#pragma unroll
        for (int u = 0; u < kFusedMaxKeys; ++u) {
            if (u < p.n_used) {
                const uint32_t w = static_cast<uint32_t>(p.want[u]);
                if (p.width[u] == 1) {
                    pass &= eq_mask_bytes(src[slot * 32], (w & 0xffu) * 0x01010101u);
                    slot += 1;
                } else {
                    pass &= eq_mask_halves(src[slot * 32], (w & 0xffffu) * 0x00010001u) |
                            (eq_mask_halves(src[(slot + 1) * 32], (w & 0xffffu) * 0x00010001u) << 8);
                    slot += 2;
                }
            }
        }
        return pass;
    };
    const int64_t n_stripes = (p.n_rows + kFusedStripe - 1) / kFusedStripe;
    const int64_t step = static_cast<int64_t>(gridDim.x) * kK1Warps;
    const int64_t st0 = static_cast<int64_t>(blockIdx.x) * kK1Warps + warp;
    for (int d = 0; d < depth; ++d) fetch(d, st0 + d * step, st0 + d * step < n_stripes);
    int it = 0;
    for (int64_t st = st0; st < n_stripes; st += step, ++it) {
        const int stage = it % depth;
        cp_async_wait_pending(depth - 1);                          // this stage's group is the oldest one pending
        __syncwarp();
        const int64_t r0 = st * kFusedStripe + 16 * lane;
        const uint32_t pass_all = evaluate(stage);
        uint32_t pass = r0 + 16 <= p.n_rows ? pass_all : 0u;      // the shard's last n_rows % 16 rows: after the loop
        __syncwarp();
        fetch(stage, st + depth * step, st + depth * step < n_stripes);    // refill the stage: `depth` stripes ahead
        if (p.row_mask && pass) {
            // the lane's 16 rows are one half of a mask word (r0 is a multiple of 16)
            pass &= (__ldg(p.row_mask + (r0 >> 5)) >> (r0 & 31)) & 0xffffu;
        }
        // queue the survivors (order is irrelevant: keys carry the row), scoring whenever R are waiting
        const int mine = __popc(pass);
        int incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int vv = __shfl_up_sync(CRB_FULL_MASK, incl, o);
            if (lane >= o) incl += vv;
        }
        const int total = __shfl_sync(CRB_FULL_MASK, incl, 31);
        if (total == 0) continue;
        int at = queued + incl - mine;
        while (pass) {
            const int j = __ffs(pass) - 1;
            pass &= pass - 1u;
            queue[at++] = static_cast<uint32_t>(r0 + j);
        }
        queued += total;
        __syncwarp();
        // drain from the front in whole batches; what is left (< R rows) moves to the front of the queue
        int head = 0;
        while (queued - head >= R) {
            score_batch(head, R);
            head += R;
        }
        const int rest = queued - head;           // < R <= 32
        const uint32_t keep = lane < rest ? queue[head + lane] : 0u;
        __syncwarp();
        if (lane < rest) queue[lane] = keep;
        queued = rest;
        __syncwarp();
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    // The shard's last n_rows % 16 rows belong to no whole 16-row lane group: one warp tests them, a row per lane.
    // (As an element-wise pass over a whole ragged stripe inside the loop this was a chain of dependent loads.)
    if (blockIdx.x == 0 && warp == 0) {
        const int rem = static_cast<int>(p.n_rows & 15);
        const int64_t r = p.n_rows - rem + lane;
        bool ok = lane < rem;
        for (int u = 0; u < p.n_used; ++u) {
            if (lane < rem) {
                const int32_t c = p.width[u] == 1 ? static_cast<int32_t>(__ldg(static_cast<const int8_t*>(p.cols[u]) + r))
                                                  : static_cast<int32_t>(__ldg(static_cast<const int16_t*>(p.cols[u]) + r));
                ok = ok && (c == p.want[u]);
            }
        }
        if (ok && p.row_mask) ok = (__ldg(p.row_mask + (r >> 5)) >> (r & 31)) & 1u;
        const uint32_t b = __ballot_sync(CRB_FULL_MASK, ok);
        if (ok) queue[queued + __popc(b & ((1u << lane) - 1u))] = static_cast<uint32_t>(r);
        queued += __popc(b);
        __syncwarp();
    }
    for (int head = 0; head < queued; head += R) score_batch(head, min(R, queued - head));

    uint64_t* out = p.cand + static_cast<int64_t>(blockIdx.x) * p.k;
    block_extract_topk<KPL>(top, p.k, out, xchg, lane, warp, kK1Warps);

    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned t = atomicAdd(p.counter, 1u);
        s_last = (t == gridDim.x - 1) ? 1 : 0;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();

    if (warp == 0) {
        const float iq = cosine ? warp_inv_qnorm(p.query, p.dim, lane) : 1.0f;
        if (lane == 0) s_invq = iq;
    }
    __syncthreads();
    TopKSink sink;
    sink.scores = p.out_scores;
    sink.rows = p.out_rows;
    sink.n_valid = p.out_nvalid;
    sink.keys = p.out_keys;
    sink.scale = s_invq;
    sink.ex = p.ex.world > 0 ? &p.ex : nullptr;
    sink.qi = p.qi;
    sink.k = p.k;
    block_merge_sorted_lists<kK1MergeLPT>(p.cand, gridDim.x, p.k, p.k, sink, xchg, lane, warp, kK1Warps);
    if (threadIdx.x == 0) *p.counter = 0u;
}

}  // namespace crb
