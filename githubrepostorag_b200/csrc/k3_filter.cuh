// K3 -- metadata filter: AND of equalities over dictionary-coded columns -> survivor lists.
//
// Replaces the SAI predicates `metadata_s['k'] = 'v' AND ...` that cassio adds to the ANN
// query (filters built at rag_worker/src/worker/services/agent_graph.py:198-239,357-364,
// 532-534; traversal edges add one more equality per adjacent search,
// graph_rag_retrievers.py:82-100).  One pass over the code columns evaluates up to 32
// filters at once (one pass-bit per filter per row) and appends the surviving local row
// indices of each filter to its list; K1/K2 then touch only surviving rows.
//
// Code columns are as narrow as their dictionary allows (1, 2 or 4 bytes per row: `namespace`,
// `branch`, `component_kind`, `scope` have a handful of values, `repo` thousands), one column per
// key, and a thread reads 16 consecutive rows of a column with 16-byte loads.
// Bound: HBM.  Algorithmic bytes = n_rows * sum(width of the used keys) + survivors * 4 (lists).
#pragma once
#include "common.cuh"

namespace crb {

constexpr int kK3Threads = 256;
constexpr int kK3Warps = kK3Threads / 32;
constexpr int kK3RowsPerThread = 16;     // consecutive rows: one 16-byte load of an int8 column

struct K3Params {
    const void* cols[kMaxKeys];   // column of key m: [n_rows] codes of width[m] bytes, -1 = key absent
    uint8_t width[kMaxKeys];      // 1, 2 or 4
    uint8_t vec_ok[kMaxKeys];     // column start is 16-byte aligned: vector loads
    uint32_t* lists;       // [n_filters, list_stride]
    uint32_t* counts;      // [n_filters], zero on entry
    uint32_t* masks;       // [n_filters, mask_stride] one pass bit per row (for K2), or nullptr
    int64_t mask_stride;
    int64_t n_rows;
    int64_t list_stride;
    int32_t n_keys;
    int32_t n_filters;
    int32_t any_of;        // 1: the filters are ALTERNATIVES -- a row that passes any of them goes to list 0 (once)
    uint32_t used_keys;    // bit m set when any filter constrains key m
    int32_t n_used;        // the constrained keys, compacted (host-filled)
    uint8_t ukeys[kMaxKeys];
    int32_t want[kMaxFilters][kMaxKeys];
    const uint32_t* extra[kMaxFilters];   // optional caller-built pass-bit mask ANDed into filter f (or nullptr)
};

// 16 consecutive codes of one column, sign-extended to int32 (the "missing" code -1 survives narrowing).
CRB_DEVINL void k3_load_codes(const void* col, int width, bool vec, int64_t row0, int n_ok, int32_t (&code)[kK3RowsPerThread]) {
    if (vec && n_ok == kK3RowsPerThread) {
        if (width == 1) {
            const uint4 v = __ldg(reinterpret_cast<const uint4*>(static_cast<const int8_t*>(col) + row0));
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int j = 0; j < 16; ++j) code[j] = static_cast<int32_t>(static_cast<int8_t>((w[j >> 2] >> (8 * (j & 3))) & 0xffu));
        } else if (width == 2) {
            const uint4* p = reinterpret_cast<const uint4*>(static_cast<const int16_t*>(col) + row0);
            const uint4 a = __ldg(p), b = __ldg(p + 1);
            const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
            for (int j = 0; j < 16; ++j) code[j] = static_cast<int32_t>(static_cast<int16_t>((w[j >> 1] >> (16 * (j & 1))) & 0xffffu));
        } else {
            const int4* p = reinterpret_cast<const int4*>(static_cast<const int32_t*>(col) + row0);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int4 v = __ldg(p + i);
                code[4 * i] = v.x; code[4 * i + 1] = v.y; code[4 * i + 2] = v.z; code[4 * i + 3] = v.w;
            }
        }
        return;
    }
#pragma unroll
    for (int j = 0; j < kK3RowsPerThread; ++j) {
        int32_t c = -1;
        if (j < n_ok) {
            if (width == 1) c = __ldg(static_cast<const int8_t*>(col) + row0 + j);
            else if (width == 2) c = __ldg(static_cast<const int16_t*>(col) + row0 + j);
            else c = __ldg(static_cast<const int32_t*>(col) + row0 + j);
        }
        code[j] = c;
    }
}

__global__ void __launch_bounds__(kK3Threads) k3_filter_kernel(const __grid_constant__ K3Params p) {
    __shared__ uint32_t s_cnt[kMaxFilters];
    __shared__ uint32_t s_base[kMaxFilters];
    __shared__ uint32_t s_wbase[kK3Warps][kMaxFilters];

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int64_t row0 = (static_cast<int64_t>(blockIdx.x) * kK3Threads + threadIdx.x) * kK3RowsPerThread;
    const int64_t left = p.n_rows - row0;
    const int n_ok = left >= kK3RowsPerThread ? kK3RowsPerThread : (left > 0 ? static_cast<int>(left) : 0);
    const uint32_t all = p.n_filters >= 32 ? 0xffffffffu : ((1u << p.n_filters) - 1u);

    if (threadIdx.x < kMaxFilters) s_cnt[threadIdx.x] = 0u;

    // pass[j]: bit f set when row row0 + j survives filter f
    uint32_t pass[kK3RowsPerThread];
#pragma unroll
    for (int j = 0; j < kK3RowsPerThread; ++j) pass[j] = j < n_ok ? all : 0u;
    // One filter (every plain filtered query): the 16 codes of a narrow column are compared inside their 32-bit
    // words (4 bytes / 2 halves at a time), all keys' loads issued before the first compare.  The general path
    // below unpacks every code and costs ~40 instructions per row: it made this kernel instruction-bound (ncu, 10M
    // rows x int8/int16/int8: 40 MB in 22 us, issue slots 61 % busy, DRAM 28 %).
    bool fast = p.n_filters == 1 && n_ok == kK3RowsPerThread;
    for (int m = 0; m < p.n_keys && fast; ++m)
        if (((p.used_keys >> m) & 1u) && (!p.vec_ok[m] || p.width[m] > 2 || p.want[0][m] < 0 ||
                                          p.want[0][m] > (p.width[m] == 1 ? 127 : 32767))) fast = false;
    fast = __all_sync(CRB_FULL_MASK, fast);
    if (fast) {
        uint32_t pm = 0xffffu;
        uint4 v[4][2];
        for (int u0 = 0; u0 < p.n_used; u0 += 4) {
            // up to four constrained keys per round: loads first, compares after
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (u0 + u < p.n_used) {
                    const int mm = p.ukeys[u0 + u];
                    if (p.width[mm] == 1) {
                        v[u][0] = __ldg(reinterpret_cast<const uint4*>(static_cast<const uint8_t*>(p.cols[mm]) + row0));
                    } else {
                        const uint4* src = reinterpret_cast<const uint4*>(static_cast<const uint16_t*>(p.cols[mm]) + row0);
                        v[u][0] = __ldg(src);
                        v[u][1] = __ldg(src + 1);
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (u0 + u < p.n_used) {
                    const int mm = p.ukeys[u0 + u];
                    const uint32_t w = static_cast<uint32_t>(p.want[0][mm]);
                    if (p.width[mm] == 1) pm &= eq_mask_bytes(v[u][0], (w & 0xffu) * 0x01010101u);
                    else pm &= eq_mask_halves(v[u][0], (w & 0xffffu) * 0x00010001u) | (eq_mask_halves(v[u][1], (w & 0xffffu) * 0x00010001u) << 8);
                }
            }
        }
#pragma unroll
        for (int j = 0; j < kK3RowsPerThread; ++j) pass[j] = (pm >> j) & 1u;
    } else {
        for (int m = 0; m < p.n_keys; ++m) {
            if (!((p.used_keys >> m) & 1u)) continue;
            int32_t code[kK3RowsPerThread];
            k3_load_codes(p.cols[m], p.width[m], p.vec_ok[m] != 0, row0, n_ok, code);
            for (int f = 0; f < p.n_filters; ++f) {
                const int32_t w = p.want[f][m];
                if (w == -1) continue;  // CRB_FILT_ANY
#pragma unroll
                for (int j = 0; j < kK3RowsPerThread; ++j)
                    if (code[j] != w) pass[j] &= ~(1u << f);
            }
        }
    }
    // caller-built row masks (membership in list-valued metadata, IN lists, OR ...): one bit per row; the
    // thread's 16 rows are one half of a mask word
    for (int f = 0; f < p.n_filters; ++f) {
        const uint32_t* ex = p.extra[f];
        if (!ex || n_ok == 0) continue;
        const uint32_t half = (__ldg(ex + (row0 >> 5)) >> (row0 & 31)) & 0xffffu;
#pragma unroll
        for (int j = 0; j < kK3RowsPerThread; ++j)
            if (!((half >> j) & 1u)) pass[j] &= ~(1u << f);
    }
    // alternatives (one query vector, the union of several edges' row sets -- the adapter's adjacent()): one list
    const int n_out = p.any_of ? 1 : p.n_filters;
    if (p.any_of) {
#pragma unroll
        for (int j = 0; j < kK3RowsPerThread; ++j) pass[j] = pass[j] ? 1u : 0u;
    }
    __syncthreads();

    // per filter: this thread's 16-bit survivor mask; the warp's total claims the warp's place in the block
    for (int f = 0; f < n_out; ++f) {
        uint32_t pm = 0u;
#pragma unroll
        for (int j = 0; j < kK3RowsPerThread; ++j) pm |= ((pass[j] >> f) & 1u) << j;
        const uint32_t cnt = __popc(pm);
        const uint32_t tot = __reduce_add_sync(CRB_FULL_MASK, cnt);
        if (lane == 0) s_wbase[warp][f] = tot ? atomicAdd(&s_cnt[f], tot) : 0u;
        // the pass-bit words of K2's masks: two lanes share a word
        if (p.masks) {
            const uint32_t hi = __shfl_down_sync(CRB_FULL_MASK, pm, 1);
            if (!(lane & 1) && row0 < p.n_rows) p.masks[static_cast<int64_t>(f) * p.mask_stride + (row0 >> 5)] = pm | (hi << 16);
        }
    }
    __syncthreads();
    if (threadIdx.x < n_out) {
        const uint32_t c = s_cnt[threadIdx.x];
        s_base[threadIdx.x] = c ? atomicAdd(&p.counts[threadIdx.x], c) : 0u;
    }
    __syncthreads();

    for (int f = 0; f < n_out; ++f) {
        uint32_t pm = 0u;
#pragma unroll
        for (int j = 0; j < kK3RowsPerThread; ++j) pm |= ((pass[j] >> f) & 1u) << j;
        if (__ballot_sync(CRB_FULL_MASK, pm != 0u) == 0u) continue;
        // exclusive prefix of the lanes' survivor counts
        const uint32_t cnt = __popc(pm);
        uint32_t incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(CRB_FULL_MASK, incl, o);
            if (lane >= o) incl += v;
        }
        uint32_t off = s_base[f] + s_wbase[warp][f] + incl - cnt;
        uint32_t* list = p.lists + static_cast<int64_t>(f) * p.list_stride;
        while (pm) {
            const int j = __ffs(pm) - 1;
            pm &= pm - 1u;
            list[off++] = static_cast<uint32_t>(row0 + j);
        }
    }
}

// ---- ingest-side conversion ------------------------------------------------------------------
// fp32 rows -> fp16 rows (round-to-nearest-even) + inverse norm of the *rounded* row.  One warp
// per row, lane-strided 8-element chunks, sequential sum of squares, butterfly.
__global__ void __launch_bounds__(256) encode_rows_kernel(const float* __restrict__ src,
                                                          __half* __restrict__ dst,
                                                          float* __restrict__ inv_norm,
                                                          int64_t n_rows, int dim) {
    const int lane = threadIdx.x & 31;
    const int64_t gw = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    const int chunks = dim >> 3;
    for (int64_t row = gw; row < n_rows; row += nw) {
        float acc = 0.f;
        for (int ch = lane; ch < chunks; ch += 32) {
            const float4* s = reinterpret_cast<const float4*>(src + row * dim + ch * 8);
            const float4 a = s[0], b = s[1];
            __half2 h[4];
            h[0] = __floats2half2_rn(a.x, a.y);
            h[1] = __floats2half2_rn(a.z, a.w);
            h[2] = __floats2half2_rn(b.x, b.y);
            h[3] = __floats2half2_rn(b.z, b.w);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float2 f = __half22float2(h[e]);
                acc = fmaf(f.x, f.x, acc);
                acc = fmaf(f.y, f.y, acc);
            }
            *reinterpret_cast<uint4*>(dst + row * dim + ch * 8) = *reinterpret_cast<const uint4*>(h);
        }
        acc = warp_butterfly_sum(acc);
        if (lane == 0) inv_norm[row] = acc > 0.f ? __fdiv_rn(1.0f, __fsqrt_rn(acc)) : 0.f;
    }
}

__global__ void __launch_bounds__(256) row_inv_norms_kernel(const __half* __restrict__ rows,
                                                            float* __restrict__ inv_norm,
                                                            int64_t n_rows, int dim) {
    const int lane = threadIdx.x & 31;
    const int64_t gw = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    const int chunks = dim >> 3;
    for (int64_t row = gw; row < n_rows; row += nw) {
        float acc = 0.f;
        for (int ch = lane; ch < chunks; ch += 32) {
            const uint4 v = ld_stream_16(rows + row * dim + ch * 8);
            const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float2 f = __half22float2(h[e]);
                acc = fmaf(f.x, f.x, acc);
                acc = fmaf(f.y, f.y, acc);
            }
        }
        acc = warp_butterfly_sum(acc);
        if (lane == 0) inv_norm[row] = acc > 0.f ? __fdiv_rn(1.0f, __fsqrt_rn(acc)) : 0.f;
    }
}

// Deleted rows: the inverse norm becomes NaN and so does the first 16-byte chunk of the row itself, so
// every score of the row is NaN under either metric (cosine: NaN scale; inner product: NaN dot) and
// fails the `s == s` / `s > threshold` admission tests of every kernel at no cost to the scan loops.
__global__ void __launch_bounds__(256) tombstone_rows_kernel(float* __restrict__ inv_norm, __half* __restrict__ rows,
                                                             int dim, const int64_t* __restrict__ idx, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        inv_norm[idx[i]] = __uint_as_float(0x7fc00000u);
        *reinterpret_cast<uint4*>(rows + idx[i] * dim) = make_uint4(0x7e007e00u, 0x7e007e00u, 0x7e007e00u, 0x7e007e00u);
    }
}

// Stored rows -> fp32 (the `embedding` member of LCCassandra's (Document, embedding, id)).
__global__ void __launch_bounds__(256) fetch_rows_kernel(const __half* __restrict__ rows,
                                                         const int64_t* __restrict__ idx, int n,
                                                         int dim, float* __restrict__ out) {
    const int64_t total = static_cast<int64_t>(n) * dim;
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int r = static_cast<int>(i / dim);
        const int c = static_cast<int>(i - static_cast<int64_t>(r) * dim);
        out[i] = __half2float(rows[idx[r] * dim + c]);
    }
}

// The stored vectors of a finished search's hits, fp32: out[i] = row (result_rows[i] - row_offset), zeros where the
// result slot is padding (-1).  Lets the (Document, embedding, id) triples leave with the search's own copy.
__global__ void __launch_bounds__(256) fetch_result_rows_kernel(const __half* __restrict__ rows,
                                                                const int64_t* __restrict__ result_rows, int n,
                                                                int64_t row_offset, int dim, float* __restrict__ out) {
    const int64_t total = static_cast<int64_t>(n) * dim;
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int r = static_cast<int>(i / dim);
        const int c = static_cast<int>(i - static_cast<int64_t>(r) * dim);
        const int64_t g = result_rows[r];
        out[i] = g >= 0 ? __half2float(rows[(g - row_offset) * dim + c]) : 0.f;
    }
}

}  // namespace crb
