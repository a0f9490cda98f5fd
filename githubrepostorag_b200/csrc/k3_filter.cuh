// K3 -- metadata filter: AND of equalities over dictionary-coded columns -> survivor lists.
//
// Replaces the SAI predicates `metadata_s['k'] = 'v' AND ...` that cassio adds to the ANN
// query (filters built at rag_worker/src/worker/services/agent_graph.py:198-239,357-364,
// 532-534; traversal edges add one more equality per adjacent search,
// graph_rag_retrievers.py:82-100).  One pass over the code columns evaluates up to 32
// filters at once (one pass-bit per filter per row) and appends the surviving local row
// indices of each filter to its list; K1/K2 then touch only surviving rows.
// Bound: HBM.  Algorithmic bytes = n_rows * used_keys * 4 (codes) + survivors * 4 (lists).
#pragma once
#include "common.cuh"

namespace crb {

constexpr int kK3Threads = 256;
constexpr int kK3Warps = kK3Threads / 32;
constexpr int kK3RowsPerThread = 8;

struct K3Params {
    const int32_t* codes;  // [n_keys, n_rows]
    uint32_t* lists;       // [n_filters, list_stride]
    uint32_t* counts;      // [n_filters], zero on entry
    uint32_t* masks;       // [n_filters, mask_stride] one pass bit per row (for K2), or nullptr
    int64_t mask_stride;
    int64_t n_rows;
    int64_t code_stride;   // elements between consecutive key columns (>= n_rows)
    int64_t list_stride;
    int32_t n_keys;
    int32_t n_filters;
    uint32_t used_keys;    // bit m set when any filter constrains key m
    int32_t want[kMaxFilters][kMaxKeys];
    const uint32_t* extra[kMaxFilters];   // optional caller-built pass-bit mask ANDed into filter f (or nullptr)
};

__global__ void __launch_bounds__(kK3Threads) k3_filter_kernel(const __grid_constant__ K3Params p) {
    __shared__ uint32_t s_cnt[kMaxFilters];
    __shared__ uint32_t s_base[kMaxFilters];
    __shared__ uint32_t s_wbase[kK3Warps][kMaxFilters];

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int64_t blk0 = static_cast<int64_t>(blockIdx.x) * (kK3Threads * kK3RowsPerThread);
    const uint32_t all = p.n_filters >= 32 ? 0xffffffffu : ((1u << p.n_filters) - 1u);

    if (threadIdx.x < kMaxFilters) s_cnt[threadIdx.x] = 0u;

    // pass[j]: bit f set when row j of this thread survives filter f
    uint32_t pass[kK3RowsPerThread];
#pragma unroll
    for (int j = 0; j < kK3RowsPerThread; ++j) {
        const int64_t row = blk0 + j * kK3Threads + threadIdx.x;
        pass[j] = row < p.n_rows ? all : 0u;
    }
    for (int m = 0; m < p.n_keys; ++m) {
        if (!((p.used_keys >> m) & 1u)) continue;
        int32_t code[kK3RowsPerThread];
#pragma unroll
        for (int j = 0; j < kK3RowsPerThread; ++j) {
            const int64_t row = blk0 + j * kK3Threads + threadIdx.x;
            code[j] = row < p.n_rows ? __ldg(p.codes + static_cast<int64_t>(m) * p.code_stride + row) : -1;
        }
        for (int f = 0; f < p.n_filters; ++f) {
            const int32_t w = p.want[f][m];
            if (w == -1) continue;  // CRB_FILT_ANY
#pragma unroll
            for (int j = 0; j < kK3RowsPerThread; ++j)
                if (code[j] != w) pass[j] &= ~(1u << f);
        }
    }
    // caller-built row masks (membership in list-valued metadata, IN lists, ...): one bit per row
    for (int f = 0; f < p.n_filters; ++f) {
        const uint32_t* ex = p.extra[f];
        if (!ex) continue;
#pragma unroll
        for (int j = 0; j < kK3RowsPerThread; ++j) {
            const int64_t row = blk0 + j * kK3Threads + threadIdx.x;
            if (row < p.n_rows && !((__ldg(ex + (row >> 5)) >> (row & 31)) & 1u)) pass[j] &= ~(1u << f);
        }
    }
    __syncthreads();

    // count per filter: warp totals -> block totals -> one global atomic per filter per block
    for (int f = 0; f < p.n_filters; ++f) {
        uint32_t cnt = 0;
#pragma unroll
        for (int j = 0; j < kK3RowsPerThread; ++j)
            cnt += __popc(__ballot_sync(CRB_FULL_MASK, (pass[j] >> f) & 1u));
        if (lane == 0) s_wbase[warp][f] = cnt ? atomicAdd(&s_cnt[f], cnt) : 0u;
    }
    __syncthreads();
    if (threadIdx.x < p.n_filters) {
        const uint32_t c = s_cnt[threadIdx.x];
        s_base[threadIdx.x] = c ? atomicAdd(&p.counts[threadIdx.x], c) : 0u;
    }
    __syncthreads();

    const uint32_t lt = (1u << lane) - 1u;
    for (int f = 0; f < p.n_filters; ++f) {
        uint32_t off = s_base[f] + s_wbase[warp][f];
        uint32_t* list = p.lists + static_cast<int64_t>(f) * p.list_stride;
#pragma unroll
        for (int j = 0; j < kK3RowsPerThread; ++j) {
            const uint32_t bit = (pass[j] >> f) & 1u;
            const uint32_t b = __ballot_sync(CRB_FULL_MASK, bit);
            const int64_t row = blk0 + j * kK3Threads + threadIdx.x;
            if (bit) list[off + __popc(b & lt)] = static_cast<uint32_t>(row);
            // the warp's 32 rows are consecutive and 32-aligned: one mask word per ballot
            if (p.masks && lane == 0 && row < p.n_rows) p.masks[static_cast<int64_t>(f) * p.mask_stride + (row >> 5)] = b;
            off += __popc(b);
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

}  // namespace crb
