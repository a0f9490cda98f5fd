// Shared device helpers for the retrieval kernels (sm_100a only).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#define CRB_DEVINL __device__ __forceinline__
#define CRB_FULL_MASK 0xffffffffu

namespace crb {

constexpr int kMaxKeys = 16;     // CRB_MAX_KEYS
constexpr int kMaxTopK = 128;    // CRB_MAX_TOPK
constexpr int kMaxGroups = 64;   // query groups per launch
constexpr int kMaxFilters = 32;  // distinct filters per launch (one u32 pass-bit each in K3)
static_assert(kMaxKeys == 16, "kMaxKeys must equal CRB_MAX_KEYS of include/coderag_b200.h");

// ---- order-preserving (score, row) keys ------------------------------------------------
// key = ord(score) << 32 | global_row.  Larger key == better hit; among equal scores the
// larger row wins, which is the row the reference heap keeps at the k-th boundary
// (oracle/simple_vector_store.py::get_top_k_embeddings).  key 0 is "empty".
CRB_DEVINL uint32_t f32_to_ord(float f) {
    uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
CRB_DEVINL float ord_to_f32(uint32_t o) {
    uint32_t b = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
    return __uint_as_float(b);
}
CRB_DEVINL uint64_t make_key(float score, uint32_t global_row) {
    return (static_cast<uint64_t>(f32_to_ord(score)) << 32) | global_row;
}
CRB_DEVINL float key_score(uint64_t key) { return ord_to_f32(static_cast<uint32_t>(key >> 32)); }
CRB_DEVINL uint32_t key_row(uint64_t key) { return static_cast<uint32_t>(key); }

// ---- warp-wide u64 max / min through redux.sync (two 32-bit reductions) ------------------
CRB_DEVINL uint64_t warp_max_u64(uint64_t v) {
    uint32_t hi = static_cast<uint32_t>(v >> 32);
    uint32_t mh = __reduce_max_sync(CRB_FULL_MASK, hi);
    uint32_t lo = (hi == mh) ? static_cast<uint32_t>(v) : 0u;
    uint32_t ml = __reduce_max_sync(CRB_FULL_MASK, lo);
    return (static_cast<uint64_t>(mh) << 32) | ml;
}
CRB_DEVINL uint64_t warp_min_u64(uint64_t v) {
    uint32_t hi = static_cast<uint32_t>(v >> 32);
    uint32_t mh = __reduce_min_sync(CRB_FULL_MASK, hi);
    uint32_t lo = (hi == mh) ? static_cast<uint32_t>(v) : 0xffffffffu;
    uint32_t ml = __reduce_min_sync(CRB_FULL_MASK, lo);
    return (static_cast<uint64_t>(mh) << 32) | ml;
}

// ---- streaming loads ---------------------------------------------------------------------
// Corpus rows are read exactly once per pass: bypass L1 allocation.
CRB_DEVINL uint4 ld_stream_16(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
CRB_DEVINL uint64_t ld_cg_u64(const uint64_t* p) {
    uint64_t r;
    asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(r) : "l"(p));
    return r;
}

// ---- in-word equality tests over narrow code columns ----------------------------------------
// 16-bit mask of the bytes (16-bit halves) of `v` equal to the replicated wanted code: bit j <-> element j.
// A zero byte of x = w ^ want becomes 0x80 exactly (no borrow between bytes), the four flags are gathered by a
// multiply whose partial products cannot collide.
CRB_DEVINL uint32_t eq_mask_bytes(const uint4& v, uint32_t want4) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    uint32_t m = 0u;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t x = w[i] ^ want4;
        const uint32_t t = ~((((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x) | 0x7f7f7f7fu);      // 0x80 where the byte of x is 0
        m |= (((t >> 7) * 0x01020408u) >> 24) << (4 * i);
    }
    return m;
}
CRB_DEVINL uint32_t eq_mask_halves(const uint4& v, uint32_t want2) {      // 8 elements -> 8 bits
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    uint32_t m = 0u;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t x = w[i] ^ want2;
        const uint32_t t = ~((((x & 0x7fff7fffu) + 0x7fff7fffu) | x) | 0x7fff7fffu);      // 0x8000 where the half of x is 0
        m |= (((t >> 15) & 1u) | ((t >> 30) & 2u)) << (2 * i);
    }
    return m;
}

// ---- the canonical exact score ------------------------------------------------------------
// Every score this library returns is produced by this one summation order, whichever kernel
// found the candidate and whichever shard holds the row: lane l of a warp accumulates 16-byte
// chunks l, l+32, l+64, ... with sequential fp32 FMAs (fp16 element upcast exactly), then the
// 32 partial sums are combined by an xor-butterfly (16, 8, 4, 2, 1).  fp16 x fp32 products are
// rounded once by the FMA.
CRB_DEVINL float chunk_fma(const uint4& v, const float* __restrict__ q8, float acc) {
    const __half2* h = reinterpret_cast<const __half2*>(&v);
    float2 a = __half22float2(h[0]);
    float2 b = __half22float2(h[1]);
    float2 c = __half22float2(h[2]);
    float2 d = __half22float2(h[3]);
    acc = fmaf(q8[0], a.x, acc);
    acc = fmaf(q8[1], a.y, acc);
    acc = fmaf(q8[2], b.x, acc);
    acc = fmaf(q8[3], b.y, acc);
    acc = fmaf(q8[4], c.x, acc);
    acc = fmaf(q8[5], c.y, acc);
    acc = fmaf(q8[6], d.x, acc);
    acc = fmaf(q8[7], d.y, acc);
    return acc;
}
CRB_DEVINL float warp_butterfly_sum(float v) {
    v += __shfl_xor_sync(CRB_FULL_MASK, v, 16);
    v += __shfl_xor_sync(CRB_FULL_MASK, v, 8);
    v += __shfl_xor_sync(CRB_FULL_MASK, v, 4);
    v += __shfl_xor_sync(CRB_FULL_MASK, v, 2);
    v += __shfl_xor_sync(CRB_FULL_MASK, v, 1);
    return v;
}

// Canonical inverse query norm (one warp): lane-strided sequential sum of squares, butterfly,
// IEEE sqrt and divide.  A zero query gives 0 (its cosine scores are reported as 0).
CRB_DEVINL float warp_inv_qnorm(const float* __restrict__ q, int dim, int lane) {
    float acc = 0.f;
    for (int i = lane; i < dim; i += 32) acc = fmaf(q[i], q[i], acc);
    acc = warp_butterfly_sum(acc);
    return acc > 0.f ? __fdiv_rn(1.0f, __fsqrt_rn(acc)) : 0.f;
}

}  // namespace crb
