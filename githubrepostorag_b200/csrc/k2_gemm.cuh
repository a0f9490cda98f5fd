// K2 -- tensor-core candidate generation: TMA-staged tcgen05.mma fp16 Q x C^T with the
// accumulators in TMEM and a fused per-query top-k' epilogue; K5 -- exact rescore.
//
// Replaces the same Cassandra ANN scoring as K1 (graph_rag_retrievers.py:104-134) for query
// batches: the graph-traversal fan-out and query expansion of the reference produce natural
// batches (agent_graph.py:249-278).  The B x N score matrix is never
// materialised: each epilogue thread owns one query row of the 128 x 256 accumulator tile,
// scales by the row's inverse norm, and keeps a private k'-entry candidate list in shared
// memory across all corpus tiles its CTA visits.
//
// K2 scores are APPROXIMATE (the query is rounded to fp16 after normalisation); K5 re-scores
// every candidate with the library's canonical fp32 order (common.cuh) and proves the result
// exact with a margin test, flagging the rare query that needs the K1 fallback:
//   any row NOT kept by a full slot list has approx <= tau_slot, hence exact <= tau_max + eps,
//   eps = ||q_hat - fp16(q_hat)||_2 (Cauchy-Schwarz) + the worst-case rounding of the two evaluations
//   being compared (k2_rounding_allowance: linear in dim; cosine units, scaled by the largest row
//   norm for inner product).  The exact top-k is proven when its k-th score exceeds tau_max + eps.
//
// Bound: tensor pipe for batch >= ~256 (2*B*N*d flop), HBM below that (N*d*2 bytes).
#pragma once
#include <cuda.h>

#include "topk.cuh"

namespace crb {

constexpr int kK2BlockM = 128;   // queries per tile (UMMA M)
constexpr int kK2BlockN = 256;   // corpus rows per tile (UMMA N)
constexpr int kK2BlockK = 64;    // fp16 elements per k-block = one 128 B swizzle row
constexpr int kK2UmmaK = 16;
constexpr int kK2Threads = 192;  // warp 0 TMA, warp 1 MMA + TMEM alloc, warps 2..5 epilogue
constexpr int kK2StageBytesA = kK2BlockM * kK2BlockK * 2;   // 16 KB
constexpr int kK2StageBytesB = kK2BlockN * kK2BlockK * 2;   // 32 KB
constexpr int kK2StageBytes = kK2StageBytesA + kK2StageBytesB;
constexpr int kK2TmemCols = 512;                            // 2 accumulator buffers x 256 columns
constexpr int kK2MaxMTiles = 16;

struct K2Params {
    const float* inv_norm;      // [n_rows]
    const uint32_t* mask;       // [ceil(n_rows/32)] pass bits or nullptr
    uint64_t* cand;             // [batch, slots, kprime] approximate keys (local row in low 32 bits)
    float* tau;                 // [batch, slots] smallest key score of each list that filled up (-inf otherwise)
    uint32_t* gthr;             // [batch, gthr_groups] shared threshold classes per query (ord(score), 0 = none yet) or nullptr
    int32_t gthr_groups;        // power of two in [16, 128], >= k
    int32_t gthr_every;         // a list re-reads its query's classes every this many corpus tiles (0 = k2_gthr_due's schedule)
    int32_t gthr_prepass;       // read the first accumulator tile of a list twice (see k2_first_visit)
    int64_t n_rows;
    int32_t batch;
    int32_t dim;
    int32_t kprime;
    int32_t metric;
    int32_t m_tiles;
    int32_t slots;              // CTAs per m-tile; grid = m_tiles * slots
    int32_t stages;
    int32_t prefetch;           // CRB_TUNING builds only: L2-prefetch the next corpus tile
    int32_t debug;              // CRB_TUNING builds only (timing experiments that change results): 1 = no TMA after the first fill, 2 = no epilogue math
    int32_t lstride;            // candidate-list stride = query rows that own a list (<= 128)
    int32_t m_inner;            // pair kernel: query-tile pairs served per corpus tile by one CTA pair (>= 1)
};

// Timing experiments (stale tiles, no epilogue, L2 prefetch) exist only in -DCRB_TUNING builds; the shipped
// library has none of those branches.
#ifdef CRB_TUNING
#define CRB_K2_DEBUG(p) ((p).debug)
#define CRB_K2_PREFETCH(p) ((p).prefetch)
#else
#define CRB_K2_DEBUG(p) 0
#define CRB_K2_PREFETCH(p) 0
#endif

// ---- PTX wrappers ------------------------------------------------------------------------------
CRB_DEVINL uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

CRB_DEVINL void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
CRB_DEVINL void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
CRB_DEVINL void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
CRB_DEVINL void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra WAIT_DONE;\n\t"
        "bra WAIT_LOOP;\n\t"
        "WAIT_DONE:\n\t"
        "}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
CRB_DEVINL void tma_load_2d(const CUtensorMap* map, void* smem_dst, uint64_t* bar, int32_t c0, int32_t c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
CRB_DEVINL void tma_prefetch_l2_2d(const CUtensorMap* map, int32_t c0, int32_t c1) {
    asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];"
                 ::"l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1) : "memory");
}
CRB_DEVINL void tmem_alloc(uint32_t* smem_dst, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(cols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
}
CRB_DEVINL void tmem_dealloc(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols));
}
CRB_DEVINL void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
CRB_DEVINL void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
CRB_DEVINL void umma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
CRB_DEVINL void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
CRB_DEVINL void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr));
}
CRB_DEVINL void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
CRB_DEVINL void epi_bar_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// K-major, 128-byte-swizzled shared-memory matrix descriptor (sm_100 format): start address and
// offsets in 16-byte units; SBO = 1024 B between 8-row groups; LBO unused for swizzled K-major;
// version 1; layout type 2 = SWIZZLE_128B.
CRB_DEVINL uint64_t umma_desc_sw128(uint32_t smem_addr) {
    return static_cast<uint64_t>((smem_addr >> 4) & 0x3fffu) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// kind::f16 instruction descriptor: D = f32, A = B = f16, both K-major, M = 128, N = 256.
constexpr uint32_t kK2IDesc = (1u << 4) | ((kK2BlockN >> 3) << 17) | ((kK2BlockM >> 4) << 24);

// Insert `key` into a thread-private k'-entry candidate list (entry i at list[i*128]); returns the
// new {admission threshold bits, count}.  The list fills as a plain array, is turned into a
// binary MIN-heap when it reaches k' entries, and from then on an admitted key replaces the root
// and sifts down (O(log k')).  Kept out of line: the epilogue has 32 call sites per chunk and the
// call is rare after warm-up.
CRB_DEVINL void k2_sift_down(uint64_t* list, int ls, int i, int n, uint64_t key) {
    for (;;) {
        int c = 2 * i + 1;
        if (c >= n) break;
        uint64_t cv = list[c * ls];
        if (c + 1 < n) {
            const uint64_t rv = list[(c + 1) * ls];
            if (rv < cv) { cv = rv; ++c; }
        }
        if (cv >= key) break;
        list[i * ls] = cv;
        i = c;
    }
    list[i * ls] = key;
}
CRB_DEVINL uint2 k2_list_insert(uint64_t* list, uint64_t key, uint32_t state, float thr, int kprime, int ls) {
    int count = static_cast<int>(state);
    if (count < kprime) {
        list[count * ls] = key;
        ++count;
        if (count == kprime) {
            for (int i = kprime / 2 - 1; i >= 0; --i) k2_sift_down(list, ls, i, kprime, list[i * ls]);
            thr = key_score(list[0]);
        }
    } else {
        k2_sift_down(list, ls, 0, kprime, key);
        thr = key_score(list[0]);
    }
    return make_uint2(__float_as_uint(thr), static_cast<uint32_t>(count));
}

CRB_DEVINL float4 lds_f4(uint32_t saddr) {
    float4 r;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(saddr));
    return r;
}

// s[j] for a run-time j without spilling the array: a 5-level select tree (31 selects).
CRB_DEVINL float k2_pick(const float (&s)[32], int j) {
    float a[16], b[8], c[4], d[2];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = (j & 16) ? s[i + 16] : s[i];
#pragma unroll
    for (int i = 0; i < 8; ++i) b[i] = (j & 8) ? a[i + 8] : a[i];
#pragma unroll
    for (int i = 0; i < 4; ++i) c[i] = (j & 4) ? b[i + 4] : b[i];
#pragma unroll
    for (int i = 0; i < 2; ++i) d[i] = (j & 2) ? c[i + 2] : c[i];
    return (j & 1) ? d[1] : d[0];
}

// Examine one 32-column chunk of this thread's accumulator row: scale, and admit columns that
// beat the thread's threshold.  `inv_saddr` is the shared-window address of the chunk's scales.
// Admissions run in warp-lockstep rounds (every lane that still has a pending column inserts one
// per round), so the heap updates of different lanes overlap instead of serialising.
CRB_DEVINL float k2_scan_chunk(const uint32_t (&v)[32], uint32_t inv_saddr, float thr, uint32_t& state,
                               uint64_t* my_list, int64_t n_first, int kprime, int ls, float& item_best) {
    float s[32];
#pragma unroll
    for (int g = 0; g < 8; ++g) {
        const float4 w = lds_f4(inv_saddr + g * 16);
        s[4 * g + 0] = __uint_as_float(v[4 * g + 0]) * w.x;
        s[4 * g + 1] = __uint_as_float(v[4 * g + 1]) * w.y;
        s[4 * g + 2] = __uint_as_float(v[4 * g + 2]) * w.z;
        s[4 * g + 3] = __uint_as_float(v[4 * g + 3]) * w.w;
    }
    // fmaxf drops NaNs (masked / out-of-range columns), so `best` is the largest real score
    float best = s[0];
#pragma unroll
    for (int j = 1; j < 32; ++j) best = fmaxf(best, s[j]);
    item_best = fmaxf(item_best, best);
    if (__any_sync(CRB_FULL_MASK, best > thr)) {
        uint32_t hit = 0u;
#pragma unroll
        for (int j = 0; j < 32; ++j) hit |= (s[j] > thr) ? (1u << j) : 0u;
        while (__any_sync(CRB_FULL_MASK, hit != 0u)) {
            if (hit != 0u) {
                const int j = __ffs(hit) - 1;
                hit &= hit - 1u;
                const float sj = k2_pick(s, j);
                if (sj > thr) {     // the threshold may have risen since the mask was built
                    const uint2 r = k2_list_insert(my_list, make_key(sj, static_cast<uint32_t>(n_first + j)), state, thr, kprime, ls);
                    thr = fmaxf(__uint_as_float(r.x), thr);   // never below the shared threshold folded in earlier
                    state = r.y;
                }
            }
        }
    }
    return thr;
}

// Shared admission threshold.  Every CTA (pair) that serves a query keeps its own k'-list, so on its own
// each list would admit ~k'(1 + ln(rows seen / k')) columns, and every admission costs its whole warp a
// round -- at B = 1024 that was a quarter of the kernel on a 1.25M-row shard.  The lists of one query
// therefore share a lower bound of the query's k-th best score: corpus tile t belongs to class t mod G
// (G >= k classes), each thread raises its tile's class to the tile's best score with a fire-and-forget
// atomic max, and the bound is the MINIMUM over the classes -- G distinct rows (one per class, tiles are
// disjoint) score at least that much, so the k-th best does too.  A thread re-reads the G words of its
// query when it starts a tile (the loads fly during the wait for the accumulator) and drops every column
// at or below the bound without touching its list.  K5 folds the final bound into tau_max, so the proof
// of exactness covers these drops exactly like the drops of a full list.
CRB_DEVINL float k2_gthr_float(uint32_t o) { return o ? ord_to_f32(o) : __uint_as_float(0xff800000u); }
CRB_DEVINL uint32_t k2_gthr_min(const uint32_t* g, int groups) {
    uint32_t m = 0xffffffffu;
    for (int i = 0; i < groups; i += 4) {
        uint32_t a, b, c, d;
        asm volatile("ld.global.cg.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "l"(g + i));
        m = min(min(m, a), min(b, min(c, d)));
    }
    return m;
}
// The first 16 classes are fetched into registers BEFORE the thread waits for its accumulator and folded
// after, so the L2 round trip overlaps the wait (exposed, it cost 10 % of an HBM-bound B = 64 search);
// classes beyond 16 (k > 12) are read in place, every groups/16-th tile only.
CRB_DEVINL void k2_gthr_issue16(const uint32_t* g, uint32_t (&r)[16]) {
#pragma unroll
    for (int i = 0; i < 16; i += 4)
        asm volatile("ld.global.cg.v4.u32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(r[i]), "=r"(r[i + 1]), "=r"(r[i + 2]), "=r"(r[i + 3]) : "l"(g + i));
}
CRB_DEVINL uint32_t k2_gthr_fold16(const uint32_t (&r)[16]) {
    uint32_t m = r[0];
#pragma unroll
    for (int i = 1; i < 16; ++i) m = min(m, r[i]);
    return m;
}
// When a list re-reads the bound.  The bound moves like an order statistic -- fast over the first tiles,
// then logarithmically -- so visits are dense at first and then four per octave (8, 10, 12, 14, 16, 20, ...):
// ~4 log2(tiles) reads instead of one per tile (per-tile reads of lines all CTAs share cost an HBM-bound
// B = 64 search 10 %, a fixed every-8th cadence cost a B = 1024 search of a 1.25M-row shard 10 %).
CRB_DEVINL bool k2_gthr_due(int visit, int every) {
    if (every > 0) return (visit % every) == 0;
    if (visit < 8) return true;
    const int hb = 31 - __clz(visit);
    return (visit & ((1 << (hb - 2)) - 1)) == 0;
}
CRB_DEVINL void k2_gthr_publish(uint32_t* g, int groups, int64_t tile, float item_best) {
    atomicMax(g + (tile & (groups - 1)), f32_to_ord(item_best));
}

// First visit of a list: every CTA starts with an empty list and no bound, and filling a list from
// arbitrary columns costs ~k'(1 + ln(256/k')) admissions per thread that are almost all evicted later.
// So the first accumulator tile is read twice: a max-only pass publishes the tile's best score, a short
// bounded poll lets the other CTAs' first tiles land (they start together; slots >= classes fills every
// class), and the normal pass then runs under a bound that only a handful of columns clear.
CRB_DEVINL float k2_first_visit(uint32_t t_row, uint32_t inv_s, int n_chunks, bool g_live, uint32_t* my_g, int groups,
                                int64_t tile, float thr) {
    float pb = __uint_as_float(0xff800000u);
    uint32_t v[32];
#pragma unroll 1
    for (int c = 0; c < n_chunks; ++c) {
        tmem_ld32(t_row + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int g = 0; g < 8; ++g) {
            const float4 w = lds_f4(inv_s + c * 128 + g * 16);
            pb = fmaxf(pb, fmaxf(fmaxf(__uint_as_float(v[4 * g + 0]) * w.x, __uint_as_float(v[4 * g + 1]) * w.y),
                                 fmaxf(__uint_as_float(v[4 * g + 2]) * w.z, __uint_as_float(v[4 * g + 3]) * w.w)));
        }
    }
    if (g_live && pb > __uint_as_float(0xff800000u)) k2_gthr_publish(my_g, groups, tile, pb);
    uint32_t gm = 0u;
#pragma unroll 1
    for (int tries = 0; tries < 8; ++tries) {
        gm = g_live ? k2_gthr_min(my_g, groups) : 1u;
        if (__all_sync(CRB_FULL_MASK, gm != 0u)) break;
        __nanosleep(256);
    }
    return g_live ? fmaxf(thr, k2_gthr_float(gm)) : thr;
}

// Per-column scale of corpus row n: inverse norm (cosine) or 1 (inner product); NaN = "this column is not a
// hit" (past the end, filtered out, deleted or zero-norm) -- NaN never passes `>`.
CRB_DEVINL float k2_column_scale(const K2Params& p, int64_t n) {
    const float qnan = __uint_as_float(0x7fc00000u);
    if (n >= p.n_rows) return qnan;
    if (p.mask && !((p.mask[n >> 5] >> (n & 31)) & 1u)) return qnan;
    const float in = __ldg(p.inv_norm + n);          // NaN = deleted row
    if (p.metric != 0) return in == in ? 1.0f : qnan;
    return in > 0.f ? in : qnan;
}

// One accumulator tile (128 query rows x 256 corpus rows) through the epilogue, for the thread that owns query
// row `qrow` (TMEM lane quarter*32 + lane): fold the shared bound, wait for the tile, scan it into the thread's
// list.  Shared by the single-CTA and the CTA-pair kernel; the caller stages the column scales in `inv`,
// releases the accumulator buffer afterwards and keeps `thr` / `state` between tiles.
//   visit       how many corpus tiles this list has seen (schedules the re-reads of the bound)
//   first_pass  true on the list's first tile (read twice, see k2_first_visit)
CRB_DEVINL void k2_epilogue_tile(const K2Params& p, const float* inv, uint64_t* tfull, uint32_t tfull_parity,
                                 uint32_t t_row, uint64_t* my_list, float& thr, uint32_t& state, int qrow,
                                 bool q_live, int visit, bool first_pass, int64_t tile) {
    const int64_t n0 = tile * kK2BlockN;
    uint32_t* my_g = p.gthr + static_cast<int64_t>(qrow) * p.gthr_groups;
    const bool g_live = p.gthr && q_live;
    const bool g_read = g_live && k2_gthr_due(visit, p.gthr_every);
    uint32_t gr[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) gr[i] = 0u;
    if (g_read) k2_gthr_issue16(my_g, gr);
    epi_bar_sync();
    mbar_wait(tfull, tfull_parity);
    tc_fence_after();
    uint32_t g_ord = k2_gthr_fold16(gr);
    if (g_read && p.gthr_groups > 16) g_ord = min(g_ord, k2_gthr_min(my_g + 16, p.gthr_groups - 16));
    thr = fmaxf(thr, k2_gthr_float(g_ord));
    float item_best = __uint_as_float(0xff800000u);
    // TMEM -> registers one 32-column chunk ahead of the chunk being examined.  The common case per chunk is:
    // 8 broadcast LDS.128 of the column scales, 32 FMUL, a max tree and ONE branch; only a chunk that holds a
    // new candidate walks its 32 columns.
    const uint32_t inv_s = smem_u32(inv);
    const int n_chunks = (CRB_K2_DEBUG(p) & 2) ? 0 : kK2BlockN / 32;
    if (p.gthr && p.gthr_prepass && first_pass)
        thr = k2_first_visit(t_row, inv_s, n_chunks, g_live, my_g, p.gthr_groups, tile, thr);
    const float thr0 = thr;
    uint32_t va[32], vb[32];
    if (n_chunks > 0) {
        tmem_ld32(t_row, va);
        tmem_ld_wait();
    }
#pragma unroll 1
    for (int c = 0; c < n_chunks; c += 2) {
        tmem_ld32(t_row + (c + 1) * 32, vb);
        thr = k2_scan_chunk(va, inv_s + c * 128, thr, state, my_list, n0 + c * 32, p.kprime, p.lstride, item_best);
        tmem_ld_wait();
        if (c + 2 < n_chunks) tmem_ld32(t_row + (c + 2) * 32, va);
        thr = k2_scan_chunk(vb, inv_s + (c + 1) * 128, thr, state, my_list, n0 + (c + 1) * 32, p.kprime, p.lstride, item_best);
        tmem_ld_wait();
    }
    tc_fence_before();
    if (g_live && item_best > thr0) k2_gthr_publish(my_g, p.gthr_groups, tile, item_best);
}

// Flush a thread's list: the candidates, and tau = the best score the list may have dropped (the root of a
// full min-heap; -inf while the list never filled).
CRB_DEVINL void k2_flush_list(const K2Params& p, const uint64_t* my_list, int count, int qrow, int slot) {
    uint64_t* out = p.cand + (static_cast<int64_t>(qrow) * p.slots + slot) * p.kprime;
    for (int i = 0; i < p.kprime; ++i) out[i] = i < count ? my_list[i * p.lstride] : 0ull;
    p.tau[static_cast<int64_t>(qrow) * p.slots + slot] =
        count == p.kprime ? key_score(my_list[0]) : __uint_as_float(0xff800000u);
}

// ---- query preparation -------------------------------------------------------------------------
// Worst-case distance between the two evaluations K5's proof compares, for unit-norm operands (|sum of
// |q_i x_i|| <= 1 by Cauchy-Schwarz), NOT counting the fp16 rounding of the query (that is the other term of eps):
//  * tensor core: the fp16 x fp16 products are exact; the fp32 accumulator takes dim/16 additions of a
//    16-term partial sum, each allowed 2 ulp (2^-22 relative: the adder aligns and TRUNCATES, so the error is
//    biased and adds up linearly, not as sqrt(steps)) of a running sum of magnitude <= 1:  (dim/16) * 2^-22;
//  * canonical fp32 score: dim/32 sequential FMAs per lane + 5 butterfly adds + the norm multiplies, each
//    rounded to nearest (2^-24 relative of a partial sum <= 1):  (dim/32 + 8) * 2^-24.
// dim = 384: 5.9e-6, 768: 1.3e-5, 1024: 1.8e-5, 4096: 6.9e-5.  tests/test_gpu_k2.py measures the observed
// distance on fp16-exact queries (where it is the ONLY error) and asserts it stays under a quarter of this.
CRB_DEVINL float k2_rounding_allowance(int dim) {
    return static_cast<float>(dim / 16 + 1) * 2.384185791015625e-07f + static_cast<float>(dim / 32 + 8) * 5.9604644775390625e-08f;
}

// q_hat = q / ||q|| (canonical norm), q16 = fp16(q_hat); eps = ||q_hat - q16||_2 + k2_rounding_allowance(dim).
__global__ void __launch_bounds__(128) k2_prep_queries_kernel(const float* __restrict__ queries,
                                                              const int32_t* __restrict__ q_map, int n,
                                                              int dim, __half* __restrict__ q16,
                                                              float* __restrict__ invq,
                                                              float* __restrict__ eps) {
    __shared__ float s_iq;
    __shared__ float s_part[4];
    const int slot = blockIdx.x;
    if (slot >= n) {
        // padding rows up to a whole query tile: zeros in memory, so the TMA never has to synthesise
        // out-of-bounds rows (a B = 8 search spent 25 % longer per corpus tile filling 120 of them)
        for (int i = threadIdx.x; i < dim; i += blockDim.x) q16[static_cast<int64_t>(slot) * dim + i] = __float2half_rn(0.f);
        return;
    }
    const float* q = queries + static_cast<int64_t>(q_map ? q_map[slot] : slot) * dim;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (warp == 0) {
        const float iq = warp_inv_qnorm(q, dim, lane);
        if (lane == 0) s_iq = iq;
    }
    __syncthreads();
    const float iq = s_iq;
    float err = 0.f;
    for (int i = threadIdx.x; i < dim; i += blockDim.x) {
        const float qh = q[i] * iq;
        const __half h = __float2half_rn(qh);
        q16[static_cast<int64_t>(slot) * dim + i] = h;
        const float d = qh - __half2float(h);
        err = fmaf(d, d, err);
    }
    err = warp_butterfly_sum(err);
    if (lane == 0) s_part[warp] = err;
    __syncthreads();
    if (threadIdx.x == 0) {
        const float e2 = s_part[0] + s_part[1] + s_part[2] + s_part[3];
        invq[slot] = iq;
        eps[slot] = sqrtf(e2) * 1.0001f + k2_rounding_allowance(dim);
    }
}

// ---- the GEMM + top-k' kernel ------------------------------------------------------------------
// grid = m_tiles * slots persistent CTAs (1 per SM).  CTA c serves query tile c % m_tiles and
// corpus tiles  c / m_tiles + i * slots.
__global__ void __launch_bounds__(kK2Threads, 1)
k2_gemm_topk_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_c,
                    const __grid_constant__ K2Params p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // carve: [stages x (A | B)] [lists kprime x 128 u64] [s_inv 2 x 256 f32] [barriers] [tmem ptr]
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* stage_base = smem;
    uint64_t* lists = reinterpret_cast<uint64_t*>(smem + static_cast<size_t>(p.stages) * kK2StageBytes);
    float* s_inv = reinterpret_cast<float*>(lists + static_cast<size_t>(p.kprime) * p.lstride);
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_inv + 2 * kK2BlockN);
    uint64_t* full_bar = bars;                       // [stages]
    uint64_t* empty_bar = bars + p.stages;           // [stages]
    uint64_t* tfull_bar = bars + 2 * p.stages;       // [2]
    uint64_t* tempty_bar = bars + 2 * p.stages + 2;  // [2]
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * p.stages + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int m_tile = blockIdx.x % p.m_tiles;
    const int slot = blockIdx.x / p.m_tiles;
    const int64_t n_tiles = (p.n_rows + kK2BlockN - 1) / kK2BlockN;
    const int k_blocks = (p.dim + kK2BlockK - 1) / kK2BlockK;

    if (threadIdx.x == 0) {
        for (int s = 0; s < p.stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&tfull_bar[b], 1);
            mbar_init(&tempty_bar[b], 128);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) tmem_alloc(tmem_ptr, kK2TmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            int st = 0;
            uint32_t ph = 0;
            for (int64_t t = slot; t < n_tiles; t += p.slots) {
                // One CTA of each slot pulls the slot's NEXT corpus tile into L2 a whole tile-time
                // ahead, so the demand loads below (and those of the other query tiles) hit L2.
                if (CRB_K2_PREFETCH(p) && m_tile == 0 && t + p.slots < n_tiles) {
                    for (int kb = 0; kb < k_blocks; ++kb)
                        tma_prefetch_l2_2d(&map_c, kb * kK2BlockK, static_cast<int32_t>((t + p.slots) * kK2BlockN));
                }
                for (int kb = 0; kb < k_blocks; ++kb) {
                    mbar_wait(&empty_bar[st], ph ^ 1u);
                    uint8_t* a = stage_base + static_cast<size_t>(st) * kK2StageBytes;
                    uint8_t* b = a + kK2StageBytesA;
                    if ((CRB_K2_DEBUG(p) & 1) && (t != slot || kb >= p.stages)) {
                        mbar_arrive(&full_bar[st]);
                    } else {
                        mbar_expect_tx(&full_bar[st], kK2StageBytes);
                        tma_load_2d(&map_q, a, &full_bar[st], kb * kK2BlockK, m_tile * kK2BlockM);
                        tma_load_2d(&map_c, b, &full_bar[st], kb * kK2BlockK, static_cast<int32_t>(t * kK2BlockN));
                    }
                    if (++st == p.stages) { st = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            int st = 0;
            uint32_t ph = 0;
            int64_t it = 0;
            for (int64_t t = slot; t < n_tiles; t += p.slots, ++it) {
                const int buf = static_cast<int>(it & 1);
                const uint32_t use = static_cast<uint32_t>(it >> 1);
                mbar_wait(&tempty_bar[buf], (use & 1u) ^ 1u);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(buf * kK2BlockN);
                for (int kb = 0; kb < k_blocks; ++kb) {
                    mbar_wait(&full_bar[st], ph);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(stage_base + static_cast<size_t>(st) * kK2StageBytes);
                    const uint32_t b_addr = a_addr + kK2StageBytesA;
#pragma unroll
                    for (int k = 0; k < kK2BlockK / kK2UmmaK; ++k) {
                        const uint64_t da = umma_desc_sw128(a_addr + k * kK2UmmaK * 2);
                        const uint64_t db = umma_desc_sw128(b_addr + k * kK2UmmaK * 2);
                        umma_f16(d_tmem, da, db, kK2IDesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    umma_commit(&empty_bar[st]);   // frees the smem stage when these MMAs retire
                    if (++st == p.stages) { st = 0; ph ^= 1u; }
                }
                umma_commit(&tfull_bar[buf]);      // accumulator tile complete
            }
        }
    } else {
        // ===== epilogue: 4 warps = 128 TMEM lanes = 128 query rows =====
        const int quarter = warp & 3;                 // TMEM lane quarter this warp may access
        const int row = quarter * 32 + lane;          // query row inside the tile = list column of this thread
        const int tid_e = threadIdx.x - 64;           // 0..127 (staging role only)
        const int qrow = m_tile * kK2BlockM + row;
        const bool q_valid = qrow < p.batch && row < p.lstride;
        uint64_t* my_list = lists + row;              // entry i at my_list[i * lstride]
        uint32_t state = 0;                           // entries in the list
        float thr = q_valid ? __uint_as_float(0xff800000u) : __uint_as_float(0x7f800000u);  // -inf / +inf
        const float qnan = __uint_as_float(0x7fc00000u);
        float pre[2] = {qnan, qnan};

        int64_t it = 0;
        for (int64_t t = slot; t < n_tiles; t += p.slots, ++it) {
            const int buf = static_cast<int>(it & 1);
            const uint32_t use = static_cast<uint32_t>(it >> 1);
            const int64_t n0 = t * kK2BlockN;
            // The column scales for THIS tile were fetched from global memory during the previous tile.
            float* inv = s_inv + buf * kK2BlockN;
            if (it == 0) { pre[0] = k2_column_scale(p, n0 + tid_e); pre[1] = k2_column_scale(p, n0 + tid_e + 128); }
            inv[tid_e] = pre[0];
            inv[tid_e + 128] = pre[1];
            if (t + p.slots < n_tiles) {
                const int64_t nn = (t + p.slots) * kK2BlockN;
                pre[0] = k2_column_scale(p, nn + tid_e);
                pre[1] = k2_column_scale(p, nn + tid_e + 128);
            }
            const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) +
                                   static_cast<uint32_t>(buf * kK2BlockN);
            k2_epilogue_tile(p, inv, &tfull_bar[buf], use & 1u, t_row, my_list, thr, state, qrow, q_valid,
                             static_cast<int>(it), it == 0, t);
            mbar_arrive(&tempty_bar[buf]);
        }
        if (q_valid) k2_flush_list(p, my_list, static_cast<int>(state), qrow, slot);
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kK2TmemCols);
    }
}

// ---- the CTA-pair variant (cta_group::2) ---------------------------------------------------------
// Two CTAs of a cluster (one TPC) share one 256 x 256 MMA: each CTA stages its own 128 query rows
// (A) and HALF of the corpus tile (128 rows of B), so a CTA moves 32 KB per k-block instead of
// 48 KB for the same flops and the ring holds 6 stages instead of 4.  The leader (cluster rank 0)
// issues tcgen05.mma.cta_group::2; each CTA's TMEM receives the accumulator rows of its own
// queries, so the epilogue is the single-CTA one.
//   full[st]   (leader's, count 2): leader arrive.expect_tx(64 KB) + peer remote arrive; both
//              CTAs' TMA loads complete_tx on it
//   empty[st]  (own, count 1):     leader's tcgen05.commit multicast to both CTAs
//   tfull[b]   (own, count 1):     leader's tcgen05.commit multicast to both CTAs
//   tempty[b]  (leader's, count 256): all epilogue threads of both CTAs
constexpr int kK2PairStageBytesB = (kK2BlockN / 2) * kK2BlockK * 2;      // 16 KB: this CTA's half of B
constexpr int kK2PairStageBytes = kK2StageBytesA + kK2PairStageBytesB;   // 32 KB
constexpr uint32_t kK2PairIDesc = (1u << 4) | ((kK2BlockN >> 3) << 17) | ((256u >> 4) << 24);

CRB_DEVINL uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
CRB_DEVINL uint32_t mapa_u32(uint32_t saddr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
CRB_DEVINL void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
CRB_DEVINL void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
CRB_DEVINL void tma_load_2d_pair(const CUtensorMap* map, void* smem_dst, uint32_t bar_cluster_addr, int32_t c0, int32_t c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
        : "memory");
}
CRB_DEVINL void tmem_alloc_pair(uint32_t* smem_dst, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(cols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::);
}
CRB_DEVINL void tmem_dealloc_pair(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols));
}
CRB_DEVINL void umma_f16_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
CRB_DEVINL void umma_commit_pair(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"(static_cast<uint16_t>(3)) : "memory");
}

// grid = 2 * m_tiles * slots CTAs in clusters of 2.  Pair P = blockIdx.x / 2 belongs to query group
// P % m_tiles and corpus slot P / m_tiles.  A group is m_inner consecutive query-tile pairs (256
// queries each); the pair walks corpus tiles  slot + i * slots  and, for EACH tile, runs the MMA once
// per query-tile pair of its group ("m-inner" order).  With m_inner covering the whole batch every
// corpus tile is fetched from HBM by exactly one pair and its re-reads for the other query tiles
// follow within microseconds, i.e. hit L2 (with m_inner = 1, pairs of different query groups share
// tiles only by luck: ncu showed 3x the corpus in DRAM reads).
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kK2Threads, 1)
k2_gemm_topk_pair_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_c,
                         const __grid_constant__ K2Params p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int MI = p.m_inner;
    uint8_t* stage_base = smem;
    uint64_t* lists = reinterpret_cast<uint64_t*>(smem + static_cast<size_t>(p.stages) * kK2PairStageBytes);
    const size_t list_elems = static_cast<size_t>(p.kprime) * p.lstride;        // one query tile's lists
    float* s_inv = reinterpret_cast<float*>(lists + static_cast<size_t>(MI) * list_elems);
    float* s_thr = s_inv + 2 * kK2BlockN;                                       // [MI][128]
    uint32_t* s_state = reinterpret_cast<uint32_t*>(s_thr + MI * 128);          // [MI][128]
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_state + MI * 128);
    uint64_t* full_bar = bars;
    uint64_t* empty_bar = bars + p.stages;
    uint64_t* tfull_bar = bars + 2 * p.stages;
    uint64_t* tempty_bar = bars + 2 * p.stages + 2;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * p.stages + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = (rank == 0);
    const int pair = blockIdx.x >> 1;
    const int m_group = pair % p.m_tiles;
    const int slot = pair / p.m_tiles;
    const int64_t n_tiles = (p.n_rows + kK2BlockN - 1) / kK2BlockN;
    const int k_blocks = (p.dim + kK2BlockK - 1) / kK2BlockK;
    // first query row of this CTA for inner index j
    auto m_row0 = [&](int j) { return (m_group * MI + j) * 256 + static_cast<int>(rank) * kK2BlockM; };

    if (threadIdx.x == 0) {
        for (int s = 0; s < p.stages; ++s) {
            mbar_init(&full_bar[s], 2);
            mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&tfull_bar[b], 1);
            mbar_init(&tempty_bar[b], 256);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) tmem_alloc_pair(tmem_ptr, kK2TmemCols);
    tc_fence_before();
    cluster_sync_all();          // barriers of BOTH CTAs are initialised before anyone signals them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        // ===== TMA producer (one per CTA): own A rows + own half of the corpus tile =====
        if (lane == 0) {
            int st = 0;
            uint32_t ph = 0;
            for (int64_t t = slot; t < n_tiles; t += p.slots) {
                for (int j = 0; j < MI; ++j) {
                    const int a_row = m_row0(j);
                    for (int kb = 0; kb < k_blocks; ++kb) {
                        mbar_wait(&empty_bar[st], ph ^ 1u);
                        uint8_t* a = stage_base + static_cast<size_t>(st) * kK2PairStageBytes;
                        uint8_t* b = a + kK2StageBytesA;
                        const uint32_t full_leader = mapa_u32(smem_u32(&full_bar[st]), 0);
                        if ((CRB_K2_DEBUG(p) & 1) && (t != slot || j != 0 || kb >= p.stages)) {
                            // timing experiment: no TMA after the first fill (MMA runs on stale tiles)
                            if (leader) mbar_arrive(&full_bar[st]); else mbar_arrive_cluster(full_leader);
                        } else {
                            if (leader) mbar_expect_tx(&full_bar[st], 2 * kK2PairStageBytes);
                            tma_load_2d_pair(&map_q, a, full_leader, kb * kK2BlockK, a_row);
                            tma_load_2d_pair(&map_c, b, full_leader, kb * kK2BlockK,
                                             static_cast<int32_t>(t * kK2BlockN + rank * (kK2BlockN / 2)));
                            if (!leader) mbar_arrive_cluster(full_leader);
                        }
                        if (++st == p.stages) { st = 0; ph ^= 1u; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: the leader CTA only =====
        if (leader && lane == 0) {
            int st = 0;
            uint32_t ph = 0;
            int64_t it = 0;
            for (int64_t t = slot; t < n_tiles; t += p.slots) {
                for (int j = 0; j < MI; ++j, ++it) {
                    const int buf = static_cast<int>(it & 1);
                    const uint32_t use = static_cast<uint32_t>(it >> 1);
                    mbar_wait(&tempty_bar[buf], (use & 1u) ^ 1u);
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(buf * kK2BlockN);
                    for (int kb = 0; kb < k_blocks; ++kb) {
                        mbar_wait(&full_bar[st], ph);
                        tc_fence_after();
                        const uint32_t a_addr = smem_u32(stage_base + static_cast<size_t>(st) * kK2PairStageBytes);
                        const uint32_t b_addr = a_addr + kK2StageBytesA;
#pragma unroll
                        for (int k = 0; k < kK2BlockK / kK2UmmaK; ++k) {
                            const uint64_t da = umma_desc_sw128(a_addr + k * kK2UmmaK * 2);
                            const uint64_t db = umma_desc_sw128(b_addr + k * kK2UmmaK * 2);
                            umma_f16_pair(d_tmem, da, db, kK2PairIDesc, (kb | k) != 0 ? 1u : 0u);
                        }
                        umma_commit_pair(&empty_bar[st]);
                        if (++st == p.stages) { st = 0; ph ^= 1u; }
                    }
                    umma_commit_pair(&tfull_bar[buf]);
                }
            }
        }
    } else {
        // ===== epilogue: this CTA's 128 query rows of each query tile of the group =====
        const int quarter = warp & 3;
        const int row = quarter * 32 + lane;
        const int tid_e = threadIdx.x - 64;
        const float qnan = __uint_as_float(0x7fc00000u);
        for (int j = 0; j < MI; ++j) {
            const bool v = (m_row0(j) + row) < p.batch && row < p.lstride;
            s_thr[j * 128 + row] = v ? __uint_as_float(0xff800000u) : __uint_as_float(0x7f800000u);   // -inf / +inf
            s_state[j * 128 + row] = 0u;
        }
        float pre[2] = {qnan, qnan};
        int64_t it = 0;
        for (int64_t t = slot; t < n_tiles; t += p.slots) {
            const int64_t n0 = t * kK2BlockN;
            for (int j = 0; j < MI; ++j, ++it) {
                const int buf = static_cast<int>(it & 1);
                const uint32_t use = static_cast<uint32_t>(it >> 1);
                float* inv = s_inv + buf * kK2BlockN;
                if (it == 0) { pre[0] = k2_column_scale(p, n0 + tid_e); pre[1] = k2_column_scale(p, n0 + tid_e + 128); }
                inv[tid_e] = pre[0];
                inv[tid_e + 128] = pre[1];
                // scales of the NEXT item's corpus tile (the same tile until j wraps)
                const int64_t nt = (j + 1 < MI) ? t : t + p.slots;
                if (nt < n_tiles && nt != t) {
                    pre[0] = k2_column_scale(p, nt * kK2BlockN + tid_e);
                    pre[1] = k2_column_scale(p, nt * kK2BlockN + tid_e + 128);
                }
                uint64_t* my_list = lists + static_cast<size_t>(j) * list_elems + row;
                float thr = s_thr[j * 128 + row];
                uint32_t state = s_state[j * 128 + row];
                const int qrow = m_row0(j) + row;
                const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) +
                                       static_cast<uint32_t>(buf * kK2BlockN);
                k2_epilogue_tile(p, inv, &tfull_bar[buf], use & 1u, t_row, my_list, thr, state, qrow,
                                 qrow < p.batch && row < p.lstride, static_cast<int>(it / MI), it < MI, t);
                mbar_arrive_cluster(mapa_u32(smem_u32(&tempty_bar[buf]), 0));
                s_thr[j * 128 + row] = thr;
                s_state[j * 128 + row] = state;
            }
        }
        for (int j = 0; j < MI; ++j) {
            const int qrow = m_row0(j) + row;
            if (qrow < p.batch && row < p.lstride)
                k2_flush_list(p, lists + static_cast<size_t>(j) * list_elems + row, static_cast<int>(s_state[j * 128 + row]),
                              qrow, slot);
        }
    }

    tc_fence_before();
    cluster_sync_all();          // the peer may still be reading this CTA's smem / signalling its barriers
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_pair(tmem_base, kK2TmemCols);
    }
}

// ---- K5: exact rescore of the candidates, final top-k, proof of exactness ------------------------
struct K5Params {
    const __half* rows;
    const float* inv_norm;
    const float* queries;       // [*, dim] fp32 originals
    const int32_t* q_map;       // slot -> query index in the caller's batch (or nullptr = identity)
    const uint32_t* remap;      // candidate row -> shard row when K2 ran on compacted survivors (or nullptr)
    const uint64_t* cand;       // [n, slots, kprime]
    const float* tau;           // [n, slots] from K2: what each full list may have dropped
    const float* invq;          // [n]
    const float* eps;           // [n]
    float* out_scores;
    int64_t* out_rows;
    int32_t* out_nvalid;
    uint64_t* out_keys;
    const uint32_t* gthr;       // [n, gthr_groups] K2's shared threshold classes (nullptr = not used)
    int32_t gthr_groups;
    int32_t* flagged;           // [n] slots whose result is not proven
    int32_t* n_flagged;
    float* stats;               // diagnostics (crb_index_set_k5_stats) or nullptr: per query slot
                                // {max |approx - exact| over re-scored candidates, margin, re-scored count, proven 1/0}
    float max_norm;             // largest row norm (inner-product margin)
    uint32_t row_offset;
    int32_t dim;
    int32_t k;
    int32_t kprime;
    int32_t slots;
    int32_t metric;
    ExchangeDev ex;             // ex.world == 0: local search only
};

constexpr int kK5Gather = 2048;   // candidates K5 can gather into shared memory

// Block-wide bitonic sort of n (a power of two >= 32) keys in shared memory, descending; 256 threads.
CRB_DEVINL void block_bitonic_desc(uint64_t* a, int n) {
    for (int size = 2; size <= n; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            __syncthreads();
            for (int t = threadIdx.x; t < (n >> 1); t += 256) {
                const int i = 2 * t - (t & (stride - 1));
                const int j = i + stride;
                const bool desc = (i & size) == 0;
                const uint64_t x = a[i], y = a[j];
                if ((x < y) == desc) { a[i] = y; a[j] = x; }
            }
        }
    }
    __syncthreads();
}

// Exact canonical score of one approximate candidate -> its exact key (0 = not a hit).  One warp.
CRB_DEVINL uint64_t k5_exact_key(const K5Params& p, uint64_t e, const float* s_q, int chunks, bool cosine, int lane) {
    if (e == 0ull) return 0ull;
    const uint32_t row = p.remap ? p.remap[key_row(e)] : key_row(e);
    const __half* rp = p.rows + static_cast<int64_t>(row) * p.dim;
    float acc = 0.f;
    for (int ch = lane; ch < chunks; ch += 32) {
        const uint4 v = ld_stream_16(rp + ch * 8);
        float q8[8];
        const float4 a = *reinterpret_cast<const float4*>(s_q + ch * 8);
        const float4 b = *reinterpret_cast<const float4*>(s_q + ch * 8 + 4);
        q8[0] = a.x; q8[1] = a.y; q8[2] = a.z; q8[3] = a.w;
        q8[4] = b.x; q8[5] = b.y; q8[6] = b.z; q8[7] = b.w;
        acc = chunk_fma(v, q8, acc);
    }
    const float in = __ldg(p.inv_norm + row);
    const float invn = cosine ? in : (in == in ? 1.0f : in);
    const float s = warp_butterfly_sum(acc) * invn;
    if ((cosine && invn == 0.f) || !(s == s)) return 0ull;
    return make_key(s, p.row_offset + row);
}

template <int KPL>
__global__ void __launch_bounds__(256) k5_rescore_kernel(const __grid_constant__ K5Params p) {
    extern __shared__ float s_q[];   // [dim]
    __shared__ uint64_t xchg[2][32];
    __shared__ uint64_t s_out[kMaxTopK];
    __shared__ uint64_t s_c[kK5Gather];
    __shared__ float s_tau[8];
    __shared__ int s_n;
    __shared__ unsigned s_err_bits, s_rescored;      // diagnostics only (p.stats)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int slot_q = blockIdx.x;
    if (threadIdx.x == 0) { s_err_bits = 0u; s_rescored = 0u; }
    const int qi = p.q_map ? p.q_map[slot_q] : slot_q;
    const bool cosine = (p.metric == 0);
    const int chunks = p.dim >> 3;
    for (int i = threadIdx.x; i < p.dim; i += 256) s_q[i] = p.queries[static_cast<int64_t>(qi) * p.dim + i];
    __syncthreads();

    const uint64_t* cand_all = p.cand + static_cast<int64_t>(slot_q) * p.slots * p.kprime;
    const int n_all = p.slots * p.kprime;
    const float margin = p.eps[slot_q] * (cosine ? 1.0f : p.max_norm);
    const float ninf = __uint_as_float(0xff800000u);

    // tau_max: the best score any FULL slot list may have dropped, and K2's shared bound (columns dropped
    // against the bound scored at most its final value)
    float tau = ninf;
    for (int s = threadIdx.x; s < p.slots; s += 256) tau = fmaxf(tau, p.tau[static_cast<int64_t>(slot_q) * p.slots + s]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tau = fmaxf(tau, __shfl_xor_sync(CRB_FULL_MASK, tau, o));
    if (lane == 0) s_tau[warp] = tau;
    const float gbound = p.gthr ? k2_gthr_float(k2_gthr_min(p.gthr + static_cast<int64_t>(slot_q) * p.gthr_groups, p.gthr_groups)) : ninf;

    // Stage 0: the bound is at most the k-th best approximate score a_k, so only candidates within
    // 2*margin of it can matter (see below); they are a few dozen of the slots*k' entries.  Gather them
    // into shared memory and run the rest on that short list (all candidates when it would not fit).
    if (threadIdx.x == 0) s_n = 0;
    __syncthreads();
    const float gcut = gbound - 2.0f * margin;          // -inf when there is no bound
    for (int c = threadIdx.x; c < n_all; c += 256) {
        const uint64_t e = cand_all[c];
        if (e != 0ull && !(key_score(e) < gcut)) {
            const int at = atomicAdd(&s_n, 1);
            if (at < kK5Gather) s_c[at] = e;
        }
    }
    __syncthreads();
    const bool gathered = s_n <= kK5Gather;
    const uint64_t* cand = gathered ? s_c : cand_all;
    const int n_cand = gathered ? s_n : n_all;

    // Prune before the expensive part: find the k-th best APPROXIMATE score a_k among the candidates.
    // A candidate with approx < a_k - 2*margin has exact < a_k - margin <= the exact score of each of
    // the k best approximate candidates, so it cannot be in the exact top-k and is not re-scored.
    float cut = ninf;
    if (gathered) {
        // ---- short list in shared memory: sort it, re-score the prefix that can matter in place, sort again ----
        int np = 32;
        while (np < n_cand) np <<= 1;
        for (int i = n_cand + threadIdx.x; i < np; i += 256) s_c[i] = 0ull;
        block_bitonic_desc(s_c, np);
        const uint64_t kth = p.k <= n_cand ? s_c[p.k - 1] : 0ull;
        if (kth != 0ull) cut = key_score(kth) - 2.0f * margin;
        if (threadIdx.x == 0) s_n = 0;
        __syncthreads();
        {   // sorted descending, so the candidates that can matter are a prefix; count it
            int mine = 0;
            for (int i = threadIdx.x; i < n_cand; i += 256) mine += !(key_score(s_c[i]) < cut);
            mine = __reduce_add_sync(CRB_FULL_MASK, mine);
            if (lane == 0 && mine) atomicAdd(&s_n, mine);
        }
        __syncthreads();
        const int m = s_n;
        for (int c = warp; c < m; c += 8) {
            const uint64_t e = s_c[c];
            const uint64_t key = k5_exact_key(p, e, s_q, chunks, cosine, lane);
            if (lane == 0) s_c[c] = key;
            if (p.stats && lane == 0 && key != 0ull) {
                atomicMax(&s_err_bits, __float_as_uint(fabsf(key_score(e) - key_score(key) * p.invq[slot_q])));
                atomicAdd(&s_rescored, 1u);
            }
        }
        int mp = 32;
        while (mp < m) mp <<= 1;
        for (int i = m + threadIdx.x; i < mp; i += 256) s_c[i] = 0ull;
        block_bitonic_desc(s_c, mp);                     // exact keys, best first (0 = dropped / padding)
        if (threadIdx.x < p.k) s_out[threadIdx.x] = threadIdx.x < mp ? s_c[threadIdx.x] : 0ull;
        __syncthreads();
    } else {
        {
            WarpTopK<KPL> atop;
            atop.init(p.k, lane);
            for (int c0 = warp * 32; c0 < n_cand; c0 += 8 * 32) {
                const int c = c0 + lane;
                const uint64_t e = c < n_cand ? cand[c] : 0ull;
                // only lanes whose candidate beats the running k-th best are visited (the bound only rises)
                unsigned hit = __ballot_sync(CRB_FULL_MASK, e > atop.thr);
                while (hit) {
                    const int l = __ffs(hit) - 1;
                    hit &= hit - 1u;
                    const uint64_t el = __shfl_sync(CRB_FULL_MASK, e, l);
                    if (el > atop.thr) atop.insert(el, lane);      // warp-uniform
                }
            }
            block_extract_topk<KPL>(atop, p.k, s_out, xchg, lane, warp, 8);
            const uint64_t kth = s_out[p.k - 1];
            if (kth != 0ull) cut = key_score(kth) - 2.0f * margin;
            __syncthreads();
        }
        // exact canonical score of every surviving candidate (warp per candidate)
        WarpTopK<KPL> top;
        top.init(p.k, lane);
        for (int c0 = warp * 32; c0 < n_cand; c0 += 8 * 32) {
            const int cl = c0 + lane;
            const uint64_t el = cl < n_cand ? cand[cl] : 0ull;
            unsigned live = __ballot_sync(CRB_FULL_MASK, el != 0ull && !(key_score(el) < cut));
            while (live) {
                const int src = __ffs(live) - 1;
                live &= live - 1u;
                const uint64_t esrc = __shfl_sync(CRB_FULL_MASK, el, src);
                const uint64_t key = k5_exact_key(p, esrc, s_q, chunks, cosine, lane);
                if (p.stats && lane == 0 && key != 0ull) {
                    atomicMax(&s_err_bits, __float_as_uint(fabsf(key_score(esrc) - key_score(key) * p.invq[slot_q])));
                    atomicAdd(&s_rescored, 1u);
                }
                if (key > top.thr) top.insert(key, lane);
            }
        }
        block_extract_topk<KPL>(top, p.k, s_out, xchg, lane, warp, 8);
    }

    // ---- proof, then every rank of the result is written by its own thread ----
    const float iq = p.invq[slot_q];
    if (threadIdx.x == 0) {
        float tau_max = s_tau[0];
        for (int w = 1; w < 8; ++w) tau_max = fmaxf(tau_max, s_tau[w]);
        tau_max = fmaxf(tau_max, gbound);
        int nv = 0;
        for (int r = 0; r < p.k; ++r) nv += (s_out[r] != 0ull);
        // proof: every non-candidate row has exact (normalised) score <= tau_max + margin
        bool proven;
        if (tau_max == ninf) {
            proven = true;                              // no slot list overflowed: nothing was dropped
        } else if (nv < p.k) {
            proven = false;
        } else {
            const float kth = key_score(s_out[p.k - 1]) * iq;   // same units as the approximate scores
            proven = kth > tau_max + margin;
        }
        s_n = proven ? nv : -1 - nv;
    }
    __syncthreads();
    const bool proven = s_n >= 0;
    const int nv = proven ? s_n : -1 - s_n;
    TopKSink sink;
    sink.scores = p.out_scores + static_cast<int64_t>(qi) * p.k;
    sink.rows = p.out_rows + static_cast<int64_t>(qi) * p.k;
    sink.n_valid = p.out_nvalid + qi;
    sink.keys = p.out_keys ? p.out_keys + static_cast<int64_t>(qi) * p.k : nullptr;
    sink.scale = cosine ? iq : 1.0f;
    sink.qi = qi;
    sink.k = p.k;
    // an unproven query is re-run by the fallback kernel, which then publishes it to the peers
    sink.ex = (proven && p.ex.world > 0) ? &p.ex : nullptr;
    if (threadIdx.x < p.k) {
        sink.emit(threadIdx.x, s_out[threadIdx.x]);
        if (sink.ex) __threadfence_system();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        *sink.n_valid = nv;
        sink.finish();
        if (p.stats) {
            float* st = p.stats + 4 * static_cast<int64_t>(qi);
            st[0] = __uint_as_float(s_err_bits);
            st[1] = margin;
            st[2] = static_cast<float>(s_rescored);
            st[3] = proven ? 1.0f : 0.0f;
        }
        if (!proven) {
            const int at = atomicAdd(p.n_flagged, 1);
            p.flagged[at] = slot_q;
        }
    }
}

// Survivor compaction for filtered tensor-core batches: copy the surviving rows (and their inverse
// norms) of one filter into a dense buffer so K2 spends flops only on survivors.  One warp per row.
__global__ void __launch_bounds__(256) compact_rows_kernel(const __half* __restrict__ rows,
                                                           const float* __restrict__ inv_norm,
                                                           const uint32_t* __restrict__ list,
                                                           const uint32_t* __restrict__ count_ptr, int64_t cap,
                                                           int dim, __half* __restrict__ out_rows,
                                                           float* __restrict__ out_inv) {
    const int lane = threadIdx.x & 31;
    const int64_t gw = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    const int64_t cnt = static_cast<int64_t>(*count_ptr);
    const int64_t n = cnt < cap ? cnt : cap;
    const int chunks = dim >> 3;
    for (int64_t i = gw; i < n; i += nw) {
        const int64_t src = static_cast<int64_t>(list[i]);
        for (int ch = lane; ch < chunks; ch += 32)
            *reinterpret_cast<uint4*>(out_rows + i * dim + ch * 8) = ld_stream_16(rows + src * dim + ch * 8);
        if (lane == 0) out_inv[i] = inv_norm[src];
    }
}

// Largest row norm = 1 / smallest positive inverse norm (inner-product margin of K5).
__global__ void __launch_bounds__(256) min_pos_inv_norm_kernel(const float* __restrict__ inv_norm, int64_t n,
                                                               unsigned* __restrict__ out_bits) {
    float m = __uint_as_float(0x7f800000u);
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const float v = inv_norm[i];
        if (v > 0.f) m = fminf(m, v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fminf(m, __shfl_xor_sync(CRB_FULL_MASK, m, o));
    if ((threadIdx.x & 31) == 0) atomicMin(out_bits, __float_as_uint(m));   // positive floats order as uints
}

}  // namespace crb
