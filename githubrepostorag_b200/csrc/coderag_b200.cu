// C ABI of the retrieval hot path (include/coderag_b200.h) and the host-side launch planning.
// Built for sm_100a only; there is no CPU or multi-backend fallback.
#include "../../include/coderag_b200.h"

#include <algorithm>
#include <array>
#include <atomic>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <ctime>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "common.cuh"
#include "k1_stream.cuh"
#include "k2_gemm.cuh"
#include "k3_filter.cuh"
#include "topk.cuh"

namespace crb {

// ---------------------------------------------------------------------------------------------
// errors, device guard, launch counter
// ---------------------------------------------------------------------------------------------
static thread_local std::string g_err;
static std::atomic<int64_t> g_launches{0};

static int fail(const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return 1;
}

#define CRB_CUDA(expr)                                                                         \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess)                                                                 \
            return crb::fail("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                             __LINE__);                                                        \
    } while (0)

#define CRB_LAUNCH_CHECK()                                                                  \
    do {                                                                                    \
        crb::g_launches.fetch_add(1, std::memory_order_relaxed);                            \
        cudaError_t _e = cudaGetLastError();                                                \
        if (_e != cudaSuccess)                                                              \
            return crb::fail("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e),    \
                             __FILE__, __LINE__);                                           \
    } while (0)

struct DeviceGuard {
    int prev = -1;
    bool switched = false;
    cudaError_t status = cudaSuccess;
    explicit DeviceGuard(int dev) {
        status = cudaGetDevice(&prev);
        if (status == cudaSuccess && prev != dev) {
            status = cudaSetDevice(dev);
            switched = (status == cudaSuccess);
        }
    }
    ~DeviceGuard() {
        if (switched) cudaSetDevice(prev);
    }
};

// Tuning knobs (CODERAG_B200_* environment variables).  They are read ONCE per index -- at crb_index_create,
// and again only when the host calls crb_index_reload_tuning -- never on the search path: the worker runs ten
// searching threads, and getenv next to a setenv elsewhere in the process is a data race.  None of them changes
// a result bit; the result-changing timing experiments exist only in -DCRB_TUNING builds.
struct Tuning {
    int lanes = 4;            // CODERAG_B200_LANES: stream + staging lanes of the host-buffer entry points
    int compact_mb = 1024;    // CODERAG_B200_COMPACT_MB: cap of the survivor-compaction buffer
    int k1_rows = 0;          // _K1_ROWS: rows in flight per warp, single-query groups (0 = default per width)
    int k1_ex = 1;            // _K1_EX: exact-width instantiation at dim 768
    int k1_half = 1;          // _K1_HALF: half-warp layout for dim <= 512
    int k1_e3 = 1;            // _K1_E3: exact three-slot half-warp instantiation at dim 384
    int k1_half_pairs = 0;    // _K1_HALF_PAIRS: row pairs in flight, half-warp layout (0 = default)
    int k1_group = 0;         // _K1_GROUP: queries per streaming group (0 = default)
    int k1_fused = -1;        // _K1_FUSED: fused filter+score kernel for one query under one filter (-1 = plan decides)
    int k1_balanced = 1;      // _K1_BALANCED: balanced kernel for fan-outs
    int k2_pair = -1;         // _K2_PAIR: force the CTA-pair kernel on/off (-1 = by batch)
    int k2_minner = 0;        // _K2_MINNER: query-tile pairs per corpus tile (0 = default)
    int k2_gthr = 1;          // _K2_GTHR: shared admission bound
    int k2_gthr_every = 0;    // _K2_GTHR_EVERY: fixed re-read cadence of the bound (0 = schedule)
    int k2_prepass = 1;       // _K2_PREPASS: first accumulator tile read twice
    int zero_copy = 1;        // _ZEROCOPY: small results are written by the kernels straight into pinned host memory
    int host_graph = 1;       // _HOST_GRAPH: the one-query unfiltered host call is replayed as a CUDA graph
    int zone_maps = 1;        // _ZONE_MAPS: a single filtered row set is searched inside the row range its codes occupy
    int drop_trivial = 1;     // _DROP_TRIVIAL: drop filter codes that every row of the shard carries
    int k1_fold = 1;          // _K1_FOLD: sharded batch-1 calls merge in the streaming kernel's last CTA (one launch per rank)
    int k2_kprime = 16;       // candidates per list for k <= 10 (CRB_TUNING builds read _K2_KPRIME)
    int k2_prefetch = 0;      // CRB_TUNING builds only
    int k2_debug = 0;         // CRB_TUNING builds only
    static int env(const char* name, int dflt) {
        const char* v = getenv(name);
        return (v && *v) ? atoi(v) : dflt;
    }
    static Tuning from_env() {
        Tuning t;
        t.lanes = std::max(1, env("CODERAG_B200_LANES", t.lanes));
        t.compact_mb = std::max(0, env("CODERAG_B200_COMPACT_MB", t.compact_mb));
        t.k1_rows = env("CODERAG_B200_K1_ROWS", t.k1_rows);
        t.k1_ex = env("CODERAG_B200_K1_EX", t.k1_ex);
        t.k1_half = env("CODERAG_B200_K1_HALF", t.k1_half);
        t.k1_e3 = env("CODERAG_B200_K1_E3", t.k1_e3);
        t.k1_half_pairs = env("CODERAG_B200_K1_HALF_PAIRS", t.k1_half_pairs);
        t.k1_group = env("CODERAG_B200_K1_GROUP", t.k1_group);
        t.k1_fused = env("CODERAG_B200_K1_FUSED", t.k1_fused);
        t.k1_balanced = env("CODERAG_B200_K1_BALANCED", t.k1_balanced);
        t.k1_fold = env("CODERAG_B200_K1_FOLD", t.k1_fold);
        t.drop_trivial = env("CODERAG_B200_DROP_TRIVIAL", t.drop_trivial);
        t.zone_maps = env("CODERAG_B200_ZONE_MAPS", t.zone_maps);
        t.host_graph = env("CODERAG_B200_HOST_GRAPH", t.host_graph);
        t.zero_copy = env("CODERAG_B200_ZEROCOPY", t.zero_copy);
        t.k2_pair = env("CODERAG_B200_K2_PAIR", t.k2_pair);
        t.k2_minner = env("CODERAG_B200_K2_MINNER", t.k2_minner);
        t.k2_gthr = env("CODERAG_B200_K2_GTHR", t.k2_gthr);
        t.k2_gthr_every = std::max(0, env("CODERAG_B200_K2_GTHR_EVERY", t.k2_gthr_every));
        t.k2_prepass = env("CODERAG_B200_K2_PREPASS", t.k2_prepass);
#ifdef CRB_TUNING
        t.k2_kprime = env("CODERAG_B200_K2_KPRIME", t.k2_kprime);
        t.k2_prefetch = env("CODERAG_B200_K2_PREFETCH", 0);
        t.k2_debug = env("CODERAG_B200_K2_DEBUG", 0);
#endif
        return t;
    }
};

// ---------------------------------------------------------------------------------------------
// workspace layout
// ---------------------------------------------------------------------------------------------
constexpr size_t kWsHeaderBytes = 4096;   // counters [kMaxGroups] u32 @0, list counts @1024
constexpr size_t kWsCountsOffset = 1024;
constexpr size_t kWsCandBytes = size_t(8) << 20;  // per-CTA candidate lists of one launch

static inline int64_t round_up(int64_t v, int64_t m) { return (v + m - 1) / m * m; }

constexpr int kK2MaxChunk = kK2MaxMTiles * kK2BlockM;   // queries per K2 launch
constexpr int64_t kK2MinRows = 16384;                   // below this K1 is always used
constexpr int64_t kK2SmallBatchRows = 131072;           // from here on even a 2-query unfiltered batch takes K2

// k' (candidates kept per query per CTA) for a requested k
// (shared memory holds k' x 128 keys next to >= 2 pipeline stages, which caps k' at 120)
constexpr int kK2MaxKPrime = 120;
static inline int k2_kprime(int k, int kprime_small = 16) {
    if (k <= 10) return std::max(k + 2, kprime_small);
    return std::min<int>(kK2MaxKPrime, static_cast<int>(round_up(k + k / 2, 8)));
}
static inline bool k2_supports_k(int k) { return k2_kprime(k) >= k + 2; }
// classes of the shared admission threshold: a power of two >= k (see k2_gemm.cuh)
static inline int k2_gthr_groups(int k) { int g = 16; while (g < k + k / 4 + 1 && g < 128) g *= 2; return g; }

struct Lane {
    cudaStream_t stream = nullptr;
    void* d_ws = nullptr;
    size_t ws_cap = 0;
    void* d_io = nullptr;
    size_t d_io_cap = 0;
    void* h_pin = nullptr;
    size_t h_cap = 0;
    volatile uint32_t* h_flag = nullptr;   // pinned word a signal kernel writes when a zero-copy result is complete
    uint32_t flag_seq = 0;
    bool busy = false;
    // The one-query unfiltered host call as a replayable graph (copy-in -> streaming kernel -> flag): one graph launch per
    // request instead of three enqueues.  Valid for one (bind generation, k, metric) and this lane's buffers.
    cudaGraphExec_t g_exec = nullptr;
    uint64_t g_gen = 0;
    int g_k = 0, g_metric = -1;
    void *g_dio = nullptr, *g_hpin = nullptr, *g_ws = nullptr;
    uint64_t seen_gen = 0;                 // (generation, k, metric) of the previous plain call: a graph is built on the second
    int seen_k = 0, seen_metric = -1;
    bool g_failed = false;                 // capture failed for seen_*: do not try again until they change
};

}  // namespace crb

struct crb_index {
    const __half* rows = nullptr;
    const float* inv_norm = nullptr;
    const int32_t* codes = nullptr;   // int32 [n_keys, code_stride] matrix, or nullptr when the columns were bound one by one
    int64_t code_stride = 0;       // elements between key columns of `codes`
    const void* cols[CRB_MAX_KEYS] = {};   // column of key m (what the kernels read): width[m] bytes per row
    uint8_t width[CRB_MAX_KEYS] = {};
    // rows per code of the narrow columns (bin code + 1; bin 0 = key absent), counted when the columns are bound:
    // the plan's selectivity estimate (product of the constrained keys' frequencies)
    std::vector<uint32_t> hist[CRB_MAX_KEYS];
    // first / last row carrying each code (same bins; valid where hist > 0): a filter's survivors lie inside the
    // intersection of its codes' row ranges, and a code whose rows are one contiguous run (hist == last - first + 1: rows
    // written repository by repository, or a shard clustered by its keys) needs no filter pass at all
    std::vector<uint32_t> zone_first[CRB_MAX_KEYS], zone_last[CRB_MAX_KEYS];
    int64_t n_rows = 0;
    uint64_t bind_gen = 1;         // bumped whenever the buffers are (re)bound or the tuning is re-read: cached launch graphs die
    int32_t dim = 0;
    int32_t n_keys = 0;
    int64_t row_offset = 0;
    int32_t device = 0;
    int32_t sm_count = 148;
    float max_norm = 0.f;          // largest row norm (K5 inner-product margin)
    bool k2_ok = false;            // corpus tensor map encoded
    CUtensorMap map_rows;          // fp16 [n_rows, dim], box 64 x 256, 128 B swizzle
    CUtensorMap map_rows_half;     // same, box 64 x 128 (one CTA's half tile in the pair kernel)
    crb::Tuning tune;
    float* k5_stats = nullptr;     // diagnostics buffer (crb_index_set_k5_stats) or nullptr
    std::mutex mu;
    std::condition_variable cv;
    std::vector<crb::Lane> lanes;
};

namespace crb {

static int64_t list_stride_of(const crb_index* ix) { return round_up(std::max<int64_t>(ix->n_rows, 1), 64); }
static int64_t mask_stride_of(const crb_index* ix) { return round_up((std::max<int64_t>(ix->n_rows, 1) + 31) / 32, 64); }

// Workspace layout: [header 4 KB][K1 candidate lists 8 MB][K2 region][survivor lists][pass masks]
struct WsLayout {
    size_t cand = kWsHeaderBytes;
    size_t k2 = 0;          // K2 region start
    size_t q16 = 0, invq = 0, eps = 0, qmap = 0, flagged = 0, nflag = 0, fbcnt = 0, gthr = 0, tau = 0, cand2 = 0;   // inside the K2 region
    size_t filt = 0;        // survivor lists start; masks follow the lists
    size_t per_filter = 0;  // bytes one more filter costs
};
static WsLayout ws_layout(const crb_index* ix, int batch, int k) {
    WsLayout L;
    L.k2 = kWsHeaderBytes + kWsCandBytes;
    const size_t nmax = static_cast<size_t>(std::min(batch, kK2MaxChunk));
    size_t o = L.k2;
    L.q16 = o;     o += round_up(round_up(nmax, 256) * ix->dim * 2, 1024);   // padded to whole query tile pairs
    L.invq = o;    o += round_up(nmax * 4, 256);
    L.eps = o;     o += round_up(nmax * 4, 256);
    L.qmap = o;    o += round_up(nmax * 4, 256);
    L.flagged = o; o += round_up(nmax * 4, 256);
    L.nflag = o;   o += 256;
    L.fbcnt = o;   o += round_up(nmax * 4, 256);   // fallback arrival counters, cleared with nflag
    L.gthr = o;    o += round_up(nmax * size_t(k2_gthr_groups(k)) * 4, 256);   // K2's shared threshold classes, cleared with nflag
    // per query: one k'-list per CTA (pair) that served its query tile -- at most sm_count lists when the
    // batch is one tile, sm_count/2 when CTA pairs serve every query tile of the batch (m-inner)
    const size_t n_lists = std::max<size_t>(std::min<size_t>(nmax, kK2BlockM) * ix->sm_count, nmax * size_t(ix->sm_count / 2 + 1)) + nmax;
    L.tau = o;     o += round_up(n_lists * 4, 256);     // one float per list
    L.cand2 = o;   o += round_up(n_lists * k2_kprime(k, ix->tune.k2_kprime) * 8, 1024);
    L.filt = o;
    L.per_filter = size_t(list_stride_of(ix)) * 4 + size_t(mask_stride_of(ix)) * 4;
    return L;
}

// Per code of one narrow column (bin = code + 1; code -1 = key absent -> bin 0): how many rows carry it, the first and
// the last such row.  One pass; int8 columns accumulate in shared memory.
__global__ void __launch_bounds__(256) code_histogram_kernel(const void* __restrict__ col, int width, int64_t n,
                                                             unsigned* __restrict__ hist, unsigned* __restrict__ first,
                                                             unsigned* __restrict__ last) {
    __shared__ unsigned s_h[130], s_lo[130], s_hi[130];
    if (width == 1) {
        for (int i = threadIdx.x; i < 130; i += blockDim.x) { s_h[i] = 0u; s_lo[i] = 0xffffffffu; s_hi[i] = 0u; }
        __syncthreads();
    }
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const unsigned r = static_cast<unsigned>(i);
        if (width == 1) {
            const int b = static_cast<int>(static_cast<const int8_t*>(col)[i]) + 1;
            atomicAdd(&s_h[b], 1u);
            atomicMin(&s_lo[b], r);
            atomicMax(&s_hi[b], r);
        } else {
            const int b = static_cast<int>(static_cast<const int16_t*>(col)[i]) + 1;
            atomicAdd(&hist[b], 1u);
            atomicMin(&first[b], r);
            atomicMax(&last[b], r);
        }
    }
    if (width == 1) {
        __syncthreads();
        for (int i = threadIdx.x; i < 130; i += blockDim.x)
            if (s_h[i]) {
                atomicAdd(&hist[i], s_h[i]);
                atomicMin(&first[i], s_lo[i]);
                atomicMax(&last[i], s_hi[i]);
            }
    }
}

// Estimated fraction of the shard's rows that pass an AND of equalities: the product of the constrained keys'
// code frequencies (keys taken as independent).  -1 when a constrained key has no histogram (a 4-byte column).
static double estimate_selectivity(const crb_index* ix, const int32_t* want) {
    if (ix->n_rows <= 0) return -1.0;
    double sel = 1.0;
    for (int m = 0; m < ix->n_keys; ++m) {
        if (want[m] == CRB_FILT_ANY) continue;
        if (want[m] < 0) return 0.0;
        const std::vector<uint32_t>& h = ix->hist[m];
        if (h.empty()) return -1.0;
        const size_t bin = size_t(want[m]) + 1;
        sel *= bin < h.size() ? double(h[bin]) / double(ix->n_rows) : 0.0;
    }
    return sel;
}

// The rows a filter can match, from the per-code row ranges taken at bind time: survivors lie in [lo, hi) -- the
// intersection of the constrained codes' ranges.  `exact`: every constrained code's rows are one contiguous run, so every
// row of the intersection passes and the search is an UNFILTERED pass over [lo, hi) (no filter pass, no gather).  Rows
// arrive repository by repository (ingest) or the shard was clustered by its keys (B200VectorStore.compact(cluster_by=));
// on rows in random order the window is the whole shard and nothing changes.
struct RowWindow {
    int64_t lo = 0, hi = 0;
    bool exact = false;
};
static RowWindow filter_window(const crb_index* ix, const int32_t* want) {
    RowWindow w;
    w.hi = ix->n_rows;
    bool exact = true, any = false;
    for (int m = 0; m < ix->n_keys; ++m) {
        if (want[m] == CRB_FILT_ANY) continue;
        any = true;
        if (want[m] < 0) { w.lo = w.hi = 0; w.exact = true; return w; }       // impossible: no row
        const std::vector<uint32_t>& h = ix->hist[m];
        if (h.empty() || ix->zone_first[m].size() != h.size()) { exact = false; continue; }   // no ranges for this key
        const size_t bin = size_t(want[m]) + 1;
        if (bin >= h.size() || h[bin] == 0) { w.lo = w.hi = 0; w.exact = true; return w; }     // no row carries the code
        const int64_t lo = ix->zone_first[m][bin], hi = int64_t(ix->zone_last[m][bin]) + 1;
        w.lo = std::max(w.lo, lo);
        w.hi = std::min(w.hi, hi);
        if (int64_t(h[bin]) != hi - lo) exact = false;
    }
    if (w.hi < w.lo) w.hi = w.lo;
    w.exact = any && exact;
    return w;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time libcuda).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = []() -> EncodeTiledFn {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            return nullptr;
        return reinterpret_cast<EncodeTiledFn>(f);
    }();
    return fn;
}
// fp16 [rows, dim] row-major -> 2-D map, box = 64 elements (128 B, one swizzle row) x box_rows
static int make_map_f16(CUtensorMap* map, const void* base, int64_t rows, int dim, int box_rows) {
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return fail("cuTensorMapEncodeTiled is unavailable in this driver");
    const cuuint64_t gdim[2] = {static_cast<cuuint64_t>(dim), static_cast<cuuint64_t>(std::max<int64_t>(rows, 1))};
    const cuuint64_t gstride[1] = {static_cast<cuuint64_t>(dim) * 2};
    const cuuint32_t box[2] = {static_cast<cuuint32_t>(kK2BlockK), static_cast<cuuint32_t>(box_rows)};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled failed with CUresult %d", static_cast<int>(r));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// K1 dispatch
// ---------------------------------------------------------------------------------------------
typedef void (*K1Fn)(const K1Params);

// Instantiated (NQ, CH, R) combinations.  R = rows in flight per warp: small rows need more of them
// to keep enough bytes in flight (Little's law: ~35 KB per SM at HBM speed); MINB = 2 caps registers
// at 128 so two CTAs stay resident when the per-lane footprint (NQ*CH query chunks + R*CH loads) is large.
template <int NQ, int CH, int R, int MINB>
static K1Fn k1_pick_kpl(int k) {
    if (k <= 32) return k1_stream_kernel<NQ, CH, R, 1, MINB>;
    return k1_stream_kernel<NQ, CH, R, 4, MINB>;
}
// exact-width instantiation (dim = 768 = 32 lanes x 3 chunks) of the single-query kernel: no per-chunk range tests,
// 118 instead of 125 registers; interleaved A/B on 10M x 768: 2.346 vs 2.460 ms median (2.298 vs 2.396 min).
// The same specialisation for dim = 1024 measured 5 % SLOWER (3.24 vs 3.08 ms) and is not instantiated.
static K1Fn k1_pick_exact(int ch, int r, int k) {
    if (ch == 3 && r == 4) return k <= 32 ? static_cast<K1Fn>(k1_stream_kernel<1, 3, 4, 1, 2, true>) : static_cast<K1Fn>(k1_stream_kernel<1, 3, 4, 4, 2, true>);
    return nullptr;
}
static K1Fn k1_pick(int nq, int ch, int r, int k) {
#define CRB_K1_CASE(NQ_, CH_, R_, MINB_) if (nq == NQ_ && ch == CH_ && r == R_) return k1_pick_kpl<NQ_, CH_, R_, MINB_>(k);
    CRB_K1_CASE(1, 1, 4, 2) CRB_K1_CASE(1, 1, 8, 2)
    CRB_K1_CASE(1, 2, 4, 2) CRB_K1_CASE(1, 2, 6, 2)
    CRB_K1_CASE(1, 3, 2, 2) CRB_K1_CASE(1, 3, 4, 2)
    CRB_K1_CASE(1, 4, 2, 2) CRB_K1_CASE(1, 4, 3, 2)
    CRB_K1_CASE(2, 1, 4, 2)
    CRB_K1_CASE(2, 2, 2, 2) CRB_K1_CASE(2, 2, 4, 2)
    CRB_K1_CASE(2, 3, 2, 2)
    CRB_K1_CASE(2, 4, 2, 1)
    CRB_K1_CASE(4, 1, 2, 2) CRB_K1_CASE(4, 1, 4, 2)
    CRB_K1_CASE(4, 2, 2, 1)
#undef CRB_K1_CASE
    return nullptr;
}
// default rows-in-flight per (NQ, CH); Tuning::k1_rows overrides for NQ == 1 (tuning sweeps)
static int k1_default_rows(int nq, int ch, int env = 0) {
    if (nq == 1) {
        const int def[5] = {0, 8, 6, 4, 3};
        if (env > 0 && k1_pick(1, ch, env, 10)) return env;
        return def[ch];
    }
    if (nq == 2) return ch == 1 ? 4 : (ch == 2 ? 4 : 2);
    return ch == 1 ? 4 : 2;
}
// half-warp kernels for dim <= 512 (R = row PAIRS in flight per warp)
template <int NQ, int R>
static K1Fn k1_half_kpl(int k) {
    if (k <= 32) return k1_stream_half_kernel<NQ, R, 1>;
    return k1_stream_half_kernel<NQ, R, 4>;
}
template <int R>
static K1Fn k1_half_e3_kpl(int k) {
    if (k <= 32) return k1_stream_half_kernel<1, R, 1, true>;
    return k1_stream_half_kernel<1, R, 4, true>;
}
static K1Fn k1_pick_half(int nq, int r, int k, bool e3 = false) {
    if (nq == 1 && e3) return r == 2 ? k1_half_e3_kpl<2>(k) : (r == 3 ? k1_half_e3_kpl<3>(k) : k1_half_e3_kpl<4>(k));
    if (nq == 1) return r == 2 ? k1_half_kpl<1, 2>(k) : (r == 3 ? k1_half_kpl<1, 3>(k) : k1_half_kpl<1, 4>(k));
    if (nq == 2) return k1_half_kpl<2, 2>(k);
    return k1_half_kpl<4, 1>(k);
}
static K1Fn k1_pick_generic(int k) {
    if (k <= 32) return k1_stream_generic_kernel<2, 1>;
    return k1_stream_generic_kernel<2, 4>;
}

// resident CTAs per SM of a kernel (cached)
static int k1_occupancy(K1Fn fn, size_t dyn_smem) {
    static std::mutex mu;
    static std::map<std::pair<const void*, size_t>, int> cache;
    std::lock_guard<std::mutex> lk(mu);
    auto key = std::make_pair(reinterpret_cast<const void*>(fn), dyn_smem);
    auto it = cache.find(key);
    if (it != cache.end()) return it->second;
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, fn, kK1Threads, dyn_smem) != cudaSuccess || n < 1) n = 1;
    cache[key] = n;
    return n;
}

// The streaming instantiation for groups of `nq` queries on this index: width class, rows in flight, the
// half-warp layout for dim <= 512 and the exact-width forms for dim 384 / 768.  One picker for every entry point.
struct K1Choice {
    K1Fn fn = nullptr;
    bool generic = false;          // dim > 1024: query in shared memory, one query per group
    size_t dyn_smem = 0;
    int rows_per_warp_iter = 0;
};
static int k1_choose(const crb_index* ix, int nq, int k, K1Choice* out) {
    const int chunks = ix->dim / 8;
    const int ch = (chunks + 31) / 32;
    const bool generic = ch > 4;
    const Tuning& tn = ix->tune;
    int r1 = generic ? 2 : k1_default_rows(nq, ch, tn.k1_rows);
    K1Fn fn = generic ? k1_pick_generic(k) : k1_pick(nq, ch, r1, k);
    if (!generic && nq == 1 && chunks == 32 * ch && tn.k1_ex) {
        if (K1Fn ex_fn = k1_pick_exact(ch, r1, k)) fn = ex_fn;
    }
    if (!generic && ch <= 2 && nq <= 2 && tn.k1_half) {
        // short rows: two rows per warp step (bit-identical scores, fewer instructions per row)
        // dim = 384 has its own instantiation (exactly three chunks per half-lane, fewer registers): 4 pairs in flight
        const bool e3 = ix->dim == 384 && nq == 1 && tn.k1_e3;
        const int rp = nq == 1 ? std::min(std::max(tn.k1_half_pairs > 0 ? tn.k1_half_pairs : (e3 ? 4 : 3), 2), 4) : (nq == 2 ? 2 : 1);
        fn = k1_pick_half(nq, rp, k, e3);
        r1 = 2 * rp;
    }
    if (!fn) return fail("internal: no K1 instantiation for nq=%d ch=%d", nq, ch);
    out->fn = fn;
    out->generic = generic;
    out->dyn_smem = generic ? size_t(ix->dim) * sizeof(float) : 0;
    out->rows_per_warp_iter = r1;
    return 0;
}

struct QueryGroup {
    std::vector<int> queries;
    int filter;  // index into the call's distinct-filter table, -1 = none
};

// Launch K1 for `groups` (all with at most `nq` queries, filter slots already remapped to the
// lists currently in the workspace).
static int launch_k1(const crb_index* ix, const float* d_queries, int k, int metric,
                     const std::vector<QueryGroup>& groups, const std::vector<int>& slot_of_filter,
                     int nq, uint8_t* ws, const uint32_t* lists, float* d_scores, int64_t* d_rows,
                     int32_t* d_n_valid, uint64_t* d_keys, cudaStream_t stream, const ExchangeDev* ex = nullptr,
                     const RowWindow* win = nullptr) {
    // win: the launch sees rows [lo, hi) as its shard (list entries, if any, are relative to lo)
    const int64_t w_lo = win ? win->lo : 0, w_rows = win ? win->hi - win->lo : ix->n_rows;
    K1Choice c;
    if (int rc = k1_choose(ix, nq, k, &c)) return rc;
    const K1Fn fn = c.fn;
    const bool generic = c.generic;
    const size_t dyn = c.dyn_smem;
    const int rows_per_warp_iter = c.rows_per_warp_iter;
    const int resident = k1_occupancy(fn, dyn) * ix->sm_count;

    for (size_t g0 = 0; g0 < groups.size(); g0 += kMaxGroups) {
        const int ng = static_cast<int>(std::min<size_t>(kMaxGroups, groups.size() - g0));
        K1Params p;
        memset(&p, 0, sizeof p);
        p.rows = ix->rows + w_lo * ix->dim;
        p.inv_norm = ix->inv_norm + w_lo;
        p.queries = d_queries;
        p.lists = lists;
        p.list_counts = reinterpret_cast<const uint32_t*>(ws + kWsCountsOffset);
        p.cand = reinterpret_cast<uint64_t*>(ws + kWsHeaderBytes);
        p.counters = reinterpret_cast<unsigned*>(ws);
        p.out_scores = d_scores;
        p.out_rows = d_rows;
        p.out_nvalid = d_n_valid;
        p.out_keys = d_keys;
        p.n_rows = w_rows;
        p.list_stride = list_stride_of(ix);
        p.row_offset = static_cast<uint32_t>(ix->row_offset + w_lo);
        p.dim = ix->dim;
        p.k = k;
        p.metric = metric;
        if (ex) p.ex = *ex;
        for (int g = 0; g < ng; ++g) {
            const QueryGroup& qg = groups[g0 + g];
            for (int q = 0; q < kK1MaxQ; ++q)
                p.groups[g].q_idx[q] = q < static_cast<int>(qg.queries.size()) ? qg.queries[q] : -1;
            p.groups[g].filter = qg.filter < 0 ? -1 : slot_of_filter[qg.filter];
        }
        // grid.x: fill the machine once across all groups, but never more CTAs than there is
        // work for, and never more per-CTA lists than the scratch region or the merge can hold.
        const int64_t work = (w_rows + int64_t(kK1Warps) * rows_per_warp_iter * 2 - 1) /
                             (int64_t(kK1Warps) * rows_per_warp_iter * 2);
        int64_t gx = std::max<int64_t>(resident / ng, 16);
        gx = std::min<int64_t>(gx, std::max<int64_t>(work, 1));
        gx = std::min<int64_t>(gx, kK1MaxGridX);
        const int64_t cand_lists = int64_t(kWsCandBytes) / (int64_t(k) * 8);
        gx = std::min<int64_t>(gx, std::max<int64_t>(cand_lists / (int64_t(ng) * (generic ? 1 : nq)), 1));
        dim3 grid(static_cast<unsigned>(gx), static_cast<unsigned>(ng));
        fn<<<grid, kK1Threads, dyn, stream>>>(p);
        CRB_LAUNCH_CHECK();
    }
    return 0;
}

// One query under one filter: filter and score in a single pass (k1_fused_filter_kernel).
// Eligible when the constrained keys are at most kFusedMaxKeys columns of 1 or 2 bytes, 16-byte aligned.
static bool fused_eligible(const crb_index* ix, const int32_t* want) {
    int used = 0;
    for (int m = 0; m < ix->n_keys; ++m) {
        if (want[m] == CRB_FILT_ANY) continue;
        if (want[m] < 0) return false;
        if (ix->width[m] > 2 || reinterpret_cast<uintptr_t>(ix->cols[m]) % 16 != 0) return false;
        if (ix->width[m] == 1 && want[m] > 127) return false;       // a code the column cannot hold matches no row:
        if (ix->width[m] == 2 && want[m] > 32767) return false;     // leave that to the general path
        ++used;
    }
    return used <= kFusedMaxKeys;
}
static int launch_k1_fused(const crb_index* ix, const float* d_query, int qi, int k, int metric,
                           const int32_t* want, const uint32_t* row_mask, uint8_t* ws, float* d_scores,
                           int64_t* d_rows, int32_t* d_n_valid, uint64_t* d_keys, cudaStream_t stream,
                           const ExchangeDev* ex) {
    typedef void (*FusedFn)(const FusedParams);
    const int ch = (ix->dim / 8 + 31) / 32;
    FusedFn fn;
#define CRB_FUSED(CH_, R_) (k <= 32 ? static_cast<FusedFn>(k1_fused_filter_kernel<CH_, R_, 1>) : static_cast<FusedFn>(k1_fused_filter_kernel<CH_, R_, 4>))
    switch (ch) {
        // rows in flight per warp: the next stripe's codes wait in shared memory (cp.async), so the registers go to rows
        case 1: fn = CRB_FUSED(1, 8); break;
        case 2: fn = CRB_FUSED(2, 6); break;
        case 3: fn = CRB_FUSED(3, 5); break;
        default: fn = CRB_FUSED(4, 4); break;
    }
#undef CRB_FUSED
    const size_t dyn = size_t(kK1Warps) * kFusedWarpSmem;
    static std::mutex mu;
    static std::map<std::pair<const void*, int>, int> cache;
    int occ = 0;
    {
        std::lock_guard<std::mutex> lk(mu);
        auto key = std::make_pair(reinterpret_cast<const void*>(fn), ix->device);
        auto it = cache.find(key);
        if (it == cache.end()) {
            CRB_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(dyn)));
            int n = 0;
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, fn, kK1Threads, dyn) != cudaSuccess || n < 1) n = 1;
            it = cache.emplace(key, n).first;
        }
        occ = it->second;
    }
    FusedParams p;
    memset(&p, 0, sizeof p);
    p.rows = ix->rows;
    p.inv_norm = ix->inv_norm;
    p.query = d_query;
    p.row_mask = row_mask;
    p.cand = reinterpret_cast<uint64_t*>(ws + kWsHeaderBytes);
    p.counter = reinterpret_cast<unsigned*>(ws);
    p.out_scores = d_scores + size_t(qi) * k;
    p.out_rows = d_rows + size_t(qi) * k;
    p.out_nvalid = d_n_valid + qi;
    p.out_keys = d_keys ? d_keys + size_t(qi) * k : nullptr;
    p.n_rows = ix->n_rows;
    p.row_offset = static_cast<uint32_t>(ix->row_offset);
    p.dim = ix->dim;
    p.k = k;
    p.metric = metric;
    p.qi = qi;
    for (int m = 0; m < ix->n_keys; ++m) {
        if (want[m] == CRB_FILT_ANY) continue;
        p.cols[p.n_used] = ix->cols[m];
        p.want[p.n_used] = want[m];
        p.width[p.n_used] = ix->width[m];
        ++p.n_used;
    }
    for (int u = p.n_used; u < kFusedMaxKeys; ++u) p.width[u] = 1;
    if (ex) p.ex = *ex;
    const int64_t stripes = (ix->n_rows + kFusedStripe - 1) / kFusedStripe;
    int64_t grid = int64_t(occ) * ix->sm_count;
    grid = std::min<int64_t>(grid, std::max<int64_t>((stripes + kK1Warps - 1) / kK1Warps, 1));
    grid = std::min<int64_t>(grid, kK1MaxGridX);
    grid = std::min<int64_t>(grid, int64_t(kWsCandBytes) / (int64_t(k) * 8));
    fn<<<static_cast<unsigned>(grid), kK1Threads, dyn, stream>>>(p);
    CRB_LAUNCH_CHECK();
    return 0;
}

// Single-query groups with unequal row sets (one query vector, many filters): the balanced kernel.
static int launch_k1_balanced(const crb_index* ix, const float* d_queries, int k, int metric,
                              const std::vector<QueryGroup>& groups, const std::vector<int>& slot_of_filter,
                              uint8_t* ws, const uint32_t* lists, float* d_scores, int64_t* d_rows,
                              int32_t* d_n_valid, uint64_t* d_keys, cudaStream_t stream, const ExchangeDev* ex) {
    typedef void (*BalFn)(const BalParams);
    const int ch = (ix->dim / 8 + 31) / 32;     // 16-byte chunks per lane
    BalFn fn;
#define CRB_BAL(CH_, R_) (k <= 32 ? static_cast<BalFn>(k1_balanced_kernel<CH_, R_, 1>) : static_cast<BalFn>(k1_balanced_kernel<CH_, R_, 4>))
    switch (ch) {
        case 1: fn = CRB_BAL(1, 8); break;
        case 2: fn = CRB_BAL(2, 6); break;
        case 3: fn = CRB_BAL(3, 4); break;
        case 4: fn = CRB_BAL(4, 3); break;
        default: fn = CRB_BAL(0, 2); break;
    }
#undef CRB_BAL
    // dim <= 512: the half-warp layout (two rows per warp step); dim = 384 has the exact three-slot instantiation
    if (ch <= 2 && ix->tune.k1_half) {
        if (ix->dim == 384)
            fn = k <= 32 ? static_cast<BalFn>(k1_balanced_kernel<2, 3, 1, 3>) : static_cast<BalFn>(k1_balanced_kernel<2, 3, 4, 3>);
        else
            fn = k <= 32 ? static_cast<BalFn>(k1_balanced_kernel<2, 3, 1, 4>) : static_cast<BalFn>(k1_balanced_kernel<2, 3, 4, 4>);
    }
    const size_t dyn = size_t(ix->dim) * sizeof(float);
    static std::mutex mu;
    static std::map<std::pair<const void*, size_t>, int> cache;
    int occ = 0;
    {
        std::lock_guard<std::mutex> lk(mu);
        auto key = std::make_pair(reinterpret_cast<const void*>(fn), dyn);
        auto it = cache.find(key);
        if (it == cache.end()) {
            int n = 0;
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, fn, kK1Threads, dyn) != cudaSuccess || n < 1) n = 1;
            it = cache.emplace(key, n).first;
        }
        occ = it->second;
    }
    for (size_t g0 = 0; g0 < groups.size(); g0 += kMaxGroups) {
        const int ng = static_cast<int>(std::min<size_t>(kMaxGroups, groups.size() - g0));
        BalParams p;
        memset(&p, 0, sizeof p);
        p.rows = ix->rows;
        p.inv_norm = ix->inv_norm;
        p.queries = d_queries;
        p.lists = lists;
        p.list_counts = reinterpret_cast<const uint32_t*>(ws + kWsCountsOffset);
        p.cand = reinterpret_cast<uint64_t*>(ws + kWsHeaderBytes);
        p.counters = reinterpret_cast<unsigned*>(ws);
        p.out_scores = d_scores;
        p.out_rows = d_rows;
        p.out_nvalid = d_n_valid;
        p.out_keys = d_keys;
        p.n_rows = ix->n_rows;
        p.list_stride = list_stride_of(ix);
        p.row_offset = static_cast<uint32_t>(ix->row_offset);
        p.dim = ix->dim;
        p.k = k;
        p.metric = metric;
        p.n_groups = ng;
        if (ex) p.ex = *ex;
        for (int g = 0; g < ng; ++g) {
            const QueryGroup& qg = groups[g0 + g];
            p.q_idx[g] = qg.queries[0];
            p.filter[g] = qg.filter < 0 ? -1 : slot_of_filter[qg.filter];
        }
        // one wave, but never more lists than the scratch region or the merge can hold
        int64_t grid = int64_t(occ) * ix->sm_count;
        grid = std::min<int64_t>(grid, kK1MaxGridX);
        grid = std::min<int64_t>(grid, int64_t(kWsCandBytes) / (int64_t(k) * 8) - 2 * kMaxGroups);
        fn<<<static_cast<unsigned>(std::max<int64_t>(grid, 1)), kK1Threads, dyn, stream>>>(p);
        CRB_LAUNCH_CHECK();
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------
// K2 launch: tensor-core candidates + K5 exact rescore for one set of queries sharing a row set
// ---------------------------------------------------------------------------------------------
// `queries` are indices into the caller's batch; `mask` is the pass-bit array of their common
// filter (nullptr = whole shard).  Queries K5 cannot prove exact are flagged on the device and re-run
// by k1_fallback_kernel, launched right behind K5: nothing here waits for the stream.
static int launch_k2(crb_index* ix, const float* d_queries, int batch, int k, int metric,
                     const std::vector<int>& queries, const uint32_t* mask, uint8_t* ws, const WsLayout& L,
                     float* d_scores, int64_t* d_rows, int32_t* d_n_valid, uint64_t* d_keys,
                     cudaStream_t stream,
                     const uint32_t* surv_list = nullptr, const uint32_t* surv_count = nullptr,
                     int64_t n_surv = -1, uint8_t* compact_buf = nullptr, size_t compact_cap = 0,
                     const ExchangeDev* ex = nullptr) {
    const int kprime = k2_kprime(k, ix->tune.k2_kprime);

    // Filtered set with a known, small survivor count: gather the survivors into a dense buffer and
    // run the GEMM on those rows only (flops and bytes scale with selectivity, not with the shard).

    const float* k2_inv = ix->inv_norm;
    int64_t k2_n = ix->n_rows;
    const CUtensorMap* map_full = &ix->map_rows;
    const CUtensorMap* map_half = &ix->map_rows_half;
    CUtensorMap cmap_full, cmap_half;
    const uint32_t* remap = nullptr;
    if (mask && surv_list && n_surv >= 0 && compact_buf) {
        const size_t row_bytes = size_t(ix->dim) * 2;
        const size_t inv_off = round_up(size_t(std::max<int64_t>(n_surv, 1)) * row_bytes, 1024);
        if (inv_off + size_t(std::max<int64_t>(n_surv, 1)) * 4 <= compact_cap && n_surv * 4 <= ix->n_rows) {
            __half* c_rows = reinterpret_cast<__half*>(compact_buf);
            float* c_inv = reinterpret_cast<float*>(compact_buf + inv_off);
            if (n_surv > 0) {
                const unsigned blocks = static_cast<unsigned>(std::min<int64_t>((n_surv + 7) / 8, int64_t(ix->sm_count) * 16));
                compact_rows_kernel<<<blocks, 256, 0, stream>>>(ix->rows, ix->inv_norm, surv_list, surv_count, n_surv,
                                                               ix->dim, c_rows, c_inv);
                CRB_LAUNCH_CHECK();
            }
            if (int rc = make_map_f16(&cmap_full, c_rows, n_surv, ix->dim, kK2BlockN)) return rc;
            if (int rc = make_map_f16(&cmap_half, c_rows, n_surv, ix->dim, kK2BlockN / 2)) return rc;
            k2_inv = c_inv; k2_n = n_surv;
            map_full = &cmap_full; map_half = &cmap_half;
            remap = surv_list;
            mask = nullptr;
        }
    }

    const size_t tail_bytes = 2 * kK2BlockN * 4 + 256;
    // opt in to 227 KB of dynamic shared memory: a per-device attribute, set once per device ordinal
    {
        static std::mutex attr_mu;
        static bool attr_done[64] = {};
        std::lock_guard<std::mutex> lk(attr_mu);
        const int dv = ix->device & 63;
        if (!attr_done[dv]) {
            CRB_CUDA(cudaFuncSetAttribute(k2_gemm_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
            CRB_CUDA(cudaFuncSetAttribute(k2_gemm_topk_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
            attr_done[dv] = true;
        }
    }
    // CTA pairs (cta_group::2) pay off once there are >= 2 query tiles; Tuning::k2_pair = 0/1 forces.
    const int pair_env = ix->tune.k2_pair;

    __half* q16 = reinterpret_cast<__half*>(ws + L.q16);
    float* invq = reinterpret_cast<float*>(ws + L.invq);
    float* eps = reinterpret_cast<float*>(ws + L.eps);
    int32_t* qmap = reinterpret_cast<int32_t*>(ws + L.qmap);
    int32_t* flagged = reinterpret_cast<int32_t*>(ws + L.flagged);
    int32_t* nflag = reinterpret_cast<int32_t*>(ws + L.nflag);
    uint64_t* cand2 = reinterpret_cast<uint64_t*>(ws + L.cand2);

    for (size_t c0 = 0; c0 < queries.size(); c0 += kK2MaxChunk) {
        const int n = static_cast<int>(std::min<size_t>(kK2MaxChunk, queries.size() - c0));
        bool identity = (c0 == 0);
        for (int i = 0; i < n && identity; ++i) identity = (queries[c0 + i] == static_cast<int>(c0) + i);
        const int32_t* d_qmap = nullptr;
        if (!identity) {
            CRB_CUDA(cudaMemcpyAsync(qmap, queries.data() + c0, size_t(n) * 4, cudaMemcpyHostToDevice, stream));
            d_qmap = qmap;
        }
        const int n_pad = static_cast<int>(round_up(n, 256));   // zero rows up to a whole tile pair: no TMA out-of-bounds fill
        k2_prep_queries_kernel<<<n_pad, 128, 0, stream>>>(d_queries, d_qmap, n, ix->dim, q16, invq, eps);
        CRB_LAUNCH_CHECK();

        // one clear for the flag count, the fallback's arrival counters and the shared thresholds
        const int groups = k2_gthr_groups(k);
        CRB_CUDA(cudaMemsetAsync(nflag, 0, (L.gthr - L.nflag) + size_t(n) * groups * 4, stream));

        CUtensorMap map_q;
        if (int rc = make_map_f16(&map_q, q16, n_pad, ix->dim, kK2BlockM)) return rc;
        K2Params kp;
        memset(&kp, 0, sizeof kp);
        kp.gthr = ix->tune.k2_gthr ? reinterpret_cast<uint32_t*>(ws + L.gthr) : nullptr;
        kp.gthr_groups = groups;
        kp.gthr_every = ix->tune.k2_gthr_every;   // 0 = dense first, then four reads per octave
        kp.gthr_prepass = ix->tune.k2_prepass;
        kp.inv_norm = k2_inv;
        kp.mask = mask;
        kp.cand = cand2;
        kp.tau = reinterpret_cast<float*>(ws + L.tau);
        kp.n_rows = k2_n;
        kp.batch = n;
        kp.dim = ix->dim;
        kp.kprime = kprime;
        kp.metric = metric;
        const int64_t n_tiles = (k2_n + kK2BlockN - 1) / kK2BlockN;
        // only query rows that exist own a candidate list: small batches leave room for more stages
        kp.lstride = n >= kK2BlockM ? kK2BlockM : static_cast<int>(round_up(n, 32));
        const size_t list_bytes = size_t(kprime) * kp.lstride * 8;
        const size_t smem_room = size_t(227) * 1024 - 1024 - list_bytes - tail_bytes;
        const int stages1 = static_cast<int>(std::min<size_t>(smem_room / kK2StageBytes, 4));
        if (stages1 < 2) return fail("internal: K2 does not fit shared memory for k'=%d", kprime);
        const bool use_pair = pair_env >= 0 ? (pair_env != 0) : (n > kK2BlockM);
        kp.prefetch = ix->tune.k2_prefetch;
        kp.debug = ix->tune.k2_debug;
        if (use_pair) {
            const int m_pairs = (n + 255) / 256;                // query tile PAIRS
            // m-inner: one CTA pair serves up to 4 query-tile pairs per corpus tile while their candidate
            // lists fit next to >= 4 pipeline stages (CODERAG_B200_K2_MINNER=1 restores one group per pair)
            // (measured: 5 pipeline stages with 2 query-tile pairs per corpus tile beat 4 stages with 4 pairs
            //  up to 4 pairs -- 12.4 vs 13.0 ms at B=1024; beyond that fewer HBM re-reads win)
            // (after the shared admission bound made the epilogue cheap the two are equal at 10M rows -- 12.89 vs 12.90 ms,
            //  interleaved A/B -- and 4 pairs read the corpus from HBM once instead of 1.4 times (ncu: 15.4 vs 21.6 GB);
            //  2 pairs stay 4 % ahead on shards of a few million rows, where HBM is idle anyway)
            const int mi_default = m_pairs <= 4 ? (n_tiles >= 20000 ? 4 : 2) : 4;
            int mi = std::min(m_pairs, std::max(1, ix->tune.k2_minner > 0 ? ix->tune.k2_minner : mi_default));
            auto pair_stages = [&](int mi_) {
                const size_t fixed_b = size_t(mi_) * list_bytes + tail_bytes + size_t(mi_) * 1024;
                return static_cast<int>(std::min<size_t>((size_t(227) * 1024 - 1024 - fixed_b) / kK2PairStageBytes, 8));
            };
            while (mi > 1 && (size_t(mi) * list_bytes + tail_bytes + size_t(mi) * 1024 + 1024 + 4 * size_t(kK2PairStageBytes) >
                              size_t(227) * 1024))
                --mi;
            kp.m_inner = mi;
            kp.m_tiles = (m_pairs + mi - 1) / mi;               // query groups
            kp.stages = pair_stages(mi);
            const size_t smem = 1024 + size_t(kp.stages) * kK2PairStageBytes + size_t(mi) * list_bytes + tail_bytes +
                                size_t(mi) * 1024;
            kp.slots = std::max(1, (ix->sm_count / 2) / kp.m_tiles);
            kp.slots = static_cast<int>(std::min<int64_t>(kp.slots, std::max<int64_t>(n_tiles, 1)));
            k2_gemm_topk_pair_kernel<<<2 * kp.m_tiles * kp.slots, kK2Threads, smem, stream>>>(map_q, *map_half, kp);
        } else {
            kp.m_tiles = (n + kK2BlockM - 1) / kK2BlockM;
            kp.slots = std::max(1, ix->sm_count / kp.m_tiles);
            kp.slots = static_cast<int>(std::min<int64_t>(kp.slots, std::max<int64_t>(n_tiles, 1)));
            kp.stages = stages1;
            const size_t smem = 1024 + size_t(stages1) * kK2StageBytes + list_bytes + tail_bytes;
            k2_gemm_topk_kernel<<<kp.m_tiles * kp.slots, kK2Threads, smem, stream>>>(map_q, *map_full, kp);
        }
        CRB_LAUNCH_CHECK();

        K5Params rp;
        memset(&rp, 0, sizeof rp);
        rp.rows = ix->rows;
        rp.inv_norm = ix->inv_norm;
        rp.queries = d_queries;
        rp.q_map = d_qmap;
        rp.remap = remap;
        rp.cand = cand2;
        rp.tau = kp.tau;
        rp.invq = invq;
        rp.eps = eps;
        rp.out_scores = d_scores;
        rp.out_rows = d_rows;
        rp.out_nvalid = d_n_valid;
        rp.out_keys = d_keys;
        rp.gthr = kp.gthr;
        rp.gthr_groups = groups;
        rp.flagged = flagged;
        rp.n_flagged = nflag;
        rp.stats = ix->k5_stats;
        rp.max_norm = ix->max_norm;
        rp.row_offset = static_cast<uint32_t>(ix->row_offset);
        rp.dim = ix->dim;
        rp.k = k;
        rp.kprime = kprime;
        rp.slots = kp.slots;
        rp.metric = metric;
        if (ex) rp.ex = *ex;
        const size_t dyn = size_t(ix->dim) * sizeof(float);
        if (k <= 32) k5_rescore_kernel<1><<<n, 256, dyn, stream>>>(rp);
        else k5_rescore_kernel<4><<<n, 256, dyn, stream>>>(rp);
        CRB_LAUNCH_CHECK();

        // the queries K5 flagged, through the streaming scan (a one-word read per CTA when there are none)
        FbParams fp;
        memset(&fp, 0, sizeof fp);
        fp.rows = ix->rows;
        fp.inv_norm = ix->inv_norm;
        fp.queries = d_queries;
        fp.q_map = d_qmap;
        fp.flagged = flagged;
        fp.n_flagged = nflag;
        fp.list = surv_list;
        fp.list_count = surv_count;
        fp.cand = reinterpret_cast<uint64_t*>(ws + kWsHeaderBytes);
        fp.counters = reinterpret_cast<unsigned*>(ws + L.fbcnt);
        fp.out_scores = d_scores;
        fp.out_rows = d_rows;
        fp.out_nvalid = d_n_valid;
        fp.out_keys = d_keys;
        fp.n_rows = ix->n_rows;
        fp.cand_lists = int64_t(kWsCandBytes) / (int64_t(k) * 8);
        fp.row_offset = static_cast<uint32_t>(ix->row_offset);
        fp.dim = ix->dim;
        fp.k = k;
        fp.metric = metric;
        if (ex) fp.ex = *ex;
        typedef void (*FbFn)(const FbParams);
        const FbFn fb = k <= 32 ? static_cast<FbFn>(k1_fallback_kernel<4, 1>) : static_cast<FbFn>(k1_fallback_kernel<4, 4>);
        static std::mutex fb_mu;
        static std::map<std::pair<const void*, size_t>, int> fb_cache;
        int fb_occ = 0;
        {
            std::lock_guard<std::mutex> lk(fb_mu);
            auto fkey = std::make_pair(reinterpret_cast<const void*>(fb), dyn);
            auto fit = fb_cache.find(fkey);
            if (fit == fb_cache.end()) {
                int occ = 0;
                if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fb, kK1Threads, dyn) != cudaSuccess || occ < 1) occ = 1;
                fit = fb_cache.emplace(fkey, occ).first;
            }
            fb_occ = fit->second;
        }
        fb<<<static_cast<unsigned>(fb_occ * ix->sm_count), kK1Threads, dyn, stream>>>(fp);
        CRB_LAUNCH_CHECK();
    }
    (void)batch;
    return 0;
}

// ---------------------------------------------------------------------------------------------
// the search plan: dedup filters, build survivor lists, group queries, choose K1 / K2, launch
// ---------------------------------------------------------------------------------------------
// The filter pass: survivor lists (and, for the tensor-core plan, pass-bit masks) of `n` filters into the workspace.
// any_of: the filters are alternatives of ONE row set -- a row that passes any of them lands in list 0, once.
static int launch_k3(const crb_index* ix, const std::array<int32_t, kMaxKeys>* filters, const uint32_t* const* filter_masks,
                     int n, bool any_of, uint8_t* ws, uint32_t* lists, uint32_t* masks, cudaStream_t stream,
                     const RowWindow* win = nullptr) {
    // win (lo a multiple of 32; no K2 masks): only rows [lo, hi) are tested, the list entries are relative to lo
    const int64_t w_lo = win ? win->lo : 0, w_rows = win ? win->hi - win->lo : ix->n_rows;
    const int M = ix->n_keys;
    K3Params kp;
    memset(&kp, 0, sizeof kp);
    for (int m = 0; m < M; ++m) {
        kp.cols[m] = ix->cols[m] ? static_cast<const uint8_t*>(ix->cols[m]) + w_lo * ix->width[m] : nullptr;
        kp.width[m] = ix->width[m];
        kp.vec_ok[m] = reinterpret_cast<uintptr_t>(kp.cols[m]) % 16 == 0 ? 1 : 0;
    }
    kp.lists = lists;
    kp.counts = reinterpret_cast<uint32_t*>(ws + kWsCountsOffset);
    kp.masks = masks;
    kp.mask_stride = mask_stride_of(ix);
    kp.n_rows = w_rows;
    kp.list_stride = list_stride_of(ix);
    kp.n_keys = M;
    kp.n_filters = n;
    kp.any_of = any_of ? 1 : 0;
    for (int f = 0; f < n; ++f) {
        kp.extra[f] = filter_masks && filter_masks[f] ? filter_masks[f] + (w_lo >> 5) : nullptr;
        for (int m = 0; m < kMaxKeys; ++m) {
            kp.want[f][m] = m < M ? filters[f][m] : CRB_FILT_ANY;
            if (m < M && filters[f][m] != CRB_FILT_ANY) kp.used_keys |= (1u << m);
        }
    }
    for (int m = 0; m < M; ++m)
        if ((kp.used_keys >> m) & 1u) kp.ukeys[kp.n_used++] = static_cast<uint8_t>(m);
    CRB_CUDA(cudaMemsetAsync(ws + kWsCountsOffset, 0, kMaxFilters * sizeof(uint32_t), stream));
    const int64_t rows_per_block = int64_t(kK3Threads) * kK3RowsPerThread;
    const unsigned blocks = static_cast<unsigned>((w_rows + rows_per_block - 1) / rows_per_block);
    if (blocks > 0) {
        k3_filter_kernel<<<blocks, kK3Threads, 0, stream>>>(kp);
        CRB_LAUNCH_CHECK();
    }
    return 0;
}

// A filter code that EVERY row of the shard carries constrains nothing (the reference's worker pins namespace = "default"
// on every request, agent_graph.py:198-239): such a search takes the unfiltered plan.  Exact -- the histograms are
// counted over all n_rows at bind time, and deleted rows are excluded by their NaN norm, not by a code.
static bool code_is_trivial(const crb_index* ix, int m, int32_t code) {
    if (code < 0 || !ix->tune.drop_trivial) return false;
    const std::vector<uint32_t>& h = ix->hist[m];
    return size_t(code) + 1 < h.size() && int64_t(h[size_t(code) + 1]) == ix->n_rows;
}

static int search_impl(crb_index* ix, const float* d_queries, int batch, int k, int metric,
                       const int32_t* h_filt, float* d_scores, int64_t* d_rows,
                       int32_t* d_n_valid, uint64_t* d_keys, void* d_workspace,
                       size_t workspace_bytes, int algo, cudaStream_t stream, const ExchangeDev* ex = nullptr,
                       const uint32_t* const* h_mask_ptrs = nullptr, bool* folded = nullptr) {
    uint8_t* ws = static_cast<uint8_t*>(d_workspace);
    const int M = ix->n_keys;
    if (algo < 0 || algo > 2) return fail("algo must be 0 (auto), 1 (streaming) or 2 (tensor core)");

    // distinct filters of the batch
    std::vector<std::array<int32_t, kMaxKeys>> filters;
    std::vector<const uint32_t*> filter_masks;       // caller-built row mask of each distinct filter (or nullptr)
    std::vector<int> fid(batch, -1);
    if ((h_filt && M > 0) || h_mask_ptrs) {
        for (int b = 0; b < batch; ++b) {
            std::array<int32_t, kMaxKeys> f;
            f.fill(CRB_FILT_ANY);
            bool any = false;
            for (int m = 0; m < M && h_filt; ++m) {
                f[m] = h_filt[size_t(b) * M + m];
                if (f[m] < CRB_FILT_IMPOSSIBLE) return fail("filter code %d out of range", f[m]);
                if (code_is_trivial(ix, m, f[m])) f[m] = CRB_FILT_ANY;
                any |= (f[m] != CRB_FILT_ANY);
            }
            const uint32_t* rm = h_mask_ptrs ? h_mask_ptrs[b] : nullptr;
            if (!any && !rm) continue;
            int id = -1;
            for (size_t i = 0; i < filters.size(); ++i)
                if (filters[i] == f && filter_masks[i] == rm) { id = static_cast<int>(i); break; }
            if (id < 0) { id = static_cast<int>(filters.size()); filters.push_back(f); filter_masks.push_back(rm); }
            fid[b] = id;
        }
    }
    const int F = static_cast<int>(filters.size());
    const WsLayout L = ws_layout(ix, batch, k);
    const int64_t stride = list_stride_of(ix);
    const int64_t mstride = mask_stride_of(ix);
    if (workspace_bytes < L.filt) return fail("workspace too small: %zu < %zu", workspace_bytes, L.filt);
    int lists_fit = static_cast<int>(std::min<size_t>((workspace_bytes - L.filt) / L.per_filter, kMaxFilters));
    if (F > 0 && lists_fit < 1)
        return fail("workspace too small for one survivor list: need %zu bytes", L.filt + L.per_filter);
    lists_fit = std::min(lists_fit, std::max(F, 1));
    // whatever lies beyond the lists and masks holds compacted survivors for filtered K2 batches
    const size_t compact_off = round_up(L.filt + size_t(lists_fit) * L.per_filter, 1024);
    uint8_t* compact_buf = workspace_bytes > compact_off ? ws + compact_off : nullptr;
    const size_t compact_cap = workspace_bytes > compact_off ? workspace_bytes - compact_off : 0;
    uint32_t* lists = reinterpret_cast<uint32_t*>(ws + L.filt);
    uint32_t* masks = reinterpret_cast<uint32_t*>(ws + L.filt + size_t(std::max(lists_fit, 0)) * size_t(stride) * 4);

    const int chunks = ix->dim / 8;
    const int ch = (chunks + 31) / 32;
    const int nq_cap = ch > 4 ? 1 : (ch <= 2 ? 4 : 2);     // queries a filtered row set needs to exceed before K2 is considered
    // queries per streaming group: two share their row reads; four in a group are slower than two groups of two
    // everywhere (1M x 384, 10 % filter, B=4: 0.132 ms vs 0.089 ms; unfiltered 12.5M x 384: 10.3 ms vs 2 x 3.9 ms),
    // because the streaming kernel is instruction-bound and the four-query variant fits one CTA per SM only
    // (profiles/group_width_probe.py, small_batch_dispatch.py).
    const int nq_group = std::max(1, std::min(nq_cap, ix->tune.k1_group > 0 ? ix->tune.k1_group : (ch > 4 ? 1 : 2)));

    // one query under one satisfiable filter: filter and score in one pass, no survivor list.  Chosen by the
    // selectivity estimate below (CODERAG_B200_K1_FUSED=1 / 0 forces it on / off).
    // (measured, profiles/fused_sweep.py, fused vs K3 + K1: 10M x 1024 at 100 / 10 / 1 / 0.1 % passing 3.03 vs 3.44 ms, 368 vs
    //  394 us, 70 vs 83 us, 36 vs 48 us; 10M x 768: 2.32 vs 2.46 ms, 289 vs 289, 56 vs 65, 27 vs 40 us; 1M x 384: 195 vs 138,
    //  38 vs 35, 19.5 vs 24.8, 20 vs 25 us.)
    // ONE filtered row set for the whole call (a filtered query, or a group of queries under one filter): search inside
    // the row range its codes occupy.  Where every row of that range passes (each code's rows are one contiguous run),
    // the search is an unfiltered streaming pass over the range: no filter pass, no gather.
    RowWindow win;
    const RowWindow* winp = nullptr;
    bool one_row_set = F == 1 && batch <= nq_group && algo != 2 && ix->n_rows > 0 && ix->tune.zone_maps;
    for (int b = 0; b < batch && one_row_set; ++b) one_row_set = fid[b] == 0;
    if (one_row_set) {
        win = filter_window(ix, filters[0].data());
        if (filter_masks[0]) win.exact = false;
        if (win.exact) {
            std::vector<QueryGroup> groups(1);
            groups[0].filter = -1;
            for (int b = 0; b < batch; ++b) groups[0].queries.push_back(b);
            ExchangeDev ex_fold;
            const ExchangeDev* ex_k1 = ex;
            if (ex && folded && ix->tune.k1_fold) {       // sharded: this one launch also waits for the peers and merges
                ex_fold = *ex;
                ex_fold.fold = 1;
                ex_k1 = &ex_fold;
                *folded = true;
            }
            return launch_k1(ix, d_queries, k, metric, groups, std::vector<int>(), batch <= 1 ? 1 : (batch <= 2 ? 2 : 4), ws,
                             lists, d_scores, d_rows, d_n_valid, d_keys, stream, ex_k1, &win);
        }
        // a range that leaves out at least half the shard (codes in random order span all of it: nothing to gain, and
        // the fused kernel below is the better plan there)
        if ((win.hi - win.lo) * 2 <= ix->n_rows) {
            win.lo &= ~int64_t(511);           // whole mask words / 16-byte column loads
            winp = &win;
        }
    }
    // (inside a narrowed row range the survivors tend to come in runs -- that is why the range is narrow -- and the fused
    //  kernel hands whole 512-row stripes to single warps: 10M x 1024 clustered by its filter keys, 1 % passing as ONE run
    //  inside a 17 % range: 258 us fused against the list path's even split.  A narrowed range takes the filter pass + gather.)
    bool use_fused = false;
    if (batch == 1 && F == 1 && ch <= 4 && algo != 2 && ix->n_rows > 0 && ix->tune.k1_fused != 0 && !winp &&
        fused_eligible(ix, filters[0].data())) {
        if (ix->tune.k1_fused > 0) use_fused = true;
        else {
            // wide rows (dim > 512): the fused kernel wins at every selectivity measured; short rows: only when
            // few of the scanned rows pass (its per-row queueing costs more than the streaming kernel's whole row step)
            const double est = estimate_selectivity(ix, filters[0].data());
            use_fused = est >= 0.0 && est <= (ch >= 3 ? 1.0 : 0.03);
        }
    }
    if (use_fused) {
        bool impossible = false;
        for (int m = 0; m < M; ++m) impossible |= (filters[0][m] == CRB_FILT_IMPOSSIBLE);
        if (!impossible)
            return launch_k1_fused(ix, d_queries, 0, k, metric, filters[0].data(), filter_masks[0], ws, d_scores, d_rows,
                                   d_n_valid, d_keys, stream, ex);
    }
    const bool k2_possible = ix->k2_ok && ix->n_rows >= kK2MinRows && algo != 1 && k2_supports_k(k);

    // rounds: round 0 serves the unfiltered queries together with the first block of filters
    int f_lo = 0;
    bool first = true;
    while (first || f_lo < F) {
        const int f_hi = std::min(F, f_lo + std::max(lists_fit, 1));
        std::vector<int> slot_of_filter(std::max(F, 1), -1);

        // queries of this round by filter
        std::map<int, std::vector<int>> by_filter;
        for (int b = 0; b < batch; ++b) {
            const int f = fid[b];
            const bool mine = (f < 0) ? first : (f >= f_lo && f < f_hi);
            if (mine) by_filter[f].push_back(b);
        }
        // which row sets go to the tensor-core kernel
        std::map<int, bool> use_k2;
        bool any_k2_filtered = false;
        for (auto& kv : by_filter) {
            // The streaming kernel is instruction-bound, so a second query in a group costs almost a second
            // pass (10M x 768: B=1 2.29 ms, B=2 3.93 ms, B=4 6.9 ms) while the tensor-core kernel serves any
            // B <= 128 in one HBM pass (2.2 ms): an unfiltered row set with two or more queries goes there as
            // soon as the shard is big enough to hide its fixed cost (200k x 384, B=2: 0.056 vs 0.076 ms).
            // A filtered set still needs more queries than one streaming pass serves, because choosing costs
            // a survivor-count read-back.
            const int sz = static_cast<int>(kv.second.size());
            const bool big = kv.first < 0 ? (sz > nq_cap || (sz >= 2 && ix->n_rows >= kK2SmallBatchRows)) : sz > nq_cap;
            use_k2[kv.first] = k2_possible && (algo == 2 || big);
            any_k2_filtered |= (use_k2[kv.first] && kv.first >= 0);
        }

        if (f_hi > f_lo) {
            for (int f = f_lo; f < f_hi; ++f) slot_of_filter[f] = f - f_lo;
            if (int rc = launch_k3(ix, filters.data() + f_lo, filter_masks.data() + f_lo, f_hi - f_lo, false, ws, lists,
                                   any_k2_filtered ? masks : nullptr, stream, winp))
                return rc;
        }

        // A filtered row set is worth the tensor-core kernel (which scans the whole shard under a
        // pass-bit mask) only if that beats gathering the survivors once per nq_cap queries.  The
        // survivor counts are on the device: read them back (one small synchronising copy -- the only
        // point where a search waits for its stream, and only for filtered tensor-core row sets).
        uint32_t h_counts[kMaxFilters] = {0};
        if (any_k2_filtered) {
            CRB_CUDA(cudaMemcpyAsync(h_counts, ws + kWsCountsOffset, sizeof h_counts, cudaMemcpyDeviceToHost, stream));
            CRB_CUDA(cudaStreamSynchronize(stream));
            for (auto& kv : by_filter) {
                if (kv.first < 0 || !use_k2[kv.first] || algo != 0) continue;
                const double nq_set = static_cast<double>(kv.second.size());
                const double surv = static_cast<double>(h_counts[slot_of_filter[kv.first]]);
                const double row_bytes = double(ix->dim) * 2.0;
                const bool can_compact = surv * (row_bytes + 4.0) + 2048.0 <= double(compact_cap) && surv * 4.0 <= double(ix->n_rows);
                const double scan_rows = can_compact ? surv : double(ix->n_rows);
                // (a group of nq_cap queries shares the row reads but not the arithmetic: charge every query its bytes)
                const double t_k1 = nq_set * (surv * row_bytes / 5.0e12) + std::ceil(nq_set / nq_group) * 10e-6;
                const double t_k2 = std::max(scan_rows * row_bytes / 3.0e12, 2.0 * nq_set * scan_rows * ix->dim / 9.0e14) +
                                    (can_compact ? 2.0 * surv * row_bytes / 5.0e12 : 0.0) + 0.3e-3;
                if (t_k1 <= t_k2) use_k2[kv.first] = false;
            }
        }

        // tensor-core row sets first (what K5 cannot prove is re-run on the device, inside launch_k2)
        std::map<int, std::vector<int>> k1_sets;
        for (auto& kv : by_filter) {
            if (!use_k2[kv.first]) { k1_sets[kv.first] = kv.second; continue; }
            const uint32_t* mask = kv.first < 0 ? nullptr : masks + size_t(slot_of_filter[kv.first]) * size_t(mstride);
            int rc;
            if (kv.first < 0) {
                rc = launch_k2(ix, d_queries, batch, k, metric, kv.second, mask, ws, L, d_scores, d_rows, d_n_valid,
                               d_keys, stream, nullptr, nullptr, -1, nullptr, 0, ex);
            } else {
                const int slot_f = slot_of_filter[kv.first];
                rc = launch_k2(ix, d_queries, batch, k, metric, kv.second, mask, ws, L, d_scores, d_rows, d_n_valid,
                               d_keys, stream, lists + size_t(slot_f) * size_t(stride),
                               reinterpret_cast<const uint32_t*>(ws + kWsCountsOffset) + slot_f,
                               static_cast<int64_t>(std::min<uint64_t>(h_counts[slot_f], uint64_t(ix->n_rows))),
                               compact_buf, compact_cap, ex);
            }
            if (rc) return rc;
        }

        std::vector<QueryGroup> groups;
        int max_group = 1;
        for (auto& kv : k1_sets) {
            const std::vector<int>& qs = kv.second;
            for (size_t i = 0; i < qs.size(); i += nq_group) {
                QueryGroup g;
                g.filter = kv.first;
                g.queries.assign(qs.begin() + i, qs.begin() + std::min(qs.size(), i + nq_group));
                max_group = std::max<int>(max_group, static_cast<int>(g.queries.size()));
                groups.push_back(std::move(g));
            }
        }
        if (!groups.empty()) {
            const int nq = max_group <= 1 ? 1 : (max_group <= 2 ? 2 : 4);
            // single-query groups over different row sets (the adjacent() fan-out): split the grid by rows, not by group
            bool any_filtered = false;
            for (const QueryGroup& g : groups) any_filtered |= (g.filter >= 0);
            const bool balanced = max_group == 1 && groups.size() >= 2 && any_filtered && ix->tune.k1_balanced;
            // A sharded call that is exactly ONE streaming launch (the batch-1 request): its last CTA also waits for
            // the peers and merges, so no second launch follows.
            ExchangeDev ex_fold;
            const ExchangeDev* ex_k1 = ex;
            if (ex && folded && !balanced && first && f_hi == F && k1_sets.size() == by_filter.size() &&
                groups.size() <= size_t(kMaxGroups) && ix->tune.k1_fold) {
                ex_fold = *ex;
                ex_fold.fold = 1;
                ex_k1 = &ex_fold;
                *folded = true;
            }
            int rc = balanced ? launch_k1_balanced(ix, d_queries, k, metric, groups, slot_of_filter, ws, lists, d_scores,
                                                   d_rows, d_n_valid, d_keys, stream, ex)
                              : launch_k1(ix, d_queries, k, metric, groups, slot_of_filter, nq, ws, lists, d_scores,
                                          d_rows, d_n_valid, d_keys, stream, ex_k1, winp);
            if (rc) return rc;
        }
        first = false;
        f_lo = f_hi;
    }
    return 0;
}

// One query under a DISJUNCTION of filters: a row passes when ANY of the n_alt alternatives (each an AND of equalities)
// passes.  This is the adapter's adjacent(): one query vector, one row set per edge, the best k of their union
// (graph_rag_retrievers.py:82-134 hands the edges to graph-retriever, whose adapter runs one filtered search per edge and
// keeps the best k of the merged, de-duplicated hits).  The top k of a union is the top k of the per-set top k's, so one
// search over the union returns the same rows -- and a row in several sets (a file's rows are its module's and its
// repository's too) is read and scored once.
static int search_any_of_impl(crb_index* ix, const float* d_query, int n_alt, const int32_t* h_filt, int k, int metric,
                              float* d_scores, int64_t* d_rows, int32_t* d_n_valid, void* d_workspace,
                              size_t workspace_bytes, cudaStream_t stream, const ExchangeDev* ex = nullptr,
                              bool* folded = nullptr) {
    const int M = ix->n_keys;
    std::vector<std::array<int32_t, kMaxKeys>> alts;
    bool everything = false;
    for (int a = 0; a < n_alt && !everything; ++a) {
        std::array<int32_t, kMaxKeys> f;
        f.fill(CRB_FILT_ANY);
        bool any = false, impossible = false;
        for (int m = 0; m < M; ++m) {
            f[m] = h_filt[size_t(a) * M + m];
            if (f[m] < CRB_FILT_IMPOSSIBLE) return fail("filter code %d out of range", f[m]);
            impossible |= (f[m] == CRB_FILT_IMPOSSIBLE);
            if (code_is_trivial(ix, m, f[m])) f[m] = CRB_FILT_ANY;
            any |= (f[m] != CRB_FILT_ANY);
        }
        if (impossible) continue;                     // matches no row: adds nothing to the union
        if (!any) { everything = true; break; }       // matches every row: the union is the shard
        if (std::find(alts.begin(), alts.end(), f) == alts.end()) alts.push_back(f);
    }
    if (everything)
        return search_impl(ix, d_query, 1, k, metric, nullptr, d_scores, d_rows, d_n_valid, nullptr, d_workspace,
                           workspace_bytes, 0, stream, ex, nullptr, folded);
    if (alts.size() <= 1) {
        std::array<int32_t, kMaxKeys> f;
        f.fill(CRB_FILT_IMPOSSIBLE);
        if (!alts.empty()) f = alts[0];
        return search_impl(ix, d_query, 1, k, metric, f.data(), d_scores, d_rows, d_n_valid, nullptr, d_workspace,
                           workspace_bytes, 0, stream, ex, nullptr, folded);
    }
    if (alts.size() > size_t(kMaxFilters))
        return fail("at most %d distinct alternatives per call (got %zu)", kMaxFilters, alts.size());
    uint8_t* ws = static_cast<uint8_t*>(d_workspace);
    const WsLayout L = ws_layout(ix, 1, k);
    if (workspace_bytes < L.filt + L.per_filter)
        return fail("workspace too small for one survivor list: need %zu bytes", L.filt + L.per_filter);
    uint32_t* lists = reinterpret_cast<uint32_t*>(ws + L.filt);
    // the union lies inside the hull of the alternatives' row ranges; when every alternative is a contiguous run and the
    // runs join up into one (a file inside its module inside its repository), the union IS that range
    RowWindow hull;
    const RowWindow* winp = nullptr;
    bool exact_union = false;
    if (ix->tune.zone_maps && ix->n_rows > 0) {
        std::vector<std::pair<int64_t, int64_t>> runs;
        bool all_exact = true;
        hull.lo = ix->n_rows;
        hull.hi = 0;
        for (const auto& a : alts) {
            const RowWindow w = filter_window(ix, a.data());
            if (w.hi <= w.lo) continue;                       // this alternative matches no row
            all_exact &= w.exact;
            runs.emplace_back(w.lo, w.hi);
            hull.lo = std::min(hull.lo, w.lo);
            hull.hi = std::max(hull.hi, w.hi);
        }
        if (runs.empty()) { hull.lo = hull.hi = 0; exact_union = true; }
        else if (all_exact) {
            std::sort(runs.begin(), runs.end());
            int64_t reach = runs[0].second;
            exact_union = true;
            for (size_t i = 1; i < runs.size() && exact_union; ++i) {
                if (runs[i].first > reach) exact_union = false;          // a gap: rows between the runs do not pass
                reach = std::max(reach, runs[i].second);
            }
        }
        if (!exact_union) hull.lo &= ~int64_t(511);
        if (exact_union || (hull.hi - hull.lo) * 2 <= ix->n_rows) winp = &hull;
    }
    if (!exact_union)
        if (int rc = launch_k3(ix, alts.data(), nullptr, static_cast<int>(alts.size()), true, ws, lists, nullptr, stream, winp))
            return rc;
    std::vector<QueryGroup> groups(1);
    groups[0].filter = exact_union ? -1 : 0;
    groups[0].queries.assign(1, 0);
    const std::vector<int> slot_of_filter(1, 0);
    // sharded: the one streaming launch also waits for the peers and merges (as a batch-1 search does)
    ExchangeDev ex_fold;
    if (ex && folded && ix->tune.k1_fold) {
        ex_fold = *ex;
        ex_fold.fold = 1;
        ex = &ex_fold;
        *folded = true;
    }
    return launch_k1(ix, d_query, k, metric, groups, slot_of_filter, 1, ws, lists, d_scores, d_rows, d_n_valid, nullptr,
                     stream, ex, winp);
}

// ---------------------------------------------------------------------------------------------
// cross-shard merge
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) merge_keys_kernel(const uint64_t* __restrict__ keys_in,
                                                         int n_lists, int batch, int k, int metric,
                                                         const float* __restrict__ queries, int dim,
                                                         float* scores, int64_t* rows,
                                                         int32_t* n_valid, uint64_t* keys_out) {
    __shared__ uint64_t xchg[2][32];
    __shared__ float s_invq;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int b = blockIdx.x;
    if (warp == 0) {
        const float iq = (metric == 0) ? warp_inv_qnorm(queries + static_cast<int64_t>(b) * dim, dim, lane) : 1.0f;
        if (lane == 0) s_invq = iq;
    }
    __syncthreads();
    TopKSink sink;
    sink.scores = scores + static_cast<int64_t>(b) * k;
    sink.rows = rows + static_cast<int64_t>(b) * k;
    sink.n_valid = n_valid + b;
    sink.keys = keys_out ? keys_out + static_cast<int64_t>(b) * k : nullptr;
    sink.scale = s_invq;
    block_merge_sorted_lists<4>(keys_in + static_cast<int64_t>(b) * k, n_lists, k,
                                static_cast<int64_t>(batch) * k, sink, xchg, lane, warp, 8);
}

// Exchange epilogue: wait until every source rank has published `expected` queries into THIS rank's
// gather buffer (the sinks of the peers' search kernels do the publishing), then merge.  Runs on
// every GPU of the box at the same time; the spin is bounded so a lost peer cannot hang the GPU
// (n_valid = -2 marks a timed-out query).
__global__ void __launch_bounds__(256) merge_wait_kernel(const uint64_t* __restrict__ gather, uint64_t cap,
                                                         const uint32_t* counters, uint32_t expected,
                                                         int world, int batch, int k, int metric,
                                                         const float* __restrict__ queries, int dim,
                                                         float* scores, int64_t* rows, int32_t* n_valid,
                                                         uint64_t* keys_out) {
    __shared__ uint64_t xchg[2][32];
    __shared__ float s_invq;
    __shared__ int s_ok;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int b = blockIdx.x;
    if (threadIdx.x == 0) {
        const volatile uint32_t* c = counters;
        int ok = 1;
        for (int r = 0; r < world; ++r) {
            long long spins = 0;
            while (static_cast<int32_t>(c[r] - expected) < 0) {
                if (++spins > (1ll << 27)) { ok = 0; break; }
                __nanosleep(64);
            }
        }
        __threadfence_system();
        s_ok = ok;
    }
    if (warp == 1) {
        const float iq = (metric == 0) ? warp_inv_qnorm(queries + static_cast<int64_t>(b) * dim, dim, lane) : 1.0f;
        if (lane == 0) s_invq = iq;
    }
    __syncthreads();
    if (!s_ok) {
        if (threadIdx.x == 0) n_valid[b] = -2;
        return;
    }
    TopKSink sink;
    sink.scores = scores + static_cast<int64_t>(b) * k;
    sink.rows = rows + static_cast<int64_t>(b) * k;
    sink.n_valid = n_valid + b;
    sink.keys = keys_out ? keys_out + static_cast<int64_t>(b) * k : nullptr;
    sink.scale = s_invq;
    block_merge_sorted_lists<1>(gather + static_cast<int64_t>(b) * k, world, k, static_cast<int64_t>(cap), sink, xchg,
                                lane, warp, 8);
}

// ---------------------------------------------------------------------------------------------
// lanes for the host-buffer entry points
// ---------------------------------------------------------------------------------------------
static Lane* checkout_lane(crb_index* ix) {
    std::unique_lock<std::mutex> lk(ix->mu);
    for (;;) {
        for (auto& l : ix->lanes)
            if (!l.busy) { l.busy = true; return &l; }
        ix->cv.wait(lk);
    }
}
static void checkin_lane(crb_index* ix, Lane* l) {
    {
        std::lock_guard<std::mutex> lk(ix->mu);
        l->busy = false;
    }
    ix->cv.notify_one();
}
struct LaneHold {
    crb_index* ix;
    Lane* lane;
    ~LaneHold() { checkin_lane(ix, lane); }
};

// Small results skip the device-to-host copy and the stream synchronisation: the search kernels write scores / rows /
// n_valid straight into the lane's pinned host buffer (device-visible under unified addressing), a one-thread kernel
// behind them on the stream writes a sequence number into a pinned word, and the host spins on that word.
constexpr size_t kZeroCopyMaxBytes = 4096;
__global__ void signal_flag_kernel(volatile uint32_t* flag, uint32_t seq) {
    __threadfence_system();
    *flag = seq;
}
// (graph form: the sequence number travels with the query in the copy-in, so the recorded launch never changes)
__global__ void signal_flag_from_kernel(volatile uint32_t* flag, const uint32_t* seq) {
    __threadfence_system();
    *flag = *seq;
}
static void lane_drop_graph(Lane* l) {
    if (l->g_exec) cudaGraphExecDestroy(l->g_exec);
    l->g_exec = nullptr;
}
static int lane_signal(Lane* l) {
    l->flag_seq += 1;
    signal_flag_kernel<<<1, 1, 0, l->stream>>>(l->h_flag, l->flag_seq);
    CRB_LAUNCH_CHECK();
    return 0;
}
static int lane_wait_flag(Lane* l) {
    const uint32_t want = l->flag_seq;
    for (uint64_t spins = 0;; ++spins) {
        if (*l->h_flag == want) return 0;
        if ((spins & 4095) == 4095) {                 // a failed launch never raises the flag: ask the stream now and then
            const cudaError_t e = cudaStreamQuery(l->stream);
            if (e == cudaSuccess) { if (*l->h_flag == want) return 0; }
            else if (e != cudaErrorNotReady) return fail("search failed on the device: %s", cudaGetErrorString(e));
        }
#if defined(__x86_64__)
        __builtin_ia32_pause();
#endif
    }
}

static int lane_reserve(Lane* l, size_t ws_bytes, size_t io_bytes, size_t pin_bytes) {
    if (!l->stream) CRB_CUDA(cudaStreamCreateWithFlags(&l->stream, cudaStreamNonBlocking));
    if (!l->h_flag) {
        void* f = nullptr;
        CRB_CUDA(cudaMallocHost(&f, 64));
        memset(f, 0, 64);
        l->h_flag = static_cast<volatile uint32_t*>(f);
    }
    if (l->ws_cap < ws_bytes) {
        if (l->d_ws) CRB_CUDA(cudaFree(l->d_ws));
        l->d_ws = nullptr; l->ws_cap = 0;
        CRB_CUDA(cudaMalloc(&l->d_ws, ws_bytes));
        CRB_CUDA(cudaMemsetAsync(l->d_ws, 0, kWsHeaderBytes, l->stream));
        l->ws_cap = ws_bytes;
    }
    if (l->d_io_cap < io_bytes) {
        if (l->d_io) CRB_CUDA(cudaFree(l->d_io));
        l->d_io = nullptr; l->d_io_cap = 0;
        CRB_CUDA(cudaMalloc(&l->d_io, io_bytes));
        l->d_io_cap = io_bytes;
    }
    if (l->h_cap < pin_bytes) {
        if (l->h_pin) CRB_CUDA(cudaFreeHost(l->h_pin));
        l->h_pin = nullptr; l->h_cap = 0;
        CRB_CUDA(cudaMallocHost(&l->h_pin, pin_bytes));
        l->h_cap = pin_bytes;
    }
    return 0;
}

// Is this host pointer page-locked (cudaHostRegister / cudaMallocHost)?  Then a copy can DMA straight from it and
// the staging memcpy into the lane's own pinned buffer is skipped (worth asking only for large batches).
static bool host_is_pinned(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { (void)cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

static int count_distinct_filters(const int32_t* h_filt, int batch, int M) {
    if (!h_filt || M <= 0) return 0;
    std::vector<std::vector<int32_t>> seen;
    for (int b = 0; b < batch; ++b) {
        std::vector<int32_t> f(h_filt + size_t(b) * M, h_filt + size_t(b + 1) * M);
        bool any = false;
        for (int32_t v : f) any |= (v != CRB_FILT_ANY);
        if (!any) continue;
        if (std::find(seen.begin(), seen.end(), f) == seen.end()) seen.push_back(f);
    }
    return static_cast<int>(seen.size());
}

static int validate_search(const crb_index* ix, int batch, int k, int metric) {
    if (!ix) return fail("index is NULL");
    if (batch < 1) return fail("batch must be >= 1 (got %d)", batch);
    if (k < 1 || k > CRB_MAX_TOPK) return fail("k must be in [1, %d] (got %d)", CRB_MAX_TOPK, k);
    if (metric != CRB_METRIC_COSINE && metric != CRB_METRIC_IP) return fail("unknown metric %d", metric);
    return 0;
}

}  // namespace crb

namespace crb {
// (Re)bind the corpus buffers of an index: pointers, row count, corpus tensor maps, largest norm.
static int bind_buffers(crb_index* ix, const void* d_rows_f16, const float* d_inv_norm, const int32_t* d_codes,
                        int64_t n_rows, int64_t code_stride) {
    ix->rows = static_cast<const __half*>(d_rows_f16);
    ix->inv_norm = d_inv_norm;
    ix->codes = d_codes;
    ix->code_stride = code_stride;
    for (int m = 0; m < ix->n_keys; ++m) {
        ix->cols[m] = d_codes ? d_codes + int64_t(m) * code_stride : nullptr;
        ix->width[m] = 4;
        ix->hist[m].clear();
        ix->zone_first[m].clear();
        ix->zone_last[m].clear();
    }
    ix->n_rows = n_rows;
    ix->bind_gen += 1;
    ix->k2_ok = false;
    ix->max_norm = 0.f;
    // tensor-core path set-up: corpus tensor maps and the largest row norm
    if (n_rows > 0 && make_map_f16(&ix->map_rows, d_rows_f16, n_rows, ix->dim, kK2BlockN) == 0 &&
        make_map_f16(&ix->map_rows_half, d_rows_f16, n_rows, ix->dim, kK2BlockN / 2) == 0) {
        unsigned* d_bits = nullptr;
        const unsigned inf_bits = 0x7f800000u;
        unsigned h_bits = inf_bits;
        cudaError_t e = cudaMalloc(&d_bits, sizeof(unsigned));
        if (e == cudaSuccess) e = cudaMemcpy(d_bits, &inf_bits, sizeof(unsigned), cudaMemcpyHostToDevice);
        if (e == cudaSuccess) {
            min_pos_inv_norm_kernel<<<static_cast<unsigned>(std::min<int64_t>((n_rows + 255) / 256, 148 * 8)), 256>>>(d_inv_norm, n_rows, d_bits);
            g_launches.fetch_add(1, std::memory_order_relaxed);
            e = cudaMemcpy(&h_bits, d_bits, sizeof(unsigned), cudaMemcpyDeviceToHost);
        }
        if (d_bits) cudaFree(d_bits);
        if (e != cudaSuccess) return fail("index set-up failed: %s", cudaGetErrorString(e));
        float min_inv;
        memcpy(&min_inv, &h_bits, 4);
        ix->max_norm = (h_bits == inf_bits || !(min_inv > 0.f)) ? 0.f : 1.0f / min_inv;
        ix->k2_ok = true;
    }
    return 0;
}
}  // namespace crb

// =============================================================================================
// extern "C"
// =============================================================================================
using namespace crb;

extern "C" int crb_abi_version(void) { return CRB_ABI_VERSION; }
extern "C" const char* crb_last_error(void) { return g_err.c_str(); }
extern "C" int64_t crb_launch_count(void) { return g_launches.load(); }

extern "C" int crb_index_create(crb_index** out, const void* d_rows_f16, const float* d_inv_norm,
                                const int32_t* d_codes, int64_t n_rows, int32_t dim, int32_t n_keys,
                                int64_t row_offset, int32_t device) {
    if (!out) return fail("out is NULL");
    *out = nullptr;
    if (n_rows < 0) return fail("n_rows must be >= 0");
    if (dim < 8 || dim % 8 != 0 || dim > CRB_MAX_DIM)
        return fail("dim must be a multiple of 8 in [8, %d] (got %d)", CRB_MAX_DIM, dim);
    if (n_keys < 0 || n_keys > CRB_MAX_KEYS) return fail("n_keys must be in [0, %d]", CRB_MAX_KEYS);
    if (n_rows > 0 && (!d_rows_f16 || !d_inv_norm)) return fail("rows / inv_norm pointers are NULL");
    if (n_keys > 0 && n_rows > 0 && !d_codes) return fail("codes pointer is NULL with n_keys > 0");
    if (reinterpret_cast<uintptr_t>(d_rows_f16) % 16 != 0) return fail("rows must be 16-byte aligned");
    if (row_offset < 0 || row_offset + n_rows > int64_t(0xffffffffll))
        return fail("row_offset + n_rows must fit in 32 bits");
    DeviceGuard guard(device);
    CRB_CUDA(guard.status);
    cudaDeviceProp prop;
    CRB_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return fail("device %d is sm_%d%d; this library is built for sm_100a (B200) only", device,
                    prop.major, prop.minor);
    crb_index* ix = new crb_index();
    ix->device = device;
    ix->sm_count = prop.multiProcessorCount;
    ix->row_offset = row_offset;
    ix->dim = dim;
    ix->n_keys = n_keys;
    ix->tune = Tuning::from_env();
    ix->lanes.resize(ix->tune.lanes);
    if (int rc = bind_buffers(ix, d_rows_f16, d_inv_norm, d_codes, n_rows, n_rows)) {
        delete ix;
        return rc;
    }
    *out = ix;
    return 0;
}

extern "C" int crb_index_rebind(crb_index* ix, const void* d_rows_f16, const float* d_inv_norm,
                                const int32_t* d_codes, int64_t n_rows, int64_t code_stride) {
    if (!ix) return fail("index is NULL");
    if (n_rows < 0) return fail("n_rows must be >= 0");
    if (n_rows > 0 && (!d_rows_f16 || !d_inv_norm)) return fail("rows / inv_norm pointers are NULL");
    if (ix->n_keys > 0 && n_rows > 0 && !d_codes) return fail("codes pointer is NULL with n_keys > 0");
    if (code_stride < n_rows) return fail("code_stride must be >= n_rows");
    if (reinterpret_cast<uintptr_t>(d_rows_f16) % 16 != 0) return fail("rows must be 16-byte aligned");
    if (ix->row_offset + n_rows > int64_t(0xffffffffll)) return fail("row_offset + n_rows must fit in 32 bits");
    DeviceGuard guard(ix->device);
    CRB_CUDA(guard.status);
    return bind_buffers(ix, d_rows_f16, d_inv_norm, d_codes, n_rows, code_stride);
}

extern "C" int crb_index_rebind_columns(crb_index* ix, const void* d_rows_f16, const float* d_inv_norm,
                                        const void* const* h_code_cols, const int32_t* h_code_widths, int64_t n_rows) {
    if (!ix) return fail("index is NULL");
    if (n_rows < 0) return fail("n_rows must be >= 0");
    if (n_rows > 0 && (!d_rows_f16 || !d_inv_norm)) return fail("rows / inv_norm pointers are NULL");
    if (ix->n_keys > 0 && (!h_code_cols || !h_code_widths)) return fail("column pointers / widths are NULL with n_keys > 0");
    for (int m = 0; m < ix->n_keys; ++m) {
        const int w = h_code_widths[m];
        if (w != 1 && w != 2 && w != 4) return fail("code column %d: width must be 1, 2 or 4 bytes (got %d)", m, w);
        if (n_rows > 0 && !h_code_cols[m]) return fail("code column %d is NULL", m);
        if (reinterpret_cast<uintptr_t>(h_code_cols[m]) % w != 0) return fail("code column %d is not aligned to its width", m);
    }
    if (reinterpret_cast<uintptr_t>(d_rows_f16) % 16 != 0) return fail("rows must be 16-byte aligned");
    if (ix->row_offset + n_rows > int64_t(0xffffffffll)) return fail("row_offset + n_rows must fit in 32 bits");
    DeviceGuard guard(ix->device);
    CRB_CUDA(guard.status);
    if (int rc = bind_buffers(ix, d_rows_f16, d_inv_norm, nullptr, n_rows, n_rows)) return rc;
    unsigned* d_hist = nullptr;            // [3][bins]: counts, first rows, last rows
    constexpr size_t kMaxBins = 32770;
    for (int m = 0; m < ix->n_keys; ++m) {
        ix->cols[m] = h_code_cols[m];
        ix->width[m] = static_cast<uint8_t>(h_code_widths[m]);
        ix->hist[m].clear();
        ix->zone_first[m].clear();
        ix->zone_last[m].clear();
        if (h_code_widths[m] > 2 || n_rows == 0) continue;
        const size_t bins = h_code_widths[m] == 1 ? 130 : kMaxBins;
        if (!d_hist) CRB_CUDA(cudaMalloc(&d_hist, 3 * kMaxBins * sizeof(unsigned)));
        cudaError_t e = cudaMemset(d_hist, 0, bins * sizeof(unsigned));
        if (e == cudaSuccess) e = cudaMemset(d_hist + kMaxBins, 0xff, bins * sizeof(unsigned));
        if (e == cudaSuccess) e = cudaMemset(d_hist + 2 * kMaxBins, 0, bins * sizeof(unsigned));
        if (e == cudaSuccess) {
            code_histogram_kernel<<<static_cast<unsigned>(std::min<int64_t>((n_rows + 255) / 256, int64_t(ix->sm_count) * 8)), 256>>>(
                h_code_cols[m], h_code_widths[m], n_rows, d_hist, d_hist + kMaxBins, d_hist + 2 * kMaxBins);
            g_launches.fetch_add(1, std::memory_order_relaxed);
            ix->hist[m].resize(bins);
            ix->zone_first[m].resize(bins);
            ix->zone_last[m].resize(bins);
            e = cudaMemcpy(ix->hist[m].data(), d_hist, bins * sizeof(unsigned), cudaMemcpyDeviceToHost);
            if (e == cudaSuccess)
                e = cudaMemcpy(ix->zone_first[m].data(), d_hist + kMaxBins, bins * sizeof(unsigned), cudaMemcpyDeviceToHost);
            if (e == cudaSuccess)
                e = cudaMemcpy(ix->zone_last[m].data(), d_hist + 2 * kMaxBins, bins * sizeof(unsigned), cudaMemcpyDeviceToHost);
        }
        if (e != cudaSuccess) {
            cudaFree(d_hist);
            ix->hist[m].clear();
            ix->zone_first[m].clear();
            ix->zone_last[m].clear();
            return fail("code histogram failed: %s", cudaGetErrorString(e));
        }
    }
    if (d_hist) cudaFree(d_hist);
    return 0;
}

extern "C" int crb_index_reload_tuning(crb_index* ix) {
    if (!ix) return fail("index is NULL");
    const int lanes = ix->tune.lanes;
    ix->tune = Tuning::from_env();
    ix->tune.lanes = lanes;          // the lane pool is sized at creation
    ix->bind_gen += 1;
    return 0;
}

extern "C" int crb_index_set_k5_stats(crb_index* ix, float* d_stats) {
    if (!ix) return fail("index is NULL");
    ix->k5_stats = d_stats;
    return 0;
}

extern "C" int crb_tombstone_rows(crb_index* ix, const int64_t* h_rows_idx, int32_t n) {
    if (!ix) return fail("index is NULL");
    if (n <= 0) return 0;
    if (!h_rows_idx) return fail("NULL pointer passed to crb_tombstone_rows");
    for (int i = 0; i < n; ++i)
        if (h_rows_idx[i] < 0 || h_rows_idx[i] >= ix->n_rows)
            return fail("row index %lld out of range [0, %lld)", (long long)h_rows_idx[i], (long long)ix->n_rows);
    DeviceGuard guard(ix->device);
    CRB_CUDA(guard.status);
    LaneHold hold{ix, checkout_lane(ix)};
    Lane* l = hold.lane;
    const size_t bytes = round_up(size_t(n) * sizeof(int64_t), 256);
    if (int rc = lane_reserve(l, 0, bytes, bytes)) return rc;
    memcpy(l->h_pin, h_rows_idx, size_t(n) * sizeof(int64_t));
    CRB_CUDA(cudaMemcpyAsync(l->d_io, l->h_pin, size_t(n) * sizeof(int64_t), cudaMemcpyHostToDevice, l->stream));
    tombstone_rows_kernel<<<(n + 255) / 256, 256, 0, l->stream>>>(const_cast<float*>(ix->inv_norm),
                                                                  const_cast<__half*>(ix->rows), ix->dim,
                                                                  static_cast<const int64_t*>(l->d_io), n);
    CRB_LAUNCH_CHECK();
    CRB_CUDA(cudaStreamSynchronize(l->stream));
    return 0;
}

extern "C" int crb_index_destroy(crb_index* ix) {
    if (!ix) return 0;
    DeviceGuard guard(ix->device);
    for (auto& l : ix->lanes) {
        if (l.stream) cudaStreamSynchronize(l.stream);
        if (l.g_exec) cudaGraphExecDestroy(l.g_exec);
        if (l.d_ws) cudaFree(l.d_ws);
        if (l.d_io) cudaFree(l.d_io);
        if (l.h_pin) cudaFreeHost(l.h_pin);
        if (l.h_flag) cudaFreeHost(const_cast<uint32_t*>(l.h_flag));
        if (l.stream) cudaStreamDestroy(l.stream);
    }
    delete ix;
    return 0;
}

extern "C" int crb_encode_rows(const float* d_src_f32, void* d_dst_f16, float* d_inv_norm,
                               int64_t n_rows, int32_t dim, int32_t device, void* stream) {
    if (dim < 8 || dim % 8 != 0) return fail("dim must be a positive multiple of 8");
    if (n_rows <= 0) return 0;
    DeviceGuard guard(device);
    CRB_CUDA(guard.status);
    const unsigned blocks = static_cast<unsigned>(std::min<int64_t>((n_rows + 7) / 8, 148 * 16));
    encode_rows_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        d_src_f32, static_cast<__half*>(d_dst_f16), d_inv_norm, n_rows, dim);
    CRB_LAUNCH_CHECK();
    return 0;
}

extern "C" int crb_row_inv_norms(const void* d_rows_f16, float* d_inv_norm, int64_t n_rows,
                                 int32_t dim, int32_t device, void* stream) {
    if (dim < 8 || dim % 8 != 0) return fail("dim must be a positive multiple of 8");
    if (n_rows <= 0) return 0;
    DeviceGuard guard(device);
    CRB_CUDA(guard.status);
    const unsigned blocks = static_cast<unsigned>(std::min<int64_t>((n_rows + 7) / 8, 148 * 16));
    row_inv_norms_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __half*>(d_rows_f16), d_inv_norm, n_rows, dim);
    CRB_LAUNCH_CHECK();
    return 0;
}

extern "C" size_t crb_workspace_bytes(const crb_index* ix, int32_t batch, int32_t k, int32_t n_filters) {
    if (!ix) return 0;
    const int lists = std::min<int>(std::max<int>(n_filters, 0), kMaxFilters);
    const WsLayout L = ws_layout(ix, std::max(batch, 1), std::min(std::max(k, 1), CRB_MAX_TOPK));
    size_t total = L.filt + size_t(lists) * L.per_filter;
    if (lists > 0 && batch > 1) {
        // room to compact the survivors of a filtered tensor-core batch (up to a quarter of the shard,
        // capped by CODERAG_B200_COMPACT_MB, default 1024)
        const size_t cap = size_t(ix->tune.compact_mb) << 20;
        const size_t quarter = size_t(ix->n_rows / 4 + 1) * (size_t(ix->dim) * 2 + 4) + 4096;
        total = round_up(total, 1024) + std::min(cap, quarter);
    }
    return total;
}

extern "C" int crb_search(crb_index* ix, const float* d_queries, int32_t batch, int32_t k,
                          int32_t metric, const int32_t* h_filt, float* d_scores, int64_t* d_rows,
                          int32_t* d_n_valid, uint64_t* d_keys, void* d_workspace,
                          size_t workspace_bytes, int32_t algo, void* stream) {
    if (int rc = validate_search(ix, batch, k, metric)) return rc;
    if (!d_queries || !d_scores || !d_rows || !d_n_valid || !d_workspace)
        return fail("NULL device pointer passed to crb_search");
    DeviceGuard guard(ix->device);
    CRB_CUDA(guard.status);
    return search_impl(ix, d_queries, batch, k, metric, h_filt, d_scores, d_rows, d_n_valid, d_keys,
                       d_workspace, workspace_bytes, algo, static_cast<cudaStream_t>(stream));
}

extern "C" int crb_search_masked(crb_index* ix, const float* d_queries, int32_t batch, int32_t k,
                                 int32_t metric, const int32_t* h_filt, const uint32_t* const* h_row_mask_ptrs,
                                 float* d_scores, int64_t* d_rows, int32_t* d_n_valid, uint64_t* d_keys,
                                 void* d_workspace, size_t workspace_bytes, int32_t algo, void* stream) {
    if (int rc = validate_search(ix, batch, k, metric)) return rc;
    if (!d_queries || !d_scores || !d_rows || !d_n_valid || !d_workspace)
        return fail("NULL device pointer passed to crb_search_masked");
    DeviceGuard guard(ix->device);
    CRB_CUDA(guard.status);
    return search_impl(ix, d_queries, batch, k, metric, h_filt, d_scores, d_rows, d_n_valid, d_keys,
                       d_workspace, workspace_bytes, algo, static_cast<cudaStream_t>(stream), nullptr,
                       h_row_mask_ptrs);
}

namespace crb {
// crb_search_host and crb_search_host_vectors: host buffers in, host buffers out, one lane per caller
static int search_host_impl(crb_index* ix, const float* h_queries, int32_t batch, int32_t k, int32_t metric,
                            const int32_t* h_filt, float* h_scores, int64_t* h_rows, int32_t* h_n_valid,
                            float* h_vectors, int32_t algo, int32_t n_alt = 0) {
    // n_alt > 0: ONE query (batch == 1) under a disjunction of n_alt filter rows (crb_search_host_any_of)
    if (int rc = validate_search(ix, batch, k, metric)) return rc;
    if (!h_queries || !h_scores || !h_rows || !h_n_valid) return fail("NULL host pointer passed to crb_search_host");
    DeviceGuard guard(ix->device);
    CRB_CUDA(guard.status);
    LaneHold hold{ix, checkout_lane(ix)};
    Lane* l = hold.lane;

    const size_t q_bytes = size_t(batch) * ix->dim * sizeof(float);
    const size_t s_bytes = size_t(batch) * k * sizeof(float);
    const size_t r_bytes = size_t(batch) * k * sizeof(int64_t);
    const size_t v_bytes = size_t(batch) * sizeof(int32_t);
    const size_t e_bytes = h_vectors ? size_t(batch) * k * ix->dim * sizeof(float) : 0;
    // device io: [queries, flag sequence word | rows (8B aligned) | scores | n_valid | vectors]; pinned mirrors the same layout
    const size_t off_seq = round_up(q_bytes, 16);
    const size_t off_rows = round_up(off_seq + 16, 256);
    const size_t off_scores = off_rows + round_up(r_bytes, 256);
    const size_t off_valid = off_scores + round_up(s_bytes, 256);
    const size_t off_vec = off_valid + round_up(v_bytes, 256);
    const size_t io_bytes = off_vec + round_up(e_bytes, 256);
    const size_t ws_bytes = crb_workspace_bytes(ix, batch, k, n_alt > 0 ? 1 : count_distinct_filters(h_filt, batch, ix->n_keys));
    if (int rc = lane_reserve(l, ws_bytes, io_bytes, io_bytes)) return rc;

    uint8_t* d_io = static_cast<uint8_t*>(l->d_io);
    uint8_t* h_io = static_cast<uint8_t*>(l->h_pin);
    memcpy(h_io, h_queries, q_bytes);
    const bool zc = ix->tune.zero_copy && !h_vectors && (io_bytes - off_rows) <= kZeroCopyMaxBytes;
    uint8_t* o_io = zc ? h_io : d_io;          // where the kernels write the result

    // One unfiltered query with a zero-copy result is always the same three enqueues (copy-in, ONE streaming launch, flag)
    // on the same addresses: from the second such call on they are replayed as one graph launch.
    const bool plain1 = batch == 1 && n_alt == 0 && zc && algo != 2 && ix->tune.host_graph && ix->n_rows > 0 &&
                        (!h_filt || ix->n_keys == 0 || count_distinct_filters(h_filt, 1, ix->n_keys) == 0);
    if (plain1) {
        const bool have = l->g_exec && l->g_gen == ix->bind_gen && l->g_k == k && l->g_metric == metric &&
                          l->g_dio == l->d_io && l->g_hpin == l->h_pin && l->g_ws == l->d_ws;
        const bool seen = l->seen_gen == ix->bind_gen && l->seen_k == k && l->seen_metric == metric;
        if (!have && seen && !l->g_failed) {
            // record (nothing runs while recording; the caches the launch path fills were filled by the previous call)
            lane_drop_graph(l);
            cudaGraph_t graph = nullptr;
            bool ok = cudaStreamBeginCapture(l->stream, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
            if (ok) {
                ok = cudaMemcpyAsync(d_io, h_io, off_seq + 4, cudaMemcpyHostToDevice, l->stream) == cudaSuccess;
                if (ok)
                    ok = search_impl(ix, reinterpret_cast<const float*>(d_io), 1, k, metric, nullptr,
                                     reinterpret_cast<float*>(o_io + off_scores), reinterpret_cast<int64_t*>(o_io + off_rows),
                                     reinterpret_cast<int32_t*>(o_io + off_valid), nullptr, l->d_ws, l->ws_cap, algo,
                                     l->stream) == 0;
                if (ok) {
                    g_launches.fetch_sub(1, std::memory_order_relaxed);      // recorded, not launched (replays are counted)
                    signal_flag_from_kernel<<<1, 1, 0, l->stream>>>(l->h_flag, reinterpret_cast<const uint32_t*>(d_io + off_seq));
                    ok = cudaGetLastError() == cudaSuccess;
                }
                ok = (cudaStreamEndCapture(l->stream, &graph) == cudaSuccess) && ok && graph;
            }
            if (ok) ok = cudaGraphInstantiate(&l->g_exec, graph, 0) == cudaSuccess;
            if (graph) cudaGraphDestroy(graph);
            if (ok) {
                l->g_gen = ix->bind_gen; l->g_k = k; l->g_metric = metric;
                l->g_dio = l->d_io; l->g_hpin = l->h_pin; l->g_ws = l->d_ws;
            } else {
                (void)cudaGetLastError();
                l->g_exec = nullptr;
                l->g_failed = true;                 // the ordinary path below serves this and later calls
            }
        }
        if (!seen) { l->seen_gen = ix->bind_gen; l->seen_k = k; l->seen_metric = metric; l->g_failed = false; }
        if (l->g_exec && l->g_gen == ix->bind_gen && l->g_k == k && l->g_metric == metric && l->g_dio == l->d_io &&
            l->g_hpin == l->h_pin && l->g_ws == l->d_ws) {
            l->flag_seq += 1;
            *reinterpret_cast<uint32_t*>(h_io + off_seq) = l->flag_seq;
            if (cudaGraphLaunch(l->g_exec, l->stream) == cudaSuccess) {
                g_launches.fetch_add(2, std::memory_order_relaxed);      // the recorded streaming launch and flag launch
                if (int rw = lane_wait_flag(l)) return rw;
                memcpy(h_rows, h_io + off_rows, r_bytes);
                memcpy(h_scores, h_io + off_scores, s_bytes);
                memcpy(h_n_valid, h_io + off_valid, v_bytes);
                return 0;
            }
            (void)cudaGetLastError();
            l->flag_seq -= 1;
            lane_drop_graph(l);
            l->g_failed = true;
        }
    }
    CRB_CUDA(cudaMemcpyAsync(d_io, h_io, q_bytes, cudaMemcpyHostToDevice, l->stream));
    int rc = n_alt > 0
        ? search_any_of_impl(ix, reinterpret_cast<const float*>(d_io), n_alt, h_filt, k, metric,
                             reinterpret_cast<float*>(o_io + off_scores), reinterpret_cast<int64_t*>(o_io + off_rows),
                             reinterpret_cast<int32_t*>(o_io + off_valid), l->d_ws, l->ws_cap, l->stream)
        : search_impl(ix, reinterpret_cast<const float*>(d_io), batch, k, metric, h_filt,
                      reinterpret_cast<float*>(o_io + off_scores),
                      reinterpret_cast<int64_t*>(o_io + off_rows),
                      reinterpret_cast<int32_t*>(o_io + off_valid), nullptr, l->d_ws, l->ws_cap,
                      algo, l->stream);
    if (rc) { cudaStreamSynchronize(l->stream); return rc; }
    if (zc) {
        if (int rs = lane_signal(l)) { cudaStreamSynchronize(l->stream); return rs; }
        if (int rw = lane_wait_flag(l)) return rw;
        memcpy(h_rows, h_io + off_rows, r_bytes);
        memcpy(h_scores, h_io + off_scores, s_bytes);
        memcpy(h_n_valid, h_io + off_valid, v_bytes);
        return 0;
    }
    if (h_vectors && ix->n_rows > 0) {
        const int n = batch * k;
        const unsigned blocks = static_cast<unsigned>(std::min<int64_t>((int64_t(n) * ix->dim + 255) / 256, int64_t(ix->sm_count) * 8));
        fetch_result_rows_kernel<<<blocks, 256, 0, l->stream>>>(ix->rows, reinterpret_cast<const int64_t*>(d_io + off_rows), n,
                                                                ix->row_offset, ix->dim, reinterpret_cast<float*>(d_io + off_vec));
        CRB_LAUNCH_CHECK();
    }
    CRB_CUDA(cudaMemcpyAsync(h_io + off_rows, d_io + off_rows, io_bytes - off_rows, cudaMemcpyDeviceToHost, l->stream));
    CRB_CUDA(cudaStreamSynchronize(l->stream));
    memcpy(h_rows, h_io + off_rows, r_bytes);
    memcpy(h_scores, h_io + off_scores, s_bytes);
    memcpy(h_n_valid, h_io + off_valid, v_bytes);
    if (h_vectors) {
        if (ix->n_rows > 0) memcpy(h_vectors, h_io + off_vec, e_bytes);
        else memset(h_vectors, 0, e_bytes);
    }
    return 0;
}
}  // namespace crb

extern "C" int crb_search_host(crb_index* ix, const float* h_queries, int32_t batch, int32_t k,
                               int32_t metric, const int32_t* h_filt, float* h_scores,
                               int64_t* h_rows, int32_t* h_n_valid, int32_t algo) {
    return search_host_impl(ix, h_queries, batch, k, metric, h_filt, h_scores, h_rows, h_n_valid, nullptr, algo);
}

extern "C" int crb_search_host_vectors(crb_index* ix, const float* h_queries, int32_t batch, int32_t k,
                                       int32_t metric, const int32_t* h_filt, float* h_scores,
                                       int64_t* h_rows, int32_t* h_n_valid, float* h_vectors, int32_t algo) {
    if (!h_vectors) return fail("NULL vector buffer passed to crb_search_host_vectors");
    return search_host_impl(ix, h_queries, batch, k, metric, h_filt, h_scores, h_rows, h_n_valid, h_vectors, algo);
}

extern "C" int crb_search_host_any_of(crb_index* ix, const float* h_query, int32_t n_alt, const int32_t* h_filt,
                                      int32_t k, int32_t metric, float* h_scores, int64_t* h_rows,
                                      int32_t* h_n_valid, float* h_vectors) {
    if (n_alt < 1) return fail("n_alt must be >= 1");
    if (!ix) return fail("index is NULL");
    if (!h_filt && ix->n_keys > 0) return fail("NULL filter pointer passed to crb_search_host_any_of");
    return search_host_impl(ix, h_query, 1, k, metric, h_filt, h_scores, h_rows, h_n_valid, h_vectors, 0, n_alt);
}

extern "C" int crb_merge_keys(const uint64_t* d_keys_in, int32_t n_lists, int32_t batch, int32_t k,
                              int32_t metric, const float* d_queries, int32_t dim, float* d_scores,
                              int64_t* d_rows, int32_t* d_n_valid, uint64_t* d_keys_out,
                              int32_t device, void* stream) {
    if (n_lists < 1 || n_lists > 1024) return fail("n_lists must be in [1, 1024]");
    if (batch < 1) return fail("batch must be >= 1");
    if (k < 1 || k > CRB_MAX_TOPK) return fail("k must be in [1, %d]", CRB_MAX_TOPK);
    if (!d_keys_in || !d_scores || !d_rows || !d_n_valid) return fail("NULL pointer passed to crb_merge_keys");
    if (metric == CRB_METRIC_COSINE && !d_queries) return fail("cosine merge needs the queries");
    DeviceGuard guard(device);
    CRB_CUDA(guard.status);
    merge_keys_kernel<<<batch, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        d_keys_in, n_lists, batch, k, metric, d_queries, dim, d_scores, d_rows, d_n_valid, d_keys_out);
    CRB_LAUNCH_CHECK();
    return 0;
}

namespace crb {
// crb_exchange -> the device-side descriptor of one call
static int make_exchange_dev(const crb_exchange* exch, int batch, int k, uint32_t call_index, uint32_t expected_total,
                             ExchangeDev* out) {
    if (!exch) return fail("exchange descriptor is NULL");
    if (exch->world < 1 || exch->world > CRB_MAX_PEERS || exch->rank < 0 || exch->rank >= exch->world)
        return fail("exchange: world must be in [1, %d] and rank in [0, world)", CRB_MAX_PEERS);
    if (uint64_t(batch) * uint64_t(k) > exch->cap_keys)
        return fail("exchange: batch*k = %lld exceeds the gather slot capacity %llu", (long long)batch * k,
                    (unsigned long long)exch->cap_keys);
    const uint32_t all = exch->world >= 32 ? 0xffffffffu : ((1u << exch->world) - 1u);
    ExchangeDev ex;
    memset(&ex, 0, sizeof ex);
    ex.world = exch->world;
    ex.rank = exch->rank;
    ex.parity = static_cast<int32_t>(call_index & 1u);
    ex.dest_mask = exch->dest_mask ? (exch->dest_mask & all) : all;
    if (!ex.dest_mask) return fail("exchange: dest_mask names no rank of the world");
    ex.cap = exch->cap_keys;
    ex.expected = expected_total;
    for (int p = 0; p < exch->world; ++p) {
        if (!exch->peer_keys[p] || !exch->peer_counters[p]) return fail("exchange: peer %d pointers are NULL", p);
        ex.keys[p] = exch->peer_keys[p];
        ex.counters[p] = exch->peer_counters[p];
    }
    *out = ex;
    return 0;
}
// the receiving half: wait for every source's arrivals and merge (only on ranks the keys are sent to)
static int launch_merge_wait(const crb_index* ix, const crb_exchange* exch, const ExchangeDev& ex, const float* d_queries,
                             int batch, int k, int metric, float* d_scores, int64_t* d_rows, int32_t* d_n_valid,
                             uint64_t* d_keys, uint32_t expected_total, cudaStream_t stream) {
    if (!((ex.dest_mask >> ex.rank) & 1u)) return 0;
    const uint64_t* my_gather = exch->peer_keys[exch->rank] + uint64_t(ex.parity) * uint64_t(ex.world) * ex.cap;
    merge_wait_kernel<<<batch, 256, 0, stream>>>(my_gather, ex.cap, exch->peer_counters[exch->rank], expected_total,
                                                 ex.world, batch, k, metric, d_queries, ix->dim, d_scores, d_rows,
                                                 d_n_valid, d_keys);
    CRB_LAUNCH_CHECK();
    return 0;
}
}  // namespace crb

extern "C" int crb_search_exchange(crb_index* ix, const float* d_queries, int32_t batch, int32_t k,
                                   int32_t metric, const int32_t* h_filt, float* d_scores, int64_t* d_rows,
                                   int32_t* d_n_valid, uint64_t* d_keys, void* d_workspace,
                                   size_t workspace_bytes, int32_t algo, void* stream_,
                                   const crb_exchange* exch, uint32_t call_index, uint32_t expected_total) {
    if (int rc = validate_search(ix, batch, k, metric)) return rc;
    if (!d_queries || !d_scores || !d_rows || !d_n_valid || !d_workspace)
        return fail("NULL device pointer passed to crb_search_exchange");
    ExchangeDev ex;
    if (int rc = make_exchange_dev(exch, batch, k, call_index, expected_total, &ex)) return rc;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    DeviceGuard guard(ix->device);
    CRB_CUDA(guard.status);
    bool folded = false;
    // (a caller that wants the merged KEYS back needs the separate merge kernel: the folded merge writes scores / rows only)
    if (int rc = search_impl(ix, d_queries, batch, k, metric, h_filt, d_scores, d_rows, d_n_valid, nullptr,
                             d_workspace, workspace_bytes, algo, stream, &ex, nullptr, d_keys ? nullptr : &folded))
        return rc;
    if (folded) return 0;
    return launch_merge_wait(ix, exch, ex, d_queries, batch, k, metric, d_scores, d_rows, d_n_valid, d_keys,
                             expected_total, stream);
}

namespace crb {
// n_alt > 0: ONE query (batch == 1) under a disjunction of n_alt filter rows (crb_search_exchange_host_any_of)
static int search_exchange_host_impl(crb_index* ix, const float* h_queries, int32_t batch, int32_t k,
                                     int32_t metric, const int32_t* h_filt,
                                     const uint32_t* const* h_row_mask_ptrs, float* h_scores, int64_t* h_rows,
                                     int32_t* h_n_valid, int32_t algo, const crb_exchange* exch,
                                     uint32_t call_index, uint32_t expected_total, int32_t n_alt) {
    if (int rc = validate_search(ix, batch, k, metric)) return rc;
    if (!h_queries) return fail("NULL query pointer passed to crb_search_exchange_host");
    ExchangeDev ex;
    if (int rc = make_exchange_dev(exch, batch, k, call_index, expected_total, &ex)) return rc;
    const bool receives = (ex.dest_mask >> ex.rank) & 1u;
    if (receives && (!h_scores || !h_rows || !h_n_valid))
        return fail("crb_search_exchange_host: this rank receives the merged result and needs output arrays");
    DeviceGuard guard(ix->device);
    CRB_CUDA(guard.status);
    LaneHold hold{ix, checkout_lane(ix)};
    Lane* l = hold.lane;

    const size_t q_bytes = size_t(batch) * ix->dim * sizeof(float);
    const size_t s_bytes = size_t(batch) * k * sizeof(float);
    const size_t r_bytes = size_t(batch) * k * sizeof(int64_t);
    const size_t v_bytes = size_t(batch) * sizeof(int32_t);
    const size_t off_rows = round_up(q_bytes, 256);
    const size_t off_scores = off_rows + round_up(r_bytes, 256);
    const size_t off_valid = off_scores + round_up(s_bytes, 256);
    const size_t io_bytes = off_valid + round_up(v_bytes, 256);
    int n_filters = n_alt > 0 ? 1 : count_distinct_filters(h_filt, batch, ix->n_keys);
    if (h_row_mask_ptrs) n_filters = std::min<int>(kMaxFilters, n_filters + batch);   // masks split filter classes further
    const size_t ws_bytes = crb_workspace_bytes(ix, batch, k, n_filters);
    if (int rc = lane_reserve(l, ws_bytes, io_bytes, io_bytes)) return rc;

    uint8_t* d_io = static_cast<uint8_t*>(l->d_io);
    uint8_t* h_io = static_cast<uint8_t*>(l->h_pin);
    if (q_bytes >= (size_t(64) << 10) && host_is_pinned(h_queries)) {
        // (the serving store registers its shared-memory command block: a 3 MB batch goes to the GPU without a second host copy)
        CRB_CUDA(cudaMemcpyAsync(d_io, h_queries, q_bytes, cudaMemcpyHostToDevice, l->stream));
    } else {
        memcpy(h_io, h_queries, q_bytes);
        CRB_CUDA(cudaMemcpyAsync(d_io, h_io, q_bytes, cudaMemcpyHostToDevice, l->stream));
    }
    const float* d_q = reinterpret_cast<const float*>(d_io);
    const bool zc = receives && ix->tune.zero_copy && (io_bytes - off_rows) <= kZeroCopyMaxBytes;
    uint8_t* o_io = zc ? h_io : d_io;          // where the kernels write the result
    float* d_scores = reinterpret_cast<float*>(o_io + off_scores);
    int64_t* d_rows = reinterpret_cast<int64_t*>(o_io + off_rows);
    int32_t* d_valid = reinterpret_cast<int32_t*>(o_io + off_valid);
    bool folded = false;
    int rc = n_alt > 0
        ? search_any_of_impl(ix, d_q, n_alt, h_filt, k, metric, d_scores, d_rows, d_valid, l->d_ws, l->ws_cap, l->stream,
                             &ex, &folded)
        : search_impl(ix, d_q, batch, k, metric, h_filt, d_scores, d_rows, d_valid, nullptr, l->d_ws, l->ws_cap,
                      algo, l->stream, &ex, h_row_mask_ptrs, &folded);
    if (!rc && !folded) rc = launch_merge_wait(ix, exch, ex, d_q, batch, k, metric, d_scores, d_rows, d_valid, nullptr,
                                               expected_total, l->stream);
    if (rc) { cudaStreamSynchronize(l->stream); return rc; }
    if (zc) {
        if (int rs = lane_signal(l)) { cudaStreamSynchronize(l->stream); return rs; }
        if (int rw = lane_wait_flag(l)) return rw;
    } else {
        if (receives)
            CRB_CUDA(cudaMemcpyAsync(h_io + off_rows, d_io + off_rows, io_bytes - off_rows, cudaMemcpyDeviceToHost, l->stream));
        // every rank waits for its own stream: the staging buffers are free again, and a sender knows its keys left
        CRB_CUDA(cudaStreamSynchronize(l->stream));
    }
    if (receives) {
        memcpy(h_rows, h_io + off_rows, r_bytes);
        memcpy(h_scores, h_io + off_scores, s_bytes);
        memcpy(h_n_valid, h_io + off_valid, v_bytes);
        for (int b = 0; b < batch; ++b)
            if (h_n_valid[b] == -2)
                return fail("sharded search: a peer did not publish query %d before the wait ran out (lost or stalled rank); "
                            "the exchange counters are out of step -- reset the exchange on every rank", b);
    }
    return 0;
}

}  // namespace crb

extern "C" int crb_search_exchange_host(crb_index* ix, const float* h_queries, int32_t batch, int32_t k,
                                        int32_t metric, const int32_t* h_filt,
                                        const uint32_t* const* h_row_mask_ptrs, float* h_scores, int64_t* h_rows,
                                        int32_t* h_n_valid, int32_t algo, const crb_exchange* exch,
                                        uint32_t call_index, uint32_t expected_total) {
    return crb::search_exchange_host_impl(ix, h_queries, batch, k, metric, h_filt, h_row_mask_ptrs, h_scores, h_rows,
                                          h_n_valid, algo, exch, call_index, expected_total, 0);
}

extern "C" int crb_search_exchange_host_any_of(crb_index* ix, const float* h_query, int32_t n_alt, const int32_t* h_filt,
                                               int32_t k, int32_t metric, float* h_scores, int64_t* h_rows,
                                               int32_t* h_n_valid, const crb_exchange* exch, uint32_t call_index,
                                               uint32_t expected_total) {
    if (n_alt < 1) return crb::fail("n_alt must be >= 1");
    if (!ix) return crb::fail("index is NULL");
    if (!h_filt && ix->n_keys > 0) return crb::fail("NULL filter pointer passed to crb_search_exchange_host_any_of");
    return crb::search_exchange_host_impl(ix, h_query, 1, k, metric, h_filt, nullptr, h_scores, h_rows, h_n_valid, 0, exch,
                                          call_index, expected_total, n_alt);
}

// ---------------------------------------------------------------------------------------------
// The serving ranks' request path in C: what ShardedB200VectorStore does per plain search, without Python between
// the shared-memory command word and the kernel launch.
// ---------------------------------------------------------------------------------------------
namespace crb {
constexpr int32_t kCmdSearch = 1;      // sharded_store.py::CMD_SEARCH
static inline uint64_t ctl_load_u64(const void* base, uint64_t off) {
    return __atomic_load_n(reinterpret_cast<const uint64_t*>(static_cast<const uint8_t*>(base) + off), __ATOMIC_ACQUIRE);
}
static inline void ctl_store_u64(void* base, uint64_t off, uint64_t v) {
    __atomic_store_n(reinterpret_cast<uint64_t*>(static_cast<uint8_t*>(base) + off), v, __ATOMIC_RELEASE);
}
static inline void cpu_relax() {
#if defined(__x86_64__)
    __builtin_ia32_pause();
#endif
}
static double now_s() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return double(ts.tv_sec) + 1e-9 * double(ts.tv_nsec);
}
}  // namespace crb

extern "C" int crb_serve_post_search(crb_index* ix, void* ctl_base, const crb_ctl* ctl, const float* h_queries,
                                     int32_t batch, int32_t k, int32_t metric, const int32_t* h_filt, float* h_scores,
                                     int64_t* h_rows, int32_t* h_n_valid, int32_t algo, const crb_exchange* exch,
                                     uint32_t* call_state, uint64_t* seq_state, double timeout_s) {
    if (!ix || !ctl_base || !ctl || !exch || !call_state || !seq_state) return fail("NULL argument passed to crb_serve_post_search");
    if (int rc = validate_search(ix, batch, k, metric)) return rc;
    uint8_t* base = static_cast<uint8_t*>(ctl_base);
    const uint64_t seq = *seq_state;
    // every rank has finished (and acknowledged) the previous command
    const double t0 = now_s();
    for (int r = 1; r < ctl->world; ++r) {
        uint64_t spins = 0;
        while (ctl_load_u64(base, ctl->ack_off + 8 * uint64_t(r)) != seq) {
            cpu_relax();
            if ((++spins & 0xfffff) == 0 && now_s() - t0 > timeout_s)
                return fail("sharded store: timed out waiting for rank %d to finish the previous command", r);
        }
        if (reinterpret_cast<const int32_t*>(base + ctl->status_off)[r] != 0) return 2;     // the host reads the rank's error reply
    }
    int32_t* hdr = reinterpret_cast<int32_t*>(base + ctl->hdr_off);
    hdr[0] = kCmdSearch; hdr[1] = batch; hdr[2] = k; hdr[3] = metric; hdr[4] = algo;
    hdr[5] = h_filt ? 1 : 0; hdr[6] = 0; hdr[7] = 0; hdr[8] = 0;
    float* q_ctl = reinterpret_cast<float*>(base + ctl->queries_off);
    memcpy(q_ctl, h_queries, size_t(batch) * ix->dim * sizeof(float));
    if (h_filt) memcpy(base + ctl->filt_off, h_filt, size_t(batch) * ix->n_keys * sizeof(int32_t));
    *seq_state = seq + 1;
    ctl_store_u64(base, ctl->seq_off, seq + 1);
    // this rank's share of the command (the queries are read from the page-locked block, like every rank does)
    call_state[1] += static_cast<uint32_t>(batch);
    const uint32_t call = call_state[0]++;
    return crb_search_exchange_host(ix, q_ctl, batch, k, metric, h_filt, nullptr, h_scores, h_rows, h_n_valid, algo, exch,
                                    call, call_state[1]);
}

extern "C" int crb_serve_next(crb_index* ix, void* ctl_base, const crb_ctl* ctl, const crb_exchange* exch,
                              uint32_t* call_state, uint64_t* last_seq) {
    if (!ix || !ctl_base || !ctl || !exch || !call_state || !last_seq) return fail("NULL argument passed to crb_serve_next"), -1;
    uint8_t* base = static_cast<uint8_t*>(ctl_base);
    for (;;) {
        // wait for the next command: spin (the hand-off is on the request's critical path), ease off when idle
        uint64_t spins = 0, seq;
        while ((seq = ctl_load_u64(base, ctl->seq_off)) == *last_seq) {
            cpu_relax();
            if (++spins > (1ull << 22)) { timespec ts{0, 200000}; nanosleep(&ts, nullptr); }
        }
        *last_seq = seq;
        const int32_t* hdr = reinterpret_cast<const int32_t*>(base + ctl->hdr_off);
        if (hdr[0] != kCmdSearch || hdr[6] != 0) return 0;       // not a plain search: the host side executes it
        const int32_t batch = hdr[1], k = hdr[2], metric = hdr[3], algo = hdr[4];
        const int32_t* filt = hdr[5] ? reinterpret_cast<const int32_t*>(base + ctl->filt_off) : nullptr;
        call_state[1] += static_cast<uint32_t>(batch);
        const uint32_t call = call_state[0]++;
        const int rc = crb_search_exchange_host(ix, reinterpret_cast<const float*>(base + ctl->queries_off), batch, k, metric,
                                                filt, nullptr, nullptr, nullptr, nullptr, algo, exch, call, call_state[1]);
        if (rc) return -1;                                        // the host reports crb_last_error() and acknowledges
        ctl_store_u64(base, ctl->ack_off + 8 * uint64_t(ctl->rank), seq);
    }
}

extern "C" int crb_host_register(void* h_ptr, size_t bytes) {
    if (!h_ptr || bytes == 0) return fail("crb_host_register: empty range");
    const cudaError_t e = cudaHostRegister(h_ptr, bytes, cudaHostRegisterDefault);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();       // not a sticky error: leave the runtime clean for the caller
        return fail("cudaHostRegister failed: %s", cudaGetErrorString(e));
    }
    return 0;
}

extern "C" int crb_host_unregister(void* h_ptr) {
    if (!h_ptr) return 0;
    const cudaError_t e = cudaHostUnregister(h_ptr);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        return fail("cudaHostUnregister failed: %s", cudaGetErrorString(e));
    }
    return 0;
}

extern "C" int crb_fetch_rows_host(crb_index* ix, const int64_t* h_rows_idx, int32_t n, float* h_out) {
    if (!ix) return fail("index is NULL");
    if (n <= 0) return 0;
    if (!h_rows_idx || !h_out) return fail("NULL pointer passed to crb_fetch_rows_host");
    for (int i = 0; i < n; ++i)
        if (h_rows_idx[i] < 0 || h_rows_idx[i] >= ix->n_rows)
            return fail("row index %lld out of range [0, %lld)", (long long)h_rows_idx[i], (long long)ix->n_rows);
    DeviceGuard guard(ix->device);
    CRB_CUDA(guard.status);
    LaneHold hold{ix, checkout_lane(ix)};
    Lane* l = hold.lane;
    const size_t idx_bytes = round_up(size_t(n) * sizeof(int64_t), 256);
    const size_t out_bytes = size_t(n) * ix->dim * sizeof(float);
    if (int rc = lane_reserve(l, 0, idx_bytes + out_bytes, idx_bytes + out_bytes)) return rc;
    uint8_t* d_io = static_cast<uint8_t*>(l->d_io);
    uint8_t* h_io = static_cast<uint8_t*>(l->h_pin);
    memcpy(h_io, h_rows_idx, size_t(n) * sizeof(int64_t));
    CRB_CUDA(cudaMemcpyAsync(d_io, h_io, size_t(n) * sizeof(int64_t), cudaMemcpyHostToDevice, l->stream));
    const unsigned blocks = static_cast<unsigned>(std::min<int64_t>((int64_t(n) * ix->dim + 255) / 256, 148 * 8));
    fetch_rows_kernel<<<blocks, 256, 0, l->stream>>>(ix->rows, reinterpret_cast<const int64_t*>(d_io), n,
                                                     ix->dim, reinterpret_cast<float*>(d_io + idx_bytes));
    CRB_LAUNCH_CHECK();
    CRB_CUDA(cudaMemcpyAsync(h_io + idx_bytes, d_io + idx_bytes, out_bytes, cudaMemcpyDeviceToHost, l->stream));
    CRB_CUDA(cudaStreamSynchronize(l->stream));
    memcpy(h_out, h_io + idx_bytes, out_bytes);
    return 0;
}

// =============================================================================================
// Allocation-free search (see the comment on crb_search_noalloc in include/coderag_b200.h)
// =============================================================================================
// The scratch itself, at file scope: part of the library image, present on every GPU that loads
// it.  In host code this name is only a stand-in for the symbol -- the address a kernel can use
// comes from cudaGetSymbolAddress on the device in question.
static __device__ __align__(256) unsigned char g_noalloc_scratch[CRB_NOALLOC_SCRATCH_BYTES];

namespace crb {

constexpr size_t kNoAllocHeaderBytes = 4096;               // arrival counters of the K1 groups
constexpr size_t kNoAllocCandBytes = size_t(1) << 20;      // per-CTA candidate lists
static_assert(kNoAllocHeaderBytes + kNoAllocCandBytes == CRB_NOALLOC_SCRATCH_BYTES, "header and source disagree");

constexpr int kNoAllocMaxDevices = 64;
static std::mutex g_noalloc_mu;
static unsigned char* g_noalloc_ptr[kNoAllocMaxDevices] = {};

// Device address of the scratch array on `device`, which must be the current device.  Looked up
// once per device; later calls return the cached pointer.
static int noalloc_scratch(int device, unsigned char** out) {
    if (device < 0 || device >= kNoAllocMaxDevices) return fail("device ordinal %d out of range", device);
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        unsigned char* ptr = nullptr;
        CRB_CUDA(cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch));
        if (!ptr) return fail("cudaGetSymbolAddress returned a null scratch address on device %d", device);
        g_noalloc_ptr[device] = ptr;
    }
    *out = g_noalloc_ptr[device];
    return 0;
}

}  // namespace crb

extern "C" int crb_search_noalloc(crb_index* ix, const float* d_queries, int32_t batch, int32_t k,
                                  int32_t metric, float* d_scores, int64_t* d_rows,
                                  int32_t* d_n_valid, void* stream_) {
    if (int rc = validate_search(ix, batch, k, metric)) return rc;
    if (batch > CRB_NOALLOC_MAX_BATCH)
        return fail("crb_search_noalloc: batch %d exceeds CRB_NOALLOC_MAX_BATCH (%d); its scratch is a fixed "
                    "%d-byte array -- use crb_search with a caller-provided workspace for larger batches",
                    batch, CRB_NOALLOC_MAX_BATCH, CRB_NOALLOC_SCRATCH_BYTES);
    if (!d_queries || !d_scores || !d_rows || !d_n_valid) return fail("NULL device pointer passed to crb_search_noalloc");
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    DeviceGuard guard(ix->device);
    CRB_CUDA(guard.status);
    unsigned char* scratch = nullptr;
    if (int rc = noalloc_scratch(ix->device, &scratch)) return rc;

    // plan: as crb_search plans its streaming groups -- consecutive queries pair up (one query per group for
    // dim > 1024), and the instantiation comes from the same picker (half-warp and exact-width forms included)
    const int ch = (ix->dim / 8 + 31) / 32;
    const int nq_cap = ch > 4 ? 1 : std::max(1, std::min(2, ix->tune.k1_group > 0 ? ix->tune.k1_group : 2));
    const int ng = (batch + nq_cap - 1) / nq_cap;
    const int nq = std::min<int>(batch, nq_cap);
    K1Choice c;
    if (int rc = k1_choose(ix, nq, k, &c)) return rc;
    const K1Fn fn = c.fn;
    const bool generic = c.generic;
    const size_t dyn = c.dyn_smem;
    const int rows_per_warp_iter = c.rows_per_warp_iter;
    const int resident = k1_occupancy(fn, dyn) * ix->sm_count;

    K1Params p;
    memset(&p, 0, sizeof p);
    p.rows = ix->rows;
    p.inv_norm = ix->inv_norm;
    p.queries = d_queries;
    p.cand = reinterpret_cast<uint64_t*>(scratch + kNoAllocHeaderBytes);
    p.counters = reinterpret_cast<unsigned*>(scratch);
    p.out_scores = d_scores;
    p.out_rows = d_rows;
    p.out_nvalid = d_n_valid;
    p.n_rows = ix->n_rows;
    p.list_stride = 0;
    p.row_offset = static_cast<uint32_t>(ix->row_offset);
    p.dim = ix->dim;
    p.k = k;
    p.metric = metric;
    for (int g = 0; g < ng; ++g) {
        for (int q = 0; q < kK1MaxQ; ++q) {
            const int b = g * nq_cap + q;
            p.groups[g].q_idx[q] = (q < nq_cap && b < batch) ? b : -1;
        }
        p.groups[g].filter = -1;
    }
    const int64_t work = (ix->n_rows + int64_t(kK1Warps) * rows_per_warp_iter * 2 - 1) /
                         (int64_t(kK1Warps) * rows_per_warp_iter * 2);
    int64_t gx = std::max<int64_t>(resident / ng, 16);
    gx = std::min<int64_t>(gx, std::max<int64_t>(work, 1));
    gx = std::min<int64_t>(gx, kK1MaxGridX);
    const int64_t cand_lists = int64_t(kNoAllocCandBytes) / (int64_t(k) * 8);
    gx = std::min<int64_t>(gx, std::max<int64_t>(cand_lists / (int64_t(ng) * (generic ? 1 : nq)), 1));
    if (int64_t(ng) * (generic ? 1 : nq) * gx * k * 8 > int64_t(kNoAllocCandBytes))
        return fail("crb_search_noalloc: batch %d with k %d does not fit the %d-byte scratch", batch, k,
                    CRB_NOALLOC_SCRATCH_BYTES);

    // The array still holds the previous call's lists and counters: the counters must start at 0.
    CRB_CUDA(cudaMemsetAsync(scratch, 0, kNoAllocHeaderBytes, stream));
    dim3 grid(static_cast<unsigned>(gx), static_cast<unsigned>(ng));
    fn<<<grid, kK1Threads, dyn, stream>>>(p);
    CRB_LAUNCH_CHECK();
    return 0;
}
