// Top-k primitives: per-warp running lists, block extraction, S-way merge of sorted lists.
#pragma once
#include "common.cuh"

namespace crb {

// Running top-k of one warp for one query.  Slot s (= lane + 32*i) is live when s < k.
// The list is unsorted; `thr` is the smallest live slot (0 while any live slot is empty), so
// "key > thr" is the warp-uniform admission test.
template <int KPL>
struct WarpTopK {
    uint64_t slot[KPL];
    uint64_t thr;

    CRB_DEVINL void init(int k, int lane) {
#pragma unroll
        for (int i = 0; i < KPL; ++i) slot[i] = (lane + 32 * i < k) ? 0ull : ~0ull;
        thr = 0ull;
    }
    // Warp-uniform call with key > thr: replace the smallest live slot.
    CRB_DEVINL void insert(uint64_t key, int lane) {
        int which = -1;
#pragma unroll
        for (int i = 0; i < KPL; ++i)
            if (slot[i] == thr) which = i;
        const unsigned holders = __ballot_sync(CRB_FULL_MASK, which >= 0);
        const int owner = __ffs(holders) - 1;
        if (lane == owner) {
#pragma unroll
            for (int i = 0; i < KPL; ++i)
                if (i == which) slot[i] = key;
        }
        uint64_t m = slot[0];
#pragma unroll
        for (int i = 1; i < KPL; ++i) m = slot[i] < m ? slot[i] : m;
        thr = warp_min_u64(m);
    }
    CRB_DEVINL void retire_dead_slots() {
#pragma unroll
        for (int i = 0; i < KPL; ++i)
            if (slot[i] == ~0ull) slot[i] = 0ull;
    }
    CRB_DEVINL uint64_t local_max() const {
        uint64_t m = slot[0];
#pragma unroll
        for (int i = 1; i < KPL; ++i) m = slot[i] > m ? slot[i] : m;
        return m;
    }
    CRB_DEVINL void remove(uint64_t key) {
#pragma unroll
        for (int i = 0; i < KPL; ++i)
            if (slot[i] == key) slot[i] = 0ull;
    }
};

// Block-wide maximum of one u64 per thread.  `xchg` is a [2][32] shared array, `round` picks
// the buffer so consecutive calls need only one barrier each.
CRB_DEVINL uint64_t block_max_u64(uint64_t v, uint64_t (*xchg)[32], int round, int lane, int warp,
                                  int nwarps) {
    const uint64_t wm = warp_max_u64(v);
    if (lane == 0) xchg[round & 1][warp] = wm;
    __syncthreads();
    const uint64_t o = (lane < nwarps) ? xchg[round & 1][lane] : 0ull;
    return warp_max_u64(o);
}

// Merge the per-warp lists of a block into k keys, descending, written by thread 0 to
// `out[0..k)` (zero padded).  Keys are unique (they embed the row) so each round's winner is
// held by exactly one thread.
template <int KPL>
CRB_DEVINL void block_extract_topk(WarpTopK<KPL>& w, int k, uint64_t* __restrict__ out,
                                   uint64_t (*xchg)[32], int lane, int warp, int nwarps) {
    w.retire_dead_slots();
    int r = 0;
    for (; r < k; ++r) {
        const uint64_t best = block_max_u64(w.local_max(), xchg, r, lane, warp, nwarps);
        if (best == 0ull) break;
        w.remove(best);
        if (threadIdx.x == 0) out[r] = best;
    }
    if (threadIdx.x == 0)
        for (; r < k; ++r) out[r] = 0ull;
    __syncthreads();
}

// Peer-memory exchange of a row-sharded search (crb_search_exchange): every finished query's k
// keys are stored straight into EVERY rank's gather buffer over NVLink, then a per-source arrival
// counter on each rank is bumped (system-scope release), so no collective launch sits between
// the shard searches and the cross-shard merge.
constexpr int kMaxPeers = 8;
struct ExchangeDev {
    int32_t world;          // 0 = no exchange
    int32_t rank;
    int32_t parity;         // gather buffers are double-buffered by call parity
    uint32_t dest_mask;     // bit p set: rank p receives the keys (serving: only the rank that owns the store face)
    uint64_t cap;           // keys per (parity, source) slot
    uint64_t* keys[kMaxPeers];      // peer p's gather buffer: [2][world][cap]
    uint32_t* counters[kMaxPeers];  // peer p's arrival counters: [world], monotonic
    uint32_t expected;      // arrivals per source once this call is complete (see crb_search_exchange)
    int32_t fold;           // 1: the streaming kernel's last CTA also waits for the peers and merges (no second launch)
};

// Wait (bounded) until every source rank has published `expected` queries into THIS rank's gather buffer.
// Called by a whole CTA; returns false when a source never arrived.
CRB_DEVINL bool exchange_wait_all(const ExchangeDev& ex) {
    int ok = 1;
    if (threadIdx.x < ex.world) {
        const volatile uint32_t* c = ex.counters[ex.rank] + threadIdx.x;
        long long spins = 0;
        while (static_cast<int32_t>(*c - ex.expected) < 0) {
            if (++spins > (1ll << 27)) { ok = 0; break; }
            __nanosleep(32);
        }
        __threadfence_system();
    }
    return __syncthreads_and(ok) != 0;
}

// Output of one finished query: scores (times inv_qnorm), global rows, n_valid, keys.
struct TopKSink {
    float* scores;      // [k]
    int64_t* rows;      // [k]
    int32_t* n_valid;   // [1]
    uint64_t* keys;     // [k] or nullptr
    float scale;        // 1/||q|| for cosine, 1 for inner product
    const ExchangeDev* ex = nullptr;   // publish to peers when set
    int32_t qi = 0;     // query index in the batch
    int32_t k = 0;
    CRB_DEVINL void emit(int r, uint64_t key) const {
        if (key != 0ull) {
            scores[r] = key_score(key) * scale;
            rows[r] = static_cast<int64_t>(key_row(key));
        } else {
            scores[r] = __uint_as_float(0xff800000u);  // -inf
            rows[r] = -1;
        }
        if (keys) keys[r] = key;
        if (ex) {
            const uint64_t off = (static_cast<uint64_t>(ex->parity) * ex->world + ex->rank) * ex->cap +
                                 static_cast<uint64_t>(qi) * k + r;
            for (int p = 0; p < ex->world; ++p)
                if ((ex->dest_mask >> p) & 1u) ex->keys[p][off] = key;
        }
    }
    // called once, by one thread, after every emit of the query is visible to it (block barrier, or the same
    // thread emitted): system-scope fence, then one arrival per destination
    CRB_DEVINL void finish() const {
        if (ex) {
            __threadfence_system();
            for (int p = 0; p < ex->world; ++p)
                if ((ex->dest_mask >> p) & 1u) atomicAdd_system(ex->counters[p] + ex->rank, 1u);
        }
    }
};

// S-way merge of `n_lists` descending lists of k keys (list l starts at lists + l*stride), in two stages of
// "best head wins" rounds that need no block barrier: (A) every warp merges its share of the lists -- warp w owns lists
// w, w + nwarps, ...; a lane holds the heads of up to LPT of them -- into one sorted list in shared memory, all warps at
// once; (B) warp 0 merges the nwarps lists.  (One block-wide stage cost a barrier per round: 6.8 us for 296 lists at
// k = 20, a tenth of a selective filtered search -- in-kernel timestamps, profiles/README.md.)  Heads are read with
// ld.cg (they were written by other CTAs of the same grid).
template <int LPT>
CRB_DEVINL void block_merge_sorted_lists(const uint64_t* __restrict__ lists, int n_lists, int k,
                                         int64_t stride, const TopKSink& sink,
                                         uint64_t (*xchg)[32], int lane, int warp, int nwarps) {
    __shared__ uint64_t s_win[kMaxTopK];
    __shared__ uint64_t s_wl[8][kMaxTopK];          // stage A results (nwarps <= 8)
    const int T = nwarps * 32;
    // ---- stage A ----
    uint64_t head[LPT], nxt[LPT];
    int pos[LPT];
#pragma unroll
    for (int j = 0; j < LPT; ++j) {
        const int l = warp + nwarps * (lane + 32 * j);
        pos[j] = 0;
        head[j] = 0ull;
        nxt[j] = 0ull;
        if (l < n_lists) {
            head[j] = ld_cg_u64(lists + l * stride);
            if (k > 1) nxt[j] = ld_cg_u64(lists + l * stride + 1);
        }
    }
    {
        int r = 0;
        for (; r < k; ++r) {
            uint64_t m = head[0];
#pragma unroll
            for (int j = 1; j < LPT; ++j) m = head[j] > m ? head[j] : m;
            const uint64_t best = warp_max_u64(m);
            if (best == 0ull) break;
#pragma unroll
            for (int j = 0; j < LPT; ++j) {
                if (head[j] == best) {
                    const int l = warp + nwarps * (lane + 32 * j);
                    head[j] = nxt[j];
                    pos[j] += 1;
                    nxt[j] = (pos[j] + 1 < k) ? ld_cg_u64(lists + l * stride + pos[j] + 1) : 0ull;
                }
            }
            if (lane == 0) s_wl[warp][r] = best;
        }
        if (lane == 0)
            for (; r < k; ++r) s_wl[warp][r] = 0ull;
    }
    __syncthreads();
    // ---- stage B ----
    if (warp == 0) {
        int p = 0;
        uint64_t h = lane < nwarps ? s_wl[lane][0] : 0ull;
        int r = 0;
        for (; r < k; ++r) {
            const uint64_t best = warp_max_u64(h);
            if (best == 0ull) break;
            if (h == best) {
                p += 1;
                h = p < k ? s_wl[lane][p] : 0ull;
            }
            if (lane == 0) s_win[r] = best;
        }
        if (lane == 0) xchg[0][0] = static_cast<uint64_t>(r);
    }
    __syncthreads();
    const int nv = static_cast<int>(xchg[0][0]);
    // every rank of the result is written (and, in a sharded search, stored to the peers) by its own thread
    for (int i = threadIdx.x; i < k; i += T) {
        sink.emit(i, i < nv ? s_win[i] : 0ull);
        if (sink.ex) __threadfence_system();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        *sink.n_valid = nv;
        sink.finish();
    }
    __syncthreads();
}

}  // namespace crb
