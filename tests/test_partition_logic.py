"""CPU check of the work-partition arithmetic the device kernels use (no GPU needed).

`k1_balanced_kernel` (csrc/k1_stream.cuh) lays the row sets of all groups end to end in CTA-step units and
gives CTA c the units [c*T/C, (c+1)*T/C).  Every CTA must agree, from closed-form integer expressions alone,
on which CTAs own a list of group g (first(g)..last(g)), and list index c + g must never collide.  The same
expressions are restated here and checked against a brute-force assignment.  `k2_gthr_due` (csrc/k2_gemm.cuh)
is the schedule on which a candidate list re-reads its query's shared bound."""
import numpy as np
from hypothesis import given, settings, strategies as st


def partition(units, C):
    """first/last CTA per group exactly as the kernel computes them."""
    upref = np.concatenate([[0], np.cumsum(units)]).astype(np.int64)
    T = int(upref[-1])
    out = []
    for g, (u0, u1) in enumerate(zip(upref[:-1], upref[1:])):
        if u1 > u0:
            first = ((int(u0) + 1) * C + T - 1) // T - 1
            last = (int(u1) * C + T - 1) // T - 1
        else:
            first = last = 0
        out.append((first, last))
    return upref, T, out


@settings(max_examples=300, deadline=None)
@given(st.lists(st.integers(min_value=0, max_value=5000), min_size=1, max_size=64), st.integers(min_value=1, max_value=700))
def test_balanced_partition_covers_every_unit_once_and_list_slots_do_not_collide(units, C):
    upref, T, fl = partition(units, C)
    seen_slots = set()
    for g, (first, last) in enumerate(fl):
        u0, u1 = int(upref[g]), int(upref[g + 1])
        if u1 == u0:
            slot = C + len(units) + g                      # empty groups: CTA 0, a list of their own
            assert slot not in seen_slots
            seen_slots.add(slot)
            continue
        assert 0 <= first <= last < C
        covered = 0
        for c in range(C):
            lo, hi = (c * T) // C, ((c + 1) * T) // C
            a, b = max(lo, u0), min(hi, u1)
            if b > a:
                assert first <= c <= last, "a CTA with work must own a list"
                covered += b - a
        assert covered == u1 - u0, "every unit of the group is served exactly once"
        for c in range(first, last + 1):
            slot = c + g
            assert slot not in seen_slots, "two (group, CTA) pairs share a list slot"
            assert slot < C + len(units)
            seen_slots.add(slot)


def gthr_due(visit, every=0):
    if every > 0:
        return visit % every == 0
    if visit < 8:
        return True
    hb = visit.bit_length() - 1
    return (visit & ((1 << (hb - 2)) - 1)) == 0


def test_bound_reread_schedule_is_dense_then_four_per_octave():
    due = [v for v in range(4096) if gthr_due(v)]
    assert due[:12] == [0, 1, 2, 3, 4, 5, 6, 7, 8, 10, 12, 14]
    for lo in (16, 64, 1024):
        assert sum(1 for v in due if lo <= v < 2 * lo) == 4
    assert len(due) == 8 + 4 * 9          # 8 dense visits, then four in each of the 9 octaves up to 4096


def bitonic_desc(a):
    """block_bitonic_desc (csrc/k2_gemm.cuh) restated: the compare-exchange schedule K5 sorts its candidates with."""
    a = list(a)
    n = len(a)
    size = 2
    while size <= n:
        stride = size >> 1
        while stride > 0:
            for t in range(n >> 1):
                i = 2 * t - (t & (stride - 1))
                j = i + stride
                desc = (i & size) == 0
                if (a[i] < a[j]) == desc:
                    a[i], a[j] = a[j], a[i]
            stride >>= 1
        size <<= 1
    return a


@settings(max_examples=60, deadline=None)
@given(st.sampled_from([32, 64, 128, 256, 512, 1024, 2048]), st.integers(min_value=0, max_value=2**32 - 1))
def test_k5_bitonic_schedule_sorts_descending(n, seed):
    rng = np.random.default_rng(seed)
    keys = rng.integers(0, 2**63 - 1, size=n, dtype=np.int64).tolist()
    keys[: n // 4] = [0] * (n // 4)                 # zero padding, as K5 pads its list to a power of two
    assert bitonic_desc(keys) == sorted(keys, reverse=True)


@settings(max_examples=300, deadline=None)
@given(st.integers(min_value=1, max_value=2048), st.integers(min_value=1, max_value=1200),
       st.integers(min_value=2048, max_value=120000))
def test_fallback_kernel_split_serves_every_flagged_query_once(nf, grid, cand_lists):
    """k1_fallback_kernel (csrc/k1_stream.cuh): `grid` CTAs split themselves over `nf` flagged queries --
    cpq CTAs per query, query f served by column f mod n_cols; every query gets exactly cpq lists and the
    lists of different queries never overlap in the scratch region."""
    cpq = min(grid // nf, cand_lists // nf, 1024)
    cpq = max(cpq, 1)
    n_cols = grid // cpq
    assert n_cols >= 1
    served = {}
    for cta in range(grid):
        col, x = cta // cpq, cta % cpq
        if col >= n_cols:
            continue
        for f in range(col, nf, n_cols):
            served.setdefault(f, []).append(x)
    assert sorted(served) == list(range(nf))
    for f, xs in served.items():
        assert sorted(xs) == list(range(cpq))        # arrival count the last CTA waits for
    assert nf * cpq <= cand_lists                    # scratch holds [flagged][cpq][k]; the library's region holds >= 8192 lists, a chunk <= 2048 queries
