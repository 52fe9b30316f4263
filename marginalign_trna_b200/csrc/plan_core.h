// plan_core.h -- the integer logic of the planner, written once for host (plan.cpp, g++) and device
// (plan_device.cu, nvcc): anchors from a guide cigar, split regions, band geometry, checkpoint schedule.
//
// Everything here is a streaming restatement of the reference's list-building code:
//   AnchorStream     convertPairwiseForwardStrandAlignmentToAnchorPairs (pairwiseAligner.c:789-813, trim 0) +
//                    cPecanRealign's matchFn filter (cPecanRealign.c:259-263, 567), or explicit anchors
//   for_each_region  getSplitPoints / getSplitPointsP (pairwiseAligner.c:1016-1067) and the per-region ragged flags
//                    of getPosteriorProbsWithBandingSplittingAlignmentsByLargeGaps (:1121-1122)
//   band_walk        band_construct + band_setCurrentDiagonal (:94-179)
//   schedule_walk    the traceback points of getPosteriorProbsWithBanding (:664-744) and the every-10th-diagonal
//                    total-probability refresh (:712-720)
// The callers supply sinks (count or write), so the same walk serves the host's vectors and the device's
// count / scan / fill passes.
#pragma once
#include <cstdint>

#include "mab_internal.h"

#ifdef __CUDACC__
#define MAB_HD __host__ __device__ __forceinline__
#else
#define MAB_HD inline
#endif

namespace plan_core {

MAB_HD uint8_t symbol_of(uint8_t c) {  // symbol_convertCharToSymbol, pairwiseAligner.c:262-279
    switch (c) {
        case 'A': case 'a': return 0;
        case 'C': case 'c': return 1;
        case 'G': case 'g': return 2;
        case 'T': case 't': return 3;
        default: return 4;
    }
}

MAB_HD uint8_t upper_of(uint8_t c) { return (c >= 'a' && c <= 'z') ? uint8_t(c - 32) : c; }

// One alignment's inputs.  Either a cigar (ops / lens, n_ops entries) or explicit anchors (anc = x,y interleaved).
struct AlnInput {
    const uint8_t *sx, *sy;
    int64_t lX, lY;
    const uint8_t *ops;
    const int32_t *lens;
    int64_t n_ops;
    const int64_t *anc;
    int64_t n_anc;
};

// Checks what the reference's checkPairwiseAlignment / band_construct asserts would: returns 0 if fine, else a code
// (1 bad op / overrun, 2 cigar does not span the sequences, 3 anchors out of order or range, 4 bad lengths).
MAB_HD int validate(const AlnInput &in) {
    if (in.lX < 0 || in.lY < 0 || in.lX > INT32_MAX / 4 || in.lY > INT32_MAX / 4) return 4;
    if (in.ops != nullptr) {
        int64_t j = 0, k = 0;
        for (int64_t o = 0; o < in.n_ops; o++) {
            const int op = in.ops[o];
            const int64_t len = in.lens[o];
            if (op > 2 || len < 0 || (op != 2 && j + len > in.lX) || (op != 1 && k + len > in.lY)) return 1;
            if (op != 2) j += len;
            if (op != 1) k += len;
        }
        if (j != in.lX || k != in.lY) return 2;
    } else {
        int64_t px = -1, py = -1;
        for (int64_t o = 0; o < in.n_anc; o++) {
            const int64_t x = in.anc[2 * o], y = in.anc[2 * o + 1];
            if (x <= px || y <= py || x >= in.lX || y >= in.lY) return 3;
            px = x;
            py = y;
        }
    }
    return 0;
}

// Streams the anchors of one alignment in order.  Copyable: a copy is a cursor that can be resumed later.
struct AnchorStream {
    int64_t o = 0, l = 0, j = 0, k = 0;  // cigar mode: op index, offset inside the op, x and y at the op's start
    MAB_HD bool next(const AlnInput &in, int64_t &x, int64_t &y) {
        if (in.ops == nullptr) {
            if (o >= in.n_anc) return false;
            x = in.anc[2 * o];
            y = in.anc[2 * o + 1];
            o++;
            return true;
        }
        while (o < in.n_ops) {
            const int op = in.ops[o];
            const int64_t len = in.lens[o];
            if (op == 0) {
                while (l < len) {
                    const uint8_t cx = upper_of(in.sx[j + l]), cy = upper_of(in.sy[k + l]);
                    l++;
                    if (cx == cy && cx != 'N') {
                        x = j + l - 1;
                        y = k + l - 1;
                        return true;
                    }
                }
            }
            if (op != 2) j += len;
            if (op != 1) k += len;
            o++;
            l = 0;
        }
        return false;
    }
};

struct Split {
    int64_t x1, y1, x2, y2;
};

// integer floor(sqrt(v)) for v >= 0: (int64_t) sqrt((double) v) in the reference (pairwiseAligner.c:1027); the
// correction steps make it exact and identical on host and device whatever the libm
MAB_HD int64_t isqrt_floor(int64_t v) {
    int64_t r = (int64_t) sqrt((double) v);
    while (r * r > v) r--;
    while ((r + 1) * (r + 1) <= v) r++;
    return r;
}

// Region walker: step() consumes one anchor (or, at the end of the stream, closes the alignment) and calls
// sink(split, cursor_at_first_anchor, ragged_left, ragged_right) for every region that becomes final, in order.
// getSplitPoints keeps, on each side of a gap box bigger than `limit`, min(gap/2, floor(sqrt(limit))) rows/columns.
// A region is only known to be the last one when the stream ends, so the sink runs one region late.
// The walk is exposed one step at a time so that a GPU caller can keep the lanes of a warp in step with each other
// (`while (__any_sync(mask, more)) if (more) more = w.step(sink);`); for_each_region is the plain loop.
struct RegionWalker {
    const AlnInput &in;
    const mab_params &p;
    int64_t limit, cap;
    AnchorStream st, region_start, pend_cur;
    int64_t x1 = 0, y1 = 0, x2 = 0, y2 = 0, i = 0, n_emitted = 0;
    bool have_pending = false;
    Split pend{};

    MAB_HD RegionWalker(const AlnInput &in_, const mab_params &p_)
        : in(in_), p(p_), limit(p_.split_matrix_bigger_than), cap(isqrt_floor(p_.split_matrix_bigger_than)) {}

    template <typename Sink>
    MAB_HD void emit(const Split &sp, const AnchorStream &cur, Sink &&sink) {
        if (have_pending) {
            sink(pend, pend_cur, p.ragged_left || n_emitted > 0, true /* not the last region */);
            n_emitted++;
        }
        pend = sp;
        pend_cur = cur;
        have_pending = true;
    }

    // getSplitPointsP (:1016-1038): cut between (x2, y2) and the next anchor / corner (x3, y3) if the box is too big
    template <typename Sink>
    MAB_HD bool cut(int64_t x3, int64_t y3, const AnchorStream &next_start, Sink &&sink) {
        const int64_t gx = x3 - x2, gy = y3 - y2;
        if (gx * gy <= limit) return false;
        const int64_t hx = gx / 2 > cap ? cap : gx / 2, hy = gy / 2 > cap ? cap : gy / 2;
        if (!(p.ragged_left && i == 0)) emit(Split{x1, y1, x2 + hx, y2 + hy}, region_start, sink);
        x1 = x3 - hx;
        y1 = y3 - hy;
        region_start = next_start;
        return true;
    }

    template <typename Sink>
    MAB_HD bool step(Sink &&sink) {
        const AnchorStream before = st;
        int64_t x3, y3;
        if (st.next(in, x3, y3)) {
            cut(x3, y3, before, sink);
            x2 = x3 + 1;
            y2 = y3 + 1;
            i++;
            return true;
        }
        // end of the stream (:1056-1066): the gap to the corner, then the trailing region unless it is cut off by a
        // ragged right end; st is exhausted, so a trailing region that starts here has no anchors
        const bool was_cut = cut(in.lX, in.lY, st, sink);
        if (!was_cut || !p.ragged_right) emit(Split{x1, y1, in.lX, in.lY}, region_start, sink);
        if (have_pending) sink(pend, pend_cur, p.ragged_left || n_emitted > 0, p.ragged_right /* the last region */);
        return false;
    }
};

template <typename Sink>
MAB_HD void for_each_region(const AlnInput &in, const mab_params &p, Sink &&sink) {
    RegionWalker w(in, p);
    while (w.step(sink)) {
    }
}

MAB_HD int64_t clamp_to(int64_t z, int64_t hi) { return z < 0 ? 0 : (z > hi ? hi : z); }

// band_setCurrentDiagonal (:104-122)
MAB_HD void band_diagonal(int64_t xay, int64_t xL, int64_t yL, int64_t xU, int64_t yU, int64_t &l, int64_t &r) {
    l = xL - yL;
    r = xU - yU;
    if ((xay + l) % 2 != 0) l += 1;  // band_avoidOffByOne
    if ((xay + r) % 2 != 0) r += 1;
    int64_t v = (xay + l) / 2;
    if (v < xL) l += 2 * (xL - v);
    v = (xay - l) / 2;
    if (yL < v) l += 2 * (v - yL);
    v = (xay + r) / 2;
    if (xU < v) r -= 2 * (v - xU);
    v = (xay - r) / 2;
    if (v < yU) r -= 2 * (yU - v);
}

// band_construct (:128-179) for the region `sp`, anchors streamed from `cur` (absolute coordinates; the anchors of the
// region are those with x + y < sp.x2 + sp.y2).  Calls diag(xay, xmyL, width) for xay = 0 .. lX + lY.
//
// The reference walks diagonal by diagonal and steps to the next anchor whenever it reaches the current one.  The
// same sequence is produced here anchor by anchor -- diagonal 0 with the all-zero initial rectangle, then for each
// successive pair (previous anchor P, next anchor N; the last N is the corner (lX, lY)) the diagonals P.xay+1 .. N.xay
// with the rectangle of (P, N) -- because on a GPU the lanes of a warp then fetch their k-th anchor together instead
// of each lane branching into the anchor fetch at its own diagonal.
template <typename DiagSink>
MAB_HD void band_walk(const AlnInput &in, const mab_params &p, const Split &sp, AnchorStream cur, DiagSink &&diag) {
    const int64_t lX = sp.x2 - sp.x1, lY = sp.y2 - sp.y1, D = lX + lY, e = p.diagonal_expansion;
    const int64_t end_xay = sp.x2 + sp.y2;
    int64_t l, r;
    band_diagonal(0, 0, 0, 0, 0, l, r);
    diag(0, l, (r - l) / 2 + 1);
    int64_t xay = 1, pxay = 0, pxmy = 0;
    bool more = true;
    while (xay <= D) {
        int64_t x = lX, y = lY, ax, ay;
        if (more && cur.next(in, ax, ay) && ax + ay < end_xay) {
            x = ax - sp.x1 + 1;
            y = ay - sp.y1 + 1;
        } else {
            more = false;
        }
        const int64_t nxay = x + y, nxmy = x - y;
        const int64_t xL = clamp_to((pxay + (pxmy - e)) / 2, lX), yL = clamp_to((nxay - (nxmy - e)) / 2, lY);
        const int64_t xU = clamp_to((nxay + (nxmy + e)) / 2, lX), yU = clamp_to((pxay - (pxmy + e)) / 2, lY);
        for (; xay <= nxay; xay++) {
            band_diagonal(xay, xL, yL, xU, yU, l, r);
            diag(xay, l, (r - l) / 2 + 1);
        }
        pxay = nxay;
        pxmy = nxmy;
    }
}

// Band span of a region (mab_internal.h, mab_band_class): feed the diagonals in order.
struct SpanTracker {
    int64_t l1 = 0, r1 = 0, l2 = 0, r2 = 0, span = 0;
    int n = 0;
    MAB_HD void add(int64_t l, int64_t w) {
        const int64_t r = l + 2 * (w - 1);
        int64_t lo = l, hi = r;
        if (n >= 1) {
            if (l1 < lo) lo = l1;
            if (r1 > hi) hi = r1;
        }
        if (n >= 2) {
            if (l2 < lo) lo = l2;
            if (r2 > hi) hi = r2;
        }
        if (hi - lo > span) span = hi - lo;
        l2 = l1, r2 = r1, l1 = l, r1 = r;
        if (n < 2) n++;
    }
    MAB_HD int32_t flag_bits() const {
        return (int32_t) ((span > MAB_SPAN_MAX ? MAB_SPAN_MAX : span) << MAB_SPAN_SHIFT);
    }
};

// Traceback schedule of one region with D = lX + lY >= 1 diagonals after diagonal 0.  width(d) returns the number of
// cells of diagonal d.  Sink methods:
//   traceback(lo, hi, from, at_end, n_tot_of_this_tb, edge_cells)   called once per traceback, after its totals
//   total(q, through_kind, width_q)                                 through_kind: 0 none (q == hi), 1 edge buffer
//                                                                   (q == from < hi), 2 backward matrix
//   diag_flags(q, total_index_in_region, is_total, no_middle)       for q = from .. lo + 1
template <typename WidthFn, typename Sink>
MAB_HD void schedule_walk(const mab_params &p, int64_t D, WidthFn &&width, Sink &&sink) {
    const int64_t e = p.diagonal_expansion;
    int64_t traced_to = 0, n_tot_region = 0, d = 1;
    while (d <= D) {
        // scan to the next traceback point: the end, or >= min_diags past the last one on a narrow diagonal (:674-675).
        // (Written as scan-then-process so that the lanes of a warp leave the cheap scan together and run the
        // per-traceback loop below side by side.)
        if (d < traced_to + p.min_diags_between_traceback) d = traced_to + p.min_diags_between_traceback;
        if (d > D) d = D;
        while (d < D && width(d) > e * 2 + 1) d++;
        const bool at_end = d == D;
        const int64_t lo = traced_to, hi = d, from = d - (at_end ? 0 : p.traceback_diagonals + 1);
        const int64_t edge_cells = at_end ? 0 : width(from + 1);
        int64_t k = 0, n_tot = 0;
        for (int64_t q = from; q > lo; q--, k++) {
            const bool is_total = k % 10 == 0;
            if (is_total) {
                sink.total(q, q == hi ? 0 : (q == from ? 1 : 2), width(q));
                n_tot++;
                n_tot_region++;
            }
            sink.diag_flags(q, n_tot_region - 1, is_total, q - 2 < lo);
        }
        sink.traceback(lo, hi, from, at_end, n_tot, edge_cells);
        traced_to = from;
        d++;
    }
}

}  // namespace plan_core
