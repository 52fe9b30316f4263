// plan.cpp -- host planner: guide alignments -> flat work plan (SURVEY 8a rows a2-a6, a16 schedule).
//
// For every alignment of a batch it derives, with integer arithmetic only:
//   anchors        convertPairwiseForwardStrandAlignmentToAnchorPairs (pairwiseAligner.c:789-813, trim 0) followed
//                  by cPecanRealign's matchFn filter (cPecanRealign.c:259-263, 567)
//   split regions  getSplitPoints (pairwiseAligner.c:1016-1067) and the per-region anchor sub-lists / ragged flags
//                  of getPosteriorProbsWithBandingSplittingAlignmentsByLargeGaps (:1091-1122)
//   bands          band_construct / band_setCurrentDiagonal (:94-179): one [xmyL, xmyR] per anti-diagonal
//   schedule       the checkpointed traceback points of getPosteriorProbsWithBanding (:666-744) and the diagonals
//                  on which the reference refreshes totalProbability (:712-720)
// and serialises them into one blob that the kernels index directly.  No floating point happens here except the
// single sqrt of getSplitPointsP (:1027).  Threads own contiguous ranges of alignments; their pieces are stitched
// with prefix sums, so the blob does not depend on the thread count.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <numeric>
#include <thread>

#include "mab_internal.h"

namespace {

inline uint8_t symbol_of(uint8_t c) {  // symbol_convertCharToSymbol, pairwiseAligner.c:262-279
    switch (c) {
        case 'A': case 'a': return 0;
        case 'C': case 'c': return 1;
        case 'G': case 'g': return 2;
        case 'T': case 't': return 3;
        default: return 4;
    }
}

inline uint8_t upper_of(uint8_t c) { return (c >= 'a' && c <= 'z') ? uint8_t(c - 32) : c; }

struct Piece {  // what one worker thread produces for its range of alignments (all indices piece-local)
    std::vector<RegionDesc> regions;
    std::vector<DiagDesc> diags;
    std::vector<TracebackDesc> tbs;
    std::vector<TotalDesc> totals;
    std::vector<int64_t> anchors, anchor_off;  // anchor_off has one entry per alignment (start), local
    std::vector<int64_t> aln_first_region;     // per alignment, local
    int64_t n_cells = 0, max_width = 0, n_edge = 0, n_full = 0;
    std::string error;
};

inline int64_t clamp_to(int64_t z, int64_t hi) { return z < 0 ? 0 : (z > hi ? hi : z); }

// band_setCurrentDiagonal (:104-122)
inline void band_diagonal(int64_t xay, int64_t xL, int64_t yL, int64_t xU, int64_t yU, int64_t &l, int64_t &r) {
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

struct Split { int64_t x1, y1, x2, y2; };

// getSplitPointsP (:1016-1038)
inline bool split_gap(int64_t &x1, int64_t &y1, int64_t x2, int64_t y2, int64_t x3, int64_t y3, std::vector<Split> &out,
                      int64_t limit, bool skip) {
    const int64_t gx = x3 - x2, gy = y3 - y2;
    if (gx * gy > limit) {
        const int64_t cap = (int64_t) std::sqrt((double) limit);
        const int64_t hx = gx / 2 > cap ? cap : gx / 2, hy = gy / 2 > cap ? cap : gy / 2;
        if (!skip) out.push_back({x1, y1, x2 + hx, y2 + hy});
        x1 = x3 - hx;
        y1 = y3 - hy;
        return true;
    }
    return false;
}

// Builds one region: band, cells, traceback schedule.  anchors are region-local.
void plan_region(Piece &pc, const mab_params &p, int32_t aln_local, const Split &sp, const int64_t *anc, int64_t n_anc,
                 bool ragged_left, bool ragged_right, int64_t sym_x_off, int64_t sym_y_off) {
    const int64_t lX = sp.x2 - sp.x1, lY = sp.y2 - sp.y1, D = lX + lY;
    if (D == 0) return;  // getPosteriorProbsWithBanding :650-652
    RegionDesc rg;
    std::memset(&rg, 0, sizeof(rg));
    rg.sym_x_off = sym_x_off + sp.x1;
    rg.sym_y_off = sym_y_off + sp.y1;
    rg.diag_base = (int64_t) pc.diags.size();
    rg.lX = (int32_t) lX;
    rg.lY = (int32_t) lY;
    rg.x1 = (int32_t) sp.x1;
    rg.y1 = (int32_t) sp.y1;
    rg.aln = aln_local;
    rg.flags = (ragged_left ? 1 : 0) | (ragged_right ? 2 : 0);
    rg.tb_base = (int32_t) pc.tbs.size();
    rg.tot_base = (int32_t) pc.totals.size();

    // band_construct (:128-179)
    const int64_t e = p.diagonal_expansion;
    int64_t ai = 0, xay = 0, pxay = 0, pxmy = 0, nxay = 0, nxmy = 0, xL = 0, yL = 0, xU = 0, yU = 0;
    const size_t d0 = pc.diags.size();
    pc.diags.resize(d0 + D + 1);
    int64_t cell = pc.n_cells;
    while (xay <= D) {
        int64_t l, r;
        band_diagonal(xay, xL, yL, xU, yU, l, r);
        DiagDesc &dd = pc.diags[d0 + xay];
        dd.cell_off = cell;
        dd.xmyL = (int32_t) l;
        dd.tot_flags = MAB_DIAG_TOT_MASK;  // "none" until the schedule below assigns one
        const int64_t w = (r - l) / 2 + 1;
        cell += w;
        if (w > pc.max_width) pc.max_width = w;
        if (nxay == xay++) {
            pxay = nxay;
            pxmy = nxmy;
            int64_t x = lX, y = lY;
            if (ai < n_anc) {
                x = anc[2 * ai] + 1;
                y = anc[2 * ai + 1] + 1;
                ai++;
            }
            nxay = x + y;
            nxmy = x - y;
            xL = clamp_to((pxay + (pxmy - e)) / 2, lX);
            yL = clamp_to((nxay - (nxmy - e)) / 2, lY);
            xU = clamp_to((nxay + (nxmy + e)) / 2, lX);
            yU = clamp_to((pxay - (pxmy + e)) / 2, lY);
        }
    }
    pc.n_cells = cell;

    // traceback schedule (:664-744)
    auto width = [&](int64_t d) {
        const int64_t next = d == D ? cell : pc.diags[d0 + d + 1].cell_off;
        return next - pc.diags[d0 + d].cell_off;
    };
    int64_t traced_to = 0;
    for (int64_t d = 1; d <= D; d++) {
        const bool at_end = d == D;
        const bool tb_point = d >= traced_to + p.min_diags_between_traceback && width(d) <= e * 2 + 1;
        if (!(at_end || tb_point)) continue;
        TracebackDesc tb;
        std::memset(&tb, 0, sizeof(tb));
        tb.region = -1;  // patched when pieces are merged
        tb.lo = (int32_t) traced_to;
        tb.hi = (int32_t) d;
        tb.from = (int32_t) (d - (at_end ? 0 : p.traceback_diagonals + 1));
        tb.at_end = at_end;
        tb.tot_base = (int32_t) pc.totals.size();
        tb.edge_off = -1;
        if (!at_end) {  // from + 1 <= hi always holds here
            tb.edge_off = (int32_t) pc.n_edge;
            pc.n_edge += width(tb.from + 1);
        }
        // posterior diagonals are from, from-1, ..., lo+1; the total is refreshed on every 10th of them (:712)
        int64_t k = 0;
        for (int64_t q = tb.from; q > tb.lo; q--, k++) {
            const bool is_total = k % 10 == 0;
            if (is_total) {
                TotalDesc td;
                td.region = -1;
                td.d = (int32_t) q;
                // backward diagonal q+1 is alive iff q < hi (:573-575); on q == from it is this traceback's own copy
                td.edge_off = q == tb.hi ? MAB_THROUGH_NONE : (q == tb.from ? tb.edge_off : MAB_THROUGH_MATRIX);
                td.full_off = (int32_t) pc.n_full;
                pc.n_full += width(q);
                pc.totals.push_back(td);
            }
            uint32_t tf = (uint32_t) (pc.totals.size() - 1);
            if (is_total) tf |= MAB_DIAG_IS_TOTAL;
            // the forward diagonal q-2 has been deleted by the previous traceback when q-2 < lo (:725-727, :737)
            if (q - 2 < tb.lo) tf |= MAB_DIAG_NO_MIDDLE;
            pc.diags[d0 + q].tot_flags = tf;
        }
        tb.n_tot = (int32_t) pc.totals.size() - tb.tot_base;
        pc.tbs.push_back(tb);
        traced_to = tb.from;
    }
    rg.n_tb = (int32_t) pc.tbs.size() - rg.tb_base;
    rg.n_tot = (int32_t) pc.totals.size() - rg.tot_base;
    pc.regions.push_back(rg);
}

struct BatchInput {
    int64_t n;
    const uint8_t *seq_x, *seq_y;
    const int64_t *off_x, *off_y;
    const uint8_t *ops;
    const int32_t *lens;
    const int64_t *off_ops;
    const int64_t *anchors, *off_anc;
};

void plan_range(Piece &pc, const mab_params &p, const BatchInput &in, int64_t a0, int64_t a1) {
    std::vector<int64_t> anc;
    std::vector<Split> splits;
    std::vector<int64_t> sub;
    char msg[256];
    for (int64_t a = a0; a < a1; a++) {
        const int64_t lX = in.off_x[a + 1] - in.off_x[a], lY = in.off_y[a + 1] - in.off_y[a];
        const uint8_t *sx = in.seq_x + in.off_x[a], *sy = in.seq_y + in.off_y[a];
        if (lX < 0 || lY < 0 || lX > INT32_MAX / 4 || lY > INT32_MAX / 4) {
            snprintf(msg, sizeof(msg), "alignment %lld: bad sequence lengths %lld %lld", (long long) a, (long long) lX,
                     (long long) lY);
            pc.error = msg;
            return;
        }
        anc.clear();
        if (in.ops != nullptr) {
            int64_t j = 0, k = 0;
            for (int64_t o = in.off_ops[a]; o < in.off_ops[a + 1]; o++) {
                const int op = in.ops[o];
                const int64_t len = in.lens[o];
                if (op > 2 || len < 0 || (op != 2 && j + len > lX) || (op != 1 && k + len > lY)) {
                    snprintf(msg, sizeof(msg), "alignment %lld: cigar op %d len %lld runs past the sequences", (long long) a,
                             op, (long long) len);
                    pc.error = msg;
                    return;
                }
                if (op == 0) {
                    for (int64_t l = 0; l < len; l++) {
                        const uint8_t cx = upper_of(sx[j + l]), cy = upper_of(sy[k + l]);
                        if (cx == cy && cx != 'N') {
                            anc.push_back(j + l);
                            anc.push_back(k + l);
                        }
                    }
                }
                if (op != 2) j += len;
                if (op != 1) k += len;
            }
            if (j != lX || k != lY) {  // checkPairwiseAlignment in the reference
                snprintf(msg, sizeof(msg), "alignment %lld: cigar spans %lld x %lld but the sequences are %lld x %lld",
                         (long long) a, (long long) j, (long long) k, (long long) lX, (long long) lY);
                pc.error = msg;
                return;
            }
        } else {
            int64_t px = -1, py = -1;
            for (int64_t o = in.off_anc[a]; o < in.off_anc[a + 1]; o++) {
                const int64_t x = in.anchors[2 * o], y = in.anchors[2 * o + 1];
                if (x <= px || y <= py || x >= lX || y >= lY) {  // band_construct's (compiled-out) asserts :159-164
                    snprintf(msg, sizeof(msg), "alignment %lld: anchor (%lld,%lld) out of order or out of range",
                             (long long) a, (long long) x, (long long) y);
                    pc.error = msg;
                    return;
                }
                anc.push_back(x);
                anc.push_back(y);
                px = x;
                py = y;
            }
        }
        const int64_t n_anc = (int64_t) anc.size() / 2;
        pc.anchor_off.push_back((int64_t) pc.anchors.size() / 2);
        pc.anchors.insert(pc.anchors.end(), anc.begin(), anc.end());
        pc.aln_first_region.push_back((int64_t) pc.regions.size());

        // getSplitPoints (:1040-1067)
        splits.clear();
        {
            int64_t x1 = 0, y1 = 0, x2 = 0, y2 = 0;
            for (int64_t i = 0; i < n_anc; i++) {
                const int64_t x3 = anc[2 * i], y3 = anc[2 * i + 1];
                split_gap(x1, y1, x2, y2, x3, y3, splits, p.split_matrix_bigger_than, p.ragged_left && i == 0);
                x2 = x3 + 1;
                y2 = y3 + 1;
            }
            if (!split_gap(x1, y1, x2, y2, lX, lY, splits, p.split_matrix_bigger_than, p.ragged_left && n_anc == 0) ||
                !p.ragged_right) {
                splits.push_back({x1, y1, lX, lY});
            }
        }
        // regions (:1091-1122)
        int64_t j = 0;
        for (size_t i = 0; i < splits.size(); i++) {
            const Split &sp = splits[i];
            sub.clear();
            while (j < n_anc && anc[2 * j] + anc[2 * j + 1] < sp.x2 + sp.y2) {
                sub.push_back(anc[2 * j] - sp.x1);
                sub.push_back(anc[2 * j + 1] - sp.y1);
                j++;
            }
            plan_region(pc, p, (int32_t) (a - a0), sp, sub.data(), (int64_t) sub.size() / 2, p.ragged_left || i > 0,
                        p.ragged_right || i + 1 < splits.size(), in.off_x[a], in.off_y[a]);
        }
    }
}

inline int64_t align16(int64_t v) { return (v + 15) & ~int64_t(15); }

int build_plan(const mab_params *p, const BatchInput &in, int n_threads, mab_plan **out) {
    if (p == nullptr || out == nullptr || in.n < 0 || in.off_x == nullptr || in.off_y == nullptr) {
        mab_set_error("mab_plan: null argument");
        return MAB_E_ARG;
    }
    if (p->diagonal_expansion < 0 || p->diagonal_expansion % 2 != 0 || p->traceback_diagonals < 1 ||
        p->min_diags_between_traceback < 2 || p->traceback_diagonals + 1 >= p->min_diags_between_traceback ||
        p->split_matrix_bigger_than < 0 || !(p->threshold >= 0.0 && p->threshold <= 1.0)) {
        // the prerequisites asserted at pairwiseAligner.c:643-647 and 588-589
        mab_set_error("mab_plan: invalid parameters (expansion must be even and >= 0, traceback_diagonals >= 1, "
                      "traceback_diagonals + 1 < min_diags_between_traceback, 0 <= threshold <= 1)");
        return MAB_E_ARG;
    }
    if (n_threads <= 0) n_threads = (int) std::thread::hardware_concurrency();
    if (n_threads <= 0) n_threads = 1;
    const int64_t per = 256;  // alignments per work item granularity
    n_threads = (int) std::max<int64_t>(1, std::min<int64_t>(n_threads, (in.n + per - 1) / per));
    std::vector<Piece> pieces(n_threads);
    std::vector<int64_t> bounds(n_threads + 1);
    {
        // balance by sequence length, which is proportional to the planning work
        const int64_t total = (in.n > 0) ? (in.off_x[in.n] - in.off_x[0]) + (in.off_y[in.n] - in.off_y[0]) : 0;
        bounds[0] = 0;
        int64_t a = 0;
        for (int t = 1; t < n_threads; t++) {
            const int64_t target = total * t / n_threads;
            while (a < in.n && (in.off_x[a] - in.off_x[0]) + (in.off_y[a] - in.off_y[0]) < target) a++;
            bounds[t] = a;
        }
        bounds[n_threads] = in.n;
    }
    if (n_threads == 1) {
        plan_range(pieces[0], *p, in, 0, in.n);
    } else {
        std::vector<std::thread> th;
        for (int t = 0; t < n_threads; t++)
            th.emplace_back([&, t]() { plan_range(pieces[t], *p, in, bounds[t], bounds[t + 1]); });
        for (auto &t : th) t.join();
    }
    for (auto &pc : pieces) {
        if (!pc.error.empty()) {
            mab_set_error("%s", pc.error.c_str());
            return MAB_E_INPUT;
        }
    }
    // stitch
    mab_plan *plan = new mab_plan();
    plan->params = *p;
    int64_t nR = 0, nD = 0, nT = 0, nTot = 0, nC = 0, nA = 0, maxW = 0, nE = 0, nF = 0;
    for (auto &pc : pieces) {
        nR += (int64_t) pc.regions.size();
        nD += (int64_t) pc.diags.size();
        nT += (int64_t) pc.tbs.size();
        nTot += (int64_t) pc.totals.size();
        nC += pc.n_cells;
        nE += pc.n_edge;
        nF += pc.n_full;
        nA += (int64_t) pc.anchors.size() / 2;
        maxW = std::max(maxW, pc.max_width);
    }
    if (nR > INT32_MAX || nT > INT32_MAX || nE > INT32_MAX || nF > INT32_MAX || nTot >= (int64_t) MAB_DIAG_TOT_MASK) {
        delete plan;
        mab_set_error("mab_plan: batch too large for 32-bit region/traceback indices; split it");
        return MAB_E_ARG;
    }
    plan->regions.reserve(nR);
    plan->diags.reserve(nD + 1);
    plan->tbs.reserve(nT);
    plan->totals.reserve(nTot);
    plan->anchors.reserve(2 * nA);
    plan->anchor_off.reserve(in.n + 1);
    plan->aln_region.reserve(in.n + 1);
    int64_t bR = 0, bD = 0, bT = 0, bTot = 0, bC = 0, bA = 0, bE = 0, bF = 0;
    for (int t = 0; t < n_threads; t++) {
        Piece &pc = pieces[t];
        for (auto v : pc.anchor_off) plan->anchor_off.push_back(v + bA);
        for (auto v : pc.aln_first_region) plan->aln_region.push_back(v + bR);
        plan->anchors.insert(plan->anchors.end(), pc.anchors.begin(), pc.anchors.end());
        for (size_t i = 0; i < pc.regions.size(); i++) {
            RegionDesc rg = pc.regions[i];
            // stamp the owning region into this region's tracebacks and totals before rebasing
            for (int32_t k = 0; k < rg.n_tb; k++) pc.tbs[rg.tb_base + k].region = (int32_t) (bR + (int64_t) i);
            for (int32_t k = 0; k < rg.n_tot; k++) pc.totals[rg.tot_base + k].region = (int32_t) (bR + (int64_t) i);
            rg.diag_base += bD;
            rg.aln += (int32_t) bounds[t];
            rg.tb_base += (int32_t) bT;
            rg.tot_base += (int32_t) bTot;
            plan->regions.push_back(rg);
        }
        for (auto dd : pc.diags) {
            dd.cell_off += bC;
            const uint32_t idx = dd.tot_flags & MAB_DIAG_TOT_MASK;
            if (idx != MAB_DIAG_TOT_MASK) dd.tot_flags = (dd.tot_flags & ~MAB_DIAG_TOT_MASK) | (uint32_t) (idx + bTot);
            plan->diags.push_back(dd);
        }
        for (auto tb : pc.tbs) {
            tb.tot_base += (int32_t) bTot;
            if (tb.edge_off >= 0) tb.edge_off += (int32_t) bE;
            plan->tbs.push_back(tb);
        }
        for (auto td : pc.totals) {
            if (td.edge_off >= 0) td.edge_off += (int32_t) bE;
            td.full_off += (int32_t) bF;
            plan->totals.push_back(td);
        }
        bR += (int64_t) pc.regions.size();
        bD += (int64_t) pc.diags.size();
        bT += (int64_t) pc.tbs.size();
        bTot += (int64_t) pc.totals.size();
        bC += pc.n_cells;
        bE += pc.n_edge;
        bF += pc.n_full;
        bA += (int64_t) pc.anchors.size() / 2;
        std::vector<DiagDesc>().swap(pc.diags);
    }
    plan->anchor_off.push_back(bA);
    plan->aln_region.push_back(bR);
    DiagDesc sentinel;
    sentinel.cell_off = nC;
    sentinel.xmyL = 0;
    sentinel.tot_flags = MAB_DIAG_TOT_MASK;
    plan->diags.push_back(sentinel);

    // symbols (a2): whole concatenated sequences, so regions index them by (alignment offset + origin)
    const int64_t lenX = in.n > 0 ? in.off_x[in.n] : 0, lenY = in.n > 0 ? in.off_y[in.n] : 0;
    plan->sym_x.resize(lenX);
    plan->sym_y.resize(lenY);
    for (int64_t i = 0; i < lenX; i++) plan->sym_x[i] = symbol_of(in.seq_x[i]);
    for (int64_t i = 0; i < lenY; i++) plan->sym_y[i] = symbol_of(in.seq_y[i]);

    // processing orders: longest first, so the dynamic work queue ends on short items
    plan->order_fwd.resize(nR);
    std::iota(plan->order_fwd.begin(), plan->order_fwd.end(), 0);
    std::stable_sort(plan->order_fwd.begin(), plan->order_fwd.end(), [&](int32_t a, int32_t b) {
        return plan->regions[a].lX + plan->regions[a].lY > plan->regions[b].lX + plan->regions[b].lY;
    });
    plan->order_bwd.resize(nT);
    std::iota(plan->order_bwd.begin(), plan->order_bwd.end(), 0);
    std::stable_sort(plan->order_bwd.begin(), plan->order_bwd.end(), [&](int32_t a, int32_t b) {
        return plan->tbs[a].hi - plan->tbs[a].lo > plan->tbs[b].hi - plan->tbs[b].lo;
    });

    BlobHeader &h = plan->hdr;
    std::memset(&h, 0, sizeof(h));
    h.magic = MAB_BLOB_MAGIC;
    h.n_alignments = in.n;
    h.n_regions = nR;
    h.n_diagonals = nD;
    h.n_cells = nC;
    h.n_tracebacks = nT;
    h.n_totals = nTot;
    h.n_anchors = nA;
    h.n_edge_cells = nE;
    h.n_full_cells = nF;
    h.max_width = maxW;
    h.len_sym_x = lenX;
    h.len_sym_y = lenY;
    int64_t off = sizeof(BlobHeader);
    auto place = [&](int64_t bytes) {
        const int64_t at = off;
        off = align16(off + bytes);
        return at;
    };
    h.off_regions = place(nR * (int64_t) sizeof(RegionDesc));
    h.off_diags = place((nD + 1) * (int64_t) sizeof(DiagDesc));
    h.off_tracebacks = place(nT * (int64_t) sizeof(TracebackDesc));
    h.off_totals = place(nTot * (int64_t) sizeof(TotalDesc));
    // one guard byte before and after each symbol array: the wavefront kernels read X[-1] / X[lX] (and ignore it)
    h.off_sym_x = place(lenX + 32) + 16;
    h.off_sym_y = place(lenY + 32) + 16;
    h.off_order_fwd = place(nR * 4);
    h.off_order_bwd = place(nT * 4);
    h.off_aln_region = place((in.n + 1) * 8);
    h.blob_bytes = off;
    *out = plan;
    return MAB_OK;
}

}  // namespace

extern "C" void mab_params_default(mab_params *p) {
    // pairwiseAlignmentBandingParameters_construct (:1144-1157) with the overrides every margin* call site applies
    p->threshold = 0.01;
    p->min_diags_between_traceback = 1000;
    p->traceback_diagonals = 40;
    p->diagonal_expansion = 10;
    p->split_matrix_bigger_than = (int64_t) 3000 * 3000;
    p->ragged_left = 1;
    p->ragged_right = 1;
}

extern "C" int mab_plan_from_cigars(const mab_params *p, int64_t n, const uint8_t *seq_x, const int64_t *off_x,
                                    const uint8_t *seq_y, const int64_t *off_y, const uint8_t *ops, const int32_t *lens,
                                    const int64_t *off_ops, int n_threads, mab_plan **out) {
    if (off_ops == nullptr || (n > 0 && off_ops[n] > off_ops[0] && (ops == nullptr || lens == nullptr))) {
        mab_set_error("mab_plan_from_cigars: null cigar arrays");
        return MAB_E_ARG;
    }
    static const uint8_t dummy_op = 0;
    BatchInput in{n, seq_x, seq_y, off_x, off_y, ops ? ops : &dummy_op, lens, off_ops, nullptr, nullptr};
    return build_plan(p, in, n_threads, out);
}

extern "C" int mab_plan_from_anchors(const mab_params *p, int64_t n, const uint8_t *seq_x, const int64_t *off_x,
                                     const uint8_t *seq_y, const int64_t *off_y, const int64_t *anchors,
                                     const int64_t *off_anc, int n_threads, mab_plan **out) {
    if (off_anc == nullptr) {
        mab_set_error("mab_plan_from_anchors: null anchor offsets");
        return MAB_E_ARG;
    }
    BatchInput in{n, seq_x, seq_y, off_x, off_y, nullptr, nullptr, nullptr, anchors, off_anc};
    return build_plan(p, in, n_threads, out);
}

extern "C" void mab_plan_destroy(mab_plan *plan) { delete plan; }

extern "C" int mab_plan_get_info(const mab_plan *plan, mab_plan_info *info) {
    if (plan == nullptr || info == nullptr) return MAB_E_ARG;
    const BlobHeader &h = plan->hdr;
    info->n_alignments = h.n_alignments;
    info->n_regions = h.n_regions;
    info->n_diagonals = h.n_diagonals;
    info->n_cells = h.n_cells;
    info->n_tracebacks = h.n_tracebacks;
    info->n_totals = h.n_totals;
    info->n_anchors = h.n_anchors;
    info->max_width = h.max_width;
    info->blob_bytes = h.blob_bytes;
    info->workspace_bytes_posteriors = mab_workspace_layout(h, MAB_MAX_S, false).total_bytes;
    info->workspace_bytes_expectations = mab_workspace_layout(h, MAB_MAX_S, true).total_bytes;
    return MAB_OK;
}

extern "C" int mab_plan_write_blob(const mab_plan *plan, void *dst, int64_t dst_bytes) {
    if (plan == nullptr || dst == nullptr) return MAB_E_ARG;
    const BlobHeader &h = plan->hdr;
    if (dst_bytes < h.blob_bytes) {
        mab_set_error("mab_plan_write_blob: need %lld bytes, got %lld", (long long) h.blob_bytes, (long long) dst_bytes);
        return MAB_E_CAPACITY;
    }
    uint8_t *b = (uint8_t *) dst;
    std::memcpy(b, &h, sizeof(h));
    std::memset(b + h.off_sym_x - 16, 4, 16);
    std::memset(b + h.off_sym_x + h.len_sym_x, 4, 16);
    std::memset(b + h.off_sym_y - 16, 4, 16);
    std::memset(b + h.off_sym_y + h.len_sym_y, 4, 16);
    auto put = [&](int64_t off, const void *src, size_t bytes) {
        if (bytes) std::memcpy(b + off, src, bytes);
    };
    put(h.off_regions, plan->regions.data(), plan->regions.size() * sizeof(RegionDesc));
    put(h.off_diags, plan->diags.data(), plan->diags.size() * sizeof(DiagDesc));
    put(h.off_tracebacks, plan->tbs.data(), plan->tbs.size() * sizeof(TracebackDesc));
    put(h.off_totals, plan->totals.data(), plan->totals.size() * sizeof(TotalDesc));
    put(h.off_sym_x, plan->sym_x.data(), plan->sym_x.size());
    put(h.off_sym_y, plan->sym_y.data(), plan->sym_y.size());
    put(h.off_order_fwd, plan->order_fwd.data(), plan->order_fwd.size() * 4);
    put(h.off_order_bwd, plan->order_bwd.data(), plan->order_bwd.size() * 4);
    put(h.off_aln_region, plan->aln_region.data(), plan->aln_region.size() * 8);
    return MAB_OK;
}

extern "C" int64_t mab_plan_anchors(const mab_plan *plan, int64_t aln, int64_t *out_xy, int64_t cap_pairs) {
    if (plan == nullptr || aln < 0 || aln >= plan->hdr.n_alignments) return MAB_E_ARG;
    const int64_t a0 = plan->anchor_off[aln], a1 = plan->anchor_off[aln + 1];
    for (int64_t i = a0; i < a1 && i - a0 < cap_pairs; i++) {
        out_xy[2 * (i - a0)] = plan->anchors[2 * i];
        out_xy[2 * (i - a0) + 1] = plan->anchors[2 * i + 1];
    }
    return a1 - a0;
}

extern "C" int64_t mab_plan_regions(const mab_plan *plan, int64_t *out5, int64_t cap_rows) {
    if (plan == nullptr) return MAB_E_ARG;
    const int64_t n = (int64_t) plan->regions.size();
    for (int64_t i = 0; i < n && i < cap_rows; i++) {
        const RegionDesc &r = plan->regions[i];
        out5[5 * i] = r.aln;
        out5[5 * i + 1] = r.x1;
        out5[5 * i + 2] = r.y1;
        out5[5 * i + 3] = r.x1 + r.lX;
        out5[5 * i + 4] = r.y1 + r.lY;
    }
    return n;
}

extern "C" int64_t mab_plan_band(const mab_plan *plan, int64_t region, int64_t *out2, int64_t cap_rows) {
    if (plan == nullptr || region < 0 || region >= (int64_t) plan->regions.size()) return MAB_E_ARG;
    const RegionDesc &r = plan->regions[region];
    const int64_t D = r.lX + r.lY;
    for (int64_t d = 0; d <= D && d < cap_rows; d++) {
        const DiagDesc &dd = plan->diags[r.diag_base + d];
        const int64_t w = plan->diags[r.diag_base + d + 1].cell_off - dd.cell_off;
        out2[2 * d] = dd.xmyL;
        out2[2 * d + 1] = dd.xmyL + 2 * (w - 1);
    }
    return D + 1;
}

extern "C" int64_t mab_plan_tracebacks(const mab_plan *plan, int64_t *out5, int64_t cap_rows) {
    if (plan == nullptr) return MAB_E_ARG;
    const int64_t n = (int64_t) plan->tbs.size();
    for (int64_t i = 0; i < n && i < cap_rows; i++) {
        const TracebackDesc &t = plan->tbs[i];
        out5[5 * i] = t.region;
        out5[5 * i + 1] = t.lo;
        out5[5 * i + 2] = t.hi;
        out5[5 * i + 3] = t.from;
        out5[5 * i + 4] = t.at_end;
    }
    return n;
}

WorkspaceLayout mab_workspace_layout(const BlobHeader &h, int S, bool expectations) {
    // Posterior mode keeps only what the posteriors read: the match plane of both matrices plus all states on the
    // diagonals that carry a total (1 in 10).  Expectation mode needs every state of every cell of both matrices.
    WorkspaceLayout w;
    int64_t off = 0;
    auto place = [&](int64_t bytes) {
        const int64_t at = off;
        off = (off + bytes + 255) & ~int64_t(255);
        return at;
    };
    w.off_counters = place(256);
    w.off_fwd = place(h.n_cells * 8 * (expectations ? S : 1));
    w.off_bwd = place(h.n_cells * 8 * (expectations ? S : 1));
    w.off_fwd_tot = place(expectations ? 0 : h.n_full_cells * S * 8);
    w.off_bwd_tot = place(expectations ? 0 : h.n_full_cells * S * 8);
    w.off_edge = place(h.n_edge_cells * 8);
    w.off_totals = place(h.n_totals * 8);
    w.off_pint = place(expectations ? 0 : h.n_cells * 4);
    w.off_counts = place(expectations ? 0 : (h.n_regions + 1) * 8);
    w.total_bytes = off;
    return w;
}
