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

#include "plan_core.h"

namespace {

using plan_core::AlnInput;
using plan_core::AnchorStream;
using plan_core::Split;
using plan_core::symbol_of;

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

// Builds one region: band, cells, traceback schedule (the walks live in plan_core.h).
void plan_region(Piece &pc, const mab_params &p, const AlnInput &in, int32_t aln_local, const Split &sp,
                 const AnchorStream &cur, bool ragged_left, bool ragged_right, int64_t sym_x_off, int64_t sym_y_off) {
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

    const size_t d0 = pc.diags.size();
    pc.diags.resize(d0 + D + 1);
    int64_t cell = pc.n_cells;
    plan_core::SpanTracker span;
    plan_core::band_walk(in, p, sp, cur, [&](int64_t xay, int64_t l, int64_t w) {
        DiagDesc &dd = pc.diags[d0 + xay];
        dd.cell_off = cell;
        dd.xmyL = (int32_t) l;
        dd.tot_flags = MAB_DIAG_TOT_MASK;  // "none" until the schedule below assigns one
        cell += w;
        if (w > pc.max_width) pc.max_width = w;
        span.add(l, w);
    });
    pc.n_cells = cell;
    rg.flags |= span.flag_bits();

    auto width = [&](int64_t d) {
        const int64_t next = d == D ? cell : pc.diags[d0 + d + 1].cell_off;
        return next - pc.diags[d0 + d].cell_off;
    };
    struct Sink {
        Piece &pc;
        size_t d0;
        int32_t tot_base_region;
        void total(int64_t q, int through, int64_t w) {
            TotalDesc td;
            td.region = -1;  // patched when pieces are merged
            td.d = (int32_t) q;
            td.edge_off = through == 0 ? MAB_THROUGH_NONE : (through == 1 ? (int32_t) pc.n_edge : MAB_THROUGH_MATRIX);
            td.full_off = (int32_t) pc.n_full;
            pc.n_full += w;
            pc.totals.push_back(td);
        }
        void diag_flags(int64_t q, int64_t tot_in_region, bool is_total, bool no_middle) {
            uint32_t tf = (uint32_t) (tot_base_region + tot_in_region);
            if (is_total) tf |= MAB_DIAG_IS_TOTAL;
            if (no_middle) tf |= MAB_DIAG_NO_MIDDLE;
            pc.diags[d0 + q].tot_flags = tf;
        }
        void traceback(int64_t lo, int64_t hi, int64_t from, bool at_end, int64_t n_tot, int64_t edge_cells) {
            TracebackDesc tb;
            std::memset(&tb, 0, sizeof(tb));
            tb.region = -1;
            tb.lo = (int32_t) lo;
            tb.hi = (int32_t) hi;
            tb.from = (int32_t) from;
            tb.at_end = at_end;
            tb.tot_base = (int32_t) ((int64_t) pc.totals.size() - n_tot);
            tb.n_tot = (int32_t) n_tot;
            tb.edge_off = at_end ? -1 : (int32_t) pc.n_edge;
            pc.n_edge += edge_cells;
            pc.tbs.push_back(tb);
        }
    } sink{pc, d0, rg.tot_base};
    plan_core::schedule_walk(p, D, width, sink);
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
    char msg[256];
    static const char *const kWhy[] = {"", "a cigar op runs past the sequences", "the cigar does not span the sequences",
                                       "an anchor is out of order or out of range", "bad sequence lengths"};
    for (int64_t a = a0; a < a1; a++) {
        AlnInput ai;
        ai.lX = in.off_x[a + 1] - in.off_x[a];
        ai.lY = in.off_y[a + 1] - in.off_y[a];
        ai.sx = in.seq_x + in.off_x[a];
        ai.sy = in.seq_y + in.off_y[a];
        if (in.ops != nullptr) {
            ai.ops = in.ops + in.off_ops[a];
            ai.lens = in.lens + in.off_ops[a];
            ai.n_ops = in.off_ops[a + 1] - in.off_ops[a];
            ai.anc = nullptr;
            ai.n_anc = 0;
        } else {
            ai.ops = nullptr;
            ai.lens = nullptr;
            ai.n_ops = 0;
            ai.anc = in.anchors + 2 * in.off_anc[a];
            ai.n_anc = in.off_anc[a + 1] - in.off_anc[a];
        }
        const int bad = plan_core::validate(ai);
        if (bad) {
            snprintf(msg, sizeof(msg), "alignment %lld (%lld x %lld): %s", (long long) a, (long long) ai.lX,
                     (long long) ai.lY, kWhy[bad]);
            pc.error = msg;
            return;
        }
        // anchors, kept for introspection
        pc.anchor_off.push_back((int64_t) pc.anchors.size() / 2);
        {
            AnchorStream st;
            int64_t x, y;
            while (st.next(ai, x, y)) {
                pc.anchors.push_back(x);
                pc.anchors.push_back(y);
            }
        }
        pc.aln_first_region.push_back((int64_t) pc.regions.size());
        plan_core::for_each_region(ai, p, [&](const Split &sp, const AnchorStream &cur, bool rl, bool rr) {
            plan_region(pc, p, ai, (int32_t) (a - a0), sp, cur, rl, rr, in.off_x[a], in.off_y[a]);
        });
    }
}

inline int64_t align16(int64_t v) { return (v + 15) & ~int64_t(15); }

int build_plan(const mab_params *p, const BatchInput &in, int n_threads, mab_plan **out) {
    if (p == nullptr || out == nullptr || in.n < 0 || in.off_x == nullptr || in.off_y == nullptr) {
        mab_set_error("mab_plan: null argument");
        return MAB_E_ARG;
    }
    {
        const int prc = mab_params_validate(p);
        if (prc) return prc;
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
    auto pack = [](const std::vector<uint8_t> &sym, std::vector<uint32_t> &out) {
        const int64_t len = (int64_t) sym.size(), cw = mab_pack_code_words(len);
        out.assign(mab_pack_words(len), 0);
        for (int64_t i = 0; i < len; i++) {
            const int64_t p = i + MAB_PACK_PAD_BASES;
            if (sym[i] < 4) out[p >> 4] |= (uint32_t) sym[i] << (2 * (p & 15));
            else out[cw + (p >> 5)] |= 1u << (p & 31);
        }
    };
    pack(plan->sym_x, plan->pack_x);
    pack(plan->sym_y, plan->pack_y);

    plan->diag_region.assign(nD + 1, -1);
    for (int64_t r = 0; r < nR; r++) {
        const RegionDesc &rg = plan->regions[r];
        for (int64_t d = 0; d <= (int64_t) rg.lX + rg.lY; d++) plan->diag_region[rg.diag_base + d] = (int32_t) r;
    }

    // processing orders: longest first, so the dynamic work queue ends on short items
    plan->order_fwd.resize(nR);
    std::iota(plan->order_fwd.begin(), plan->order_fwd.end(), 0);
    auto region_class = [&](int32_t r) { return mab_band_class(mab_region_span(plan->regions[r].flags)); };
    std::stable_sort(plan->order_fwd.begin(), plan->order_fwd.end(), [&](int32_t a, int32_t b) {
        const int ca = region_class(a), cb = region_class(b);
        if (ca != cb) return ca > cb;
        return plan->regions[a].lX + plan->regions[a].lY > plan->regions[b].lX + plan->regions[b].lY;
    });
    plan->order_bwd.resize(nT);
    std::iota(plan->order_bwd.begin(), plan->order_bwd.end(), 0);
    std::stable_sort(plan->order_bwd.begin(), plan->order_bwd.end(), [&](int32_t a, int32_t b) {
        const int ca = region_class(plan->tbs[a].region), cb = region_class(plan->tbs[b].region);
        if (ca != cb) return ca > cb;
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
    for (int64_t r = 0; r < nR; r++) h.n_class_regions[region_class((int32_t) r)]++;
    for (int64_t t = 0; t < nT; t++) h.n_class_tbs[region_class(plan->tbs[t].region)]++;
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
    h.off_pack_x = place(mab_pack_words(lenX) * 4);
    h.off_pack_y = place(mab_pack_words(lenY) * 4);
    h.off_order_fwd = place(nR * 4);
    h.off_order_bwd = place(nT * 4);
    h.off_aln_region = place((in.n + 1) * 8);
    h.off_diag_region = place((nD + 1) * 4);
    h.blob_bytes = off;
    *out = plan;
    return MAB_OK;
}

}  // namespace

int mab_params_validate(const mab_params *p) {
    if (p->diagonal_expansion < 0 || p->diagonal_expansion % 2 != 0 || p->traceback_diagonals < 1 ||
        p->min_diags_between_traceback < 2 || p->traceback_diagonals + 1 >= p->min_diags_between_traceback ||
        p->split_matrix_bigger_than < 0 || !(p->threshold >= 0.0 && p->threshold <= 1.0)) {
        // the prerequisites asserted at pairwiseAligner.c:643-647 and 588-589
        mab_set_error("mab_plan: invalid parameters (expansion must be even and >= 0, traceback_diagonals >= 1, "
                      "traceback_diagonals + 1 < min_diags_between_traceback, 0 <= threshold <= 1)");
        return MAB_E_ARG;
    }
    return MAB_OK;
}

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
    if (plan->d_blob != nullptr) {
        mab_set_error("mab_plan_write_blob: this plan was built on the device; its blob is mab_plan_device_blob()");
        return MAB_E_ARG;
    }
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
    put(h.off_pack_x, plan->pack_x.data(), plan->pack_x.size() * 4);
    put(h.off_pack_y, plan->pack_y.data(), plan->pack_y.size() * 4);
    put(h.off_order_fwd, plan->order_fwd.data(), plan->order_fwd.size() * 4);
    put(h.off_order_bwd, plan->order_bwd.data(), plan->order_bwd.size() * 4);
    put(h.off_aln_region, plan->aln_region.data(), plan->aln_region.size() * 8);
    put(h.off_diag_region, plan->diag_region.data(), plan->diag_region.size() * 4);
    return MAB_OK;
}

extern "C" int64_t mab_plan_anchors(const mab_plan *plan, int64_t aln, int64_t *out_xy, int64_t cap_pairs) {
    if (plan == nullptr || plan->d_blob != nullptr || aln < 0 || aln >= plan->hdr.n_alignments) return MAB_E_ARG;
    const int64_t a0 = plan->anchor_off[aln], a1 = plan->anchor_off[aln + 1];
    for (int64_t i = a0; i < a1 && i - a0 < cap_pairs; i++) {
        out_xy[2 * (i - a0)] = plan->anchors[2 * i];
        out_xy[2 * (i - a0) + 1] = plan->anchors[2 * i + 1];
    }
    return a1 - a0;
}

extern "C" int64_t mab_plan_regions(const mab_plan *plan, int64_t *out5, int64_t cap_rows) {
    if (plan == nullptr || plan->d_blob != nullptr) return MAB_E_ARG;
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
    if (plan == nullptr || plan->d_blob != nullptr || region < 0 || region >= (int64_t) plan->regions.size())
        return MAB_E_ARG;
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
    if (plan == nullptr || plan->d_blob != nullptr) return MAB_E_ARG;
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
    // per-diagonal pair counts in emission-slot order, scanned in place into output positions (+ sentinel)
    w.off_counts = place(expectations ? 0 : (h.n_diagonals + 2) * 8);
    w.scan_tmp_bytes = expectations ? 0 : h.n_diagonals / 32 + (1 << 20);
    w.off_scan_tmp = place(w.scan_tmp_bytes);
    w.total_bytes = off;
    return w;
}
