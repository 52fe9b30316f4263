/*
 * pairhmm_oracle.c -- TEST INFRASTRUCTURE ONLY: CPU restatement of the reference hot path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this; the
 * product library (libmargin_b200.so) never links or calls it.  Parity status: PINNED -- checked bit-for-bit
 * against outputs of the reference's own C code (oracle/_ref; fixtures in tests/golden/) and against the
 * known answers in the reference's unit tests (cPecan/tests/pairwiseAlignerTest.c).
 *
 * It is a restatement, not a copy: the DP lives in flat arrays indexed by a per-diagonal cell offset, the state
 * machines are data (ordered transition lists) rather than function-pointer objects, and the driver marks
 * diagonals alive/dead instead of malloc/free.  What is kept *exactly* is what the numbers depend on:
 *   - the approximate logAdd polynomial and its branch thresholds      cPecan/impl/pairwiseAligner.c:232-252
 *   - the order in which terms are logAdd-ed into a cell (scatter order for the backward pass)   :312-330, 539-554
 *   - the band geometry                                                :94-179
 *   - the checkpoint / traceback schedule and the stale total-probability refresh every 10th diagonal  :638-759
 *   - the split-by-large-gap rule                                      :1016-1136
 *   - the state machines' transition order, boundary vectors and emission tables  cPecan/impl/stateMachine.c:210-764
 * Build: gcc -std=c99 -O2 -ffp-contract=off (no FMA contraction, round-to-nearest doubles, like the reference build).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "pairhmm_oracle.h"

#define LOGZERO (-INFINITY)

/* ------------------------------------------------------------------ logAdd (pairwiseAligner.c:232-252) */

static inline double ph_lookup(double x) {
    /* float-suffixed literals in the reference: each coefficient is rounded to float, then promoted. */
    if (x <= 1.00f)
        return ((-0.009350833524763f * x + 0.130659527668286f) * x + 0.498799810682272f) * x + 0.693203116424741f;
    if (x <= 2.50f)
        return ((-0.014532321752540f * x + 0.139942324101744f) * x + 0.495635523139337f) * x + 0.692140569840976f;
    if (x <= 4.50f)
        return ((-0.004605031767994f * x + 0.063427417320019f) * x + 0.695956496475118f) * x + 0.514272634594009f;
    return ((-0.000458661602210f * x + 0.009695946122598f) * x + 0.930734667215156f) * x + 0.168037164329057f;
}

double ph_logadd(double x, double y) {
    if (x < y)
        return (x == LOGZERO || y - x >= 7.5) ? y : ph_lookup(y - x) + x;
    return (y == LOGZERO || x - y >= 7.5) ? x : ph_lookup(x - y) + y;
}

/* ------------------------------------------------------------------ symbols (pairwiseAligner.c:262-279) */

int ph_symbol(char c) {
    switch (c) {
    case 'A': case 'a': return 0;
    case 'C': case 'c': return 1;
    case 'G': case 'g': return 2;
    case 'T': case 't': return 3;
    default: return 4;
    }
}

/* ------------------------------------------------------------------ state machines (stateMachine.c) */

enum { M = 0, SX = 1, SY = 2, LX = 3, LY = 4 };

static void add_trans(ph_model *m, int group, int from, int to, double tp) {
    int k = m->n_trans[group]++;
    m->from[group][k] = from;
    m->to[group][k] = to;
    m->tp[group][k] = tp;
}

/* named log transitions; the 3-state machine uses the "short" slots only */
typedef struct {
    double match_continue;
    double match_from_short_x, match_from_long_x, short_open_x, short_extend_x, short_switch_to_x, long_open_x,
            long_extend_x, long_switch_to_x;
    double match_from_short_y, match_from_long_y, short_open_y, short_extend_y, short_switch_to_y, long_open_y,
            long_extend_y, long_switch_to_y;
} named_t;

static void finish_model(ph_model *m, const named_t *t, const double *emMatch, const double *emGapX,
        const double *emGapY) {
    memset(m->n_trans, 0, sizeof(m->n_trans));
    for (int s = 0; s < PH_MAX_S; s++) {
        m->start[s] = m->ragged_start[s] = m->end[s] = m->ragged_end[s] = 0.0;
    }
    if (m->S == 5) {
        /* stateMachine5_cellCalculate, stateMachine.c:399-429 (the switch transitions are commented out there) */
        add_trans(m, 0, M, SX, t->short_open_x);
        add_trans(m, 0, SX, SX, t->short_extend_x);
        add_trans(m, 0, M, LX, t->long_open_x);
        add_trans(m, 0, LX, LX, t->long_extend_x);
        add_trans(m, 1, M, M, t->match_continue);
        add_trans(m, 1, SX, M, t->match_from_short_x);
        add_trans(m, 1, SY, M, t->match_from_short_y);
        add_trans(m, 1, LX, M, t->match_from_long_x);
        add_trans(m, 1, LY, M, t->match_from_long_y);
        add_trans(m, 2, M, SY, t->short_open_y);
        add_trans(m, 2, SY, SY, t->short_extend_y);
        add_trans(m, 2, M, LY, t->long_open_y);
        add_trans(m, 2, LY, LY, t->long_extend_y);
        /* :350-397 */
        for (int s = 0; s < 5; s++) {
            m->start[s] = s == M ? 0 : LOGZERO;
            m->ragged_start[s] = (s == LX || s == LY) ? 0 : LOGZERO;
        }
        m->end[M] = t->match_continue;
        m->end[SX] = t->match_from_short_x;
        m->end[SY] = t->match_from_short_y;
        m->end[LX] = t->match_from_long_x;
        m->end[LY] = t->match_from_long_y;
        m->ragged_end[M] = t->long_open_x;
        m->ragged_end[SX] = t->long_open_x;
        m->ragged_end[SY] = t->long_open_y;
        m->ragged_end[LX] = t->long_extend_x;
        m->ragged_end[LY] = t->long_extend_y;
    } else {
        /* stateMachine3_cellCalculate, stateMachine.c:636-661 */
        add_trans(m, 0, M, SX, t->short_open_x);
        add_trans(m, 0, SX, SX, t->short_extend_x);
        add_trans(m, 0, SY, SX, t->short_switch_to_x);
        add_trans(m, 1, M, M, t->match_continue);
        add_trans(m, 1, SX, M, t->match_from_short_x);
        add_trans(m, 1, SY, M, t->match_from_short_y);
        add_trans(m, 2, M, SY, t->short_open_y);
        add_trans(m, 2, SY, SY, t->short_extend_y);
        add_trans(m, 2, SX, SY, t->short_switch_to_y);
        /* :595-634 */
        for (int s = 0; s < 3; s++) {
            m->start[s] = s == M ? 0 : LOGZERO;
            m->ragged_start[s] = (s == SX || s == SY) ? 0 : LOGZERO;
        }
        m->end[M] = t->match_continue;
        m->end[SX] = t->match_from_short_x;
        m->end[SY] = t->match_from_short_y;
        m->ragged_end[M] = (t->short_open_x + t->short_open_y) / 2.0;
        m->ragged_end[SX] = t->short_extend_x;
        m->ragged_end[SY] = t->short_extend_y;
    }
    /* emission_getGapProb / emission_getMatchProb, stateMachine.c:300-315: N uses rounded literals */
    for (int x = 0; x < 5; x++) {
        m->em_gap_x[x] = x == 4 ? -1.386294361 : emGapX[x];
        m->em_gap_y[x] = x == 4 ? -1.386294361 : emGapY[x];
        for (int y = 0; y < 5; y++) {
            m->em_match[x * 5 + y] = (x == 4 || y == 4) ? -2.772588722 : emMatch[x * 4 + y];
        }
    }
}

int ph_model_default(int type, ph_model *m) {
    /* stateMachine5_construct :431-468, stateMachine3_construct :663-690, emission defaults :218-241 */
    if (type < 0 || type > 3) return -1;
    m->type = type;
    m->S = type <= 1 ? 5 : 3;
    named_t t;
    memset(&t, 0, sizeof(t));
    t.match_continue = -0.030064059121770816;
    t.match_from_short_x = t.match_from_short_y = -1.272871422049609;
    t.short_extend_x = t.short_extend_y = -0.3388262689231553;
    t.short_switch_to_x = t.short_switch_to_y = -4.910694825551255;
    if (m->S == 5) {
        t.match_from_long_x = t.match_from_long_y = -5.673280173170473;
        t.short_open_x = t.short_open_y = -4.34381910900448;
        t.long_open_x = t.long_open_y = -6.30810595366929;
        t.long_extend_x = t.long_extend_y = -0.003442492794189331;
        t.long_switch_to_x = t.long_switch_to_y = -6.30810595366929;
    } else {
        t.short_open_x = t.short_open_y = -4.21256642;
    }
    const double em = -2.1149196655034745, tv = -4.5691014376830479, ti = -3.9833860032220842;
    const double emMatch[16] = { em, tv, ti, tv, tv, em, tv, ti, ti, tv, em, tv, tv, ti, tv, em };
    const double gap = -1.6094379124341003;
    const double emGap[4] = { gap, gap, gap, gap };
    finish_model(m, &t, emMatch, emGap, emGap);
    return 0;
}

static inline double hmmT(const double *T, int S, int from, int to) {
    return T[from * S + to];
}

static void swapd(double *a, double *b) {
    double c = *a;
    *a = *b;
    *b = c;
}

/* emissions_loadGapProbs, stateMachine.c:268-298 */
static void load_gap_probs(double *out, const double *E, const int *xStates, int nX, const int *yStates, int nY) {
    for (int i = 0; i < 4; i++) out[i] = 0.0;
    for (int i = 0; i < nX; i++)
        for (int x = 0; x < 4; x++)
            for (int y = 0; y < 4; y++) out[x] += E[xStates[i] * 16 + x * 4 + y];
    for (int i = 0; i < nY; i++)
        for (int x = 0; x < 4; x++)
            for (int y = 0; y < 4; y++) out[y] += E[yStates[i] * 16 + x * 4 + y];
    double total = 0.0;
    for (int i = 0; i < 4; i++) total += out[i];
    for (int i = 0; i < 4; i++) out[i] = log(out[i] / total);
}

int ph_model_from_hmm(int type, const double *T, const double *E, ph_model *m) {
    if (type < 0 || type > 3) return -1;
    m->type = type;
    const int S = m->S = type <= 1 ? 5 : 3;
    named_t t;
    memset(&t, 0, sizeof(t));
    double emMatch[16], emGapX[4], emGapY[4];
    const int symmetric = (type == 0 || type == 2);

    /* match emissions: emissions_loadMatchProbs[Symmetrically] :247-266 */
    for (int x = 0; x < 4; x++) {
        for (int y = 0; y < 4; y++) emMatch[x * 4 + y] = log(E[M * 16 + x * 4 + y]);
    }
    if (symmetric) {
        for (int x = 0; x < 4; x++) {
            for (int y = x + 1; y < 4; y++) {
                double d = log((E[M * 16 + x * 4 + y] + E[M * 16 + y * 4 + x]) / 2.0);
                emMatch[x * 4 + y] = d;
                emMatch[y * 4 + x] = d;
            }
        }
    }
    t.match_continue = log(hmmT(T, S, M, M));
    if (S == 5) {
        const int xs[2] = { SX, LX }, ys[2] = { SY, LY };
        if (!symmetric) { /* stateMachine5_loadAsymmetric :476-521 */
            t.match_from_short_x = log(hmmT(T, S, SX, M));
            t.match_from_long_x = log(hmmT(T, S, LX, M));
            t.short_open_x = log(hmmT(T, S, M, SX));
            t.short_extend_x = log(hmmT(T, S, SX, SX));
            t.short_switch_to_x = log(hmmT(T, S, SY, SX));
            t.long_open_x = log(hmmT(T, S, M, LX));
            t.long_extend_x = log(hmmT(T, S, LX, LX));
            t.long_switch_to_x = log(hmmT(T, S, LY, LX));
            if (t.short_extend_x > t.long_extend_x) {
                swapd(&t.short_extend_x, &t.long_extend_x);
                swapd(&t.match_from_short_x, &t.match_from_long_x);
                swapd(&t.short_open_x, &t.long_open_x);
                swapd(&t.short_switch_to_x, &t.long_switch_to_x);
            }
            t.match_from_short_y = log(hmmT(T, S, SY, M));
            t.match_from_long_y = log(hmmT(T, S, LY, M));
            t.short_open_y = log(hmmT(T, S, M, SY));
            t.short_extend_y = log(hmmT(T, S, SY, SY));
            t.short_switch_to_y = log(hmmT(T, S, SX, SY));
            t.long_open_y = log(hmmT(T, S, M, LY));
            t.long_extend_y = log(hmmT(T, S, LY, LY));
            t.long_switch_to_y = log(hmmT(T, S, LX, LY));
            if (t.short_extend_y > t.long_extend_y) {
                swapd(&t.short_extend_y, &t.long_extend_y);
                swapd(&t.match_from_short_y, &t.match_from_long_y);
                swapd(&t.short_open_y, &t.long_open_y);
                swapd(&t.short_switch_to_y, &t.long_switch_to_y);
            }
            load_gap_probs(emGapX, E, xs, 2, NULL, 0);
            load_gap_probs(emGapY, E, NULL, 0, ys, 2);
        } else { /* stateMachine5_loadSymmetric :523-567 */
            t.match_from_short_x = log((hmmT(T, S, SX, M) + hmmT(T, S, SY, M)) / 2);
            t.match_from_long_x = log((hmmT(T, S, LX, M) + hmmT(T, S, LY, M)) / 2);
            t.short_open_x = log((hmmT(T, S, M, SX) + hmmT(T, S, M, SY)) / 2);
            t.short_extend_x = log((hmmT(T, S, SX, SX) + hmmT(T, S, SY, SY)) / 2);
            t.short_switch_to_x = log((hmmT(T, S, SX, SY) + hmmT(T, S, SY, SX)) / 2);
            t.long_open_x = log((hmmT(T, S, M, LX) + hmmT(T, S, M, LY)) / 2);
            t.long_extend_x = log((hmmT(T, S, LX, LX) + hmmT(T, S, LY, LY)) / 2);
            t.long_switch_to_x = log((hmmT(T, S, LX, LY) + hmmT(T, S, LY, LX)) / 2);
            if (t.short_extend_x > t.long_extend_x) {
                swapd(&t.short_extend_x, &t.long_extend_x);
                swapd(&t.match_from_short_x, &t.match_from_long_x);
                swapd(&t.short_open_x, &t.long_open_x);
                swapd(&t.short_switch_to_x, &t.long_switch_to_x);
            }
            t.match_from_short_y = t.match_from_short_x;
            t.match_from_long_y = t.match_from_long_x;
            t.short_open_y = t.short_open_x;
            t.short_extend_y = t.short_extend_x;
            t.short_switch_to_y = t.short_switch_to_x;
            t.long_open_y = t.long_open_x;
            t.long_extend_y = t.long_extend_x;
            t.long_switch_to_y = t.long_switch_to_x;
            load_gap_probs(emGapX, E, xs, 2, ys, 2);
            load_gap_probs(emGapY, E, xs, 2, ys, 2);
        }
    } else {
        const int xs[1] = { SX }, ys[1] = { SY };
        if (!symmetric) { /* stateMachine3_loadAsymmetric :692-710 */
            t.match_from_short_x = log(hmmT(T, S, SX, M));
            t.match_from_short_y = log(hmmT(T, S, SY, M));
            t.short_open_x = log(hmmT(T, S, M, SX));
            t.short_open_y = log(hmmT(T, S, M, SY));
            t.short_extend_x = log(hmmT(T, S, SX, SX));
            t.short_extend_y = log(hmmT(T, S, SY, SY));
            t.short_switch_to_x = log(hmmT(T, S, SY, SX));
            t.short_switch_to_y = log(hmmT(T, S, SX, SY));
            load_gap_probs(emGapX, E, xs, 1, NULL, 0);
            load_gap_probs(emGapY, E, NULL, 0, ys, 1);
        } else { /* stateMachine3_loadSymmetric :712-734 */
            t.match_from_short_x = log((hmmT(T, S, SX, M) + hmmT(T, S, SY, M)) / 2.0);
            t.match_from_short_y = t.match_from_short_x;
            t.short_open_x = log((hmmT(T, S, M, SX) + hmmT(T, S, M, SY)) / 2.0);
            t.short_open_y = t.short_open_x;
            t.short_extend_x = log((hmmT(T, S, SX, SX) + hmmT(T, S, SY, SY)) / 2.0);
            t.short_extend_y = t.short_extend_x;
            t.short_switch_to_x = log((hmmT(T, S, SY, SX) + hmmT(T, S, SX, SY)) / 2.0);
            t.short_switch_to_y = t.short_switch_to_x;
            load_gap_probs(emGapX, E, xs, 1, ys, 1);
            load_gap_probs(emGapY, E, xs, 1, ys, 1);
        }
    }
    finish_model(m, &t, emMatch, emGapX, emGapY);
    return 0;
}

/* ------------------------------------------------------------------ band (pairwiseAligner.c:94-179) */

static int64_t avoid_off_by_one(int64_t xay, int64_t xmy) {
    return (xay + xmy) % 2 == 0 ? xmy : xmy + 1;
}

static int64_t clampz(int64_t z, int64_t lZ) {
    return z < 0 ? 0 : (z > lZ ? lZ : z);
}

static void band_diagonal(int64_t xay, int64_t xL, int64_t yL, int64_t xU, int64_t yU, int64_t *outL, int64_t *outR) {
    int64_t l = avoid_off_by_one(xay, xL - yL);
    int64_t r = avoid_off_by_one(xay, xU - yU);
    int64_t v;
    /* pull each end inside the rectangle [xL,xU] x [yU,yL] */
    v = (xay + l) / 2;
    if (v < xL) l += 2 * (xL - v);
    v = (xay - l) / 2;
    if (yL < v) l += 2 * (v - yL);
    v = (xay + r) / 2;
    if (xU < v) r -= 2 * (v - xU);
    v = (xay - r) / 2;
    if (v < yU) r -= 2 * (yU - v);
    *outL = l;
    *outR = r;
}

int64_t ph_band(const int64_t *anchors, int64_t n, int64_t lX, int64_t lY, int64_t e, int64_t *xmyL, int64_t *xmyR) {
    int64_t ai = 0, xay = 0, pxay = 0, pxmy = 0, nxay = 0, nxmy = 0, xL = 0, yL = 0, xU = 0, yU = 0;
    const int64_t D = lX + lY;
    while (xay <= D) {
        band_diagonal(xay, xL, yL, xU, yU, &xmyL[xay], &xmyR[xay]);
        if (nxay == xay++) {
            pxay = nxay;
            pxmy = nxmy;
            int64_t x = lX, y = lY;
            if (ai < n) {
                x = anchors[2 * ai] + 1;
                y = anchors[2 * ai + 1] + 1;
                ai++;
            }
            nxay = x + y;
            nxmy = x - y;
            xL = clampz((pxay + (pxmy - e)) / 2, lX);
            yL = clampz((nxay - (nxmy - e)) / 2, lY);
            xU = clampz((nxay + (nxmy + e)) / 2, lX);
            yU = clampz((pxay - (pxmy + e)) / 2, lY);
        }
    }
    return D + 1;
}

/* ------------------------------------------------------------------ split points (pairwiseAligner.c:1016-1067) */

typedef struct {
    int64_t *out, cap, n;
} splits_t;

static void push_split(splits_t *s, int64_t a, int64_t b, int64_t c, int64_t d) {
    if (s->n < s->cap) {
        s->out[4 * s->n] = a;
        s->out[4 * s->n + 1] = b;
        s->out[4 * s->n + 2] = c;
        s->out[4 * s->n + 3] = d;
    }
    s->n++;
}

static int split_gap(int64_t *x1, int64_t *y1, int64_t x2, int64_t y2, int64_t x3, int64_t y3, splits_t *s,
        int64_t split, int skip) {
    int64_t gX = x3 - x2, gY = y3 - y2;
    if (gX * gY > split) {
        int64_t maxLen = (int64_t) sqrt((double) split);
        int64_t hX = gX / 2 > maxLen ? maxLen : gX / 2;
        int64_t hY = gY / 2 > maxLen ? maxLen : gY / 2;
        if (!skip) push_split(s, *x1, *y1, x2 + hX, y2 + hY);
        *x1 = x3 - hX;
        *y1 = y3 - hY;
        return 1;
    }
    return 0;
}

int64_t ph_split_points(const int64_t *anchors, int64_t n, int64_t lX, int64_t lY, int64_t split, int raggedLeft,
        int raggedRight, int64_t *out, int64_t cap) {
    splits_t s = { out, cap, 0 };
    int64_t x1 = 0, y1 = 0, x2 = 0, y2 = 0;
    for (int64_t i = 0; i < n; i++) {
        int64_t x3 = anchors[2 * i], y3 = anchors[2 * i + 1];
        split_gap(&x1, &y1, x2, y2, x3, y3, &s, split, raggedLeft && i == 0);
        x2 = x3 + 1;
        y2 = y3 + 1;
    }
    if (!split_gap(&x1, &y1, x2, y2, lX, lY, &s, split, raggedLeft && n == 0) || !raggedRight) {
        push_split(&s, x1, y1, lX, lY);
    }
    return s.n;
}

/* ------------------------------------------------------------------ cigar -> anchors */

int64_t ph_anchors_from_cigar(const int32_t *ops, const int64_t *lens, int64_t nOps, const char *sX, const char *sY,
        int64_t *out, int64_t cap) {
    /* convertPairwiseForwardStrandAlignmentToAnchorPairs (pairwiseAligner.c:789-813, trim 0) followed by the
     * matchFn filter of cPecanRealign.c:259-263,567; coordinates are relative to the aligned sub-sequences. */
    int64_t j = 0, k = 0, n = 0;
    for (int64_t i = 0; i < nOps; i++) {
        if (ops[i] == 0) {
            for (int64_t l = 0; l < lens[i]; l++) {
                /* toupper(x)==toupper(y) && x != 'N': compares raw upper-cased chars (so 'R'=='R' anchors too) */
                char cx = sX[j + l], cy = sY[k + l];
                if (cx >= 'a' && cx <= 'z') cx -= 32;
                if (cy >= 'a' && cy <= 'z') cy -= 32;
                if (cx == cy && cx != 'N') {
                    if (n < cap) {
                        out[2 * n] = j + l;
                        out[2 * n + 1] = k + l;
                    }
                    n++;
                }
            }
        }
        if (ops[i] != 2) j += lens[i];
        if (ops[i] != 1) k += lens[i];
    }
    return n;
}

/* ------------------------------------------------------------------ the banded DP (pairwiseAligner.c:527-759) */

typedef struct {
    const ph_model *m;
    int S;
    int64_t lX, lY, D;
    const int *cx, *cy;     /* symbols of the region */
    int64_t *xmyL, *xmyR;   /* band */
    int64_t *off;           /* first cell of each diagonal */
    double *F, *B;          /* [cell][S] */
    char *falive, *balive;  /* mirrors dpMatrix create/delete */
} dp_t;

static inline double *cell_at(const dp_t *d, double *base, const char *alive, int64_t xay, int64_t xmy) {
    if (xay < 0 || xay > d->D || !alive[xay]) return NULL;
    if (xmy < d->xmyL[xay] || xmy > d->xmyR[xay]) return NULL;
    return base + (d->off[xay] + (xmy - d->xmyL[xay]) / 2) * d->S;
}

static inline int64_t width_of(const dp_t *d, int64_t xay) {
    return (d->xmyR[xay] - d->xmyL[xay]) / 2 + 1;
}

static void fill_diag(const dp_t *d, double *base, int64_t xay, const double *vals) {
    for (int64_t j = 0; j < width_of(d, xay); j++)
        for (int s = 0; s < d->S; s++) base[(d->off[xay] + j) * d->S + s] = vals ? vals[s] : LOGZERO;
}

static inline int sym_x(const dp_t *d, int64_t x) {
    return x > 0 ? d->cx[x - 1] : 4;
}
static inline int sym_y(const dp_t *d, int64_t y) {
    return y > 0 ? d->cy[y - 1] : 4;
}

static inline double emission(const ph_model *m, int group, int cX, int cY) {
    return group == 0 ? m->em_gap_x[cX] : (group == 1 ? m->em_match[cX * 5 + cY] : m->em_gap_y[cY]);
}

/* diagonalCalculationForward :556 -> cell_calculateForward :317 -> doTransitionForward :312 */
static void forward_diag(dp_t *d, int64_t xay) {
    const ph_model *m = d->m;
    for (int64_t xmy = d->xmyL[xay]; xmy <= d->xmyR[xay]; xmy += 2) {
        int64_t x = (xay + xmy) / 2, y = (xay - xmy) / 2;
        int cX = sym_x(d, x), cY = sym_y(d, y);
        double *cur = cell_at(d, d->F, d->falive, xay, xmy);
        double *nb[3] = { cell_at(d, d->F, d->falive, xay - 1, xmy - 1), cell_at(d, d->F, d->falive, xay - 2, xmy),
                cell_at(d, d->F, d->falive, xay - 1, xmy + 1) };
        for (int g = 0; g < 3; g++) {
            if (nb[g] == NULL) continue;
            double eP = emission(m, g, cX, cY);
            for (int k = 0; k < m->n_trans[g]; k++) {
                int to = m->to[g][k];
                cur[to] = ph_logadd(cur[to], nb[g][m->from[g][k]] + (eP + m->tp[g][k]));
            }
        }
    }
}

/* diagonalCalculationBackward :561 -> doTransitionBackward :322 (a scatter into diagonals xay-1 and xay-2) */
static void backward_diag(dp_t *d, int64_t xay) {
    const ph_model *m = d->m;
    for (int64_t xmy = d->xmyL[xay]; xmy <= d->xmyR[xay]; xmy += 2) {
        int64_t x = (xay + xmy) / 2, y = (xay - xmy) / 2;
        int cX = sym_x(d, x), cY = sym_y(d, y);
        double *cur = cell_at(d, d->B, d->balive, xay, xmy);
        double *nb[3] = { cell_at(d, d->B, d->balive, xay - 1, xmy - 1), cell_at(d, d->B, d->balive, xay - 2, xmy),
                cell_at(d, d->B, d->balive, xay - 1, xmy + 1) };
        for (int g = 0; g < 3; g++) {
            if (nb[g] == NULL) continue;
            double eP = emission(m, g, cX, cY);
            for (int k = 0; k < m->n_trans[g]; k++) {
                int from = m->from[g][k];
                nb[g][from] = ph_logadd(nb[g][from], cur[m->to[g][k]] + (eP + m->tp[g][k]));
            }
        }
    }
}

/* cell_dotProduct :332 / dpDiagonal_dotProduct :443 */
static double dot_cells(int S, const double *a, const double *b) {
    double t = a[0] + b[0];
    for (int s = 1; s < S; s++) t = ph_logadd(t, a[s] + b[s]);
    return t;
}

/* diagonalCalculationTotalProbability :566-583 */
static double total_probability(dp_t *d, int64_t xay) {
    const ph_model *m = d->m;
    const int S = d->S;
    double total = LOGZERO;
    for (int64_t j = 0; j < width_of(d, xay); j++) {
        total = ph_logadd(total, dot_cells(S, d->F + (d->off[xay] + j) * S, d->B + (d->off[xay] + j) * S));
    }
    if (xay + 1 <= d->D && d->balive[xay + 1] && xay - 1 >= 0 && d->falive[xay - 1]) {
        /* matches that jump over xay: forward xay-1 -> (match transitions) -> backward xay+1 */
        double through = LOGZERO;
        double tmp[PH_MAX_S];
        for (int64_t xmy = d->xmyL[xay + 1]; xmy <= d->xmyR[xay + 1]; xmy += 2) {
            int64_t x = (xay + 1 + xmy) / 2, y = (xay + 1 - xmy) / 2;
            for (int s = 0; s < S; s++) tmp[s] = LOGZERO;
            double *mid = cell_at(d, d->F, d->falive, xay - 1, xmy);
            if (mid != NULL) {
                double eP = emission(m, 1, sym_x(d, x), sym_y(d, y));
                for (int k = 0; k < m->n_trans[1]; k++) {
                    int to = m->to[1][k];
                    tmp[to] = ph_logadd(tmp[to], mid[m->from[1][k]] + (eP + m->tp[1][k]));
                }
            }
            through = ph_logadd(through, dot_cells(S, tmp, cell_at(d, d->B, d->balive, xay + 1, xmy)));
        }
        total = ph_logadd(total, through);
    }
    return total;
}

typedef struct {
    int mode;         /* 0 posteriors, 1 expectations */
    double threshold;
    int64_t *pairs, cap, n; /* region-local rows (p_int, x, y) in emission order */
    double *T, *E, *lik;
} sink_t;

/* diagonalCalculationPosteriorMatchProbs :585-615 */
static void posterior_diag(dp_t *d, int64_t xay, double total, sink_t *s) {
    const int S = d->S;
    for (int64_t xmy = d->xmyL[xay]; xmy <= d->xmyR[xay]; xmy += 2) {
        int64_t x = (xay + xmy) / 2, y = (xay - xmy) / 2;
        if (x > 0 && y > 0) {
            int64_t c = d->off[xay] + (xmy - d->xmyL[xay]) / 2;
            double p = exp((d->F[c * S + M] + d->B[c * S + M]) - total);
            if (p >= s->threshold) {
                if (p > 1.0) p = 1.0;
                p = floor(p * PH_PROB_1);
                if (s->n < s->cap) {
                    s->pairs[3 * s->n] = (int64_t) p;
                    s->pairs[3 * s->n + 1] = x - 1;
                    s->pairs[3 * s->n + 2] = y - 1;
                }
                s->n++;
            }
        }
    }
}

/* diagonalCalculationExpectations :617-628 -> updateExpectations :348-362 */
static void expectation_diag(dp_t *d, int64_t xay, double total, sink_t *s) {
    const ph_model *m = d->m;
    const int S = d->S;
    *s->lik += total;
    for (int64_t xmy = d->xmyL[xay]; xmy <= d->xmyR[xay]; xmy += 2) {
        int64_t x = (xay + xmy) / 2, y = (xay - xmy) / 2;
        int cX = sym_x(d, x), cY = sym_y(d, y);
        double *cur = cell_at(d, d->B, d->balive, xay, xmy);
        double *nb[3] = { cell_at(d, d->F, d->falive, xay - 1, xmy - 1), cell_at(d, d->F, d->falive, xay - 2, xmy),
                cell_at(d, d->F, d->falive, xay - 1, xmy + 1) };
        for (int g = 0; g < 3; g++) {
            if (nb[g] == NULL) continue;
            double eP = emission(m, g, cX, cY);
            for (int k = 0; k < m->n_trans[g]; k++) {
                int from = m->from[g][k], to = m->to[g][k];
                double p = exp(nb[g][from] + cur[to] + (eP + m->tp[g][k]) - total);
                s->T[from * S + to] += p;
                if (cX < 4 && cY < 4) s->E[to * 16 + cX * 4 + cY] += p;
            }
        }
    }
}

/* getPosteriorProbsWithBanding :638-759 */
static void banded_dp(const ph_model *m, const int *cx, int64_t lX, const int *cy, int64_t lY, const int64_t *anchors,
        int64_t nAnchors, const ph_params *p, int raggedLeft, int raggedRight, sink_t *sink) {
    const int64_t D = lX + lY;
    if (D == 0) return;
    dp_t d;
    d.m = m;
    d.S = m->S;
    d.lX = lX;
    d.lY = lY;
    d.D = D;
    d.cx = cx;
    d.cy = cy;
    d.xmyL = malloc(sizeof(int64_t) * (D + 1));
    d.xmyR = malloc(sizeof(int64_t) * (D + 1));
    d.off = malloc(sizeof(int64_t) * (D + 2));
    ph_band(anchors, nAnchors, lX, lY, p->diagonal_expansion, d.xmyL, d.xmyR);
    d.off[0] = 0;
    for (int64_t i = 0; i <= D; i++) d.off[i + 1] = d.off[i] + width_of(&d, i);
    d.F = malloc(sizeof(double) * d.S * d.off[D + 1]);
    d.B = malloc(sizeof(double) * d.S * d.off[D + 1]);
    d.falive = calloc(D + 1, 1);
    d.balive = calloc(D + 1, 1);

    d.falive[0] = 1;
    fill_diag(&d, d.F, 0, raggedLeft ? m->ragged_start : m->start);

    int64_t tracedBackTo = 0;
    for (int64_t xay = 1; xay <= D; xay++) {
        d.falive[xay] = 1;
        fill_diag(&d, d.F, xay, NULL);
        forward_diag(&d, xay);

        const int atEnd = xay == D;
        const int tracebackPoint = xay >= tracedBackTo + p->min_diags_between_traceback
                && width_of(&d, xay) <= p->diagonal_expansion * 2 + 1;
        if (!(atEnd || tracebackPoint)) continue;

        d.balive[xay] = 1;
        fill_diag(&d, d.B, xay, (atEnd && raggedRight) ? m->ragged_end : m->end);
        if (xay > tracedBackTo + 1) {
            d.balive[xay - 1] = 1;
            fill_diag(&d, d.B, xay - 1, NULL);
        }
        const int64_t tracedBackFrom = xay - (atEnd ? 0 : p->traceback_diagonals + 1);
        double total = LOGZERO;
        int64_t calcs = 0;
        int64_t d2 = xay;
        while (d2 > tracedBackTo) {
            if (d2 > tracedBackTo + 2) {
                d.balive[d2 - 2] = 1;
                fill_diag(&d, d.B, d2 - 2, NULL);
            }
            if (d2 > tracedBackTo + 1) backward_diag(&d, d2);
            if (d2 <= tracedBackFrom) {
                if (calcs++ % 10 == 0) total = total_probability(&d, d2);
                if (sink->mode == 0) posterior_diag(&d, d2, total, sink);
                else expectation_diag(&d, d2, total, sink);
                if (d2 < tracedBackFrom || atEnd) d.falive[d2] = 0;
            }
            if (d2 + 1 <= D) d.balive[d2 + 1] = 0;
            d2--;
        }
        tracedBackTo = tracedBackFrom;
        d.balive[d2 + 1] = 0;
        d.falive[d2] = 0;
    }
    free(d.xmyL);
    free(d.xmyR);
    free(d.off);
    free(d.F);
    free(d.B);
    free(d.falive);
    free(d.balive);
}

/* getPosteriorProbsWithBandingSplittingAlignmentsByLargeGaps :1083-1136 */
typedef struct {
    int64_t *out, cap, n;
} pairlist_t;

static void run_regions(const ph_model *m, const char *sX, int64_t lX, const char *sY, int64_t lY,
        const int64_t *anchors, int64_t n, const ph_params *p, int raggedLeft, int raggedRight, int mode,
        pairlist_t *pl, double *T, double *E, double *lik, int64_t *cellCount) {
    int64_t nSplit = ph_split_points(anchors, n, lX, lY, p->split_matrix_bigger_than, raggedLeft, raggedRight, NULL, 0);
    int64_t *sp = malloc(sizeof(int64_t) * 4 * (nSplit > 0 ? nSplit : 1));
    ph_split_points(anchors, n, lX, lY, p->split_matrix_bigger_than, raggedLeft, raggedRight, sp, nSplit);
    int64_t j = 0;
    for (int64_t i = 0; i < nSplit; i++) {
        const int64_t x1 = sp[4 * i], y1 = sp[4 * i + 1], x2 = sp[4 * i + 2], y2 = sp[4 * i + 3];
        const int64_t rX = x2 - x1, rY = y2 - y1;
        int *cx = malloc(sizeof(int) * (rX > 0 ? rX : 1)), *cy = malloc(sizeof(int) * (rY > 0 ? rY : 1));
        for (int64_t k = 0; k < rX; k++) cx[k] = ph_symbol(sX[x1 + k]);
        for (int64_t k = 0; k < rY; k++) cy[k] = ph_symbol(sY[y1 + k]);
        int64_t j0 = j;
        while (j < n && anchors[2 * j] + anchors[2 * j + 1] < x2 + y2) j++;
        int64_t nSub = j - j0;
        int64_t *sub = malloc(sizeof(int64_t) * 2 * (nSub > 0 ? nSub : 1));
        for (int64_t k = 0; k < nSub; k++) {
            sub[2 * k] = anchors[2 * (j0 + k)] - x1;
            sub[2 * k + 1] = anchors[2 * (j0 + k) + 1] - y1;
        }
        const int rl = raggedLeft || i > 0, rr = raggedRight || i < nSplit - 1;
        if (cellCount != NULL) {
            int64_t D = rX + rY;
            if (D > 0) {
                int64_t *l = malloc(sizeof(int64_t) * (D + 1)), *r = malloc(sizeof(int64_t) * (D + 1));
                ph_band(sub, nSub, rX, rY, p->diagonal_expansion, l, r);
                for (int64_t k = 0; k <= D; k++) *cellCount += (r[k] - l[k]) / 2 + 1;
                free(l);
                free(r);
            }
        } else {
            sink_t sink;
            sink.mode = mode;
            sink.threshold = p->threshold;
            sink.T = T;
            sink.E = E;
            sink.lik = lik;
            sink.n = 0;
            /* every in-band cell can emit at most one pair */
            int64_t maxPairs = 0;
            if (mode == 0) {
                int64_t D = rX + rY;
                int64_t *l = malloc(sizeof(int64_t) * (D + 1)), *r = malloc(sizeof(int64_t) * (D + 1));
                ph_band(sub, nSub, rX, rY, p->diagonal_expansion, l, r);
                for (int64_t k = 0; k <= D; k++) maxPairs += (r[k] - l[k]) / 2 + 1;
                free(l);
                free(r);
            }
            sink.cap = maxPairs;
            sink.pairs = mode == 0 ? malloc(sizeof(int64_t) * 3 * (maxPairs > 0 ? maxPairs : 1)) : NULL;
            banded_dp(m, cx, rX, cy, rY, sub, nSub, p, rl, rr, &sink);
            if (mode == 0) {
                /* alignedPairCoordinateCorrectionFn :1163-1170: shift, then pop from the tail (reverses the region) */
                for (int64_t k = sink.n - 1; k >= 0; k--) {
                    if (pl->n < pl->cap) {
                        pl->out[3 * pl->n] = sink.pairs[3 * k];
                        pl->out[3 * pl->n + 1] = sink.pairs[3 * k + 1] + x1;
                        pl->out[3 * pl->n + 2] = sink.pairs[3 * k + 2] + y1;
                    }
                    pl->n++;
                }
                free(sink.pairs);
            }
        }
        free(sub);
        free(cx);
        free(cy);
    }
    free(sp);
}

int64_t ph_aligned_pairs(const ph_model *m, const char *sX, int64_t lX, const char *sY, int64_t lY,
        const int64_t *anchors, int64_t n, const ph_params *p, int raggedLeft, int raggedRight, int64_t *out,
        int64_t cap) {
    pairlist_t pl = { out, cap, 0 };
    run_regions(m, sX, lX, sY, lY, anchors, n, p, raggedLeft, raggedRight, 0, &pl, NULL, NULL, NULL, NULL);
    return pl.n;
}

void ph_expectations(const ph_model *m, const char *sX, int64_t lX, const char *sY, int64_t lY,
        const int64_t *anchors, int64_t n, const ph_params *p, int raggedLeft, int raggedRight, double *T, double *E,
        double *lik) {
    run_regions(m, sX, lX, sY, lY, anchors, n, p, raggedLeft, raggedRight, 1, NULL, T, E, lik, NULL);
}

int64_t ph_count_cells(int64_t lX, int64_t lY, const int64_t *anchors, int64_t n, const ph_params *p, int raggedLeft,
        int raggedRight) {
    int64_t cells = 0;
    /* sequences are not needed to count cells; pass dummies of the right length */
    char *sX = malloc(lX + 1), *sY = malloc(lY + 1);
    memset(sX, 'A', lX);
    memset(sY, 'A', lY);
    sX[lX] = sY[lY] = 0;
    run_regions(NULL, sX, lX, sY, lY, anchors, n, p, raggedLeft, raggedRight, 0, NULL, NULL, NULL, NULL, &cells);
    free(sX);
    free(sY);
    return cells;
}

/* ------------------------------------------------------------------------------------------------------------------
 * Realign tail: what cPecanRealign does with the posterior pairs of one alignment
 * (cPecanRealign.c:590-631).  Restated over flat arrays; libc srand()/rand() supply the jitter exactly as the
 * reference's st_random() does (sonLib/C/impl/sonLibRandom.c:39-43).
 *
 *   1. AMAP reweight: getIndelProbabilities (pairwiseAligner.c:1216-1231), reweightAlignedPairs (:1233-1249),
 *      reweightAlignedPairs2 (:1251-1261).
 *   2. Maximum-expected-accuracy chain: filterPairwiseAlignmentToMakePairsOrdered (multipleAligner.c:941-967) ->
 *      convertAlignedPairsToMultipleAlignedPairs (:622-629, pops from the tail: the list is reversed) ->
 *      makeAlignmentWeightAdjacencyLists (:154-169, two st_random() draws per pair: the first for the weight filed
 *      under the Y column, the second for the weight filed under the X column, which is the one the DP reads) ->
 *      pairwiseAlignColumns (:355-489).
 *   3. Sort by (x, y) (cPecanRealign.c:611) and build the cigar: convertAlignedPairsToPairwiseAlignment (:50-92).
 * ------------------------------------------------------------------------------------------------------------------ */

typedef struct {
    int64_t y, pred;   /* pred: index into the candidate array, -1 = the (-1,-1) sentinel */
    int64_t x, src;    /* src: index of the input pair */
    double score;
} ph_cand;

static int cand_cmp_y(const void *a, const void *b) {
    int64_t i = ((const ph_cand *) a)->y, j = ((const ph_cand *) b)->y;
    return i > j ? 1 : (i < j ? -1 : 0);
}

/* pairs: n rows (p_int, x, y) in the order getAlignedPairsUsingAnchors returns them.  seed < 0: leave the libc
 * stream alone; else srand(seed) (a fresh cPecanRealign process has seed 1).  out_xy: chain pairs sorted, cap n rows.
 * out_w (may be NULL): reweighted integer weight of every INPUT pair.  Returns the chain length. */
int64_t ph_realign_chain(const int64_t *pairs, int64_t n, int64_t lX, int64_t lY, double gapGamma, double matchGamma,
        int seed, int64_t *out_xy, int64_t *out_w) {
    if (seed >= 0) srand((unsigned) seed);
    int64_t *w = malloc(sizeof(int64_t) * (n > 0 ? n : 1));
    for (int64_t i = 0; i < n; i++) w[i] = pairs[3 * i];
    if (gapGamma > 0.0) { /* reweightAlignedPairs2 :1252 */
        int64_t *ix = malloc(sizeof(int64_t) * (lX > 0 ? lX : 1)), *iy = malloc(sizeof(int64_t) * (lY > 0 ? lY : 1));
        for (int64_t i = 0; i < lX; i++) ix[i] = PH_PROB_1;
        for (int64_t i = 0; i < lY; i++) iy[i] = PH_PROB_1;
        for (int64_t i = 0; i < n; i++) {
            ix[pairs[3 * i + 1]] -= pairs[3 * i];
            iy[pairs[3 * i + 2]] -= pairs[3 * i];
        }
        for (int64_t i = 0; i < lX; i++) if (ix[i] < 0) ix[i] = 0;
        for (int64_t i = 0; i < lY; i++) if (iy[i] < 0) iy[i] = 0;
        for (int64_t i = 0; i < n; i++) { /* :1244: int64 - double * int64, truncated back to int64 */
            int64_t updated = pairs[3 * i] - gapGamma * (ix[pairs[3 * i + 1]] + iy[pairs[3 * i + 2]]);
            w[i] = updated;
        }
        free(ix);
        free(iy);
    }
    if (out_w != NULL) memcpy(out_w, w, sizeof(int64_t) * n);
    /* jitter: element k of the reversed list is input pair n-1-k; draws 2k (unused by the DP) and 2k+1 */
    double *aw = malloc(sizeof(double) * (n > 0 ? n : 1));
    for (int64_t k = 0; k < n; k++) {
        const double r0 = rand() / (RAND_MAX + 1.0);
        (void) r0;
        double r1 = rand() / (RAND_MAX + 1.0);
        r1 = r1 >= 1.0 ? 0.9999 : (r1 < 0.0 ? 0.0 : r1);
        aw[n - 1 - k] = ((double) w[n - 1 - k]) / PH_PROB_1 + r1 * 0.00001;
    }
    /* pairs by X column */
    int64_t *colStart = calloc(lX + 2, sizeof(int64_t)), *byCol = malloc(sizeof(int64_t) * (n > 0 ? n : 1));
    for (int64_t i = 0; i < n; i++) colStart[pairs[3 * i + 1] + 1]++;
    for (int64_t i = 0; i < lX; i++) colStart[i + 1] += colStart[i];
    int64_t *fill = malloc(sizeof(int64_t) * (lX + 1));
    memcpy(fill, colStart, sizeof(int64_t) * (lX + 1));
    for (int64_t i = 0; i < n; i++) byCol[fill[pairs[3 * i + 1]]++] = i;
    free(fill);
    /* the DP: `cands` holds every ColumnPair ever constructed, `front` the sorted set bestScoringAlignments */
    ph_cand *cands = malloc(sizeof(ph_cand) * (n + 2));
    int64_t nCands = 0;
    int64_t *front = malloc(sizeof(int64_t) * (n + 2)); /* candidate indices; -1 = min sentinel, -2 = max sentinel */
    int64_t nFront = 2;
    front[0] = -1;
    front[1] = -2;
#define F_Y(i) ((i) == -1 ? -1 : ((i) == -2 ? lY : cands[i].y))
#define F_SCORE(i) ((i) == -1 ? 0.0 : ((i) == -2 ? (double) INT64_MAX : cands[i].score))
    ph_cand *l = malloc(sizeof(ph_cand) * (lY + 1));
    for (int64_t x = 0; x < lX; x++) {
        int64_t nl = 0;
        for (int64_t k = colStart[x]; k < colStart[x + 1]; k++) {
            const int64_t i = byCol[k];
            if (aw[i] >= matchGamma && aw[i] > 0.0) {
                const int64_t y = pairs[3 * i + 2];
                int64_t f = 0; /* searchLessThan: last frontier element with yIndex < y */
                while (f + 1 < nFront && F_Y(front[f + 1]) < y) f++;
                l[nl].x = x;
                l[nl].y = y;
                l[nl].src = i;
                l[nl].pred = front[f];
                l[nl].score = F_SCORE(front[f]) + aw[i] * 1.0;
                nl++;
            }
        }
        qsort(l, nl, sizeof(ph_cand), cand_cmp_y);
        while (nl > 0) {
            const ph_cand cP = l[--nl];
            int64_t g = 0; /* searchGreaterThanOrEqual */
            while (F_Y(front[g]) < cP.y) g++;
            if (cP.score >= F_SCORE(front[g]) || F_Y(front[g]) > cP.y) {
                int64_t e = g;
                while (cP.score >= F_SCORE(front[e])) e++;
                cands[nCands] = cP;
                /* replace front[g..e) by the new candidate */
                memmove(front + g + 1, front + e, sizeof(int64_t) * (nFront - e));
                front[g] = nCands++;
                nFront += 1 - (e - g);
            }
        }
    }
    /* traceback from the right-most real point */
    int64_t m = 0;
    for (int64_t c = front[nFront - 2]; c >= 0; c = cands[c].pred) m++;
    int64_t k = m;
    for (int64_t c = front[nFront - 2]; c >= 0; c = cands[c].pred) {
        k--;
        out_xy[2 * k] = cands[c].x;
        out_xy[2 * k + 1] = cands[c].y;
    }
#undef F_Y
#undef F_SCORE
    free(l);
    free(front);
    free(cands);
    free(byCol);
    free(colStart);
    free(aw);
    free(w);
    return m;
}

/* convertAlignedPairsToPairwiseAlignment (cPecanRealign.c:50-92).  xy: m sorted chain pairs.  out_ops rows
 * (op, length) with sonLib op codes 0 match, 1 indel-X ('D'), 2 indel-Y ('I'); cap rows.  Returns the op count. */
int64_t ph_cigar_from_pairs(const int64_t *xy, int64_t m, int64_t lX, int64_t lY, int64_t *out_ops, int64_t cap) {
    int64_t pX = -1, pY = -1, mL = 0, nOps = 0;
#define PUSH(op, len) do { if (nOps < cap) { out_ops[2 * nOps] = (op); out_ops[2 * nOps + 1] = (len); } nOps++; } while (0)
    for (int64_t i = 0; i <= m; i++) {
        const int64_t x = i < m ? xy[2 * i] : lX, y = i < m ? xy[2 * i + 1] : lY;
        if (x - pX > 0 && y - pY > 0) {
            if (x - pX > 1) {
                if (mL > 0) { PUSH(0, mL); mL = 0; }
                PUSH(1, x - pX - 1);
            }
            if (y - pY > 1) {
                if (mL > 0) { PUSH(0, mL); mL = 0; }
                PUSH(2, y - pY - 1);
            }
            mL++;
            pX = x;
            pY = y;
        }
    }
    if (mL > 1) PUSH(0, mL - 1);
#undef PUSH
    return nOps;
}
