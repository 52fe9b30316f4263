/*
 * ref_shim.c -- TEST INFRASTRUCTURE ONLY.
 *
 * A thin ctypes-friendly wrapper (ours) around the UNMODIFIED reference C library (cPecan + sonLib, compiled
 * from /root/reference by oracle/Makefile into oracle/_ref/libcpecan_ref.so).  It marshals flat arrays into the
 * reference's stList/stIntTuple containers, calls the reference's public entry points, and copies the results
 * back out at full precision (the reference CLI's "%f" text loses digits).  No algorithm lives here.
 *
 * Entry points wrapped (reference file:line):
 *   getAlignedPairsUsingAnchors   cPecan/impl/pairwiseAligner.c:1172
 *   getExpectationsUsingAnchors   cPecan/impl/pairwiseAligner.c:1201
 *   band_construct / bandIterator cPecan/impl/pairwiseAligner.c:128,208
 *   getSplitPoints                cPecan/impl/pairwiseAligner.c:1040
 *   logAdd                        cPecan/impl/pairwiseAligner.c:248
 *   hmm_loadFromFile / hmm_getStateMachine   cPecan/impl/stateMachine.c:145,742
 *   reweightAlignedPairs2         cPecan/impl/pairwiseAligner.c:1251
 *   filterPairwiseAlignmentToMakePairsOrdered  cPecan/impl/multipleAligner.c:941
 */
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include "sonLib.h"
#include "pairwiseAligner.h"
#include "multipleAligner.h"

/* ---------------------------------------------------------------- models */

void *ref_model_create(int type, const double *T, const double *E) {
    Hmm *hmm = hmm_constructEmpty(0.0, (StateMachineType) type);
    int64_t S = hmm->stateNumber;
    memcpy(hmm->transitions, T, sizeof(double) * S * S);
    memcpy(hmm->emissions, E, sizeof(double) * S * 16);
    StateMachine *sM = hmm_getStateMachine(hmm);
    hmm_destruct(hmm);
    return sM;
}

void *ref_model_load(const char *file) {
    Hmm *hmm = hmm_loadFromFile(file);
    StateMachine *sM = hmm_getStateMachine(hmm);
    hmm_destruct(hmm);
    return sM;
}

/* The model cPecanRealign uses when no --loadHmm is given (cPecanRealign.c:526). */
void *ref_model_default5(void) {
    return stateMachine5_construct(fiveState);
}

void ref_model_destroy(void *sM) {
    stateMachine_destruct((StateMachine *) sM);
}

int64_t ref_model_state_number(void *sM) {
    return ((StateMachine *) sM)->stateNumber;
}

int ref_model_type(void *sM) {
    return (int) ((StateMachine *) sM)->type;
}

/* which: 0 start, 1 raggedStart, 2 end, 3 raggedEnd */
void ref_model_boundary(void *sM_, int which, double *out) {
    StateMachine *sM = sM_;
    for (int64_t s = 0; s < sM->stateNumber; s++) {
        out[s] = which == 0 ? sM->startStateProb(sM, s) : which == 1 ? sM->raggedStartStateProb(sM, s) :
                 which == 2 ? sM->endStateProb(sM, s) : sM->raggedEndStateProb(sM, s);
    }
}

typedef struct {
    int64_t n, cap;
    double *out; /* rows of (group, from, to, eP, tP) */
    int group;
    double *lower, *middle, *upper;
} TransitionRecorder;

static void recordTransition(double *fromCells, double *toCells, int64_t from, int64_t to, double eP, double tP,
        void *extra) {
    TransitionRecorder *r = extra;
    (void) toCells;
    if (r->n < r->cap) {
        double *row = r->out + 5 * r->n;
        row[0] = fromCells == r->lower ? 0 : (fromCells == r->middle ? 1 : 2);
        row[1] = (double) from;
        row[2] = (double) to;
        row[3] = eP;
        row[4] = tP;
    }
    r->n++;
}

/* Lists the ordered transitions of cellCalculate for symbols (cX, cY): rows (group, from, to, eP, tP). */
int64_t ref_model_transitions(void *sM_, int cX, int cY, double *out, int64_t cap) {
    StateMachine *sM = sM_;
    double cur[8], lo[8], mi[8], up[8];
    TransitionRecorder r = { 0, cap, out, 0, lo, mi, up };
    sM->cellCalculate(sM, cur, lo, mi, up, (Symbol) cX, (Symbol) cY, recordTransition, &r);
    return r.n;
}

/* ---------------------------------------------------------------- small pieces */

double ref_logadd(double x, double y) {
    return logAdd(x, y);
}

int ref_symbol(char ch) {
    return (int) symbol_convertCharToSymbol(ch);
}

static stList *makeAnchorList(const int64_t *anchors, int64_t n) {
    stList *l = stList_construct3(0, (void (*)(void *)) stIntTuple_destruct);
    for (int64_t i = 0; i < n; i++) {
        stList_append(l, stIntTuple_construct2(anchors[2 * i], anchors[2 * i + 1]));
    }
    return l;
}

/* out: (xay, xmyL, xmyR) per diagonal, lX+lY+1 rows. */
int64_t ref_band(const int64_t *anchors, int64_t n, int64_t lX, int64_t lY, int64_t expansion, int64_t *out) {
    stList *l = makeAnchorList(anchors, n);
    Band *band = band_construct(l, lX, lY, expansion);
    BandIterator *it = bandIterator_construct(band);
    for (int64_t i = 0; i <= lX + lY; i++) {
        Diagonal d = bandIterator_getNext(it);
        out[3 * i] = diagonal_getXay(d);
        out[3 * i + 1] = diagonal_getMinXmy(d);
        out[3 * i + 2] = diagonal_getMaxXmy(d);
    }
    bandIterator_destruct(it);
    band_destruct(band);
    stList_destruct(l);
    return lX + lY + 1;
}

int64_t ref_split_points(const int64_t *anchors, int64_t n, int64_t lX, int64_t lY, int64_t splitMatrixBiggerThanThis,
        int raggedLeft, int raggedRight, int64_t *out, int64_t cap) {
    stList *l = makeAnchorList(anchors, n);
    stList *sp = getSplitPoints(l, lX, lY, splitMatrixBiggerThanThis, raggedLeft, raggedRight);
    int64_t m = stList_length(sp);
    for (int64_t i = 0; i < m && i < cap; i++) {
        stIntTuple *t = stList_get(sp, i);
        for (int k = 0; k < 4; k++) {
            out[4 * i + k] = stIntTuple_get(t, k);
        }
    }
    stList_destruct(sp);
    stList_destruct(l);
    return m;
}

/* ---------------------------------------------------------------- the two public entry points */

static PairwiseAlignmentParameters *makeParams(double threshold, int64_t minDiagsBetweenTraceBack,
        int64_t traceBackDiagonals, int64_t diagonalExpansion, int64_t splitMatrixBiggerThanThis) {
    PairwiseAlignmentParameters *p = pairwiseAlignmentBandingParameters_construct();
    p->threshold = threshold;
    p->minDiagsBetweenTraceBack = minDiagsBetweenTraceBack;
    p->traceBackDiagonals = traceBackDiagonals;
    p->diagonalExpansion = diagonalExpansion;
    p->splitMatrixBiggerThanThis = splitMatrixBiggerThanThis;
    p->constraintDiagonalTrim = 0;
    return p;
}

/* out rows (p_int, x, y) in the reference's list order; returns the number of pairs (may exceed cap). */
int64_t ref_aligned_pairs(void *sM, const char *sX, const char *sY, const int64_t *anchors, int64_t nAnchors,
        double threshold, int64_t minDiagsBetweenTraceBack, int64_t traceBackDiagonals, int64_t diagonalExpansion,
        int64_t splitMatrixBiggerThanThis, int raggedLeft, int raggedRight, int64_t *out, int64_t cap) {
    PairwiseAlignmentParameters *p = makeParams(threshold, minDiagsBetweenTraceBack, traceBackDiagonals,
            diagonalExpansion, splitMatrixBiggerThanThis);
    stList *l = makeAnchorList(anchors, nAnchors);
    stList *pairs = getAlignedPairsUsingAnchors(sM, sX, sY, l, p, raggedLeft, raggedRight);
    int64_t m = stList_length(pairs);
    for (int64_t i = 0; i < m && i < cap; i++) {
        stIntTuple *t = stList_get(pairs, i);
        out[3 * i] = stIntTuple_get(t, 0);
        out[3 * i + 1] = stIntTuple_get(t, 1);
        out[3 * i + 2] = stIntTuple_get(t, 2);
    }
    stList_destruct(pairs);
    stList_destruct(l);
    pairwiseAlignmentBandingParameters_destruct(p);
    return m;
}

/* Accumulates into T (S*S), E (S*16), lik (1) which the caller has initialised (cPecanRealign.c:531 uses 1e-12). */
void ref_expectations(void *sM_, const char *sX, const char *sY, const int64_t *anchors, int64_t nAnchors,
        double threshold, int64_t minDiagsBetweenTraceBack, int64_t traceBackDiagonals, int64_t diagonalExpansion,
        int64_t splitMatrixBiggerThanThis, int raggedLeft, int raggedRight, double *T, double *E, double *lik) {
    StateMachine *sM = sM_;
    PairwiseAlignmentParameters *p = makeParams(threshold, minDiagsBetweenTraceBack, traceBackDiagonals,
            diagonalExpansion, splitMatrixBiggerThanThis);
    stList *l = makeAnchorList(anchors, nAnchors);
    Hmm *hmm = hmm_constructEmpty(0.0, sM->type);
    int64_t S = hmm->stateNumber;
    memcpy(hmm->transitions, T, sizeof(double) * S * S);
    memcpy(hmm->emissions, E, sizeof(double) * S * 16);
    hmm->likelihood = *lik;
    getExpectationsUsingAnchors(sM, hmm, sX, sY, l, p, raggedLeft, raggedRight);
    memcpy(T, hmm->transitions, sizeof(double) * S * S);
    memcpy(E, hmm->emissions, sizeof(double) * S * 16);
    *lik = hmm->likelihood;
    hmm_destruct(hmm);
    stList_destruct(l);
    pairwiseAlignmentBandingParameters_destruct(p);
}

/* ---------------------------------------------------------------- downstream of the posteriors */

/*
 * The realign tail of cPecanRealign.c:590-611: AMAP reweight, MEA chain filter, sort.  pairs rows (p_int,x,y) in the
 * order getAlignedPairsUsingAnchors returned them.  seed < 0 leaves the libc rand() stream alone, else srand(seed)
 * first (a fresh cPecanRealign process has the default seed 1).  out rows (x, y) sorted.  Returns count.
 */
int64_t ref_realign_tail(const int64_t *pairs, int64_t nPairs, const char *sX, const char *sY, double gapGamma,
        double matchGamma, int seed, int64_t *out, int64_t cap) {
    if (seed >= 0) {
        srand((unsigned) seed);
    }
    stList *aligned = stList_construct3(0, (void (*)(void *)) stIntTuple_destruct);
    for (int64_t i = 0; i < nPairs; i++) {
        stList_append(aligned, stIntTuple_construct3(pairs[3 * i], pairs[3 * i + 1], pairs[3 * i + 2]));
    }
    /* gapGamma and matchGamma are floats in the reference (pairwiseAligner.h:38, cPecanRealign.c:390). */
    aligned = reweightAlignedPairs2(aligned, strlen(sX), strlen(sY), (float) gapGamma);
    aligned = filterPairwiseAlignmentToMakePairsOrdered(aligned, sX, sY, (float) matchGamma);
    int64_t m = stList_length(aligned);
    stList *xy = stList_construct3(0, (void (*)(void *)) stIntTuple_destruct);
    for (int64_t i = 0; i < m; i++) {
        stIntTuple *t = stList_get(aligned, i);
        stList_append(xy, stIntTuple_construct2(stIntTuple_get(t, 1), stIntTuple_get(t, 2)));
    }
    stList_sort(xy, (int (*)(const void *, const void *)) stIntTuple_cmpFn);
    for (int64_t i = 0; i < m && i < cap; i++) {
        stIntTuple *t = stList_get(xy, i);
        out[2 * i] = stIntTuple_get(t, 0);
        out[2 * i + 1] = stIntTuple_get(t, 1);
    }
    stList_destruct(xy);
    stList_destruct(aligned);
    return m;
}
