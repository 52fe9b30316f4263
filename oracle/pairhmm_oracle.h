/*
 * pairhmm_oracle.h -- TEST INFRASTRUCTURE ONLY (never linked into the product library).
 *
 * Plain-C CPU restatement of the reference hot path (cPecan's anchor-banded pair-HMM forward-backward):
 * used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg as the checker for the CUDA path.
 * Parity of this restatement is PINNED: tests/test_oracle_golden.py checks it bit-for-bit against fixtures
 * generated from the reference's own C code (oracle/_ref, tests/golden/make_golden.py) and against the
 * reference's unit-test known answers (test_bands, test_getSplitPoints, test_diagonalDPCalculations).
 */
#ifndef PAIRHMM_ORACLE_H_
#define PAIRHMM_ORACLE_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PH_MAX_S 5
#define PH_PROB_1 10000000 /* PAIR_ALIGNMENT_PROB_1, cPecan/inc/pairwiseAligner.h:26 */

/* Log-space parameter block of one state machine: the ordered transition lists of cellCalculate, grouped by the
 * neighbour cell they read (0 lower=(x-1,y), 1 middle=(x-1,y-1), 2 upper=(x,y-1)). */
typedef struct {
    int32_t type;         /* 0 fiveState 1 fiveStateAsymmetric 2 threeState 3 threeStateAsymmetric */
    int32_t S;            /* 5 or 3 */
    int32_t n_trans[3];
    int32_t from[3][5];
    int32_t to[3][5];
    double tp[3][5];
    double start[PH_MAX_S], ragged_start[PH_MAX_S], end[PH_MAX_S], ragged_end[PH_MAX_S];
    double em_match[25];  /* [cX*5+cY]; any N -> literal -2.772588722 */
    double em_gap_x[5];   /* [cX]; N -> literal -1.386294361 */
    double em_gap_y[5];
} ph_model;

typedef struct {
    double threshold;                 /* 0.01 */
    int64_t min_diags_between_traceback; /* 1000 */
    int64_t traceback_diagonals;      /* 40 */
    int64_t diagonal_expansion;       /* even */
    int64_t split_matrix_bigger_than; /* already squared */
} ph_params;

int ph_model_from_hmm(int type, const double *T, const double *E, ph_model *out);
int ph_model_default(int type, ph_model *out);
double ph_logadd(double x, double y);
int ph_symbol(char c);
/* fills xmyL/xmyR for xay = 0..lX+lY, returns lX+lY+1 */
int64_t ph_band(const int64_t *anchors, int64_t n, int64_t lX, int64_t lY, int64_t expansion, int64_t *xmyL,
        int64_t *xmyR);
int64_t ph_split_points(const int64_t *anchors, int64_t n, int64_t lX, int64_t lY, int64_t split, int raggedLeft,
        int raggedRight, int64_t *out, int64_t cap);
/* rows (p_int, x, y) in the reference's list order; returns count (may exceed cap) */
int64_t ph_aligned_pairs(const ph_model *m, const char *sX, int64_t lX, const char *sY, int64_t lY,
        const int64_t *anchors, int64_t n, const ph_params *p, int raggedLeft, int raggedRight, int64_t *out,
        int64_t cap);
/* accumulates into T[S*S], E[S*16], lik[1] */
void ph_expectations(const ph_model *m, const char *sX, int64_t lX, const char *sY, int64_t lY,
        const int64_t *anchors, int64_t n, const ph_params *p, int raggedLeft, int raggedRight, double *T, double *E,
        double *lik);
/* number of in-band cells summed over regions (forward cell visits; the throughput unit) */
int64_t ph_count_cells(int64_t lX, int64_t lY, const int64_t *anchors, int64_t n, const ph_params *p, int raggedLeft,
        int raggedRight);
/* cigar (ops: 0 M, 1 INDEL_X (ref only), 2 INDEL_Y (read only)) -> matching-base anchors; returns count */
int64_t ph_anchors_from_cigar(const int32_t *ops, const int64_t *lens, int64_t nOps, const char *sX, const char *sY,
        int64_t *out, int64_t cap);

/* Realign tail of one alignment (cPecanRealign.c:590-611): AMAP reweight + maximum-expected-accuracy chain.  pairs:
 * n rows (p_int, x, y) in list order; seed >= 0 -> srand(seed) first (the reference process starts at 1); out_xy:
 * sorted chain pairs (cap n rows); out_w (or NULL): reweighted weight of every input pair.  Returns chain length. */
int64_t ph_realign_chain(const int64_t *pairs, int64_t n, int64_t lX, int64_t lY, double gapGamma, double matchGamma,
        int seed, int64_t *out_xy, int64_t *out_w);
/* convertAlignedPairsToPairwiseAlignment (cPecanRealign.c:50-92): rows (op, length); returns the op count */
int64_t ph_cigar_from_pairs(const int64_t *xy, int64_t m, int64_t lX, int64_t lY, int64_t *out_ops, int64_t cap);

#ifdef __cplusplus
}
#endif
#endif
