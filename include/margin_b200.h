/*
 * margin_b200.h -- C ABI of libmargin_b200.so: the B200-native replacement for the cPecanRealign hot path.
 *
 * The reference reaches this path through a process boundary: `cPecanRealign [flags] ref.fa read.fa < cigars`
 * (submodules/cPecan/cPecanRealign.c:388-661), launched once per read from
 *     src/margin/marginAlignLib.py:315-318   (realign)
 *     src/margin/marginCallerLib.py:49-57    (posteriors)
 *     submodules/cPecan/cPecanEm.py:178-180  (EM E-step, --outputExpectations)
 * A maintainer replacing those `system()` calls binds the entry points below (ctypes stub in INTEGRATION.md).
 * Everything is extern "C", plain pointers and sizes; no torch / STL types cross the boundary.  Device pointers
 * are raw CUDA device addresses (the Python host passes torch tensors' data_ptr()), streams are cudaStream_t cast
 * to void*.
 *
 * Conventions (same as the reference): X = reference window (cigar contig1), Y = read (contig2); cigar op codes
 * 0 = M, 1 = D (consumes X only, PAIRWISE_INDEL_X), 2 = I (consumes Y only, PAIRWISE_INDEL_Y)
 * (submodules/sonLib/C/inc/pairwiseAlignment.h:17-19); states match=0, shortGapX=1, shortGapY=2, longGapX=3,
 * longGapY=4 (cPecan/impl/stateMachine.c:210-212); hmm types 0 fiveState, 1 fiveStateAsymmetric, 2 threeState,
 * 3 threeStateAsymmetric (cPecan/inc/stateMachine.h:28-33).
 *
 * All functions return 0 on success or a negative MAB_E_* code; mab_last_error() gives the message of the last
 * failure on the calling thread (the reference's convention is st_errAbort -> exit(1) -> RuntimeError in Python;
 * the Python host raises RuntimeError on any non-zero return).
 */
#ifndef MARGIN_B200_H_
#define MARGIN_B200_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MAB_OK 0
#define MAB_E_ARG (-1)      /* bad argument (null pointer, bad hmm type, odd diagonalExpansion, ...) */
#define MAB_E_INPUT (-2)    /* malformed alignment (cigar does not span the sequences, coordinates out of range) */
#define MAB_E_CUDA (-3)     /* CUDA runtime error */
#define MAB_E_CAPACITY (-4) /* caller-provided buffer too small */
#define MAB_E_NOGPU (-5)    /* no CUDA device: there is NO CPU fallback on this path */

typedef struct mab_model mab_model; /* replaces StateMachine* (cPecan/inc/stateMachine.h:59-75) */
typedef struct mab_plan mab_plan;   /* host-side work plan of one batch: regions, bands, traceback schedule */

const char *mab_last_error(void);
const char *mab_version(void);
/* number of visible CUDA devices (0 when none / driver missing); never fails */
int mab_device_count(void);

/* ---------------------------------------------------------------- models
 * mab_model_from_hmm replaces hmm_loadFromFile + hmm_getStateMachine (stateMachine.c:145-202, 742-764): T is the
 * S*S row-major transition matrix (from*S+to) and E the S*16 emission table (state*16 + x*4 + y) exactly as stored
 * in a `.hmm` file (probabilities, not logs; NOT normalised on load, cPecanRealign.c:522).
 * mab_model_default replaces stateMachine5_construct / stateMachine3_construct (the built-in model cPecanRealign
 * uses without --loadHmm, cPecanRealign.c:526). */
int mab_model_from_hmm(int hmm_type, const double *T, const double *E, mab_model **out);
int mab_model_default(int hmm_type, mab_model **out);
void mab_model_destroy(mab_model *m);
int mab_model_state_number(const mab_model *m);
int mab_model_type(const mab_model *m);
/* Log-space parameter block for inspection/tests.  transitions: rows (group, from, to, tP), group 0 lower (gap X),
 * 1 middle (match), 2 upper (gap Y), in cellCalculate order (stateMachine.c:399-429, 636-661); returns row count.
 * boundary: which = 0 start, 1 raggedStart, 2 end, 3 raggedEnd, S doubles.  emissions: em_match[25] (5x5 incl. N),
 * em_gap_x[5], em_gap_y[5]. */
int mab_model_transitions(const mab_model *m, double *rows4, int cap_rows);
int mab_model_boundary(const mab_model *m, int which, double *out);
int mab_model_emissions(const mab_model *m, double *em_match25, double *em_gap_x5, double *em_gap_y5);

/* ---------------------------------------------------------------- parameters
 * PairwiseAlignmentParameters (cPecan/inc/pairwiseAligner.h:28-39) restricted to what the anchored entry points
 * read, plus the two ragged-end flags that cPecanRealign always passes as (1,1) (cPecanRealign.c:570-576). */
typedef struct {
    double threshold;                     /* 0.01 */
    int64_t min_diags_between_traceback;  /* 1000 */
    int64_t traceback_diagonals;          /* 40 */
    int64_t diagonal_expansion;           /* even; 10 from every margin* call site */
    int64_t split_matrix_bigger_than;     /* ALREADY SQUARED (cPecanRealign.c:459 squares the flag value) */
    int32_t ragged_left;                  /* 1 */
    int32_t ragged_right;                 /* 1 */
} mab_params;
void mab_params_default(mab_params *p);

/* ---------------------------------------------------------------- planning (host only, no CUDA needed)
 * One batch = n guide alignments, concatenated:
 *   seq_x/off_x, seq_y/off_y : ASCII bases of the aligned sub-sequences (cPecanRealign.c:557-558), off_*[n+1]
 *   either  ops/lens/off_ops : the guide cigar relative to those sub-sequences -> anchors are derived as
 *                              convertPairwiseForwardStrandAlignmentToAnchorPairs(trim 0) + matchFn filter
 *                              (pairwiseAligner.c:789-813, cPecanRealign.c:259-263, 567)
 *   or      anchors/off_anc  : explicit anchor pairs (x,y interleaved int64), strictly increasing in x and y
 *                              (getAlignedPairsUsingAnchors' own contract, pairwiseAligner.c:1172)
 * The plan holds: split regions (getSplitPoints, :1040-1067), per-region bands (band_construct, :128-179),
 * the checkpoint/traceback schedule of getPosteriorProbsWithBanding (:666-744) and the every-10th-diagonal
 * total-probability schedule (:712-720), all as flat arrays in one blob that is copied to the device verbatim. */
int mab_plan_from_cigars(const mab_params *p, int64_t n, const uint8_t *seq_x, const int64_t *off_x,
        const uint8_t *seq_y, const int64_t *off_y, const uint8_t *ops, const int32_t *lens, const int64_t *off_ops,
        int n_threads, mab_plan **out);
int mab_plan_from_anchors(const mab_params *p, int64_t n, const uint8_t *seq_x, const int64_t *off_x,
        const uint8_t *seq_y, const int64_t *off_y, const int64_t *anchors, const int64_t *off_anc, int n_threads,
        mab_plan **out);
void mab_plan_destroy(mab_plan *plan);

typedef struct {
    int64_t n_alignments;
    int64_t n_regions;     /* split regions with at least one diagonal */
    int64_t n_diagonals;   /* sum over regions of (lX + lY + 1) */
    int64_t n_cells;       /* in-band cells: the throughput unit */
    int64_t n_tracebacks;
    int64_t n_totals;      /* total-probability evaluations */
    int64_t n_anchors;
    int64_t max_width;     /* widest diagonal in the batch */
    int64_t blob_bytes;    /* size of the plan blob (host == device layout) */
    int64_t workspace_bytes_posteriors;   /* device scratch needed by mab_posteriors_* */
    int64_t workspace_bytes_expectations; /* device scratch needed by mab_expectations */
} mab_plan_info;
int mab_plan_get_info(const mab_plan *plan, mab_plan_info *info);
/* Writes the blob into caller memory (typically a pinned torch tensor); dst_bytes >= blob_bytes. */
int mab_plan_write_blob(const mab_plan *plan, void *dst, int64_t dst_bytes);
/* Plan introspection for tests (host): anchors of alignment i; regions rows (aln, x1, y1, x2, y2); band of region r
 * rows (xmyL, xmyR) for xay = 0..lX+lY; tracebacks rows (region, lo, hi, from, at_end). Return counts. */
int64_t mab_plan_anchors(const mab_plan *plan, int64_t aln, int64_t *out_xy, int64_t cap_pairs);
int64_t mab_plan_regions(const mab_plan *plan, int64_t *out5, int64_t cap_rows);
int64_t mab_plan_band(const mab_plan *plan, int64_t region, int64_t *out2, int64_t cap_rows);
int64_t mab_plan_tracebacks(const mab_plan *plan, int64_t *out5, int64_t cap_rows);

/* ---------------------------------------------------------------- planning on the device
 * Same plan, built in HBM from raw inputs that are ALREADY on the device (the arrays of mab_plan_from_cigars /
 * mab_plan_from_anchors as device pointers; pass d_off_ops = NULL to use anchors).  Only bases and cigars cross PCIe
 * (~0.4 kB per tRNA read instead of ~4.8 kB of descriptors) and nothing is planned on host cores.  Replaces the same
 * reference code as the host planner (cigarRead .. band_construct .. the checkpoint schedule, see above).
 * d_arena: caller-owned device memory, 256-byte aligned; the blob is laid out inside it and stays valid as long as
 * the arena does.  mab_plan_device_arena_hint gives a first size; if the arena is too small the call returns
 * MAB_E_CAPACITY with the exact requirement in *needed_bytes (retry with that).  The call synchronises `stream`
 * twice (two 64-byte count read-backs).  The returned plan is used like a host-built one, with
 * d_blob = mab_plan_device_blob(plan); its introspection getters and mab_plan_write_blob return MAB_E_ARG. */
int64_t mab_plan_device_arena_hint(const mab_params *p, int64_t n, int64_t len_x, int64_t len_y);
int mab_plan_build_device(const mab_params *p, int64_t n, const uint8_t *d_seq_x, const int64_t *d_off_x, int64_t len_x,
        const uint8_t *d_seq_y, const int64_t *d_off_y, int64_t len_y, const uint8_t *d_ops, const int32_t *d_lens,
        const int64_t *d_off_ops, const int64_t *d_anchors, const int64_t *d_off_anc, void *d_arena,
        int64_t arena_bytes, void *stream, mab_plan **out, int64_t *needed_bytes);
const void *mab_plan_device_blob(const mab_plan *plan);

/* ---------------------------------------------------------------- device work
 * d_blob: device copy of the plan blob.  d_ws: device scratch of at least the size mab_plan_get_info reported.
 * All launches are asynchronous on `stream`.
 *
 * Posterior mode replaces getAlignedPairsUsingAnchors (pairwiseAligner.c:1172-1190):
 *   mab_posteriors_compute : forward, backward, totals, per-cell posteriors, per-region pair counts and their scan.
 *                            Writes the number of pairs to d_n_pairs[0] and per-alignment offsets (n+1) to
 *                            d_pair_off.
 *   mab_posteriors_emit    : writes the pairs (x, y, p_int = floor(min(p,1)*1e7)), alignment by alignment, in the
 *                            reference's list order (:594-613, :1167-1169), into d_x/d_y/d_p (capacity entries).
 * Expectation mode replaces getExpectationsUsingAnchors (pairwiseAligner.c:1201-1206) applied to every alignment of
 * the batch and summed: d_out[S*S + S*16 + 1] = transitions, emissions, likelihood, ADDED to what d_out holds
 * (cPecanRealign.c:531 initialises with the 1e-12 pseudo-count; the host does the same). */
int mab_posteriors_compute(const mab_model *m, const mab_plan *plan, const void *d_blob, void *d_ws, int64_t ws_bytes,
        int64_t *d_pair_off, int64_t *d_n_pairs, void *stream);
int mab_posteriors_emit(const mab_plan *plan, const void *d_blob, void *d_ws, int64_t ws_bytes, int32_t *d_x,
        int32_t *d_y, int32_t *d_p, int64_t capacity, void *stream);
int mab_expectations(const mab_model *m, const mab_plan *plan, const void *d_blob, void *d_ws, int64_t ws_bytes,
        double *d_out, void *stream);
/* mab_posteriors_emit with 16-bit coordinates: 8 instead of 12 bytes per pair leave the GPU.  Only valid when every
 * sub-sequence of the batch is shorter than 32768 bases (the caller checks: coordinates are relative to the
 * sub-sequences); the pairs and their order are the same. */
int mab_posteriors_emit_i16(const mab_plan *plan, const void *d_blob, void *d_ws, int64_t ws_bytes, int16_t *d_x,
        int16_t *d_y, int32_t *d_p, int64_t capacity, void *stream);
/* mab_posteriors_compute for callers that may not allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments, same results as mab_posteriors_compute; the difference is the scratch of the over-wide
 * band kernels (diagonals wider than a shared-memory ring, ~1850 cells for five states), which mab_posteriors_compute
 * takes from cudaMallocAsync and this call takes from a static device array of MAB_NOALLOC_SCRATCH_BYTES that every
 * GPU loading the library holds until exit.  The call itself issues no cudaMalloc / cudaFree or allocating library
 * routine.  Limit: one block of the over-wide kernels needs 3 * max_width * S * 8 * 4 bytes of ring, so the widest
 * band it takes is MAB_NOALLOC_SCRATCH_BYTES / (96 * S) cells (69905 for five states, 116508 for three); beyond that
 * it returns MAB_E_CAPACITY with a message and runs nothing.  With fewer rings than mab_posteriors_compute would
 * allocate, the over-wide items run on fewer blocks (slower, same results).  All calls on one GPU share the array
 * (nothing in it has to be cleared: every ring cell is written before it is read), so two calls that overlap on the
 * GPU -- different streams or host threads -- race; serialise them. */
#define MAB_NOALLOC_SCRATCH_BYTES (32u << 20)
int mab_posteriors_compute_noalloc(const mab_model *m, const mab_plan *plan, const void *d_blob, void *d_ws,
        int64_t ws_bytes, int64_t *d_pair_off, int64_t *d_n_pairs, void *stream);
/* ---- realign tail --------------------------------------------------------------------------------
 * Replaces what cPecanRealign does with the posterior pairs of an alignment before it prints the new cigar
 * (cPecanRealign.c:590-631): reweightAlignedPairs2 (pairwiseAligner.c:1251-1261; AMAP gap reweighting),
 * filterPairwiseAlignmentToMakePairsOrdered (multipleAligner.c:941-967; maximum-expected-accuracy chain with the
 * reference's rand() jitter) and convertAlignedPairsToPairwiseAlignment (cPecanRealign.c:50-92; the cigar).  Works on
 * the device-resident output of mab_posteriors_emit, for all n alignments of the batch at once; only cigars leave
 * the GPU.
 *   d_off_x / d_off_y [n+1] : sequence offsets of the batch (alignment i has lX = off_x[i+1] - off_x[i]); len_x/len_y
 *                             their last entries.  d_pair_off [n+1], d_x, d_y, d_p: as written by mab_posteriors_*;
 *                             n_pairs = capacity of d_x/d_y/d_p (the same value in both calls).  An alignment whose
 *                             pairs were clipped by that capacity is skipped (no operations, chain length 0): the
 *                             caller compares the pair count of mab_posteriors_compute with its capacity.
 *   gap_gamma / match_gamma : cPecanRealign --gapGamma / --matchGamma (rounded to float like the reference).
 *   d_rand [n_rand]         : the jitter stream, mab_rand_sequence(1, n_rand, ...) copied to the device; n_rand must
 *                             be at least twice the largest pair count of one alignment (every reference process
 *                             restarts rand() at seed 1 and draws twice per pair).
 *   d_ws                    : scratch of mab_realign_workspace_bytes(...) bytes, 256-byte aligned; it carries the
 *                             chains from mab_realign_chain to mab_realign_emit.
 * mab_realign_chain writes d_op_off[n+1] / d_chain_off[n+1] (cigar operations / chain pairs per alignment, exclusive
 * scans, totals in entry n); the caller sizes the outputs from them and calls mab_realign_emit, which writes the
 * operations (op codes 0 M, 1 D, 2 I as in the input cigars) and, unless d_chain_x is NULL, the chain pairs. */
int64_t mab_realign_workspace_bytes(int64_t n, int64_t n_pairs, int64_t len_x, int64_t len_y);
int mab_realign_chain(int64_t n, const int64_t *d_off_x, const int64_t *d_off_y, int64_t len_x, int64_t len_y,
        const int64_t *d_pair_off, const int32_t *d_x, const int32_t *d_y, const int32_t *d_p, int64_t n_pairs,
        double gap_gamma, double match_gamma, const int32_t *d_rand, int64_t n_rand, void *d_ws, int64_t ws_bytes,
        int64_t *d_op_off, int64_t *d_chain_off, void *stream);
int mab_realign_emit(int64_t n, const int64_t *d_off_x, const int64_t *d_off_y, int64_t len_x, int64_t len_y,
        int64_t n_pairs, void *d_ws, int64_t ws_bytes, const int64_t *d_op_off, const int64_t *d_chain_off,
        uint8_t *d_ops, int32_t *d_lens, int32_t *d_chain_x, int32_t *d_chain_y, void *stream);
/* The first n values glibc's rand() returns after srand(seed) (host memory): the jitter stream of st_random()
 * (sonLib/C/impl/sonLibRandom.c:39-43) without touching the process-wide libc generator. */
int mab_rand_sequence(uint32_t seed, int64_t n, int32_t *out);

/* ---- marginCaller columns -----------------------------------------------------------------------
 * Replaces the consumer of --outputAllPosteriorProbs in posteriorProbabilityCalculationTargetFn
 * (src/margin/marginCallerLib.py:58-68) and the per-column calculation of variantCallSamFileTargetFn (:189-209,
 * calcBasePosteriorProbs :81-92), on the device-resident output of mab_posteriors_emit.
 *   mab_columns_accumulate: table[col_base[i] + x][base of read[y]] += round6(p_int / 1e7) for every pair (x, y,
 *       p_int) of alignment i, in units of 1e-6 (the "%f" the reference writes and re-reads, cPecanRealign.c:346-351).
 *       d_table: int64 [n_columns][4] (A, C, G, T), zeroed by the caller before the first batch; d_col_base[i] =
 *       column of the first base of alignment i's reference window (offset of its reference sequence in the
 *       concatenated column space + the alignment's reference start).  d_seq_y / d_off_y: the read sub-sequences.
 *   mab_columns_call: d_post[n_columns][4] = posterior of each true base given the normalised observations;
 *       d_log_err[true][observed], d_log_evo[reference][true] are log substitution matrices; d_ref the reference
 *       bases (ASCII) of the column space.  Columns with no observation or a non-ACGT reference base give zeros. */
int mab_columns_accumulate(int64_t n, const int64_t *d_pair_off, const int32_t *d_x, const int32_t *d_y,
        const int32_t *d_p, const uint8_t *d_seq_y, const int64_t *d_off_y, const int64_t *d_col_base,
        int64_t *d_table, int64_t n_columns, void *stream);
int mab_columns_call(const int64_t *d_table, int64_t n_columns, const uint8_t *d_ref, const double *d_log_err,
        const double *d_log_evo, double *d_post, void *stream);

/* Debug/test hook: copies the forward / backward matrices ([cell][S] doubles, cells in plan order) and the totals
 * out of the workspace after a compute call.  Any pointer may be NULL. */
int mab_debug_copy_matrices(const mab_plan *plan, const void *d_ws, double *h_fwd, double *h_bwd, double *h_totals,
        void *stream);
/* Debug/test hook: warps that share one 5-state work item (region / traceback) in the wavefront kernels.
 * 0 = chosen from the mean band width (default), 1, 2, 4 or 8 = forced.  Results do not depend on it. */
int mab_debug_set_warp_group(int warps);
/* Debug/test hook: work items are dispatched by band class (0: a half-warp per item, 1: a warp per item, both on the
 * register-resident wavefront kernels; 2: wide bands on the shared-memory-ring kernels).  cls = 1 runs class-0 items as
 * class 1, cls = 2 runs everything as class 2; 0 restores the default.  Results do not depend on it. */
int mab_debug_set_min_band_class(int cls);
/* Posterior mode can run backward sweep, total probabilities and posterior match probabilities fused in one kernel for
 * band classes 0 and 1 (the reference's own loop structure, pairwiseAligner.c:695-733): on = 1.  Default 0: separate
 * kernels, which measured faster.  Results do not depend on it (bit-identical pairs). */
int mab_debug_set_fused_posterior(int on);
/* Kernel launches issued by this library on the calling thread since the last reset (bench.py's gpu_launches). */
int64_t mab_launch_count(int reset);
/* Per-kernel timing for the roofline report: when enabled every launch is bracketed by CUDA events on its stream.
 * mab_kernel_times synchronises on the recorded events and returns, per kernel id (0 .. mab_kernel_count()-1), the
 * summed device time in milliseconds and the number of launches since the last reset. */
int mab_kernel_timing(int enable);
int mab_kernel_count(void);
const char *mab_kernel_name(int kernel);
int mab_kernel_times(double *ms, int64_t *launches, int reset);

#ifdef __cplusplus
}
#endif
#endif /* MARGIN_B200_H_ */
