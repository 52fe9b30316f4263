// mab_internal.h -- layouts shared by the host planner (plan.cpp), the model builder (model.cpp) and the
// CUDA kernels (kernels.cu).  Not part of the public ABI (include/margin_b200.h is).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/margin_b200.h"

#define MAB_MAX_S 5
#define MAB_BLOB_MAGIC 0x4d41423230304231LL /* "MAB200B1" */
#define MAB_PROB_1 10000000.0               /* PAIR_ALIGNMENT_PROB_1, cPecan/inc/pairwiseAligner.h:26 */

void mab_set_error(const char *fmt, ...);

// kernel ids of realign.cu inside the per-kernel timing table of kernels.cu (mab_kernel_times)
#define MAB_K_RL_WEIGHTS 7
#define MAB_K_RL_CHAIN 8
#define MAB_K_RL_SCAN 9
#define MAB_K_RL_EMIT 10
#define MAB_K_COL_ACC 11   // columns.cu
#define MAB_K_COL_CALL 12
void *mab_launch_timer_begin(int kernel, void *stream);
void mab_launch_timer_end(void *timer);

// ---------------------------------------------------------------------------------------------- model

// Host model: the generic ordered-transition-list form (mirrors what stateMachine{5,3}_cellCalculate hard-code).
struct mab_model {
    int32_t type;
    int32_t S;
    int32_t n_trans[3];
    int32_t from[3][5];
    int32_t to[3][5];
    double tp[3][5];
    double start[MAB_MAX_S], ragged_start[MAB_MAX_S], end[MAB_MAX_S], ragged_end[MAB_MAX_S];
    double em_match[25];
    double em_gap_x[5];
    double em_gap_y[5];
};

// Device model block, passed to kernels by value (kernel parameter space) and staged into shared memory.
// etp_* = (eP + tP), the exact sub-expression the reference forms before adding the cell value
// (pairwiseAligner.c:314,324,356), so precomputing it keeps the DP bit-identical.
struct DevModel {
    int32_t S;
    int32_t n_lower, n_middle, n_upper;  // 4,5,4 (5-state) or 3,3,3 (3-state)
    double etp_lower[5][5];              // [cX][k]
    double etp_middle[25][5];            // [cX*5+cY][k]
    double etp_upper[5][5];              // [cY][k]
    double start[MAB_MAX_S], ragged_start[MAB_MAX_S], end[MAB_MAX_S], ragged_end[MAB_MAX_S];
    int32_t t_from[13], t_to[13];        // flattened transition list (lower, middle, upper) for the expectation kernel
    // the two addends of etp_* on their own (wavefront_reg.cuh forms eP + tP in the kernel)
    double tp[3][5];                     // [lower | middle | upper][k]
    double em_match[25], em_gap_x[5], em_gap_y[5];
};
void mab_model_to_device_block(const mab_model *m, DevModel *out);

// ---------------------------------------------------------------------------------------------- plan blob

struct BlobHeader {  // 256 bytes
    int64_t magic;
    int64_t n_alignments, n_regions, n_diagonals, n_cells, n_tracebacks, n_totals, n_anchors, max_width;
    int64_t len_sym_x, len_sym_y;
    // byte offsets of the arrays inside the blob (each 16-byte aligned)
    int64_t off_regions;      // RegionDesc[n_regions]
    int64_t off_diags;        // DiagDesc[n_diagonals + 1]   (last entry is a sentinel carrying n_cells)
    int64_t off_tracebacks;   // TracebackDesc[n_tracebacks]
    int64_t off_totals;       // TotalDesc[n_totals]
    int64_t off_sym_x;        // uint8 symbol codes 0..4 of all X sequences, concatenated like seq_x
    int64_t off_sym_y;
    int64_t off_order_fwd;    // int32[n_regions]   regions, longest first
    int64_t off_order_bwd;    // int32[n_tracebacks] tracebacks, longest first
    int64_t off_aln_region;   // int64[n_alignments + 1] first region of each alignment
    int64_t blob_bytes;
    int64_t n_edge_cells;     // cells of the diagonals from+1 of all intermediate tracebacks (see TracebackDesc)
    int64_t n_full_cells;     // cells of all diagonals that carry a total (posterior mode keeps all states there)
    int64_t off_diag_region;  // int32[n_diagonals + 1] owning region of every diagonal (sentinel: -1)
    // Work items by band class (mab_band_class): order_fwd / order_bwd list class 2 first, then 1, then 0, each class
    // longest first.  The register-resident wavefront kernels take classes 0 and 1, the shared-memory-ring kernels 2.
    int64_t n_class_regions[3];
    int64_t n_class_tbs[3];
    // 2-bit packed symbols + N mask of the concatenated X / Y sequences (symbol_convertStringToSymbols,
    // pairwiseAligner.c:281-289, as bits): uint32 words, mab_pack_words(len) of them per sequence -- first the 2-bit
    // codes, 16 bases per word, then the N mask, 32 bases per word (1 = not ACGT); base i at position i + 256
    int64_t off_pack_x, off_pack_y;
};
static_assert(sizeof(BlobHeader) == 256, "BlobHeader must be 256 bytes");

struct RegionDesc {  // 64 bytes
    int64_t sym_x_off, sym_y_off;  // start of this region's symbols inside sym_x / sym_y
    int64_t diag_base;             // global index of this region's xay = 0 diagonal
    int32_t lX, lY;
    int32_t x1, y1;                // origin inside the alignment (alignedPairCoordinateCorrectionFn)
    int32_t aln;
    int32_t flags;                 // bit0 ragged left, bit1 ragged right (after the split rule of :1121-1122);
                                   // bits 8..30: band span (plan_core::SpanTracker), see mab_band_class
    int32_t tb_base, n_tb;
    int32_t tot_base, n_tot;
};
static_assert(sizeof(RegionDesc) == 64, "RegionDesc must be 64 bytes");

// Band span of a region: the largest x-y extent (max xmyR - min xmyL) over any three consecutive diagonals.  The
// register-resident kernels map x-y to lanes circularly (lane = (xmy >> 1) mod LPI); a region fits LPI lanes iff its
// span is at most 2 * LPI - 2 (no two live cells of three consecutive diagonals then share a lane, and an out-of-band
// neighbour is always a lane that holds -inf).
#define MAB_SPAN_SHIFT 8
#define MAB_SPAN_MAX 0x7fffff
#ifdef __CUDACC__
#define MAB_HDI __host__ __device__ static inline
#else
#define MAB_HDI static inline
#endif
MAB_HDI int mab_region_span(int32_t flags) { return (int) (((uint32_t) flags >> MAB_SPAN_SHIFT) & MAB_SPAN_MAX); }
// 0: fits a half-warp (16 lanes), 1: fits a warp (32 lanes), 2: wider (shared-memory ring kernels)
MAB_HDI int mab_band_class(int span) { return span <= 30 ? 0 : (span <= 62 ? 1 : 2); }

// packed-symbol arrays: code words, then mask words; base i sits at bit position i + MAB_PACK_PAD_BASES and the same
// pad follows the last base, so that a 128-base window may start anywhere around the sequence
#define MAB_PACK_PAD_BASES 256
MAB_HDI int64_t mab_pack_code_words(int64_t len) { return (len + 2 * MAB_PACK_PAD_BASES + 15) / 16; }
MAB_HDI int64_t mab_pack_mask_words(int64_t len) { return (len + 2 * MAB_PACK_PAD_BASES + 31) / 32; }
MAB_HDI int64_t mab_pack_words(int64_t len) { return mab_pack_code_words(len) + mab_pack_mask_words(len); }

#define MAB_DIAG_NO_MIDDLE 0x80000000u  // forward diagonal xay-2 was already freed when the reference visits xay
#define MAB_DIAG_IS_TOTAL 0x40000000u   // a total probability is evaluated ON this diagonal (index = its own total)
#define MAB_DIAG_TOT_MASK 0x3fffffffu
struct DiagDesc {  // 16 bytes
    int64_t cell_off;   // global index of the first cell of this diagonal
    int32_t xmyL;       // smallest x-y on the diagonal; cell j has x-y = xmyL + 2j
    uint32_t tot_flags; // index of the total-probability value the reference uses on this diagonal | flags
};
static_assert(sizeof(DiagDesc) == 16, "DiagDesc must be 16 bytes");

struct TracebackDesc {  // 32 bytes
    int32_t region;
    int32_t lo;     // tracedBackTo before this traceback
    int32_t hi;     // diagonal the backward pass starts from
    int32_t from;   // tracedBackFrom: posteriors are produced for diagonals (lo, from]
    int32_t at_end;
    int32_t tot_base;  // global index of this traceback's first total (at diagonal `from`)
    int32_t n_tot;
    // Intermediate tracebacks only: the first total (on diagonal `from`) needs THIS traceback's backward match values
    // on diagonal from+1, which lies in the overlap the reference discards and the next traceback recomputes with
    // different values.  They are parked in a side buffer at this cell offset (-1 for the final traceback).
    int32_t edge_off;
};
static_assert(sizeof(TracebackDesc) == 32, "TracebackDesc must be 32 bytes");

#define MAB_THROUGH_NONE (-2)   // d == hi: backward diagonal d+1 does not exist, no "jump over d" term
#define MAB_THROUGH_MATRIX (-1) // take backward diagonal d+1 from the backward matrix
struct TotalDesc {  // 16 bytes
    int32_t region;
    int32_t d;         // diagonal (region-local xay) the total is evaluated on
    int32_t edge_off;  // >= 0: backward diagonal d+1 comes from the side buffer at this offset (d == from < hi)
    int32_t full_off;  // posterior mode: first cell of this diagonal inside the all-states side arrays FT / BT
};
static_assert(sizeof(TotalDesc) == 16, "TotalDesc must be 16 bytes");

// ---------------------------------------------------------------------------------------------- host plan object

struct mab_plan {
    mab_params params;
    BlobHeader hdr;
    std::vector<RegionDesc> regions;
    std::vector<DiagDesc> diags;
    std::vector<TracebackDesc> tbs;
    std::vector<TotalDesc> totals;
    std::vector<uint8_t> sym_x, sym_y;
    std::vector<uint32_t> pack_x, pack_y;  // see BlobHeader::off_pack_x
    std::vector<int32_t> order_fwd, order_bwd, diag_region;
    std::vector<int64_t> aln_region;
    // kept for introspection only
    std::vector<int64_t> anchors;      // x,y interleaved, alignment-local
    std::vector<int64_t> anchor_off;   // [n+1]
    // plans built by mab_plan_build_device: the blob lives in the caller's device arena and the host vectors above
    // are empty (no introspection)
    void *d_blob = nullptr;
};
int mab_params_validate(const mab_params *p);

// workspace carving (device scratch); all offsets 256-byte aligned
struct WorkspaceLayout {
    int64_t off_fwd, off_bwd, off_fwd_tot, off_bwd_tot, off_edge, off_totals, off_pint, off_counts, off_counters;
    int64_t off_scan_tmp, scan_tmp_bytes;  // posterior mode: scratch of the device-wide scan over off_counts
    int64_t total_bytes;
};
WorkspaceLayout mab_workspace_layout(const BlobHeader &h, int S, bool expectations);
