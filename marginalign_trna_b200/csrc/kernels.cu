// kernels.cu -- sm_100a kernels of the anchor-banded pair-HMM forward-backward and
// their launchers (the device half of include/margin_b200.h).
//
// Arithmetic contract (DESIGN.md "numerics"): everything the reference computes in double with its approximate
// logAdd is reproduced in FP64 with the same polynomial, the same branch thresholds and the same accumulation
// order, using explicit round-to-nearest intrinsics (__dadd_rn/__dmul_rn, never contracted into FMA), so forward,
// backward and total-probability values are bit-identical to the reference's; only exp() (CUDA libm, <= 1 ulp)
// can differ from glibc's in the last place.
//
// Decomposition (not the reference's interleaved loop):
//   k_forward    one warp per split region, anti-diagonal wavefront, lanes = cells of the diagonal; the two previous
//                diagonals live in a shared-memory ring; every diagonal is also written to HBM ([cell][S]).
//   k_backward   one warp per *traceback* (checkpoint segment).  The backward pass of a segment does not depend on
//                forward values, so all segments of all reads run concurrently.  The reference's scatter
//                (doTransitionBackward) is turned into a gather that adds the terms in the scatter's order.
//   k_totals     one thread per total-probability evaluation (every 10th diagonal of a traceback): the serial
//                logAdd chain over the cells of a diagonal, in cell order.
//   k_posterior_diag / scan / k_emit_diag   one thread per diagonal: p_int per cell and pair count, device-wide
//                exclusive scan over the diagonals in emission order, ordered write of the pairs.
//   k_expect     per-cell transition/emission expectations with in-block reduction.
#include <cuda_runtime.h>

#include <cub/cub.cuh>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "mab_internal.h"

#define WARPS_PER_BLOCK 4
#define THREADS_PER_BLOCK (WARPS_PER_BLOCK * 32)
#define FULL_MASK 0xffffffffu

// scratch of mab_posteriors_compute_noalloc (see static_ring_address below); MAB_NOALLOC_SCRATCH_BYTES in margin_b200.h
#define MAB_STATIC_RING_BYTES ((size_t) 32 << 20)
static __device__ __align__(256) double g_static_ring[MAB_STATIC_RING_BYTES / sizeof(double)];

namespace {

thread_local int64_t g_launches = 0;

// Optional per-kernel timing with CUDA events on the launching stream (bench.py's roofline numbers).
enum KernelId { K_FORWARD = 0, K_BACKWARD, K_TOTALS, K_POSTERIOR, K_SCAN, K_EMIT, K_EXPECT,
                K_RL_WEIGHTS = MAB_K_RL_WEIGHTS, K_RL_CHAIN = MAB_K_RL_CHAIN, K_RL_SCAN = MAB_K_RL_SCAN,
                K_RL_EMIT = MAB_K_RL_EMIT, K_COL_ACC = MAB_K_COL_ACC, K_COL_CALL = MAB_K_COL_CALL, K_COUNT };
const char *const kKernelNames[K_COUNT] = {"k_forward", "k_backward", "k_totals", "k_posterior_diag", "scan+offsets",
                                           "k_emit_diag", "k_expect", "k_rl_weights", "k_rl_chain", "k_rl_scan",
                                           "k_rl_emit", "k_col_accumulate", "k_col_call"};
struct TimedLaunch {
    cudaEvent_t start, stop;
    int kernel;
};
thread_local bool g_timing = false;
thread_local std::vector<TimedLaunch> g_timed;
thread_local std::vector<cudaEvent_t> g_event_pool;

cudaEvent_t take_event() {
    if (!g_event_pool.empty()) {
        cudaEvent_t e = g_event_pool.back();
        g_event_pool.pop_back();
        return e;
    }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
}

struct LaunchTimer {  // records an event pair around one kernel launch when timing is on
    cudaStream_t st;
    TimedLaunch t;
    bool on;
    LaunchTimer(int kernel, cudaStream_t s) : st(s), on(g_timing) {
        g_launches++;
        if (on) {
            t.kernel = kernel;
            t.start = take_event();
            t.stop = take_event();
            cudaEventRecord(t.start, st);
        }
    }
    ~LaunchTimer() {
        if (on) {
            cudaEventRecord(t.stop, st);
            g_timed.push_back(t);
        }
    }
};

// ------------------------------------------------------------------------------------------------ logAdd

// lookup() of pairwiseAligner.c:235-246.  The literals carry an `f` suffix in the reference: each coefficient is
// the float nearest to the decimal, promoted to double.  Horner steps are separate multiply and add (no FMA).
__device__ __forceinline__ double la_lookup(double x) {
    double c3, c2, c1, c0;
    if (x <= 1.00) {
        c3 = (double) -0.009350833524763f; c2 = (double) 0.130659527668286f;
        c1 = (double) 0.498799810682272f;  c0 = (double) 0.693203116424741f;
    } else if (x <= 2.50) {
        c3 = (double) -0.014532321752540f; c2 = (double) 0.139942324101744f;
        c1 = (double) 0.495635523139337f;  c0 = (double) 0.692140569840976f;
    } else if (x <= 4.50) {
        c3 = (double) -0.004605031767994f; c2 = (double) 0.063427417320019f;
        c1 = (double) 0.695956496475118f;  c0 = (double) 0.514272634594009f;
    } else {
        c3 = (double) -0.000458661602210f; c2 = (double) 0.009695946122598f;
        c1 = (double) 0.930734667215156f;  c0 = (double) 0.168037164329057f;
    }
    double r = __dadd_rn(__dmul_rn(c3, x), c2);
    r = __dadd_rn(__dmul_rn(r, x), c1);
    r = __dadd_rn(__dmul_rn(r, x), c0);
    return r;
}

// logAdd of pairwiseAligner.c:248-252.  With hi = max, lo = min: d = hi - lo is +inf when lo is -inf and NaN when
// both are -inf; in both cases !(d < 7.5) selects hi, which is what the reference's LOG_ZERO tests return.
__device__ __forceinline__ double la(double x, double y) {
    const bool lt = x < y;
    const double hi = lt ? y : x;
    const double lo = lt ? x : y;
    const double d = __dsub_rn(hi, lo);
    if (!(d < 7.5)) return hi;
    return __dadd_rn(la_lookup(d), lo);
}

__device__ __forceinline__ double neg_inf() { return __longlong_as_double(0xfff0000000000000LL); }

template <int S>
struct Cell {
    double v[S];
};

template <int S>
__device__ __forceinline__ Cell<S> cell_fill(double x) {
    Cell<S> c;
#pragma unroll
    for (int s = 0; s < S; s++) c.v[s] = x;
    return c;
}

template <int S>
__device__ __forceinline__ Cell<S> cell_load(const double *p) {
    Cell<S> c;
#pragma unroll
    for (int s = 0; s < S; s++) c.v[s] = p[s];
    return c;
}

template <int S>
__device__ __forceinline__ Cell<S> cell_load_cg(const double *p) {
    Cell<S> c;
#pragma unroll
    for (int s = 0; s < S; s++) c.v[s] = __ldcg(p + s);
    return c;
}

template <int S>
__device__ __forceinline__ void cell_store(double *p, const Cell<S> &c) {
#pragma unroll
    for (int s = 0; s < S; s++) p[s] = c.v[s];
}

enum { M = 0, SX = 1, SY = 2, LX = 3, LY = 4 };

// Transition order of the cells (cell_calculateForward / cell_calculateBackward over stateMachine{5,3}_cellCalculate,
// stateMachine.c:399-429, 636-661).  A missing neighbour is passed as all -inf, which leaves the affected states at -inf
// exactly as skipping the group does; the first term of each chain replaces logAdd(-inf, t) == t.
//   forward 5-state : SX = la(lo.M+eL0, lo.SX+eL1); LX = la(lo.M+eL2, lo.LX+eL3);
//                     M = la(la(la(la(mi.M+eM0, mi.SX+eM1), mi.SY+eM2), mi.LX+eM3), mi.LY+eM4);
//                     SY = la(up.M+eU0, up.SY+eU1); LY = la(up.M+eU2, up.LY+eU3)
//   forward 3-state : SX = la(la(lo.M+eL0, lo.SX+eL1), lo.SY+eL2); M, SY likewise from mi / up
// Backward is a gather.  The reference scatters (doTransitionBackward, pairwiseAligner.c:322-325): while it processes
// diagonal d+2 the target cell T=(d,xmy) receives the match group of (d+2,xmy); while it processes d+1 in ascending xmy,
// T receives the upper group of (d+1,xmy-1) [T is that cell's "upper"] and then the lower group of (d+1,xmy+1) [T is
// its "lower"].  nMi/nUp/nLo are those three cells; the terms are added in exactly that order.

// ------------------------------------------------------------------------------------------------ blob access

struct Blob {
    const BlobHeader *h;
    const RegionDesc *regions;
    const DiagDesc *diags;
    const TracebackDesc *tbs;
    const TotalDesc *totals;
    const uint8_t *sym_x, *sym_y;
    const int32_t *order_fwd, *order_bwd, *diag_region;
    const int64_t *aln_region;
    const uint32_t *pack_x, *pack_y;      // 2-bit codes + N mask (BlobHeader::off_pack_x)
    int64_t h_len_sym_x, h_len_sym_y;
};

__host__ __device__ inline Blob blob_view(const void *base, const BlobHeader &h) {
    const uint8_t *b = (const uint8_t *) base;
    Blob v;
    v.h = (const BlobHeader *) b;
    v.regions = (const RegionDesc *) (b + h.off_regions);
    v.diags = (const DiagDesc *) (b + h.off_diags);
    v.tbs = (const TracebackDesc *) (b + h.off_tracebacks);
    v.totals = (const TotalDesc *) (b + h.off_totals);
    v.sym_x = b + h.off_sym_x;
    v.sym_y = b + h.off_sym_y;
    v.order_fwd = (const int32_t *) (b + h.off_order_fwd);
    v.order_bwd = (const int32_t *) (b + h.off_order_bwd);
    v.aln_region = (const int64_t *) (b + h.off_aln_region);
    v.diag_region = (const int32_t *) (b + h.off_diag_region);
    v.pack_x = (const uint32_t *) (b + h.off_pack_x);
    v.pack_y = (const uint32_t *) (b + h.off_pack_y);
    v.h_len_sym_x = h.len_sym_x;
    v.h_len_sym_y = h.len_sym_y;
    return v;
}

__device__ __forceinline__ int next_item(int *counter, int lane) {
    int item = 0;
    if (lane == 0) item = atomicAdd(counter, 1);
    return __shfl_sync(FULL_MASK, item, 0);
}
// the same for a group of LPI lanes (32 or 16) of a warp; `mask` names the lanes of the group
template <int LPI>
__device__ __forceinline__ int next_item_sub(int *counter, int lane, unsigned mask) {
    int item = 0;
    if ((lane & (LPI - 1)) == 0) item = atomicAdd(counter, 1);
    return __shfl_sync(mask, item, lane & ~(LPI - 1));
}

// ------------------------------------------------------------------------------------------------ forward
// Generic one-lane-per-cell wavefront (any state count; the 3-state path).  The 5-state path uses wavefront5.cuh.
// FULL = true : F holds all S states of every cell ([cell][S]) -- expectation mode.
// FULL = false: posterior mode -- F holds the match state of every cell, FT all states of the cells on diagonals that
//               carry a total probability (TotalDesc::full_off).

#include "wavefront5.cuh"

// 3-state cells with the branch-free logAdd of wavefront5.cuh (coefficients from a replicated shared-memory table):
// the lanes of a (half-)warp sit on different polynomial pieces, and the branchy la() serialises them.
// One-lane-per-cell kernels: coefficient table of the branch-free logAdd of wavefront5.cuh, replicated so that a
// 128-bit load is conflict-free whatever rows the lanes of a quarter-warp index.
// (Measured and dropped, profiles/r1_ncu_summary.md r1h: a 16-byte packed row per piece -- the coefficients are float
// literals, so a word can carry each one exactly -- indexed without compares through RU(d + 2^44 - 2^-8): bit-exact,
// halves the coefficient bytes, but its 8 extra LOP3 per logAdd move the kernel from LSU-bound to issue-bound, 5 %
// slower; mixing the two forms per logAdd was slower still.)
struct CoefTable {
    double2 coef[4][2][W5_COPIES];  // [piece][(c3,c2) | (c1,c0)][replica]
};
__device__ __forceinline__ void fill_coef_table(CoefTable &t, int tid, int nthreads) {
    for (int i = tid; i < 4 * 2 * W5_COPIES; i += nthreads) {
        const int piece = i / (2 * W5_COPIES), half = (i / W5_COPIES) % 2, rep = i % W5_COPIES;
        t.coef[piece][half][rep] = make_double2(c_la_coef[piece * 4 + half * 2], c_la_coef[piece * 4 + half * 2 + 1]);
    }
}
typedef unsigned CoefAddr;  // shared address of coef[0][0][replica of the lane]
__device__ __forceinline__ double la_gap(double x, double y, CoefAddr c) { return la_bf<COEF_SMEM>(x, y, c); }
__device__ __forceinline__ double la_chain(double x, double y, CoefAddr c) { return la_bf<COEF_SMEM>(x, y, c); }
#define LG(x, y) la_gap((x), (y), cf)
#define LC(x, y) la_chain((x), (y), cf)
__device__ __forceinline__ Cell<3> forward_cell3_bf(const Cell<3> &lo, const Cell<3> &mi, const Cell<3> &up,
                                                    const double *eL, const double *eM, const double *eU,
                                                    const CoefAddr &cf) {
    Cell<3> c;
    c.v[SX] = LC(LG(__dadd_rn(lo.v[M], eL[0]), __dadd_rn(lo.v[SX], eL[1])), __dadd_rn(lo.v[SY], eL[2]));
    c.v[M] = LC(LG(__dadd_rn(mi.v[M], eM[0]), __dadd_rn(mi.v[SX], eM[1])), __dadd_rn(mi.v[SY], eM[2]));
    c.v[SY] = LC(LG(__dadd_rn(up.v[M], eU[0]), __dadd_rn(up.v[SY], eU[1])), __dadd_rn(up.v[SX], eU[2]));
    return c;
}
__device__ __forceinline__ Cell<3> backward_cell3_bf(const Cell<3> &nMi, const Cell<3> &nUp, const Cell<3> &nLo,
                                                     const double *eM, const double *eU, const double *eL,
                                                     const CoefAddr &cf) {
    Cell<3> c;
    // lower: M->X(0) X->X(1) Y->X(2); middle: M->M(0) X->M(1) Y->M(2); upper: M->Y(0) Y->Y(1) X->Y(2)
    c.v[M] = LC(LG(__dadd_rn(nMi.v[M], eM[0]), __dadd_rn(nUp.v[SY], eU[0])), __dadd_rn(nLo.v[SX], eL[0]));
    c.v[SX] = LC(LG(__dadd_rn(nMi.v[M], eM[1]), __dadd_rn(nUp.v[SY], eU[2])), __dadd_rn(nLo.v[SX], eL[1]));
    c.v[SY] = LC(LG(__dadd_rn(nMi.v[M], eM[2]), __dadd_rn(nUp.v[SY], eU[1])), __dadd_rn(nLo.v[SX], eL[2]));
    return c;
}
__device__ __forceinline__ Cell<5> forward_cell5_bf(const Cell<5> &lo, const Cell<5> &mi, const Cell<5> &up,
                                                    const double *eL, const double *eM, const double *eU,
                                                    const CoefAddr &cf) {
    Cell<5> c;
    c.v[SX] = LG(__dadd_rn(lo.v[M], eL[0]), __dadd_rn(lo.v[SX], eL[1]));
    c.v[LX] = LG(__dadd_rn(lo.v[M], eL[2]), __dadd_rn(lo.v[LX], eL[3]));
    double m = LC(__dadd_rn(mi.v[M], eM[0]), __dadd_rn(mi.v[SX], eM[1]));
    m = LC(m, __dadd_rn(mi.v[SY], eM[2]));
    m = LC(m, __dadd_rn(mi.v[LX], eM[3]));
    c.v[M] = LC(m, __dadd_rn(mi.v[LY], eM[4]));
    c.v[SY] = LG(__dadd_rn(up.v[M], eU[0]), __dadd_rn(up.v[SY], eU[1]));
    c.v[LY] = LG(__dadd_rn(up.v[M], eU[2]), __dadd_rn(up.v[LY], eU[3]));
    return c;
}
__device__ __forceinline__ Cell<5> backward_cell5_bf(const Cell<5> &nMi, const Cell<5> &nUp, const Cell<5> &nLo,
                                                     const double *eM, const double *eU, const double *eL,
                                                     const CoefAddr &cf) {
    Cell<5> c;
    double m = LC(__dadd_rn(nMi.v[M], eM[0]), __dadd_rn(nUp.v[SY], eU[0]));
    m = LC(m, __dadd_rn(nUp.v[LY], eU[2]));
    m = LC(m, __dadd_rn(nLo.v[SX], eL[0]));
    c.v[M] = LC(m, __dadd_rn(nLo.v[LX], eL[2]));
    c.v[SX] = LG(__dadd_rn(nMi.v[M], eM[1]), __dadd_rn(nLo.v[SX], eL[1]));
    c.v[SY] = LG(__dadd_rn(nMi.v[M], eM[2]), __dadd_rn(nUp.v[SY], eU[1]));
    c.v[LX] = LG(__dadd_rn(nMi.v[M], eM[3]), __dadd_rn(nLo.v[LX], eL[3]));
    c.v[LY] = LG(__dadd_rn(nMi.v[M], eM[4]), __dadd_rn(nUp.v[LY], eU[3]));
    return c;
}
#undef LG
#undef LC
template <int S>
__device__ __forceinline__ Cell<S> forward_cell_sel(const Cell<S> &lo, const Cell<S> &mi, const Cell<S> &up,
                                                    const double *eL, const double *eM, const double *eU,
                                                    const CoefAddr &cf) {
    if constexpr (S == 3) return forward_cell3_bf(lo, mi, up, eL, eM, eU, cf);
    else return forward_cell5_bf(lo, mi, up, eL, eM, eU, cf);  // only the over-wide fallback (global-memory ring)
}
template <int S>
__device__ __forceinline__ Cell<S> backward_cell_sel(const Cell<S> &nMi, const Cell<S> &nUp, const Cell<S> &nLo,
                                                     const double *eM, const double *eU, const double *eL,
                                                     const CoefAddr &cf) {
    if constexpr (S == 3) return backward_cell3_bf(nMi, nUp, nLo, eM, eU, eL, cf);
    else return backward_cell5_bf(nMi, nUp, nLo, eM, eU, eL, cf);
}

#include "wavefront_reg.cuh"

// Ring accesses of the two kernels below.  GR = false: the ring is shared memory.  GR = true: the ring is a slice of a
// global scratch buffer (bands wider than 227 KB of shared memory can hold, up to the widest gap box the split rule
// allows); a warp's lanes exchange cells through it, so the accesses go to L2 (ld.cg / st.cg) and the __syncwarp
// between diagonals orders them.
template <int S, bool GR>
__device__ __forceinline__ Cell<S> ring_load(const double *p) {
    return GR ? cell_load_cg<S>(p) : cell_load<S>(p);
}
template <int S, bool GR>
__device__ __forceinline__ void ring_store(double *p, const Cell<S> &c) {
    if (GR) {
#pragma unroll
        for (int s = 0; s < S; s++) __stcg(p + s, c.v[s]);
    } else {
        cell_store<S>(p, c);
    }
}

template <int S, bool FULL, int LPI, bool GR = false>
__global__ void __launch_bounds__(THREADS_PER_BLOCK)
k_forward(const __grid_constant__ DevModel model, Blob bl, int n_regions, double *__restrict__ F,
          double *__restrict__ FT, int *counter, int ring_w, double *g_ring = nullptr) {
    extern __shared__ __align__(16) unsigned char dyn_smem_raw[];
    __shared__ DevModel sm;
    __shared__ __align__(16) CoefTable ct;
    for (int i = threadIdx.x; i < (int) (sizeof(DevModel) / 4); i += blockDim.x)
        ((int32_t *) &sm)[i] = ((const int32_t *) &model)[i];
    fill_coef_table(ct, threadIdx.x, blockDim.x);
    __syncthreads();
    // LPI lanes (a warp, or a half-warp when the band is at most 16 cells wide: tRNA) share one work item
    const int lane32 = threadIdx.x & 31, lane = lane32 & (LPI - 1);
    const CoefAddr cf = (unsigned) __cvta_generic_to_shared(&ct.coef[0][0][lane32 & (W5_COPIES - 1)]);
    const int slot_id = (threadIdx.x >> 5) * (32 / LPI) + (lane32 / LPI);
    const unsigned gmask = LPI == 32 ? FULL_MASK : (0xffffu << (lane32 & 16));
    double *ring = GR ? g_ring + ((size_t) blockIdx.x * (THREADS_PER_BLOCK / LPI) + slot_id) * 3 * ring_w * S
                      : reinterpret_cast<double *>(dyn_smem_raw) + (size_t) slot_id * 3 * ring_w * S;
    const double NINF = neg_inf();

    for (;;) {
        const int item = next_item_sub<LPI>(counter, lane32, gmask);
        if (item >= n_regions) break;
        const RegionDesc rg = bl.regions[bl.order_fwd[item]];
        const DiagDesc *dd = bl.diags + rg.diag_base;
        const uint8_t *sx = bl.sym_x + rg.sym_x_off, *sy = bl.sym_y + rg.sym_y_off;
        const int D = rg.lX + rg.lY;
        // diagonal 0: dpDiagonal_initialiseValues with (ragged)startStateProb, pairwiseAligner.c:658-659
        int64_t c_prev1 = dd[0].cell_off, c_cur = dd[1].cell_off;
        int L2 = 0, L1 = dd[0].xmyL, w2 = 0, w1 = (int) (c_cur - c_prev1);
        {
            const double *init = (rg.flags & 1) ? sm.ragged_start : sm.start;
            for (int j = lane; j < w1; j += LPI) {
                Cell<S> c;
#pragma unroll
                for (int s = 0; s < S; s++) c.v[s] = init[s];
                if (FULL) cell_store<S>(F + (c_prev1 + j) * S, c);
                else F[c_prev1 + j] = c.v[M];
                ring_store<S, GR>(ring + (size_t) j * S, c);
            }
        }
        __syncwarp(gmask);
        for (int d = 1; d <= D; d++) {
            const DiagDesc cur = dd[d];
            const int L = cur.xmyL;
            const int64_t c_next = dd[d + 1].cell_off;
            const int w = (int) (c_next - c_cur);
            const int oLo = (L - 1 - L1) >> 1;  // index shift to the lower neighbour (d-1, xmy-1)
            const int oMi = (L - L2) >> 1;      // index shift to the middle neighbour (d-2, xmy)
            double *slot = ring + (size_t) (d % 3) * ring_w * S;
            const double *slot1 = ring + (size_t) ((d + 2) % 3) * ring_w * S;
            const double *slot2 = ring + (size_t) ((d + 1) % 3) * ring_w * S;
            double *ft = nullptr;
            if (!FULL && (cur.tot_flags & MAB_DIAG_IS_TOTAL))
                ft = FT + (int64_t) bl.totals[cur.tot_flags & MAB_DIAG_TOT_MASK].full_off * S;
            for (int j = lane; j < w; j += LPI) {
                const int xmy = L + 2 * j;
                const int x = (d + xmy) >> 1, y = (d - xmy) >> 1;
                const int cX = x > 0 ? sx[x - 1] : 4, cY = y > 0 ? sy[y - 1] : 4;
                const int iLo = j + oLo, iUp = iLo + 1, iMi = j + oMi;
                const Cell<S> lo = (iLo >= 0 && iLo < w1) ? ring_load<S, GR>(slot1 + (size_t) iLo * S) : cell_fill<S>(NINF);
                const Cell<S> up = (iUp >= 0 && iUp < w1) ? ring_load<S, GR>(slot1 + (size_t) iUp * S) : cell_fill<S>(NINF);
                const Cell<S> mi = (d >= 2 && iMi >= 0 && iMi < w2) ? ring_load<S, GR>(slot2 + (size_t) iMi * S)
                                                                     : cell_fill<S>(NINF);
                const Cell<S> c = forward_cell_sel<S>(lo, mi, up, sm.etp_lower[cX], sm.etp_middle[cX * 5 + cY],
                                                      sm.etp_upper[cY], cf);
                if (FULL) cell_store<S>(F + (c_cur + j) * S, c);
                else {
                    F[c_cur + j] = c.v[M];
                    if (ft != nullptr) cell_store<S>(ft + (int64_t) j * S, c);
                }
                ring_store<S, GR>(slot + (size_t) j * S, c);
            }
            __syncwarp(gmask);
            c_cur = c_next;
            L2 = L1;
            L1 = L;
            w2 = w1;
            w1 = w;
        }
    }
}

// ------------------------------------------------------------------------------------------------ backward

template <int S, bool FULL, int LPI, bool GR = false>
__global__ void __launch_bounds__(THREADS_PER_BLOCK)
k_backward(const __grid_constant__ DevModel model, Blob bl, int n_tbs, double *__restrict__ B,
           double *__restrict__ BT, double *__restrict__ edge, int *counter, int ring_w, double *g_ring = nullptr) {
    extern __shared__ __align__(16) unsigned char dyn_smem_raw[];
    __shared__ DevModel sm;
    __shared__ __align__(16) CoefTable ct;
    for (int i = threadIdx.x; i < (int) (sizeof(DevModel) / 4); i += blockDim.x)
        ((int32_t *) &sm)[i] = ((const int32_t *) &model)[i];
    fill_coef_table(ct, threadIdx.x, blockDim.x);
    __syncthreads();
    // LPI lanes (a warp, or a half-warp when the band is at most 16 cells wide: tRNA) share one work item
    const int lane32 = threadIdx.x & 31, lane = lane32 & (LPI - 1);
    const CoefAddr cf = (unsigned) __cvta_generic_to_shared(&ct.coef[0][0][lane32 & (W5_COPIES - 1)]);
    const int slot_id = (threadIdx.x >> 5) * (32 / LPI) + (lane32 / LPI);
    const unsigned gmask = LPI == 32 ? FULL_MASK : (0xffffu << (lane32 & 16));
    // The ring is private to the warp: the overlap diagonals (from, hi] that the reference discards must not reach
    // B, where they would collide with the next traceback's cells.
    double *ring = GR ? g_ring + ((size_t) blockIdx.x * (THREADS_PER_BLOCK / LPI) + slot_id) * 3 * ring_w * S
                      : reinterpret_cast<double *>(dyn_smem_raw) + (size_t) slot_id * 3 * ring_w * S;
    const double NINF = neg_inf();

    for (;;) {
        const int item = next_item_sub<LPI>(counter, lane32, gmask);
        if (item >= n_tbs) break;
        const TracebackDesc tb = bl.tbs[bl.order_bwd[item]];
        const RegionDesc rg = bl.regions[tb.region];
        const DiagDesc *dd = bl.diags + rg.diag_base;
        const uint8_t *sx = bl.sym_x + rg.sym_x_off, *sy = bl.sym_y + rg.sym_y_off;
        // diagonal hi: (ragged)endStateProb on every cell, pairwiseAligner.c:680-681
        int L1 = dd[tb.hi].xmyL, L2 = 0;
        int64_t c1 = dd[tb.hi].cell_off;
        int w1 = (int) (dd[tb.hi + 1].cell_off - c1), w2 = 0;
        {
            const double *init = (tb.at_end && (rg.flags & 2)) ? sm.ragged_end : sm.end;
            double *slot = ring + (size_t) (tb.hi % 3) * ring_w * S;
            const bool keep = tb.hi <= tb.from;
            double *bt = nullptr;
            if (!FULL && keep && (dd[tb.hi].tot_flags & MAB_DIAG_IS_TOTAL))
                bt = BT + (int64_t) bl.totals[dd[tb.hi].tot_flags & MAB_DIAG_TOT_MASK].full_off * S;
            for (int j = lane; j < w1; j += LPI) {
                Cell<S> c;
#pragma unroll
                for (int s = 0; s < S; s++) c.v[s] = init[s];
                ring_store<S, GR>(slot + (size_t) j * S, c);
                if (keep) {
                    if (FULL) cell_store<S>(B + (c1 + j) * S, c);
                    else {
                        B[c1 + j] = c.v[M];
                        if (bt != nullptr) cell_store<S>(bt + (int64_t) j * S, c);
                    }
                }
            }
        }
        __syncwarp(gmask);
        for (int d = tb.hi - 1; d > tb.lo; d--) {
            const DiagDesc cur = dd[d];
            const int L = cur.xmyL;
            const int64_t c0 = cur.cell_off;
            const int w = (int) (c1 - c0);
            const int oUp = (L - 1 - L1) >> 1;  // (d+1, xmy-1): the cell whose "upper" is T
            const int oMi = (L - L2) >> 1;      // (d+2, xmy)
            const bool have2 = d + 2 <= tb.hi;
            double *slot = ring + (size_t) (d % 3) * ring_w * S;
            const double *slot1 = ring + (size_t) ((d + 1) % 3) * ring_w * S;
            const double *slot2 = ring + (size_t) ((d + 2) % 3) * ring_w * S;
            const bool keep = d <= tb.from;
            double *bt = nullptr;
            if (!FULL && keep && (cur.tot_flags & MAB_DIAG_IS_TOTAL))
                bt = BT + (int64_t) bl.totals[cur.tot_flags & MAB_DIAG_TOT_MASK].full_off * S;
            for (int j = lane; j < w; j += LPI) {
                const int xmy = L + 2 * j;
                const int x = (d + xmy) >> 1, y = (d - xmy) >> 1;
                // the neighbours consume X[x] and/or Y[y] (0-based); past the end they are out of band anyway
                const int cX = x < rg.lX ? sx[x] : 4, cY = y < rg.lY ? sy[y] : 4;
                const int iUp = j + oUp, iLo = iUp + 1, iMi = j + oMi;
                const Cell<S> nUp = (iUp >= 0 && iUp < w1) ? ring_load<S, GR>(slot1 + (size_t) iUp * S) : cell_fill<S>(NINF);
                const Cell<S> nLo = (iLo >= 0 && iLo < w1) ? ring_load<S, GR>(slot1 + (size_t) iLo * S) : cell_fill<S>(NINF);
                const Cell<S> nMi = (have2 && iMi >= 0 && iMi < w2) ? ring_load<S, GR>(slot2 + (size_t) iMi * S)
                                                                     : cell_fill<S>(NINF);
                const Cell<S> c = backward_cell_sel<S>(nMi, nUp, nLo, sm.etp_middle[cX * 5 + cY], sm.etp_upper[cY],
                                                       sm.etp_lower[cX], cf);
                ring_store<S, GR>(slot + (size_t) j * S, c);
                if (keep) {
                    if (FULL) cell_store<S>(B + (c0 + j) * S, c);
                    else {
                        B[c0 + j] = c.v[M];
                        if (bt != nullptr) cell_store<S>(bt + (int64_t) j * S, c);
                    }
                } else if (d == tb.from + 1) {
                    edge[tb.edge_off + j] = c.v[M];
                }
            }
            __syncwarp(gmask);
            L2 = L1;
            L1 = L;
            w2 = w1;
            w1 = w;
            c1 = c0;
        }
    }
}

// ------------------------------------------------------------------------------------------------ totals

// diagonalCalculationTotalProbability (pairwiseAligner.c:566-583).  The "matches that jump over d" term rebuilds, in
// the reference, a temporary diagonal d+1 holding only the match state reached from forward diagonal d-1; that value
// is by construction identical to the match state of forward diagonal d+1, which we already have.
// Four lanes per total: the per-cell dot products (S adds + S-1 logAdds each, independent) are computed four cells at a
// time, then chained in cell order -- the chain itself stays serial because logAdd is not associative.
// Every lane of the group carries the same chain (no broadcast needed); groups of a warp diverge freely.
template <int S, bool FULL>
__global__ void k_totals(Blob bl, int n_totals, const double *__restrict__ F, const double *__restrict__ B,
                         const double *__restrict__ FT, const double *__restrict__ BT,
                         const double *__restrict__ edge, double *__restrict__ totals, int skip_class_le) {
    const int gt = blockIdx.x * blockDim.x + threadIdx.x;
    const int t = gt >> 2, q = gt & 3;
    if (t >= n_totals) return;
    const unsigned gmask = 0xFu << (threadIdx.x & 28);
    const TotalDesc td = bl.totals[t];
    const RegionDesc &rg = bl.regions[td.region];
    // regions whose totals the fused backward sweep computes itself (k_backward_post_rr); uniform per 4-lane group
    if (mab_band_class(mab_region_span(rg.flags)) <= skip_class_le) return;
    const DiagDesc *dd = bl.diags + rg.diag_base;
    const int64_t c0 = dd[td.d].cell_off, c1 = dd[td.d + 1].cell_off;
    // all states of diagonal d: from the full matrices, or from the side arrays in posterior mode
    const double *f = FULL ? F + c0 * S : FT + (int64_t) td.full_off * S;
    const double *b = FULL ? B + c0 * S : BT + (int64_t) td.full_off * S;
    const int w = (int) (c1 - c0);
    double total = neg_inf();
    for (int base = 0; base < w; base += 4) {
        const int c = base + q;
        double cd = neg_inf();
        if (c < w) {
            const double *fc = f + (int64_t) c * S, *bc = b + (int64_t) c * S;
            cd = __dadd_rn(fc[0], bc[0]);
#pragma unroll
            for (int s = 1; s < S; s++) cd = la(cd, __dadd_rn(fc[s], bc[s]));
        }
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const double v = __shfl_sync(gmask, cd, k, 4);
            if (base + k < w) total = la(total, v);
        }
    }
    if (td.edge_off != MAB_THROUGH_NONE) {
        const int64_t c2 = dd[td.d + 2].cell_off;
        const int w2 = (int) (c2 - c1);
        constexpr int STRIDE = FULL ? S : 1;  // match state of cell c: F[c*S] or the match plane F[c]
        const double *e = td.edge_off >= 0 ? edge + td.edge_off : nullptr;  // see TracebackDesc::edge_off
        double through = neg_inf();
        for (int base = 0; base < w2; base += 4) {
            const int64_t c = c1 + base + q;
            double v = neg_inf();
            if (base + q < w2) v = __dadd_rn(F[c * STRIDE], e != nullptr ? e[base + q] : B[c * STRIDE]);
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const double u = __shfl_sync(gmask, v, k, 4);
                if (base + k < w2) through = la(through, u);
            }
        }
        total = la(total, through);
    }
    if (q == 0) totals[t] = total;
}

// ------------------------------------------------------------------------------------------------ posteriors
// diagonalCalculationPosteriorMatchProbs (pairwiseAligner.c:585-615) and the ordered pair list, as three flat passes
// with one thread per DIAGONAL (a diagonal's cells are contiguous; its owner region comes from diag_region):
//   k_posterior_diag   p_int (or -1) of every cell; number of emitted pairs of the diagonal -> counts[slot]
//   (device-wide exclusive scan of counts, in place)
//   k_pair_offsets     per-alignment CSR offsets; k_emit_diag writes the pairs of the diagonal at counts[slot].
// `slot` orders the diagonals the way the reference's list is ordered: it appends per traceback (ascending), per
// diagonal descending, per cell in ascending x-y (:594-613, :695-733); alignedPairCoordinateCorrectionFn then pops each
// region's list from the tail (:1167-1169), i.e. reverses it.  Net order inside a region: tracebacks descending,
// diagonals ascending, x-y descending; regions (and alignments) ascending.  So slot(d) = diag_base + 1 +
// (D - from_t) + (d - lo_t - 1) for the traceback t that owns d >= 1, and slot(0) = diag_base (never emits).

struct DiagWork {
    int64_t c0;    // first cell
    int w;         // cells
    int d;         // region-local diagonal
    int xmyL;
    int64_t slot;
    uint32_t tot;  // index of the total the reference uses on this diagonal
    int x1, y1;    // region origin inside the alignment
    int cls;       // band class of the owning region
};

__device__ __forceinline__ DiagWork diag_work(const Blob &bl, int64_t g) {
    DiagWork k;
    const RegionDesc &rg = bl.regions[bl.diag_region[g]];
    const DiagDesc cur = bl.diags[g];
    k.c0 = cur.cell_off;
    k.w = (int) (bl.diags[g + 1].cell_off - cur.cell_off);
    k.d = (int) (g - rg.diag_base);
    k.xmyL = cur.xmyL;
    k.tot = cur.tot_flags & MAB_DIAG_TOT_MASK;
    k.x1 = rg.x1;
    k.y1 = rg.y1;
    k.cls = mab_band_class(mab_region_span(rg.flags));
    k.slot = rg.diag_base;
    if (k.d >= 1) {
        const int D = rg.lX + rg.lY;
        int t = rg.tb_base + rg.n_tb - 1;  // the last traceback owns (lo, D]; walk down to the owner of d
        TracebackDesc tb = bl.tbs[t];
        while (k.d <= tb.lo) tb = bl.tbs[--t];
        k.slot = rg.diag_base + 1 + (D - tb.from) + (k.d - tb.lo - 1);
    }
    return k;
}

// One warp per group of 32 consecutive diagonals: lane i owns diagonal g0 + i (its DiagWork, its total, its pair count)
// while the cells of the whole group -- contiguous in memory -- are swept 32 at a time, every lane finding the diagonal
// of its cell among the 32 starts by a shuffle binary search.  Loads of F / B and the store of pint are coalesced
// whatever the band width (a thread per diagonal walked 13 ... 500 cells serially with a stride between lanes).
__global__ void __launch_bounds__(256)
k_posterior_diag(Blob bl, int64_t n_diagonals, const double *__restrict__ F, const double *__restrict__ B,
                 const double *__restrict__ totals, double threshold, double log_reject, int32_t *__restrict__ pint,
                 int64_t *__restrict__ counts, int skip_class_le) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t g0 = warp * 32;
    if (g0 > n_diagonals) return;
    const int64_t g = g0 + lane;
    if (g == n_diagonals) counts[g] = 0;  // sentinel: the scan leaves the total number of pairs here
    if (g0 == n_diagonals) return;        // a warp that holds only the sentinel has no cells to sweep
    const bool mine = g < n_diagonals;
    DiagWork k;
    k.c0 = 0, k.w = 0, k.d = 0, k.xmyL = 0, k.slot = 0, k.tot = 0, k.x1 = 0, k.y1 = 0, k.cls = 0;
    if (mine) k = diag_work(bl, g);
    // diagonals of regions the fused backward sweep has already served (it wrote their p_int and counts): d = 0 marks
    // a diagonal this kernel leaves alone
    const bool served = mine && k.cls <= skip_class_le;
    if (__all_sync(FULL_MASK, served || !mine)) return;
    if (served) k.d = 0;
    const double my_total = (mine && k.d >= 1) ? totals[k.tot] : 0.0;
    const int64_t base = __shfl_sync(FULL_MASK, k.c0, 0);
    const int64_t g_last = min(g0 + 32, n_diagonals);  // one past the last diagonal of the group
    const int64_t end = bl.diags[g_last].cell_off;
    // start of my diagonal relative to the group; lanes past the last diagonal hold "infinity"
    const int my_start = mine ? (int) (k.c0 - base) : 0x7fffffff;
    int count = 0;
    const int n_iter = (int) ((end - base + 31) >> 5);  // uniform: the shuffles need every lane in every round
    for (int it = 0; it < n_iter; it++) {
        const int64_t c = base + (int64_t) it * 32 + lane;
        const bool active = c < end;
        const int rel = active ? (int) (c - base) : 0;
        int q = 0;  // largest lane whose diagonal starts at or before my cell
#pragma unroll
        for (int step = 16; step > 0; step >>= 1) {
            const int t = __shfl_sync(FULL_MASK, my_start, (q + step) & 31);
            if (q + step < 32 && t <= rel) q += step;
        }
        const int d = __shfl_sync(FULL_MASK, k.d, q);
        const int xmyL = __shfl_sync(FULL_MASK, k.xmyL, q);
        const int start = __shfl_sync(FULL_MASK, my_start, q);
        const double total = __shfl_sync(FULL_MASK, my_total, q);
        bool hit = false;
        if (active && d >= 1) {
            const int xmy = xmyL + 2 * (rel - start);
            const int x = (d + xmy) >> 1, y = (d - xmy) >> 1;
            int32_t v = -1;
            if (x > 0 && y > 0) {
                const double s = __dsub_rn(__dadd_rn(F[c], B[c]), total);
                // exp() only where the pair can pass: s < log(threshold) - 0.1 implies exp(s) < threshold
                if (s >= log_reject) {
                    double p = exp(s);
                    if (p >= threshold) {
                        if (p > 1.0) p = 1.0;
                        v = (int32_t) floor(__dmul_rn(p, MAB_PROB_1));
                        hit = true;
                    }
                }
            }
            pint[c] = v;
        }
        // pairs per diagonal: every lane counts the hits that landed on ITS diagonal
        const unsigned hits = __ballot_sync(FULL_MASK, hit);
        if (hits) {
#pragma unroll 1
            for (unsigned m = hits; m; m &= m - 1) {
                const int src = __ffs(m) - 1;
                if (__shfl_sync(FULL_MASK, q, src) == lane) count++;
            }
        }
    }
    if (mine && !served) counts[k.slot] = count;
}

__global__ void k_pair_offsets(Blob bl, int64_t n_aln, int64_t n_regions, int64_t n_diagonals,
                               const int64_t *__restrict__ pos, int64_t *__restrict__ pair_off,
                               int64_t *__restrict__ n_pairs) {
    const int64_t a = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (a > n_aln) return;
    const int64_t r0 = bl.aln_region[a];
    pair_off[a] = r0 < n_regions ? pos[bl.regions[r0].diag_base] : pos[n_diagonals];
    if (a == 0) n_pairs[0] = pos[n_diagonals];
}

template <typename XY>  // int32_t, or int16_t when every sub-sequence of the batch is shorter than 32768 (8 B per pair)
__global__ void __launch_bounds__(256)
k_emit_diag(Blob bl, int64_t n_diagonals, const int32_t *__restrict__ pint, const int64_t *__restrict__ pos,
            XY *__restrict__ out_x, XY *__restrict__ out_y, int32_t *__restrict__ out_p, int64_t capacity) {
    const int64_t g = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n_diagonals) return;
    const DiagWork k = diag_work(bl, g);
    if (k.d < 1) return;
    int64_t at = pos[k.slot];
    if (pos[k.slot + 1] == at) return;  // nothing to emit on this diagonal
    for (int j = k.w - 1; j >= 0; j--) {
        const int32_t v = pint[k.c0 + j];
        if (v >= 0) {
            if (at < capacity) {
                const int xmy = k.xmyL + 2 * j;
                out_x[at] = (XY) (k.x1 + ((k.d + xmy) >> 1) - 1);
                out_y[at] = (XY) (k.y1 + ((k.d - xmy) >> 1) - 1);
                out_p[at] = v;
            }
            at++;
        }
    }
}

// ------------------------------------------------------------------------------------------------ expectations
// diagonalCalculationExpectations + updateExpectations (pairwiseAligner.c:348-368, 617-628).
//
// One warp per region (dynamic queue), but the lanes do not follow diagonals: a region's cells are contiguous in
// memory (diagonal after diagonal), so the warp sweeps them 32 at a time -- every lane busy whatever the band width --
// and each lane finds the diagonal of its cell among the few diagonal starts the chunk touches.  Per cell: 13 (9)
// exp(f_from + b_to + (eP + tP) - total), in the reference's operand order.  Transition sums live in registers;
// emission sums in warp-private shared-memory bins with one column per lane pair [S*16 bins][16], added with plain
// read-modify-writes (lanes 0-15, then 16-31), so there is no atomic and no conflict.  One global atomic per bin per
// warp at the very end.
#define EXP_WARPS 4
#define EXP_COLS 8     // bin columns per warp: lanes with the same (lane % EXP_COLS) take turns
#define EXP_SEGMENT 4096  // cells per work item of k_expect
#define EXP_PREFETCH 64   // cells ahead of the current one whose F / B lines are prefetched (two chunks; 128: no better)

template <int S>
struct TransTable;
template <>
struct TransTable<5> {  // stateMachine5 cellCalculate order (stateMachine.c:399-429): group, from, to
    static constexpr int N = 13, NL = 4, NM = 5;
    __device__ static constexpr int from(int k) {
        constexpr int f[13] = {M, SX, M, LX, M, SX, SY, LX, LY, M, SY, M, LY};
        return f[k];
    }
    __device__ static constexpr int to(int k) {
        constexpr int t[13] = {SX, SX, LX, LX, M, M, M, M, M, SY, SY, LY, LY};
        return t[k];
    }
};
template <>
struct TransTable<3> {  // stateMachine3 cellCalculate order (stateMachine.c:636-661)
    static constexpr int N = 9, NL = 3, NM = 3;
    __device__ static constexpr int from(int k) {
        constexpr int f[9] = {M, SX, SY, M, SX, SY, M, SY, SX};
        return f[k];
    }
    __device__ static constexpr int to(int k) {
        constexpr int t[9] = {SX, SX, SX, M, M, M, SY, SY, SY};
        return t[k];
    }
};

// exp(x) for the expectation kernel: x = log of a posterior transition probability (<= ~0).  Table-driven
// (exp(x) = 2^k 2^(j/64) e^r, |r| <= ln2/128, degree-5 polynomial; max relative error 2.2e-16 against libm over
// [-700, 1], tools/exp_tab_check.c): 10 FP64 operations in a dependent chain of 9 instead of the ~20 of the library
// routine, which is what k_expect waits on.  Below -700 (and for -inf) the result is 0: such terms cannot change a sum
// of probabilities.  The expectation sums are not bit-exact quantities (their order differs from the reference's
// anyway); the posterior kernel, whose p_int must match bit for bit, keeps the library exp.
__device__ __forceinline__ double exp_tab(double x, const double *__restrict__ pow2_64) {
    const double INV = 0x1.71547652b82fep+6, C1 = 0x1.62e42fe000000p-7, C2 = 0x1.f473de6af278fp-36;  // 64/ln2, ln2/64 hi + lo
    const double MAGIC = 6755399441055744.0;  // 1.5 * 2^52: the integer lands in the low word
    const double t = __fma_rn(x, INV, MAGIC);
    const int n = __double2loint(t);
    const double tn = __dsub_rn(t, MAGIC);
    double r = __fma_rn(tn, -C1, x);
    r = __fma_rn(tn, -C2, r);
    double q = __fma_rn(r, 1.0 / 120, 1.0 / 24);
    q = __fma_rn(q, r, 1.0 / 6);
    q = __fma_rn(q, r, 0.5);
    q = __fma_rn(q, r, 1.0);
    const double p = __dmul_rn(q, r);
    const double tj = pow2_64[n & 63];
    const double res = __fma_rn(tj, p, tj);
    const double scaled = __hiloint2double(__double2hiint(res) + ((n >> 6) << 20), __double2loint(res));
    return x < -700.0 ? 0.0 : scaled;
}

template <int S, int NB>
__global__ void __launch_bounds__(EXP_WARPS * 32, NB)
k_expect(const __grid_constant__ DevModel model, Blob bl, int64_t n_cells, int64_t n_diagonals,
         const double *__restrict__ F, const double *__restrict__ B, const double *__restrict__ totals,
         double *__restrict__ out, int *counter) {
    using TT = TransTable<S>;
    constexpr int NT = TT::N, NL = TT::NL, NM = TT::NM;
    constexpr int NE = S * 16, NOUT = S * S + NE + 1;
    __shared__ DevModel sm;
    __shared__ double bins_all[EXP_WARPS][NE][EXP_COLS];
    __shared__ double pow2_64[64];
    if (threadIdx.x < 64) pow2_64[threadIdx.x] = exp2((double) threadIdx.x / 64.0);
    for (int i = threadIdx.x; i < (int) (sizeof(DevModel) / 4); i += blockDim.x)
        ((int32_t *) &sm)[i] = ((const int32_t *) &model)[i];
    for (int i = threadIdx.x; i < EXP_WARPS * NE * EXP_COLS; i += blockDim.x) (&bins_all[0][0][0])[i] = 0.0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double(*bins)[EXP_COLS] = bins_all[warp];
    const int col = lane & (EXP_COLS - 1);
    const double NINF = neg_inf();
    double tacc[NT];
#pragma unroll
    for (int k = 0; k < NT; k++) tacc[k] = 0.0;
    double lik = 0.0;

    // Work item = a segment of EXP_SEGMENT consecutive cells of the global cell numbering (cells are numbered region by
    // region, diagonal by diagonal), so that a batch of a few long reads spreads over the whole GPU as well as a
    // batch of many short ones.  Inside a segment the cells are walked region by region in flat 32-cell chunks.
    const int64_t n_segments = (n_cells + EXP_SEGMENT - 1) / EXP_SEGMENT;
    for (;;) {
        const int item = next_item(counter, lane);
        if (item >= n_segments) break;
        int64_t c0 = (int64_t) item * EXP_SEGMENT;
        const int64_t seg_end = min(n_cells, c0 + EXP_SEGMENT);
        // global diagonal that holds cell c0: the last one whose first cell is <= c0 (warp-uniform binary search)
        int64_t g;
        {
            int64_t lo = 0, hi = n_diagonals;  // diags[n_diagonals] is the sentinel (cell_off = n_cells > c0)
            while (hi - lo > 1) {
                const int64_t mid = (lo + hi) >> 1;
                if (bl.diags[mid].cell_off <= c0) lo = mid;
                else hi = mid;
            }
            g = lo;
        }
        while (c0 < seg_end) {
        const RegionDesc rg = bl.regions[bl.diag_region[g]];
        const DiagDesc *dd = bl.diags + rg.diag_base;
        const uint8_t *sx = bl.sym_x + rg.sym_x_off, *sy = bl.sym_y + rg.sym_y_off;
        const int D = rg.lX + rg.lY;
        const int64_t region_end = dd[D + 1].cell_off;
        const int64_t c_end = min(region_end, seg_end);  // diagonals 1 .. D of this region, cut at the segment end
        int g0 = (int) (g - rg.diag_base);  // diagonal of the first cell of the current chunk
        if (g0 == 0) {  // diagonal 0 has no incoming transitions (the reference starts at diagonal 1, :617-628)
            g0 = 1;
            c0 = dd[1].cell_off;
        }
        g = rg.diag_base + D + 1;  // where the walk continues: diagonal 0 of the next region
        for (; c0 < c_end; c0 += 32) {
            // starts of the diagonals after g0: lane i holds the start of diagonal g0 + 1 + i (or "infinity")
            const int gi = g0 + 1 + lane;
            const int64_t start_i = gi <= D ? dd[gi].cell_off : c_end + 64;
            const int64_t c = c0 + lane;
            int k_adv = 0, n_starts = 0;  // starts <= my cell; starts inside the chunk (uniform)
            bool next_starts_new = false;
            for (int i = 0; i < 32; i++) {
                const int64_t si = __shfl_sync(FULL_MASK, start_i, i);
                if (si > c0 + 31) {  // uniform: no later diagonal starts inside this chunk (starts strictly increase,
                    next_starts_new = si == c0 + 32;  // so this always triggers by i = 31)
                    break;
                }
                k_adv += si <= c ? 1 : 0;
                n_starts++;
            }
            const bool active = c < c_end;
            const int d = g0 + k_adv;
            g0 += n_starts + (next_starts_new ? 1 : 0);  // diagonal of the first cell of the next chunk
            // The backward cells are touched once, front to back, and the forward cells a chunk reads lie one and two
            // diagonals behind them: both streams are sequential, so the lines of the chunk after the next one are
            // requested now and arrive under this chunk's 13 exp chains (the kernel waited on these loads: 3.9
            // long-scoreboard stalls per issue before, 2.5 after; profiles/r2_ncu_summary.md sections 6 and 9).
            if (c + EXP_PREFETCH < c_end) {
                asm volatile("prefetch.global.L1 [%0];" ::"l"(B + (c + EXP_PREFETCH) * S));
                asm volatile("prefetch.global.L1 [%0];" ::"l"(F + (c + EXP_PREFETCH) * S));
            }
            double e_to[S];
#pragma unroll
            for (int s = 0; s < S; s++) e_to[s] = 0.0;
            int row = -1;  // emission row cX*4+cY, or -1: no emission (inactive lane, or a symbol is N, :359)
            if (active) {
                const DiagDesc cur = dd[d], p1 = dd[d - 1];
                const int j = (int) (c - cur.cell_off);
                const int w1 = (int) (cur.cell_off - p1.cell_off);
                const double total = totals[cur.tot_flags & MAB_DIAG_TOT_MASK];
                if (j == 0) lik += total;  // hmmExpectations->likelihood += totalProbability, once per diagonal (:625)
                const bool mid_ok = d >= 2 && !(cur.tot_flags & MAB_DIAG_NO_MIDDLE);
                int64_t c2 = 0;
                int L2 = 0, w2 = 0;
                if (d >= 2) {
                    const DiagDesc p2 = dd[d - 2];
                    c2 = p2.cell_off;
                    L2 = p2.xmyL;
                    w2 = (int) (p1.cell_off - p2.cell_off);
                }
                const int xmy = cur.xmyL + 2 * j;
                const int x = (d + xmy) >> 1, y = (d - xmy) >> 1;
                const int cX = x > 0 ? sx[x - 1] : 4, cY = y > 0 ? sy[y - 1] : 4;
                if (cX < 4 && cY < 4) row = cX * 4 + cY;
                const int iLo = j + ((cur.xmyL - 1 - p1.xmyL) >> 1), iUp = iLo + 1, iMi = j + ((cur.xmyL - L2) >> 1);
                const Cell<S> b = cell_load<S>(B + c * S);
                const Cell<S> lo = (iLo >= 0 && iLo < w1) ? cell_load<S>(F + (p1.cell_off + iLo) * S) : cell_fill<S>(NINF);
                const Cell<S> up = (iUp >= 0 && iUp < w1) ? cell_load<S>(F + (p1.cell_off + iUp) * S) : cell_fill<S>(NINF);
                const Cell<S> mi = (mid_ok && iMi >= 0 && iMi < w2) ? cell_load<S>(F + (c2 + iMi) * S) : cell_fill<S>(NINF);
                const double *eL = sm.etp_lower[cX], *eM = sm.etp_middle[cX * 5 + cY], *eU = sm.etp_upper[cY];
#pragma unroll
                for (int k = 0; k < NT; k++) {
                    const double f = k < NL ? lo.v[TT::from(k)] : (k < NL + NM ? mi.v[TT::from(k)] : up.v[TT::from(k)]);
                    const double e = k < NL ? eL[k] : (k < NL + NM ? eM[k - NL] : eU[k - NL - NM]);
                    // exp(fromCells[from] + toCells[to] + (eP + tP) - totalProbability), left to right (:356)
                    const double p = exp_tab(__dsub_rn(__dadd_rn(__dadd_rn(f, b.v[TT::to(k)]), e), total), pow2_64);
                    tacc[k] += p;
                    e_to[TT::to(k)] += p;
                }
            }
            // emission bins: the lanes that share a column take turns (conflict-free: one lane per column per turn)
#pragma unroll
            for (int q = 0; q < 32 / EXP_COLS; q++) {
                if (row >= 0 && lane / EXP_COLS == q) {
#pragma unroll
                    for (int s = 0; s < S; s++) bins[s * 16 + row][col] += e_to[s];
                }
                __syncwarp();
            }
        }
        if (c0 >= region_end) c0 = region_end;  // region finished: go on with the next one (else the segment ended)
        }
    }
    // per-warp reduction: transitions and likelihood by shuffles, emission bins over their 16 columns
#pragma unroll
    for (int k = 0; k < NT; k++) {
        double v = tacc[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL_MASK, v, o);
        if (lane == 0 && v != 0.0) atomicAdd(&out[TT::from(k) * S + TT::to(k)], v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) lik += __shfl_xor_sync(FULL_MASK, lik, o);
    if (lane == 0 && lik != 0.0) atomicAdd(&out[NOUT - 1], lik);
    __syncwarp();
    for (int i = lane; i < NE; i += 32) {
        double v = 0.0;
#pragma unroll
        for (int q = 0; q < EXP_COLS; q++) v += bins[i][q];
        if (v != 0.0) atomicAdd(&out[S * S + i], v);
    }
}

// ------------------------------------------------------------------------------------------------ launch helpers

int check(cudaError_t e, const char *what) {
    if (e != cudaSuccess) {
        mab_set_error("CUDA error in %s: %s", what, cudaGetErrorString(e));
        return MAB_E_CUDA;
    }
    return MAB_OK;
}

struct DeviceInfo {
    int sm_count = 0;
    int max_smem_optin = 0;
    bool ok = false;
};

const DeviceInfo &device_info() {
    static thread_local DeviceInfo info;
    static thread_local int cached_dev = -1;
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) {
        info.ok = false;
        return info;
    }
    if (dev != cached_dev) {
        cudaDeviceProp prop;
        info.ok = cudaGetDeviceProperties(&prop, dev) == cudaSuccess;
        info.sm_count = prop.multiProcessorCount;
        info.max_smem_optin = (int) prop.sharedMemPerBlockOptin;
        cached_dev = dev;
    }
    return info;
}

struct Workspace {
    double *fwd, *bwd, *fwd_tot, *bwd_tot, *edge, *totals;
    int32_t *pint;
    int64_t *counts;
    void *scan_tmp;
    int64_t scan_tmp_bytes;
    int *counters;
};

int carve(const mab_plan *plan, void *d_ws, int64_t ws_bytes, int S, bool expectations, Workspace *w) {
    const WorkspaceLayout lay = mab_workspace_layout(plan->hdr, S, expectations);
    if (ws_bytes < lay.total_bytes) {
        mab_set_error("workspace too small: need %lld bytes, got %lld", (long long) lay.total_bytes, (long long) ws_bytes);
        return MAB_E_CAPACITY;
    }
    if (((uintptr_t) d_ws & 255) != 0) {
        mab_set_error("workspace must be 256-byte aligned");
        return MAB_E_ARG;
    }
    uint8_t *b = (uint8_t *) d_ws;
    w->counters = (int *) (b + lay.off_counters);
    w->fwd = (double *) (b + lay.off_fwd);
    w->bwd = (double *) (b + lay.off_bwd);
    w->fwd_tot = (double *) (b + lay.off_fwd_tot);
    w->bwd_tot = (double *) (b + lay.off_bwd_tot);
    w->edge = (double *) (b + lay.off_edge);
    w->totals = (double *) (b + lay.off_totals);
    w->pint = (int32_t *) (b + lay.off_pint);
    w->counts = (int64_t *) (b + lay.off_counts);
    w->scan_tmp = b + lay.off_scan_tmp;
    w->scan_tmp_bytes = lay.scan_tmp_bytes;
    return MAB_OK;
}

int grid_for(int64_t items, int blocks_per_sm, int warps_per_block = WARPS_PER_BLOCK) {
    const DeviceInfo &di = device_info();
    const int64_t want = (items + warps_per_block - 1) / warps_per_block;
    const int64_t cap = (int64_t) di.sm_count * blocks_per_sm;
    return (int) (want < 1 ? 1 : (want < cap ? want : cap));
}

// warps per 5-state work item: 0 = chosen from the mean band width (run_dp), 1/2/4/8 = forced (tests, profiling)
int g_w5_group = 0;

template <typename K>
int allow_smem(K kernel, int64_t bytes) {
    return check(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) bytes), "smem attribute");
}

// Launch shape of a wavefront kernel.  G warps share one work item and its shared-memory ring (wavefront5.cuh); the
// choice maximises the number of busy warps: few items with wide diagonals want a large G (expansion 50, forward:
// 333 regions x 4 chunks of 16 cells), many items want G = 1 unless the ring is so big that only a few one-warp
// items fit an SM (expansion 200).  Measured on 10 kb reads (profiles/r1_ncu_summary.md, r1c).
struct WaveLaunch {
    int G, groups, warps, blocks_per_sm;
    int64_t ring_bytes;
};
WaveLaunch wave_launch_for(int S, int G, const BlobHeader &h, bool wave5 = true) {
    wave5 = wave5 && S == 5;
    const int w5_warps = 8;
    const DeviceInfo &di = device_info();
    WaveLaunch l;
    l.G = G;
    const int want_warps = wave5 ? std::max(w5_warps / G, 1) * G : WARPS_PER_BLOCK;
    const int64_t static_bytes = wave5 ? (int64_t) sizeof(Wave5Shared) + 1024 : 10 * 1024;
    const int64_t ring_per_group = (int64_t) 3 * (h.max_width + 2) * S * 8;
    const int64_t budget = (int64_t) di.max_smem_optin - static_bytes;
    l.groups = (int) std::min<int64_t>(want_warps / G, budget / std::max<int64_t>(1, ring_per_group));
    if (G > 1) l.groups = std::min(l.groups, 15);  // one named barrier (1..15) per group
    l.warps = l.groups * G;
    l.ring_bytes = ring_per_group * l.groups;
    l.blocks_per_sm = l.warps < 1 ? 0 : (int) std::max<int64_t>(
        1, std::min<int64_t>(64 / l.warps, (220 * 1024) / (l.ring_bytes + static_bytes + 1024)));
    return l;
}
WaveLaunch wave_launch(int S, int64_t items, const BlobHeader &h) {
    if (S != 5) return wave_launch_for(S, 1, h);
    if (g_w5_group == 1 || g_w5_group == 2 || g_w5_group == 4 || g_w5_group == 8) return wave_launch_for(S, g_w5_group, h);
    const double mean_w = h.n_diagonals > 0 ? (double) h.n_cells / (double) h.n_diagonals : 1.0;
    const int chunks = std::max(1, (int) std::ceil(mean_w / 16.0));
    const double saturated = 24.0 * device_info().sm_count;  // warps beyond this add no throughput
    WaveLaunch best = wave_launch_for(S, 1, h);
    double best_score = -1.0;
    for (int G = 1; G <= 8; G *= 2) {
        const WaveLaunch l = wave_launch_for(S, G, h);
        if (l.warps < 1) continue;
        const double resident = (double) device_info().sm_count * l.blocks_per_sm * l.groups;
        const double busy = std::min<double>((double) items, resident) * std::min(G, chunks);
        const double rounds = std::ceil((double) chunks / G) * G / chunks;  // idle warps in the last round of chunks
        const double score = std::min(busy, saturated) / (rounds * (G > 1 ? 1.1 : 1.0));
        if (score > best_score * 1.0001) best = l, best_score = score;
    }
    return best;
}

// test hook (mab_debug_set_min_band_class): 1 runs class-0 items on the warp-per-item kernel, 2 everything on the
// shared-memory-ring kernels; results must not depend on it
int g_min_band_class = 0;
// mab_debug_set_fused_posterior: 1 runs posterior mode through the fused sweep (backward + totals + posteriors in one
// kernel, k_backward_post_rr) for band classes 0 / 1; 0 (default) through the separate kernels.  Results are
// bit-identical (tests/test_gpu_parity.py).  Measured on 1 M tRNA reads (profiles/r2_ncu_summary.md): fused sweep
// 92.3 ms against 52.4 (backward) + 8.8 (k_totals) + 16.1 (k_posterior_diag) = 77.3 ms -- the serial total-probability
// chains and the exp() of the ~7 % candidate cells cost more issue slots inside the latency-bound sweep (554 against
// 330 warp instructions per diagonal) than the two streaming kernels they replace, so the separate path stays default.
int g_fuse_posterior = 0;
#define RR_LA_DEFAULT 1    // |x - y| logAdd (wavefront_reg.cuh); measured against la_bf and four other coefficient
#define RR_MINB_POST 6     // fused sweep: 92.3 ms per 1 M tRNA reads at 6 blocks per SM, 95.9 at 5, 99.6 at 4
#ifndef RR_MINB_DEFAULT
#define RR_MINB_DEFAULT 5  // fetches (profiles/r2_ncu_summary.md); 4 ... 7 resident blocks per SM are within 2 % of each other
#endif

// register-resident kernels (wavefront_reg.cuh): resident blocks per SM from the occupancy calculator, once per kernel
template <typename K>
int rr_grid(K kernel, int64_t items, int items_per_block) {
    int blocks_per_sm = 0;  // (K is a function-pointer type shared by every instantiation: no static cache here)
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, kernel, RR_THREADS, 0) != cudaSuccess ||
        blocks_per_sm < 1)
        blocks_per_sm = 1;
    const int64_t want = (items + items_per_block - 1) / items_per_block;
    const int64_t cap = (int64_t) device_info().sm_count * blocks_per_sm;
    return (int) std::max<int64_t>(1, std::min(want, cap));
}

// ---- mab_posteriors_compute_noalloc: over-wide ring scratch that is never allocated ------------------------------
// The one place this library allocates device memory on its own is the ring scratch of the over-wide kernels
// (k_forward / k_backward with GR = true, below).  Callers that cannot allocate (CUDA graph capture, serving loops,
// harnesses that forbid cudaMalloc) get the same kernels with the ring in this file-scope array instead: 32 MiB, held
// by every GPU that loads the library until the process exits.  In host code `g_static_ring` is only a stand-in for the
// symbol, not a device address: static_ring_address() asks the runtime once per device (with that device current) and
// that pointer is what the kernels receive (the array itself is defined at file scope, top of this file).  The kernels initialise every ring cell they read (the same code runs on
// uninitialised cudaMallocAsync memory), so nothing has to be cleared between calls; all calls on one GPU share the
// array, so calls that overlap on the GPU (two streams) race.
#define MAB_MAX_DEVICES 64
double *g_static_ring_addr[MAB_MAX_DEVICES];  // per device, from cudaGetSymbolAddress
std::mutex g_static_ring_mutex;

int static_ring_address(double **out) {
    int dev = -1, rc;
    if ((rc = check(cudaGetDevice(&dev), "cudaGetDevice"))) return rc;
    if (dev < 0 || dev >= MAB_MAX_DEVICES) {
        mab_set_error("static ring scratch: device ordinal %d outside 0..%d", dev, MAB_MAX_DEVICES - 1);
        return MAB_E_ARG;
    }
    std::lock_guard<std::mutex> lock(g_static_ring_mutex);
    if (!g_static_ring_addr[dev]) {
        void *ptr = nullptr;
        if ((rc = check(cudaGetSymbolAddress(&ptr, g_static_ring), "cudaGetSymbolAddress(static ring scratch)")))
            return rc;
        g_static_ring_addr[dev] = static_cast<double *>(ptr);
    }
    *out = g_static_ring_addr[dev];
    return MAB_OK;
}

// set by mab_posteriors_compute_noalloc for the duration of its call (per host thread): run_dp takes the over-wide
// ring from here instead of cudaMallocAsync
thread_local double *t_ext_ring = nullptr;
thread_local size_t t_ext_ring_bytes = 0;

// forward + backward + totals: shared by posterior mode (FULL = false) and expectation mode (FULL = true).
// The order arrays list the work items by band class (mab_internal.h): class 2 (wide bands) first -- they run on the
// shared-memory-ring kernels, G warps per item -- then class 1 (a warp per item) and class 0 (a half-warp per item) on
// the register-resident kernels.
// posterior mode: what the fused backward sweep (k_backward_post_rr) needs besides the DP buffers
struct PostArgs {
    double threshold, log_reject;
};

template <int S, bool FULL>
int run_dp(const DevModel &dm, const mab_plan *plan, const Blob &bl, const Workspace &w, cudaStream_t st,
           const PostArgs *post = nullptr, bool *unfused_items = nullptr) {
    const BlobHeader &h = plan->hdr;
    int nR2 = (int) h.n_class_regions[2], nR1 = (int) h.n_class_regions[1], nR0 = (int) h.n_class_regions[0];
    int nT2 = (int) h.n_class_tbs[2], nT1 = (int) h.n_class_tbs[1], nT0 = (int) h.n_class_tbs[0];
    if (g_min_band_class >= 1) nR1 += nR0, nT1 += nT0, nR0 = nT0 = 0;
    if (g_min_band_class >= 2) nR2 += nR1, nT2 += nT1, nR1 = nT1 = 0;
    if ((int64_t) nR2 + nR1 + nR0 != h.n_regions || (int64_t) nT2 + nT1 + nT0 != h.n_tracebacks) {
        mab_set_error("plan header: band classes do not add up (%d + %d + %d regions of %lld)", nR2, nR1, nR0,
                      (long long) h.n_regions);
        return MAB_E_ARG;
    }
    int rc;
    if ((rc = check(cudaMemsetAsync(w.counters, 0, 256, st), "memset counters"))) return rc;
    if (nR2 > 0) {
        // shared-memory ring per work item: 3 diagonals x (max_width + 2 guard) cells x S doubles.  The 5-state kernels
        // also hold ~27 KB of replicated tables per block and may give one item to G warps (wave_launch).
        const bool wave5 = S == 5;
        const WaveLaunch lf = wave5 ? wave_launch(S, nR2, h) : wave_launch_for(S, 1, h, false),
                         lb = wave5 ? wave_launch(S, nT2, h) : wave_launch_for(S, 1, h, false);
        const int ring_w = (int) h.max_width;
        if (lf.warps < 1 || lb.warps < 1) {
            // A diagonal wider than a shared-memory ring can hold (~1850 cells for 5 states): the reference handles any
            // gap box up to splitMatrixBiggerThanThis (3000 x 3000 at marginAlign's realign call site).  The class-2
            // items of such a batch run on the one-lane ring kernels with the ring in a global scratch buffer (L2):
            // slow, correct, and only the batches that contain such a box pay.
            int gf = grid_for(nR2, 2), gb = grid_for(nT2, 2);
            const size_t block_bytes = (size_t) 3 * ring_w * S * sizeof(double) * WARPS_PER_BLOCK;
            const size_t bytes = block_bytes * std::max(gf, gb);
            double *g_ring = nullptr;
            const bool ext_ring = t_ext_ring != nullptr;
            if (ext_ring) {
                // no-allocation variant: as many blocks as the static scratch has rings for (the work queue hands
                // every block item after item, so a smaller grid only takes longer)
                const size_t fit = t_ext_ring_bytes / block_bytes;
                if (fit < 1) {
                    mab_set_error("mab_posteriors_compute_noalloc: a band of %d cells needs %zu bytes of ring scratch per "
                                  "block, the static scratch holds %zu (widest band it can take: %zu cells for %d states); "
                                  "use mab_posteriors_compute",
                                  ring_w, block_bytes, t_ext_ring_bytes,
                                  t_ext_ring_bytes / ((size_t) 3 * S * sizeof(double) * WARPS_PER_BLOCK), S);
                    return MAB_E_CAPACITY;
                }
                gf = (int) std::min<size_t>(gf, fit), gb = (int) std::min<size_t>(gb, fit);
                g_ring = t_ext_ring;
            } else if ((rc = check(cudaMallocAsync((void **) &g_ring, bytes, st), "over-wide ring scratch")))
                return rc;
            {
                LaunchTimer lt(K_FORWARD, st);
                k_forward<S, FULL, 32, true><<<gf, THREADS_PER_BLOCK, 0, st>>>(dm, bl, nR2, w.fwd, w.fwd_tot,
                                                                                w.counters + 0, ring_w, g_ring);
            }
            {
                LaunchTimer lt(K_BACKWARD, st);
                k_backward<S, FULL, 32, true><<<gb, THREADS_PER_BLOCK, 0, st>>>(dm, bl, nT2, w.bwd, w.bwd_tot, w.edge,
                                                                                 w.counters + 1, ring_w, g_ring);
            }
            if (!ext_ring && (rc = check(cudaFreeAsync(g_ring, st), "over-wide ring scratch"))) return rc;
        } else if constexpr (S == 5) {
            auto kf = k_forward5<FULL, COEF_SMEM, 1>;
            auto kb = k_backward5<FULL, COEF_SMEM, 1>;
            if (lf.G == 2) kf = k_forward5<FULL, COEF_SMEM, 2>;
            if (lf.G == 4) kf = k_forward5<FULL, COEF_SMEM, 4>;
            if (lf.G == 8) kf = k_forward5<FULL, COEF_SMEM, 8>;
            if (lb.G == 2) kb = k_backward5<FULL, COEF_SMEM, 2>;
            if (lb.G == 4) kb = k_backward5<FULL, COEF_SMEM, 4>;
            if (lb.G == 8) kb = k_backward5<FULL, COEF_SMEM, 8>;
            if ((rc = allow_smem(kf, lf.ring_bytes)) || (rc = allow_smem(kb, lb.ring_bytes))) return rc;
            {
                LaunchTimer lt(K_FORWARD, st);
                kf<<<grid_for(nR2, lf.blocks_per_sm, lf.groups), lf.warps * 32, lf.ring_bytes, st>>>(
                    dm, bl, nR2, w.fwd, w.fwd_tot, w.counters + 0, ring_w);
            }
            {
                LaunchTimer lt(K_BACKWARD, st);
                kb<<<grid_for(nT2, lb.blocks_per_sm, lb.groups), lb.warps * 32, lb.ring_bytes, st>>>(
                    dm, bl, nT2, w.bwd, w.bwd_tot, w.edge, w.counters + 1, ring_w);
            }
        } else {
            auto kf = k_forward<S, FULL, 32>;
            auto kb = k_backward<S, FULL, 32>;
            if ((rc = allow_smem(kf, lf.ring_bytes)) || (rc = allow_smem(kb, lb.ring_bytes))) return rc;
            {
                LaunchTimer lt(K_FORWARD, st);
                kf<<<grid_for(nR2, lf.blocks_per_sm, lf.warps), lf.warps * 32, lf.ring_bytes, st>>>(
                    dm, bl, nR2, w.fwd, w.fwd_tot, w.counters + 0, ring_w, nullptr);
            }
            {
                LaunchTimer lt(K_BACKWARD, st);
                kb<<<grid_for(nT2, lb.blocks_per_sm, lb.warps), lb.warps * 32, lb.ring_bytes, st>>>(
                    dm, bl, nT2, w.bwd, w.bwd_tot, w.edge, w.counters + 1, ring_w, nullptr);
            }
        }
    }
    constexpr int WPB = RR_THREADS / 32;
    constexpr int LA = RR_LA_DEFAULT, MB = RR_MINB_DEFAULT;
    if (nR1 > 0) {
        LaunchTimer lt(K_FORWARD, st);
        auto k = k_forward_rr<S, FULL, 32, LA, MB>;
        k<<<rr_grid(k, nR1, WPB), RR_THREADS, 0, st>>>(dm, bl, nR2, nR2 + nR1, w.fwd, w.fwd_tot, w.counters + 8);
    }
    if (nR0 > 0) {
        LaunchTimer lt(K_FORWARD, st);
        auto k = k_forward_rr<S, FULL, 16, LA, MB>;
        k<<<rr_grid(k, nR0, 2 * WPB), RR_THREADS, 0, st>>>(dm, bl, nR2 + nR1, nR2 + nR1 + nR0, w.fwd, w.fwd_tot,
                                                          w.counters + 9);
    }
    // Posterior mode: classes 0 / 1 run the fused sweep (backward + totals + posteriors, wavefront_reg.cuh); only
    // class-2 items still go through B / BT / edge, k_totals and k_posterior_diag.
    const bool fused = !FULL && post != nullptr;
    if (unfused_items) *unfused_items = !fused || nT2 > 0;
    if constexpr (!FULL) {
        if (fused) {
            if (nT1 > 0) {
                LaunchTimer lt(K_BACKWARD, st);
                auto k = k_backward_post_rr<S, 32, LA, RR_MINB_POST>;
                k<<<rr_grid(k, nT1, WPB), RR_THREADS, 0, st>>>(dm, bl, nT2, nT2 + nT1, w.fwd, w.fwd_tot, post->threshold,
                                                               post->log_reject, w.pint, w.counts, w.counters + 10);
            }
            if (nT0 > 0) {
                LaunchTimer lt(K_BACKWARD, st);
                auto k = k_backward_post_rr<S, 16, LA, RR_MINB_POST>;
                k<<<rr_grid(k, nT0, 2 * WPB), RR_THREADS, 0, st>>>(dm, bl, nT2 + nT1, nT2 + nT1 + nT0, w.fwd, w.fwd_tot,
                                                                   post->threshold, post->log_reject, w.pint, w.counts,
                                                                   w.counters + 11);
            }
        }
    }
    if (!fused) {
        if (nT1 > 0) {
            LaunchTimer lt(K_BACKWARD, st);
            auto k = k_backward_rr<S, FULL, 32, LA, MB>;
            k<<<rr_grid(k, nT1, WPB), RR_THREADS, 0, st>>>(dm, bl, nT2, nT2 + nT1, w.bwd, w.bwd_tot, w.edge, w.counters + 10);
        }
        if (nT0 > 0) {
            LaunchTimer lt(K_BACKWARD, st);
            auto k = k_backward_rr<S, FULL, 16, LA, MB>;
            k<<<rr_grid(k, nT0, 2 * WPB), RR_THREADS, 0, st>>>(dm, bl, nT2 + nT1, nT2 + nT1 + nT0, w.bwd, w.bwd_tot, w.edge,
                                                              w.counters + 11);
        }
    }
    if (h.n_totals > 0 && (!fused || nT2 > 0)) {
        LaunchTimer lt(K_TOTALS, st);
        k_totals<S, FULL><<<(int) ((4 * h.n_totals + 127) / 128), 128, 0, st>>>(
            bl, (int) h.n_totals, w.fwd, w.bwd, w.fwd_tot, w.bwd_tot, w.edge, w.totals,
            fused && g_min_band_class < 2 ? 1 : -1);
    }
    return check(cudaGetLastError(), "dp launch");
}

template <int S>
int run_posteriors(const DevModel &dm, const mab_plan *plan, const void *d_blob, const Workspace &w,
                   int64_t *d_pair_off, int64_t *d_n_pairs, cudaStream_t st) {
    const BlobHeader &h = plan->hdr;
    const Blob bl = blob_view(d_blob, h);
    const int64_t nD = h.n_diagonals;
    const double thr = plan->params.threshold;
    // exp() is skipped where it cannot reach the threshold (log(0) = -inf: never skipped)
    const PostArgs post{thr, thr > 0.0 ? std::log(thr) - 0.1 : -INFINITY};
    const bool fuse = g_fuse_posterior != 0;
    int rc;
    // pair counts per emission slot: the fused sweep writes the slots of its kept diagonals, k_posterior_diag those of
    // the others; diagonal 0 of every region and the sentinel stay 0
    if (fuse && (rc = check(cudaMemsetAsync(w.counts, 0, (size_t) (nD + 1) * sizeof(int64_t), st), "memset counts")))
        return rc;
    bool unfused_items = true;
    if ((rc = run_dp<S, false>(dm, plan, bl, w, st, fuse ? &post : nullptr, &unfused_items))) return rc;
    if (unfused_items) {
        LaunchTimer lt(K_POSTERIOR, st);
        k_posterior_diag<<<(unsigned) ((nD + 1 + 255) / 256), 256, 0, st>>>(bl, nD, w.fwd, w.bwd, w.totals, post.threshold,
                                                                          post.log_reject, w.pint, w.counts,
                                                                          fuse && g_min_band_class < 2 ? 1 : -1);  // a warp per 32 diagonals
    }
    {
        LaunchTimer lt(K_SCAN, st);
        size_t need = 0;
        cub::DeviceScan::ExclusiveSum(nullptr, need, w.counts, w.counts, nD + 1, st);
        if ((int64_t) need > w.scan_tmp_bytes) {
            mab_set_error("scan scratch too small: need %lld bytes, reserved %lld", (long long) need,
                          (long long) w.scan_tmp_bytes);
            return MAB_E_CAPACITY;
        }
        cub::DeviceScan::ExclusiveSum(w.scan_tmp, need, w.counts, w.counts, nD + 1, st);
        k_pair_offsets<<<(unsigned) ((h.n_alignments + 1 + 255) / 256), 256, 0, st>>>(
            bl, h.n_alignments, h.n_regions, nD, w.counts, d_pair_off, d_n_pairs);
    }
    return check(cudaGetLastError(), "posterior launch");
}

template <int S>
int run_expectations(const DevModel &dm, const mab_plan *plan, const void *d_blob, const Workspace &w, double *d_out,
                     cudaStream_t st) {
    const BlobHeader &h = plan->hdr;
    const Blob bl = blob_view(d_blob, h);
    int rc = run_dp<S, true>(dm, plan, bl, w, st);
    if (rc) return rc;
    {
        LaunchTimer lt(K_EXPECT, st);
        const int64_t n_segments = (h.n_cells + EXP_SEGMENT - 1) / EXP_SEGMENT;
        // resident blocks per SM (= the register budget the kernel is compiled for): measured on 300 k tRNA reads, 4 / 5 / 6
        // blocks give 38.4 / 37.9 / 40.4 ms for S = 5 (10 kb reads: 80.9 / 83.0 / 89.6) and 30.5 / 27.5 / 28.8 ms for S = 3
        // resident blocks per SM (= the register budget the kernel is compiled for).  Measured with the table-driven exp
        // on 300 k tRNA reads, 4 / 5 / 6 blocks: S = 5: 29.2 / 26.5 / 27.1 ms (10 kb reads 67.4 / 63.1 / 65.0), S = 3:
        // 24.7 / 23.2 / 18.5 ms (80 registers without spills)
        constexpr int NB = S == 3 ? 6 : 5;
        k_expect<S, NB><<<grid_for(n_segments, NB, EXP_WARPS), EXP_WARPS * 32, 0, st>>>(
            dm, bl, h.n_cells, h.n_diagonals, w.fwd, w.bwd, w.totals, d_out, w.counters + 4);
    }
    return check(cudaGetLastError(), "expectation launch");
}

int require_gpu() {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
        cudaGetLastError();
        mab_set_error("no CUDA device: libmargin_b200 has no CPU fallback");
        return MAB_E_NOGPU;
    }
    return MAB_OK;
}

}  // namespace

extern "C" int mab_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

extern "C" int mab_kernel_timing(int enable) {
    g_timing = enable != 0;
    return MAB_OK;
}

extern "C" int mab_debug_set_warp_group(int warps) {
    if (warps != 0 && warps != 1 && warps != 2 && warps != 4 && warps != 8) {
        mab_set_error("warp group must be 0 (auto), 1, 2, 4 or 8");
        return MAB_E_ARG;
    }
    g_w5_group = warps;
    return MAB_OK;
}

extern "C" int mab_debug_set_min_band_class(int cls) {
    if (cls < 0 || cls > 2) {
        mab_set_error("band class must be 0, 1 or 2");
        return MAB_E_ARG;
    }
    g_min_band_class = cls;
    return MAB_OK;
}

extern "C" int mab_debug_set_fused_posterior(int on) {
    g_fuse_posterior = on != 0;
    return MAB_OK;
}

extern "C" int mab_kernel_count(void) { return K_COUNT; }

extern "C" const char *mab_kernel_name(int kernel) {
    return (kernel >= 0 && kernel < K_COUNT) ? kKernelNames[kernel] : "";
}

extern "C" int mab_kernel_times(double *ms, int64_t *launches, int reset) {
    for (int k = 0; k < K_COUNT; k++) {
        if (ms) ms[k] = 0.0;
        if (launches) launches[k] = 0;
    }
    for (const TimedLaunch &t : g_timed) {
        if (cudaEventSynchronize(t.stop) != cudaSuccess) return check(cudaGetLastError(), "event sync");
        float v = 0.f;
        if (cudaEventElapsedTime(&v, t.start, t.stop) != cudaSuccess) return check(cudaGetLastError(), "event time");
        if (ms) ms[t.kernel] += v;
        if (launches) launches[t.kernel]++;
    }
    if (reset) {
        for (const TimedLaunch &t : g_timed) {
            g_event_pool.push_back(t.start);
            g_event_pool.push_back(t.stop);
        }
        g_timed.clear();
    }
    return MAB_OK;
}

// launch accounting / event timing for the kernels of realign.cu
void *mab_launch_timer_begin(int kernel, void *stream) { return new LaunchTimer(kernel, (cudaStream_t) stream); }
void mab_launch_timer_end(void *timer) { delete static_cast<LaunchTimer *>(timer); }

extern "C" int64_t mab_launch_count(int reset) {
    const int64_t v = g_launches;
    if (reset) g_launches = 0;
    return v;
}

extern "C" int mab_posteriors_compute(const mab_model *m, const mab_plan *plan, const void *d_blob, void *d_ws,
                                      int64_t ws_bytes, int64_t *d_pair_off, int64_t *d_n_pairs, void *stream) {
    if (!m || !plan || !d_blob || !d_ws || !d_pair_off || !d_n_pairs) {
        mab_set_error("mab_posteriors_compute: null argument");
        return MAB_E_ARG;
    }
    int rc = require_gpu();
    if (rc) return rc;
    Workspace w;
    if ((rc = carve(plan, d_ws, ws_bytes, MAB_MAX_S, false, &w))) return rc;
    DevModel dm;
    mab_model_to_device_block(m, &dm);
    cudaStream_t st = (cudaStream_t) stream;
    if (plan->hdr.n_regions == 0) {
        // nothing to align: all offsets zero
        if ((rc = check(cudaMemsetAsync(d_pair_off, 0, (plan->hdr.n_alignments + 1) * 8, st), "memset"))) return rc;
        return check(cudaMemsetAsync(d_n_pairs, 0, 8, st), "memset");
    }
    return m->S == 5 ? run_posteriors<5>(dm, plan, d_blob, w, d_pair_off, d_n_pairs, st)
                     : run_posteriors<3>(dm, plan, d_blob, w, d_pair_off, d_n_pairs, st);
}

// mab_posteriors_compute without any allocation (margin_b200.h): the over-wide ring comes from g_static_ring.
extern "C" int mab_posteriors_compute_noalloc(const mab_model *m, const mab_plan *plan, const void *d_blob, void *d_ws,
                                              int64_t ws_bytes, int64_t *d_pair_off, int64_t *d_n_pairs, void *stream) {
    int rc = require_gpu();
    if (rc) return rc;
    double *ring = nullptr;
    if ((rc = static_ring_address(&ring))) return rc;
    struct Scope {  // run_dp sees the static ring only inside this call
        Scope(double *p) { t_ext_ring = p, t_ext_ring_bytes = MAB_STATIC_RING_BYTES; }
        ~Scope() { t_ext_ring = nullptr, t_ext_ring_bytes = 0; }
    } scope(ring);
    return mab_posteriors_compute(m, plan, d_blob, d_ws, ws_bytes, d_pair_off, d_n_pairs, stream);
}

static_assert(MAB_STATIC_RING_BYTES == MAB_NOALLOC_SCRATCH_BYTES, "header and scratch array disagree");

namespace {
template <typename XY>
int posteriors_emit(const mab_plan *plan, const void *d_blob, void *d_ws, int64_t ws_bytes, XY *d_x, XY *d_y, int32_t *d_p,
                    int64_t capacity, void *stream) {
    if (!plan || !d_blob || !d_ws || (capacity > 0 && (!d_x || !d_y || !d_p))) {
        mab_set_error("mab_posteriors_emit: null argument");
        return MAB_E_ARG;
    }
    int rc = require_gpu();
    if (rc) return rc;
    Workspace w;
    if ((rc = carve(plan, d_ws, ws_bytes, MAB_MAX_S, false, &w))) return rc;
    const BlobHeader &h = plan->hdr;
    if (h.n_regions == 0 || capacity == 0) return MAB_OK;
    cudaStream_t st = (cudaStream_t) stream;
    const Blob bl = blob_view(d_blob, h);
    {
        LaunchTimer lt(K_EMIT, st);
        k_emit_diag<XY><<<(unsigned) ((h.n_diagonals + 255) / 256), 256, 0, st>>>(bl, h.n_diagonals, w.pint, w.counts, d_x,
                                                                                d_y, d_p, capacity);
    }
    return check(cudaGetLastError(), "emit launch");
}
}  // namespace

extern "C" int mab_posteriors_emit(const mab_plan *plan, const void *d_blob, void *d_ws, int64_t ws_bytes, int32_t *d_x,
                                   int32_t *d_y, int32_t *d_p, int64_t capacity, void *stream) {
    return posteriors_emit<int32_t>(plan, d_blob, d_ws, ws_bytes, d_x, d_y, d_p, capacity, stream);
}

extern "C" int mab_posteriors_emit_i16(const mab_plan *plan, const void *d_blob, void *d_ws, int64_t ws_bytes,
                                       int16_t *d_x, int16_t *d_y, int32_t *d_p, int64_t capacity, void *stream) {
    return posteriors_emit<int16_t>(plan, d_blob, d_ws, ws_bytes, d_x, d_y, d_p, capacity, stream);
}

extern "C" int mab_expectations(const mab_model *m, const mab_plan *plan, const void *d_blob, void *d_ws,
                                int64_t ws_bytes, double *d_out, void *stream) {
    if (!m || !plan || !d_blob || !d_ws || !d_out) {
        mab_set_error("mab_expectations: null argument");
        return MAB_E_ARG;
    }
    int rc = require_gpu();
    if (rc) return rc;
    Workspace w;
    if ((rc = carve(plan, d_ws, ws_bytes, MAB_MAX_S, true, &w))) return rc;
    if (plan->hdr.n_regions == 0) return MAB_OK;
    DevModel dm;
    mab_model_to_device_block(m, &dm);
    cudaStream_t st = (cudaStream_t) stream;
    return m->S == 5 ? run_expectations<5>(dm, plan, d_blob, w, d_out, st)
                     : run_expectations<3>(dm, plan, d_blob, w, d_out, st);
}

extern "C" int mab_debug_copy_matrices(const mab_plan *plan, const void *d_ws, double *h_fwd, double *h_bwd,
                                       double *h_totals, void *stream) {
    // Expectation-mode layout: forward / backward hold [cell][5-stride] doubles (a 3-state run packs stride 3 at the
    // start of the same space).  Only meaningful after mab_expectations.
    if (!plan || !d_ws) return MAB_E_ARG;
    const WorkspaceLayout lay = mab_workspace_layout(plan->hdr, MAB_MAX_S, true);
    const uint8_t *b = (const uint8_t *) d_ws;
    cudaStream_t st = (cudaStream_t) stream;
    int rc;
    const int64_t cell_bytes = plan->hdr.n_cells * MAB_MAX_S * 8;
    if (h_fwd && (rc = check(cudaMemcpyAsync(h_fwd, b + lay.off_fwd, cell_bytes, cudaMemcpyDeviceToHost, st), "copy fwd")))
        return rc;
    if (h_bwd && (rc = check(cudaMemcpyAsync(h_bwd, b + lay.off_bwd, cell_bytes, cudaMemcpyDeviceToHost, st), "copy bwd")))
        return rc;
    if (h_totals && (rc = check(cudaMemcpyAsync(h_totals, b + lay.off_totals, plan->hdr.n_totals * 8,
                                                cudaMemcpyDeviceToHost, st), "copy totals")))
        return rc;
    return check(cudaStreamSynchronize(st), "sync");
}
