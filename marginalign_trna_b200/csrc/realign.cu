// realign.cu -- the realign tail on the GPU: everything cPecanRealign does with the posterior pairs of an
// alignment before it prints the new cigar (cPecanRealign.c:590-631), for every alignment of a batch whose posterior
// pairs are still resident in HBM.  Only the cigars cross PCIe (a few ops per read instead of ~190 pairs per read).
//
//   k_rl_weights  one warp per alignment
//       AMAP reweight: getIndelProbabilities (pairwiseAligner.c:1216-1231: 1e7 minus the sum of the pair weights on
//       every X / Y position, clamped at 0), reweightAlignedPairs (:1233-1249: w' = (int64)(w - gapGamma * (iX + iY)),
//       evaluated in double, truncated), then the weight the chain DP reads: avgWeight = w' / 1e7 + st_random() * 1e-5
//       (multipleAligner.c:140-147).  The jitter draws are glibc's rand() stream from seed 1, two per pair, taken in
//       the order makeAlignmentWeightAdjacencyLists sees the pairs: the list reversed by
//       convertAlignedPairsToMultipleAlignedPairs (:622-629); the X-column weight is the SECOND draw of its pair
//       (:158-159).  Pairs are then bucketed by X column (counting sort inside the warp).
//   k_rl_chain    one thread per alignment
//       pairwiseAlignColumns (multipleAligner.c:355-489): X columns ascending; every admissible weight of a column
//       (avgWeight >= matchGamma and > 0) is scored against the frontier as it stood before the column, then the
//       column's candidates enter the frontier from right to left, evicting points that score no better.  The
//       frontier (the reference's sorted set bestScoringAlignments) is a sorted array searched from its right end:
//       the band makes every query land within a few slots of it.  The traceback leaves the chain, ascending, in the
//       scratch and counts the cigar operations of convertAlignedPairsToPairwiseAlignment (cPecanRealign.c:50-92).
//   k_rl_scan     one block: exclusive scans of the per-alignment op and chain counts
//   k_rl_emit     one thread per alignment: writes the ops (sonLib codes 0 M, 1 D = X only, 2 I = Y only) and, if asked,
//       the chain pairs, densely.
//
// Integer / FP64 work, bit-exact with the reference: additions and multiplications are explicit _rn intrinsics (no
// FMA contraction), the division is IEEE.  Bound by memory latency (dependent loads of a sequential DP), not by a pipe.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstring>

#include "mab_internal.h"

namespace {

struct Front {  // one point of the frontier: 16 bytes
    double score;
    int32_t cand;  // sorted-pair index of the candidate, -1 = the (-1, -1) sentinel, -2 = the (lX, lY) sentinel
    int32_t y;
};

struct Rec {  // one pair as the chain DP reads it: 16 bytes, one load
    double aw;
    int32_t y, x;
};

struct RlLayout {  // device scratch, all offsets 256-byte aligned
    int64_t off_ix;     // int64[len_x]          indel weights of X positions, later int32 column starts / cursors
    int64_t off_iy;     // int64[len_y]          indel weights of Y positions
    int64_t off_front;  // Front[len_y + 2 n]    frontier of every alignment (lY + 2 slots each), later its chain
    int64_t off_aw;     // double[n_pairs]       avgWeight of every input pair
    int64_t off_rec;    // Rec[n_pairs]          pairs bucketed by X column: (avgWeight, y, x), 16 bytes each
    int64_t off_score;  // double[n_pairs]       candidate score, by sorted position
    int64_t off_pred;   // int32[n_pairs]        candidate predecessor (sorted position, -1 = start)
    int64_t off_cnt;    // int32[2 n]            per alignment: op count, chain length
    int64_t total;
};

int64_t align256(int64_t v) { return (v + 255) & ~(int64_t) 255; }

RlLayout rl_layout(int64_t n, int64_t n_pairs, int64_t len_x, int64_t len_y) {
    RlLayout l;
    int64_t o = 0;
    l.off_ix = o, o = align256(o + 8 * std::max<int64_t>(len_x, 1));
    l.off_iy = o, o = align256(o + 8 * std::max<int64_t>(len_y, 1));
    l.off_front = o, o = align256(o + 16 * std::max<int64_t>(len_y + 2 * n, 1));
    l.off_aw = o, o = align256(o + 8 * std::max<int64_t>(n_pairs, 1));
    l.off_rec = o, o = align256(o + 16 * std::max<int64_t>(n_pairs, 1));
    l.off_score = o, o = align256(o + 8 * std::max<int64_t>(n_pairs, 1));
    l.off_pred = o, o = align256(o + 4 * std::max<int64_t>(n_pairs, 1));
    l.off_cnt = o, o = align256(o + 8 * std::max<int64_t>(n, 1));
    l.total = o;
    return l;
}

struct RlArgs {
    int64_t n;
    int64_t n_pairs;  // capacity of x / y / p: alignments whose pairs do not fit are skipped (chain length -1)
    const int64_t *off_x, *off_y, *pair_off;
    const int32_t *x, *y, *p;
    const int32_t *rnd;
    int64_t n_rand;
    double gap_gamma, match_gamma;
    long long *ix, *iy;
    Front *front;
    double *aw, *score;
    Rec *rec;
    int32_t *pred, *cnt;
};

#define RL_FULL 0xffffffffu
#define RL_LOOK 4  // columns whose loads are issued together in k_rl_chain

// ------------------------------------------------------------------------------------------------ weights + buckets
__global__ void __launch_bounds__(256) k_rl_weights(RlArgs a) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = ((int64_t) gridDim.x * blockDim.x) >> 5;
    for (int64_t al = warp; al < a.n; al += n_warps) {
        const int64_t ox = a.off_x[al], oy = a.off_y[al], p0 = a.pair_off[al];
        const int lX = (int) (a.off_x[al + 1] - ox), lY = (int) (a.off_y[al + 1] - oy);
        const int np = (int) (a.pair_off[al + 1] - p0);
        if (a.pair_off[al + 1] > a.n_pairs || 2 * (int64_t) np > a.n_rand) continue;  // clipped by the caller's capacity, or
                                                                    // more pairs than the jitter stream holds: see k_rl_chain
        long long *ix = a.ix + ox, *iy = a.iy + oy;
        const int32_t *px = a.x + p0, *py = a.y + p0, *pp = a.p + p0;
        if (a.gap_gamma > 0.0) {  // reweightAlignedPairs2 returns the list untouched otherwise (:1252)
            for (int i = lane; i < lX; i += 32) ix[i] = (long long) MAB_PROB_1;
            for (int i = lane; i < lY; i += 32) iy[i] = (long long) MAB_PROB_1;
            __syncwarp();
            for (int i = lane; i < np; i += 32) {
                const unsigned long long minus = (unsigned long long) (-(long long) pp[i]);
                atomicAdd((unsigned long long *) &ix[px[i]], minus);
                atomicAdd((unsigned long long *) &iy[py[i]], minus);
            }
            __syncwarp();
        }
        for (int i = lane; i < np; i += 32) {
            long long w = pp[i];
            if (a.gap_gamma > 0.0) {
                const long long sx = max(ix[px[i]], 0LL), sy = max(iy[py[i]], 0LL);
                // int64 - double * int64, evaluated in double and truncated back to int64 (:1244)
                w = (long long) __dsub_rn((double) w, __dmul_rn(a.gap_gamma, (double) (sx + sy)));
            }
            const int64_t k = np - 1 - i;  // position in the reversed list; draws 2k (Y column) and 2k+1 (X column)
            const double r = (double) a.rnd[(2 * k + 1) % a.n_rand] / 2147483648.0;  // st_random(): rand() / (RAND_MAX + 1.0)
            a.aw[p0 + i] = __dadd_rn(__ddiv_rn((double) w, MAB_PROB_1), __dmul_rn(r, 0.00001));
        }
        __syncwarp();
        if (lX == 0) continue;
        // bucket the pairs by X column: counts -> exclusive scan -> scatter; afterwards col[x] = END of column x
        int32_t *col = reinterpret_cast<int32_t *>(ix);
        for (int i = lane; i < lX; i += 32) col[i] = 0;
        __syncwarp();
        for (int i = lane; i < np; i += 32) atomicAdd(&col[px[i]], 1);
        __syncwarp();
        int carry = 0;
        for (int base = 0; base < lX; base += 32) {
            const int v = base + lane < lX ? col[base + lane] : 0;
            int s = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(RL_FULL, s, o);
                if (lane >= o) s += t;
            }
            if (base + lane < lX) col[base + lane] = s - v + carry;
            carry += __shfl_sync(RL_FULL, s, 31);
        }
        __syncwarp();
        Rec *rec = a.rec + p0;
        for (int i = lane; i < np; i += 32) rec[atomicAdd(&col[px[i]], 1)] = Rec{a.aw[p0 + i], py[i], px[i]};
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------------------ chain DP
// cigar operations of the chain `xy[0..m)` closed by the (lX, lY) end pair: convertAlignedPairsToPairwiseAlignment
// (cPecanRealign.c:50-92).  EMIT = false only counts.
template <bool EMIT>
__device__ __forceinline__ int cigar_walk(const int2 *xy, int m, int lX, int lY, uint8_t *ops, int32_t *lens) {
    int pX = -1, pY = -1, mL = 0, n = 0;
    for (int i = 0; i <= m; i++) {
        const int x = i < m ? xy[i].x : lX, y = i < m ? xy[i].y : lY;
        if (x - pX > 1) {
            if (mL > 0) {
                if (EMIT) ops[n] = 0, lens[n] = mL;
                n++;
                mL = 0;
            }
            if (EMIT) ops[n] = 1, lens[n] = x - pX - 1;
            n++;
        }
        if (y - pY > 1) {
            if (mL > 0) {
                if (EMIT) ops[n] = 0, lens[n] = mL;
                n++;
                mL = 0;
            }
            if (EMIT) ops[n] = 2, lens[n] = y - pY - 1;
            n++;
        }
        mL++;
        pX = x;
        pY = y;
    }
    if (mL > 1) {  // the trailing match run, without the end pair
        if (EMIT) ops[n] = 0, lens[n] = mL - 1;
        n++;
    }
    return n;
}

// The frontier holds the (-1, -1) start sentinel in slot 0 and the real points after it; the (lX, lY) end sentinel of
// the reference (score INT64_MAX) is virtual: "slot nf".  `top` is a register copy of the right-most point: in a band
// most candidates lie to its right, so their predecessor is `top` itself and they are appended with a single store.
struct Frontier {
    Front *fr;
    int nf;
    Front top;
    __device__ __forceinline__ Front before(int y) const {  // stSortedSet_searchLessThan: last point with y' < y
        if (top.y < y) return top;
        int f = nf - 2;
        while (fr[f].y >= y) f--;  // fr[0].y = -1 stops it
        return fr[f];
    }
    __device__ __forceinline__ void insert(double s, int cand, int y) {  // :412-431
        if (top.y < y) {  // lands just before the end sentinel, evicts nothing
            top = Front{s, cand, y};
            fr[nf++] = top;
            return;
        }
        int g = nf - 1;
        while (fr[g - 1].y >= y) g--;  // stSortedSet_searchGreaterThanOrEqual (exists: top.y >= y)
        if (s >= fr[g].score || fr[g].y > y) {
            int e = g;
            while (e < nf && s >= fr[e].score) e++;  // points that score no better are evicted
            if (e == g) {
                for (int k = nf; k > g; k--) fr[k] = fr[k - 1];
                nf++;
            } else if (e > g + 1) {
                const int drop = e - g - 1;
                for (int k = e; k < nf; k++) fr[k - drop] = fr[k];
                nf -= drop;
            }
            fr[g] = Front{s, cand, y};
            top = fr[nf - 1];
        }
    }
};

// One thread per alignment; `spread` > 0 gives every alignment 2^spread threads of which the first one works, so that
// the few long alignments of a long-read batch each get an instruction stream of their own instead of taking turns
// inside a handful of diverged warps (rl_spread below).
__global__ void __launch_bounds__(128) k_rl_chain(RlArgs a, int spread) {
    const int64_t gt = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t al = gt >> spread;
    if (al >= a.n || (gt & ((1 << spread) - 1)) != 0) return;
    const int64_t ox = a.off_x[al], oy = a.off_y[al], p0 = a.pair_off[al];
    const int lX = (int) (a.off_x[al + 1] - ox), lY = (int) (a.off_y[al + 1] - oy);
    // the pairs of this alignment were clipped by the capacity, or it has more pairs than half the jitter stream (the
    // draw index must never wrap: every reference process draws 2 np fresh values): no operations, chain length -1 --
    // the caller sees an empty cigar for a non-empty alignment and re-runs with a larger buffer / table
    if (a.pair_off[al + 1] > a.n_pairs || 2 * (a.pair_off[al + 1] - p0) > a.n_rand) {
        a.cnt[2 * al] = 0;
        a.cnt[2 * al + 1] = -1;
        return;
    }
    const int32_t *col = reinterpret_cast<const int32_t *>(a.ix + ox);
    Rec *rec = a.rec + p0;
    int32_t *pred = a.pred + p0;
    double *score = a.score + p0;
    Frontier F;
    F.fr = a.front + oy + 2 * al;  // lY + 2 slots
    F.nf = 1;
    F.top = Front{0.0, -1, -1};
    F.fr[0] = F.top;
    // Columns are taken RL_LOOK at a time: their ends and the first record of each are loaded together (independent
    // loads, one round trip for the block) before any of them is processed, so the sequential walk does not pay a
    // dependent column-end -> record -> frontier chain per column.  A first record may turn out to be stale only in
    // a column with several records, which re-reads them anyway.
    int start = 0;
    for (int x0 = 0; x0 < lX; x0 += RL_LOOK) {
        int ends[RL_LOOK];
        Rec firsts[RL_LOOK];
#pragma unroll
        for (int k = 0; k < RL_LOOK; k++) ends[k] = x0 + k < lX ? col[x0 + k] : -1;
#pragma unroll
        for (int k = 0; k < RL_LOOK; k++) {
            const int st_k = k == 0 ? start : ends[k - 1];
            firsts[k] = (ends[k] > st_k) ? rec[st_k] : Rec{0.0, 0, 0};
        }
#pragma unroll
        for (int k = 0; k < RL_LOOK; k++) {
            const int end = ends[k];
            if (end < 0) break;
            if (end == start + 1) {  // one weight in the column (the common case): score it and insert it
                const Rec r = firsts[k];
                if (r.aw >= a.match_gamma && r.aw > 0.0) {
                    const Front pr = F.before(r.y);
                    pred[start] = pr.cand;
                    F.insert(__dadd_rn(pr.score, r.aw), start, r.y);
                }
            } else if (end > start) {
                for (int i = start + 1; i < end; i++) {  // the column's weights by Y (they arrive in atomic order)
                    const Rec v = rec[i];
                    int j = i;
                    while (j > start && rec[j - 1].y > v.y) {
                        rec[j] = rec[j - 1];
                        j--;
                    }
                    rec[j] = v;
                }
                // every admissible weight against the frontier as it stood before this column (:386-407)
                for (int i = start; i < end; i++) {
                    const Rec r = rec[i];
                    if (r.aw >= a.match_gamma && r.aw > 0.0) {
                        const Front pr = F.before(r.y);
                        score[i] = __dadd_rn(pr.score, r.aw);
                        pred[i] = pr.cand;
                    } else {
                        pred[i] = -3;
                    }
                }
                // the new points, right to left along Y (:410-432)
                for (int i = end - 1; i >= start; i--)
                    if (pred[i] != -3) F.insert(score[i], i, rec[i].y);
            }
            start = end;
        }
    }
    // traceback from the right-most real point (:436-463); the chain, ascending, replaces the frontier
    const int last = F.top.cand;
    int m = 0;
    for (int t = last; t >= 0; t = pred[t]) m++;
    int2 *chain = reinterpret_cast<int2 *>(F.fr);
    int k = m;
    for (int t = last; t >= 0; t = pred[t]) {
        const Rec r = rec[t];
        chain[--k] = make_int2(r.x, r.y);
    }
    a.cnt[2 * al] = cigar_walk<false>(chain, m, lX, lY, nullptr, nullptr);
    a.cnt[2 * al + 1] = m;
}

inline int rl_spread(int64_t n) {  // same rule as the planner's spread_for (plan_device.cu)
    int s = 5;
    while (s > 0 && (n << s) > 300000) s--;
    return s;
}

// exclusive scans of the per-alignment op counts and chain lengths (one block)
__global__ void __launch_bounds__(1024) k_rl_scan(const int32_t *cnt, int64_t n, int64_t *op_off, int64_t *chain_off) {
    __shared__ long long warp_sums[2][32];
    __shared__ long long carry[2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x < 2) carry[threadIdx.x] = 0;
    __syncthreads();
    for (int64_t base = 0; base < n; base += 1024) {
        const int64_t i = base + threadIdx.x;
        long long v[2] = {i < n ? cnt[2 * i] : 0, i < n ? max(cnt[2 * i + 1], 0) : 0};
        long long s[2] = {v[0], v[1]};
#pragma unroll
        for (int q = 0; q < 2; q++) {
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const long long t = __shfl_up_sync(RL_FULL, s[q], o);
                if (lane >= o) s[q] += t;
            }
            if (lane == 31) warp_sums[q][warp] = s[q];
        }
        __syncthreads();
        if (warp == 0) {
#pragma unroll
            for (int q = 0; q < 2; q++) {
                long long t = warp_sums[q][lane];
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const long long u = __shfl_up_sync(RL_FULL, t, o);
                    if (lane >= o) t += u;
                }
                warp_sums[q][lane] = t;  // inclusive over warps
            }
        }
        __syncthreads();
        const long long before0 = (warp ? warp_sums[0][warp - 1] : 0) + carry[0];
        const long long before1 = (warp ? warp_sums[1][warp - 1] : 0) + carry[1];
        if (i < n) {
            op_off[i] = before0 + s[0] - v[0];
            chain_off[i] = before1 + s[1] - v[1];
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            carry[0] += warp_sums[0][31];
            carry[1] += warp_sums[1][31];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        op_off[n] = carry[0];
        chain_off[n] = carry[1];
    }
}

__global__ void __launch_bounds__(128)
k_rl_emit(RlArgs a, const int64_t *op_off, const int64_t *chain_off, uint8_t *ops, int32_t *lens, int32_t *chain_x,
          int32_t *chain_y, int spread) {
    const int64_t gt = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t al = gt >> spread;
    if (al >= a.n || (gt & ((1 << spread) - 1)) != 0) return;
    const int64_t oy = a.off_y[al];
    const int lX = (int) (a.off_x[al + 1] - a.off_x[al]), lY = (int) (a.off_y[al + 1] - oy);
    const int2 *chain = reinterpret_cast<const int2 *>(a.front + oy + 2 * al);
    const int m = a.cnt[2 * al + 1];
    if (m < 0) return;  // skipped in k_rl_chain
    cigar_walk<true>(chain, m, lX, lY, ops + op_off[al], lens + op_off[al]);
    if (chain_x != nullptr) {
        const int64_t c0 = chain_off[al];
        for (int i = 0; i < m; i++) {
            chain_x[c0 + i] = chain[i].x;
            chain_y[c0 + i] = chain[i].y;
        }
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------------ C ABI

static int rl_check(cudaError_t e, const char *what) {
    if (e == cudaSuccess) return MAB_OK;
    mab_set_error("%s: %s", what, cudaGetErrorString(e));
    return MAB_E_CUDA;
}


/* glibc's rand() (stdlib/random_r.c, TYPE_3: x[i] = x[i-3] + x[i-31], seeded by the 16807 Lehmer generator, first
 * 310 outputs discarded), restated so that the jitter stream does not depend on, or disturb, the process-wide libc
 * state.  sonLib's st_random() (sonLibRandom.c:39-43) is rand() / (RAND_MAX + 1.0). */
extern "C" int mab_rand_sequence(uint32_t seed, int64_t n, int32_t *out) {
    if (n < 0 || (n > 0 && out == nullptr)) {
        mab_set_error("mab_rand_sequence: bad arguments");
        return MAB_E_ARG;
    }
    if (seed == 0) seed = 1;  // srandom_r does the same
    uint32_t r[34 + 310];
    int32_t word = (int32_t) seed;
    r[0] = (uint32_t) word;
    for (int i = 1; i < 31; i++) {
        const long hi = word / 127773, lo = word % 127773;
        long w = 16807 * lo - 2836 * hi;
        if (w < 0) w += 2147483647;
        word = (int32_t) w;
        r[i] = (uint32_t) word;
    }
    for (int i = 31; i < 34; i++) r[i] = r[i - 31];
    for (int i = 34; i < 344; i++) r[i] = r[i - 31] + r[i - 3];
    uint32_t ring[31];
    for (int i = 0; i < 31; i++) ring[i] = r[344 - 31 + i];  // ring[j] = x[344 - 31 + j]
    int pos = 0;                                            // slot of x[i - 31]
    for (int64_t k = 0; k < n; k++) {
        const uint32_t v = ring[pos] + ring[(pos + 28) % 31];  // x[i-31] + x[i-3]
        ring[pos] = v;
        pos = (pos + 1) % 31;
        out[k] = (int32_t) (v >> 1);
    }
    return MAB_OK;
}

extern "C" int64_t mab_realign_workspace_bytes(int64_t n, int64_t n_pairs, int64_t len_x, int64_t len_y) {
    if (n < 0 || n_pairs < 0 || len_x < 0 || len_y < 0) return -1;
    return rl_layout(n, n_pairs, len_x, len_y).total;
}

static int rl_args(RlArgs &a, int64_t n, const int64_t *d_off_x, const int64_t *d_off_y, int64_t len_x, int64_t len_y,
                   const int64_t *d_pair_off, const int32_t *d_x, const int32_t *d_y, const int32_t *d_p,
                   int64_t n_pairs, void *d_ws, int64_t ws_bytes) {
    if (n < 0 || n_pairs < 0 || len_x < 0 || len_y < 0 || d_ws == nullptr || (n > 0 && (!d_off_x || !d_off_y || !d_pair_off))) {
        mab_set_error("realign: bad arguments");
        return MAB_E_ARG;
    }
    if (n_pairs >= ((int64_t) 1 << 31) || len_x >= ((int64_t) 1 << 31) || len_y >= ((int64_t) 1 << 31)) {
        mab_set_error("realign: batch too large for 32-bit pair indices; split it");
        return MAB_E_ARG;
    }
    const RlLayout l = rl_layout(n, n_pairs, len_x, len_y);
    if (ws_bytes < l.total || ((uintptr_t) d_ws & 255)) {
        mab_set_error("realign: workspace too small or misaligned (%lld < %lld)", (long long) ws_bytes, (long long) l.total);
        return MAB_E_CAPACITY;
    }
    char *b = (char *) d_ws;
    a.n = n;
    a.n_pairs = n_pairs;
    a.off_x = d_off_x, a.off_y = d_off_y, a.pair_off = d_pair_off;
    a.x = d_x, a.y = d_y, a.p = d_p;
    a.ix = (long long *) (b + l.off_ix), a.iy = (long long *) (b + l.off_iy);
    a.front = (Front *) (b + l.off_front);
    a.aw = (double *) (b + l.off_aw), a.score = (double *) (b + l.off_score);
    a.rec = (Rec *) (b + l.off_rec), a.pred = (int32_t *) (b + l.off_pred), a.cnt = (int32_t *) (b + l.off_cnt);
    a.rnd = nullptr, a.n_rand = 0, a.gap_gamma = 0.0, a.match_gamma = 0.0;
    return MAB_OK;
}

extern "C" int mab_realign_chain(int64_t n, const int64_t *d_off_x, const int64_t *d_off_y, int64_t len_x, int64_t len_y,
                                 const int64_t *d_pair_off, const int32_t *d_x, const int32_t *d_y, const int32_t *d_p,
                                 int64_t n_pairs, double gap_gamma, double match_gamma, const int32_t *d_rand,
                                 int64_t n_rand, void *d_ws, int64_t ws_bytes, int64_t *d_op_off, int64_t *d_chain_off,
                                 void *stream) {
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev < 1) {
        mab_set_error("no CUDA device: the realign tail has no CPU fallback");
        return MAB_E_NOGPU;
    }
    RlArgs a;
    int rc = rl_args(a, n, d_off_x, d_off_y, len_x, len_y, d_pair_off, d_x, d_y, d_p, n_pairs, d_ws, ws_bytes);
    if (rc) return rc;
    if (d_rand == nullptr || n_rand < 2 || d_op_off == nullptr || d_chain_off == nullptr) {
        mab_set_error("realign: missing jitter table or output offsets");
        return MAB_E_ARG;
    }
    a.rnd = d_rand, a.n_rand = n_rand;
    // both gammas are floats in the reference (pairwiseAligner.h:38, cPecanRealign.c:390, multipleAligner.c:941)
    a.gap_gamma = (double) (float) gap_gamma, a.match_gamma = (double) (float) match_gamma;
    cudaStream_t st = (cudaStream_t) stream;
    if (n > 0) {
        int sms = 148;
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        const int64_t blocks_w = std::min<int64_t>((n + 7) / 8, (int64_t) sms * 8);
        void *t = mab_launch_timer_begin(MAB_K_RL_WEIGHTS, st);
        k_rl_weights<<<(unsigned) blocks_w, 256, 0, st>>>(a);
        mab_launch_timer_end(t);
        t = mab_launch_timer_begin(MAB_K_RL_CHAIN, st);
        const int spread = rl_spread(n);
        k_rl_chain<<<(unsigned) (((n << spread) + 127) / 128), 128, 0, st>>>(a, spread);
        mab_launch_timer_end(t);
    }
    void *t = mab_launch_timer_begin(MAB_K_RL_SCAN, st);
    k_rl_scan<<<1, 1024, 0, st>>>(a.cnt, n, d_op_off, d_chain_off);
    mab_launch_timer_end(t);
    return rl_check(cudaGetLastError(), "realign launch");
}

extern "C" int mab_realign_emit(int64_t n, const int64_t *d_off_x, const int64_t *d_off_y, int64_t len_x, int64_t len_y,
                                int64_t n_pairs, void *d_ws, int64_t ws_bytes, const int64_t *d_op_off,
                                const int64_t *d_chain_off, uint8_t *d_ops, int32_t *d_lens, int32_t *d_chain_x,
                                int32_t *d_chain_y, void *stream) {
    RlArgs a;
    int rc = rl_args(a, n, d_off_x, d_off_y, len_x, len_y, d_op_off, nullptr, nullptr, nullptr, n_pairs, d_ws, ws_bytes);
    if (rc) return rc;
    if (n == 0) return MAB_OK;
    if (d_ops == nullptr || d_lens == nullptr || d_chain_off == nullptr || (d_chain_x == nullptr) != (d_chain_y == nullptr)) {
        mab_set_error("realign emit: missing outputs");
        return MAB_E_ARG;
    }
    void *t = mab_launch_timer_begin(MAB_K_RL_EMIT, stream);
    const int spread = rl_spread(n);
    k_rl_emit<<<(unsigned) (((n << spread) + 127) / 128), 128, 0, (cudaStream_t) stream>>>(a, d_op_off, d_chain_off, d_ops,
                                                                                          d_lens, d_chain_x, d_chain_y, spread);
    mab_launch_timer_end(t);
    return rl_check(cudaGetLastError(), "realign emit launch");
}
