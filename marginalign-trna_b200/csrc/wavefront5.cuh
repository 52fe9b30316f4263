// wavefront5.cuh -- the 5-state forward / backward wavefront kernels, "pair-lane" mapping.
//
// Why this shape (measured, profiles/r1_v1_ncu_summary.md): the first kernels mapped one lane to one cell and let
// the compiler branch inside logAdd; ncu showed them issue-bound (85 % issue slots busy, FP64 pipe 30 %) with 9.4
// of 32 lanes active.  Here
//   * every cell of a diagonal is handled by TWO adjacent lanes that execute the SAME instruction stream of four
//     logAdds: the "match lane" runs the 5-term chain into the match state, the "gap lane" the four independent
//     2-term sums into the gap states (5-state cellCalculate has exactly 4 + 4 logAdds, stateMachine.c:399-429).
//     A 12-13 cell diagonal (expansion 10) therefore fills 24-26 lanes and a warp issues 4 logAdds per diagonal
//     instead of 8;
//   * logAdd is branch-free: the polynomial piece is picked by three compares and its coefficients come from a
//     128-byte shared-memory table (2 x LDS.128) instead of select chains over 64-bit immediates;
//   * out-of-band neighbours are read from -inf guard cells at both ends of every ring slot (index clamp) instead of
//     per-state selects;
//   * the (eP + tP) rows are pre-arranged per lane role in shared memory so each lane fetches its 8 addends with
//     4 x LDS.128.
// Bit-exactness: the same doubles are added in the same order as the reference (see forward_cell<5> /
// backward_cell<5> in kernels.cu, which remain the readable statement and are used by the 3-state path).
#pragma once

// ---- branch-free logAdd ------------------------------------------------------------------------------------------
// coef: shared-memory table [4 pieces][c3, c2, c1, c0] (16-byte aligned).
__device__ __forceinline__ double la_bf(double x, double y, const double *coef) {
    const bool lt = x < y;
    const double hi = lt ? y : x;
    const double lo = lt ? x : y;
    const double d = __dsub_rn(hi, lo);  // +inf if lo = -inf, NaN if both are -inf: both fail (d < 7.5) below
    const int piece = (d > 1.00 ? 1 : 0) + (d > 2.50 ? 1 : 0) + (d > 4.50 ? 1 : 0);
    const double2 c32 = *reinterpret_cast<const double2 *>(coef + piece * 4);
    const double2 c10 = *reinterpret_cast<const double2 *>(coef + piece * 4 + 2);
    double r = __dadd_rn(__dmul_rn(c32.x, d), c32.y);
    r = __dadd_rn(__dmul_rn(r, d), c10.x);
    r = __dadd_rn(__dmul_rn(r, d), c10.y);
    r = __dadd_rn(r, lo);
    return (d < 7.5) ? r : hi;
}

__device__ __forceinline__ void fill_logadd_coefficients(double *coef, int tid) {
    // float-suffixed literals of lookup(), pairwiseAligner.c:239-245, promoted to double
    if (tid == 0) {
        coef[0] = (double) -0.009350833524763f; coef[1] = (double) 0.130659527668286f;
        coef[2] = (double) 0.498799810682272f;  coef[3] = (double) 0.693203116424741f;
        coef[4] = (double) -0.014532321752540f; coef[5] = (double) 0.139942324101744f;
        coef[6] = (double) 0.495635523139337f;  coef[7] = (double) 0.692140569840976f;
        coef[8] = (double) -0.004605031767994f; coef[9] = (double) 0.063427417320019f;
        coef[10] = (double) 0.695956496475118f; coef[11] = (double) 0.514272634594009f;
        coef[12] = (double) -0.000458661602210f; coef[13] = (double) 0.009695946122598f;
        coef[14] = (double) 0.930734667215156f; coef[15] = (double) 0.168037164329057f;
    }
}

// Four logAdds, same instruction stream for both lane roles.  Slots t1..t8 hold the addends.
//   match lane (chain = true):  r = la(la(la(la(t1,t2),t4),t6),t8)      -> r4 is the match state
//   gap lane   (chain = false): r1 = la(t1,t2) r2 = la(t3,t4) r3 = la(t5,t6) r4 = la(t7,t8)
struct Quad {
    double r1, r2, r3, r4;
};
__device__ __forceinline__ Quad four_logadds(bool chain, double t1, double t2, double t3, double t4, double t5, double t6,
                                             double t7, double t8, const double *coef) {
    Quad q;
    q.r1 = la_bf(t1, t2, coef);
    q.r2 = la_bf(chain ? q.r1 : t3, t4, coef);
    q.r3 = la_bf(chain ? q.r2 : t5, t6, coef);
    q.r4 = la_bf(chain ? q.r3 : t7, t8, coef);
    return q;
}

#define RING_CELL_BYTES 40  // 5 doubles per cell, [cell][state]

struct Wave5Shared {
    double coef[16];
    double etp[2][25][8];  // [lane role][cX*5+cY][slot]
    double boundary[4][5]; // start, ragged_start, end, ragged_end
};

// Forward tables.  Match lane: t1 = mi.M+eM0, t2 = mi.SX+eM1, t4 = mi.SY+eM2, t6 = mi.LX+eM3, t8 = mi.LY+eM4.
// Gap lane:   t1 = lo.M+eL0, t2 = lo.SX+eL1, t3 = lo.M+eL2, t4 = lo.LX+eL3, t5 = up.M+eU0, t6 = up.SY+eU1,
//             t7 = up.M+eU2, t8 = up.LY+eU3.
__device__ __forceinline__ void fill_forward_tables(Wave5Shared &sh, const DevModel &m, int tid, int nthreads) {
    for (int i = tid; i < 2 * 25 * 8; i += nthreads) {
        const int role = i / 200, code = (i / 8) % 25, slot = i % 8, cX = code / 5, cY = code % 5;
        double v = 0.0;
        if (role == 0) {
            const int map[8] = {0, 1, -1, 2, -1, 3, -1, 4};
            if (map[slot] >= 0) v = m.etp_middle[code][map[slot]];
        } else {
            v = slot < 4 ? m.etp_lower[cX][slot] : m.etp_upper[cY][slot - 4];
        }
        sh.etp[role][code][slot] = v;
    }
}

// Backward tables (gather form, see backward_cell<5>).  code = X[x]*5 + Y[y], the symbols the neighbours consume.
// Match lane: t1 = nMi.M+eM0, t2 = nUp.SY+eU0, t4 = nUp.LY+eU2, t6 = nLo.SX+eL0, t8 = nLo.LX+eL2.
// Gap lane:   t1 = nMi.M+eM1, t2 = nLo.SX+eL1 (SX); t3 = nMi.M+eM2, t4 = nUp.SY+eU1 (SY);
//             t5 = nMi.M+eM3, t6 = nLo.LX+eL3 (LX); t7 = nMi.M+eM4, t8 = nUp.LY+eU3 (LY).
__device__ __forceinline__ void fill_backward_tables(Wave5Shared &sh, const DevModel &m, int tid, int nthreads) {
    for (int i = tid; i < 2 * 25 * 8; i += nthreads) {
        const int role = i / 200, code = (i / 8) % 25, slot = i % 8, cX = code / 5, cY = code % 5;
        const double *eM = m.etp_middle[code], *eL = m.etp_lower[cX], *eU = m.etp_upper[cY];
        double v = 0.0;
        if (role == 0) {
            const double a[8] = {eM[0], eU[0], 0.0, eU[2], 0.0, eL[0], 0.0, eL[2]};
            v = a[slot];
        } else {
            const double a[8] = {eM[1], eL[1], eM[2], eU[1], eM[3], eL[3], eM[4], eU[3]};
            v = a[slot];
        }
        sh.etp[role][code][slot] = v;
    }
}

__device__ __forceinline__ void fill_boundary(Wave5Shared &sh, const DevModel &m, int tid) {
    if (tid < 20) {
        const int which = tid / 5, s = tid % 5;
        const double *src = which == 0 ? m.start : which == 1 ? m.ragged_start : which == 2 ? m.end : m.ragged_end;
        sh.boundary[which][s] = src[s];
    }
}

// One ring slot = (ring_w + 2) cells: position 0 and position w+1 are -inf guards, cell i lives at position i+1.
__device__ __forceinline__ int clamp_pos(int i, int w) {  // cell index -> ring position, out-of-band -> a guard
    return min(max(i, -1), w) + 1;
}

__device__ __forceinline__ void write_guards(double *slot, int w, int lane) {
    // lanes 0..4 write the low guard, lanes 5..9 the high guard
    if (lane < 10) {
        const int s = lane < 5 ? lane : lane - 5;
        const int pos = lane < 5 ? 0 : w + 1;
        slot[pos * 5 + s] = neg_inf();
    }
}

// ------------------------------------------------------------------------------------------------ forward
// FULL = true : F holds all 5 states of every cell ([cell][5]); required by expectation mode.
// FULL = false: posterior mode -- FM holds the match state of every cell, FT all states of the cells on the
//               diagonals where a total probability is evaluated (indexed through TotalDesc::full_off).
template <bool FULL>
__global__ void __launch_bounds__(THREADS_PER_BLOCK)
k_forward5(const __grid_constant__ DevModel model, Blob bl, int n_regions, double *__restrict__ F,
           double *__restrict__ FT, int *counter, int ring_w) {
    extern __shared__ __align__(16) unsigned char dyn_smem_raw[];
    __shared__ __align__(16) Wave5Shared sh;
    fill_logadd_coefficients(sh.coef, threadIdx.x);
    fill_forward_tables(sh, model, threadIdx.x, blockDim.x);
    fill_boundary(sh, model, threadIdx.x);
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int role = lane & 1;       // 0 match lane, 1 gap lane
    const bool chain = role == 0;
    const int jl = lane >> 1;        // cell within a chunk of 16
    const int slot_doubles = (ring_w + 2) * 5;
    double *ring = reinterpret_cast<double *>(dyn_smem_raw) + (size_t) warp * 3 * slot_doubles;
    // per-lane state picks (see table comment): v4 = c1[role ? LX : SY], v6 = c2[role ? SY : LX]
    const int s4 = role ? LX : SY, s6 = role ? SY : LX;
    const int s_out4 = role ? LY : M;  // where r4 goes

    for (;;) {
        const int item = next_item(counter, lane);
        if (item >= n_regions) break;
        const RegionDesc rg = bl.regions[bl.order_fwd[item]];
        const DiagDesc *dd = bl.diags + rg.diag_base;
        const uint8_t *sx = bl.sym_x + rg.sym_x_off, *sy = bl.sym_y + rg.sym_y_off;
        const int D = rg.lX + rg.lY;
        DiagDesc cur = dd[0], nxt = dd[1];
        int L1 = cur.xmyL, L2 = 0, w1 = (int) (nxt.cell_off - cur.cell_off), w2 = 0;
        // diagonal 0 (pairwiseAligner.c:658-659); slot (-1 mod 3) = 2 is the empty "d-2" of diagonal 1
        {
            const double *init = sh.boundary[(rg.flags & 1) ? 1 : 0];
            double *slot0 = ring;
            for (int i = lane; i < w1 * 5; i += 32) {
                const double v = init[i % 5];
                slot0[5 + i] = v;
                if (FULL) F[cur.cell_off * 5 + i] = v;
                else if (i % 5 == 0) F[cur.cell_off + i / 5] = v;
            }
            write_guards(slot0, w1, lane);
            write_guards(ring + 2 * slot_doubles, 0, lane);
        }
        __syncwarp();
        for (int d = 1; d <= D; d++) {
            cur = nxt;
            nxt = dd[d + 1];
            const int L = cur.xmyL;
            const int w = (int) (nxt.cell_off - cur.cell_off);
            const int oLo = (L - 1 - L1) >> 1, oMi = (L - L2) >> 1;
            double *slot = ring + (d % 3) * slot_doubles;
            const double *slot1 = ring + ((d + 2) % 3) * slot_doubles;
            const double *slot2 = ring + ((d + 1) % 3) * slot_doubles;
            // posterior mode: does this diagonal carry a total?  (then all states go to FT)
            double *ft = nullptr;
            if (!FULL && (cur.tot_flags & MAB_DIAG_IS_TOTAL))
                ft = FT + (int64_t) bl.totals[cur.tot_flags & MAB_DIAG_TOT_MASK].full_off * 5;
            for (int j0 = 0; j0 < w; j0 += 16) {
                const int j = j0 + jl;
                if (j < w) {
                    const int xmy = L + 2 * j;
                    const int x = (d + xmy) >> 1, y = (d - xmy) >> 1;
                    const int cX = x > 0 ? __ldg(sx + x - 1) : 4, cY = y > 0 ? __ldg(sy + y - 1) : 4;
                    const double *e = sh.etp[role][cX * 5 + cY];
                    // neighbour cells: match lane reads (d-2, xmy) twice; gap lane reads lower and upper of d-1
                    const int iLo = j + oLo;
                    const double *c1 = chain ? slot2 + clamp_pos(j + oMi, w2) * 5 : slot1 + clamp_pos(iLo, w1) * 5;
                    const double *c2 = chain ? c1 : slot1 + clamp_pos(iLo + 1, w1) * 5;
                    const double v1 = c1[M], v2 = c1[SX], v4 = c1[s4], v5 = c2[M], v6 = c2[s6], v8 = c2[LY];
                    const double2 e12 = *reinterpret_cast<const double2 *>(e);
                    const double2 e34 = *reinterpret_cast<const double2 *>(e + 2);
                    const double2 e56 = *reinterpret_cast<const double2 *>(e + 4);
                    const double2 e78 = *reinterpret_cast<const double2 *>(e + 6);
                    const Quad q = four_logadds(chain, __dadd_rn(v1, e12.x), __dadd_rn(v2, e12.y), __dadd_rn(v1, e34.x),
                                                __dadd_rn(v4, e34.y), __dadd_rn(v5, e56.x), __dadd_rn(v6, e56.y),
                                                __dadd_rn(v5, e78.x), __dadd_rn(v8, e78.y), sh.coef);
                    double *out = slot + (j + 1) * 5;
                    out[s_out4] = q.r4;
                    if (!chain) {
                        out[SX] = q.r1;
                        out[LX] = q.r2;
                        out[SY] = q.r3;
                    }
                    if (FULL) {
                        double *g = F + (cur.cell_off + j) * 5;
                        g[s_out4] = q.r4;
                        if (!chain) {
                            g[SX] = q.r1;
                            g[LX] = q.r2;
                            g[SY] = q.r3;
                        }
                    } else {
                        if (chain) F[cur.cell_off + j] = q.r4;
                        if (ft != nullptr) {
                            double *g = ft + (int64_t) j * 5;
                            g[s_out4] = q.r4;
                            if (!chain) {
                                g[SX] = q.r1;
                                g[LX] = q.r2;
                                g[SY] = q.r3;
                            }
                        }
                    }
                }
            }
            write_guards(slot, w, lane);
            __syncwarp();
            L2 = L1;
            L1 = L;
            w2 = w1;
            w1 = w;
        }
    }
}

// ------------------------------------------------------------------------------------------------ backward
// One warp per traceback.  FULL as above (B / BT); `edge` receives the match state of diagonal from+1 of
// intermediate tracebacks (see TracebackDesc::edge_off).
template <bool FULL>
__global__ void __launch_bounds__(THREADS_PER_BLOCK)
k_backward5(const __grid_constant__ DevModel model, Blob bl, int n_tbs, double *__restrict__ B,
            double *__restrict__ BT, double *__restrict__ edge, int *counter, int ring_w) {
    extern __shared__ __align__(16) unsigned char dyn_smem_raw[];
    __shared__ __align__(16) Wave5Shared sh;
    fill_logadd_coefficients(sh.coef, threadIdx.x);
    fill_backward_tables(sh, model, threadIdx.x, blockDim.x);
    fill_boundary(sh, model, threadIdx.x);
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int role = lane & 1;
    const bool chain = role == 0;
    const int jl = lane >> 1;
    const int slot_doubles = (ring_w + 2) * 5;
    double *ring = reinterpret_cast<double *>(dyn_smem_raw) + (size_t) warp * 3 * slot_doubles;
    const int s_out4 = role ? LY : M;

    for (;;) {
        const int item = next_item(counter, lane);
        if (item >= n_tbs) break;
        const TracebackDesc tb = bl.tbs[bl.order_bwd[item]];
        const RegionDesc rg = bl.regions[tb.region];
        const DiagDesc *dd = bl.diags + rg.diag_base;
        const uint8_t *sx = bl.sym_x + rg.sym_x_off, *sy = bl.sym_y + rg.sym_y_off;
        DiagDesc up1 = dd[tb.hi + 1], cur = dd[tb.hi];  // up1 = descriptor of diagonal d+1 while processing d
        int L1 = cur.xmyL, L2 = 0, w1 = (int) (up1.cell_off - cur.cell_off), w2 = 0;
        // diagonal hi: (ragged)endStateProb on every cell (pairwiseAligner.c:680-681)
        {
            const double *init = sh.boundary[(tb.at_end && (rg.flags & 2)) ? 3 : 2];
            double *slot = ring + (tb.hi % 3) * slot_doubles;
            const bool keep = tb.hi <= tb.from;  // only the final traceback keeps its top diagonal
            double *bt = nullptr;
            if (!FULL && keep && (cur.tot_flags & MAB_DIAG_IS_TOTAL))
                bt = BT + (int64_t) bl.totals[cur.tot_flags & MAB_DIAG_TOT_MASK].full_off * 5;
            for (int i = lane; i < w1 * 5; i += 32) {
                const double v = init[i % 5];
                slot[5 + i] = v;
                if (keep) {
                    if (FULL) B[cur.cell_off * 5 + i] = v;
                    else {
                        if (i % 5 == 0) B[cur.cell_off + i / 5] = v;
                        if (bt != nullptr) bt[i] = v;
                    }
                }
            }
            write_guards(slot, w1, lane);
            write_guards(ring + ((tb.hi + 1) % 3) * slot_doubles, 0, lane);  // the empty "d+2" of diagonal hi-1
        }
        __syncwarp();
        for (int d = tb.hi - 1; d > tb.lo; d--) {
            up1 = cur;
            cur = dd[d];
            const int L = cur.xmyL;
            const int w = (int) (up1.cell_off - cur.cell_off);
            const int oUp = (L - 1 - L1) >> 1;  // (d+1, xmy-1): the cell whose "upper" is the target
            const int oMi = (L - L2) >> 1;      // (d+2, xmy)
            double *slot = ring + (d % 3) * slot_doubles;
            const double *slot1 = ring + ((d + 1) % 3) * slot_doubles;
            const double *slot2 = ring + ((d + 2) % 3) * slot_doubles;
            const bool keep = d <= tb.from;
            const bool to_edge = d == tb.from + 1;
            double *bt = nullptr;
            if (!FULL && keep && (cur.tot_flags & MAB_DIAG_IS_TOTAL))
                bt = BT + (int64_t) bl.totals[cur.tot_flags & MAB_DIAG_TOT_MASK].full_off * 5;
            for (int j0 = 0; j0 < w; j0 += 16) {
                const int j = j0 + jl;
                if (j < w) {
                    const int xmy = L + 2 * j;
                    const int x = (d + xmy) >> 1, y = (d - xmy) >> 1;
                    // the neighbours consume X[x] and/or Y[y] (0-based); past the end they are out of band anyway
                    const int cX = x < rg.lX ? __ldg(sx + x) : 4, cY = y < rg.lY ? __ldg(sy + y) : 4;
                    const double *e = sh.etp[role][cX * 5 + cY];
                    const int iUp = j + oUp;
                    const double *nUp = slot1 + clamp_pos(iUp, w1) * 5;
                    const double *nLo = slot1 + clamp_pos(iUp + 1, w1) * 5;
                    const double *nMi = slot2 + clamp_pos(j + oMi, w2) * 5;
                    // match lane: (nMi.M, nUp.SY, -, nUp.LY, -, nLo.SX, -, nLo.LX)
                    // gap lane:   (nMi.M, nLo.SX, nMi.M, nUp.SY, nMi.M, nLo.LX, nMi.M, nUp.LY)
                    const double vm = nMi[M];
                    const double v2 = chain ? nUp[SY] : nLo[SX];
                    const double v4 = chain ? nUp[LY] : nUp[SY];
                    const double v6 = chain ? nLo[SX] : nLo[LX];
                    const double v8 = chain ? nLo[LX] : nUp[LY];
                    const double2 e12 = *reinterpret_cast<const double2 *>(e);
                    const double2 e34 = *reinterpret_cast<const double2 *>(e + 2);
                    const double2 e56 = *reinterpret_cast<const double2 *>(e + 4);
                    const double2 e78 = *reinterpret_cast<const double2 *>(e + 6);
                    const Quad q = four_logadds(chain, __dadd_rn(vm, e12.x), __dadd_rn(v2, e12.y), __dadd_rn(vm, e34.x),
                                                __dadd_rn(v4, e34.y), __dadd_rn(vm, e56.x), __dadd_rn(v6, e56.y),
                                                __dadd_rn(vm, e78.x), __dadd_rn(v8, e78.y), sh.coef);
                    // gap lane results: r1 = SX, r2 = SY, r3 = LX, r4 = LY
                    double *out = slot + (j + 1) * 5;
                    out[s_out4] = q.r4;
                    if (!chain) {
                        out[SX] = q.r1;
                        out[SY] = q.r2;
                        out[LX] = q.r3;
                    }
                    if (keep) {
                        if (FULL) {
                            double *g = B + (cur.cell_off + j) * 5;
                            g[s_out4] = q.r4;
                            if (!chain) {
                                g[SX] = q.r1;
                                g[SY] = q.r2;
                                g[LX] = q.r3;
                            }
                        } else {
                            if (chain) B[cur.cell_off + j] = q.r4;
                            if (bt != nullptr) {
                                double *g = bt + (int64_t) j * 5;
                                g[s_out4] = q.r4;
                                if (!chain) {
                                    g[SX] = q.r1;
                                    g[SY] = q.r2;
                                    g[LX] = q.r3;
                                }
                            }
                        }
                    } else if (to_edge && chain) {
                        edge[tb.edge_off + j] = q.r4;
                    }
                }
            }
            write_guards(slot, w, lane);
            __syncwarp();
            L2 = L1;
            L1 = L;
            w2 = w1;
            w1 = w;
        }
    }
}
