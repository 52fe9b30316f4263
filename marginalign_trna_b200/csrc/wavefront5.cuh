// wavefront5.cuh -- the 5-state forward / backward wavefront kernels, "pair-lane" mapping.
//
// Why this shape (measured; profiles/r1_ncu_summary.md has the numbers of each step):
//   v1  one lane per cell, compiler-branched logAdd: issue-bound (85 % issue slots, FP64 pipe 30 %), 9.4 of 32 lanes
//       active.
//   v2  pair-lane mapping + branch-free logAdd with a shared-memory coefficient table: shared-memory data pipe at
//       97 % (44 % of the load wavefronts were bank-conflict replays on the 64-byte (eP+tP) rows).
//   v3  rows padded to 80 B, coefficients moved to the constant bank, rotating ring-slot pointers: the indexed
//       constant cache became the limiter (IDC 92 % busy: a per-lane index makes every distinct address a request),
//       and 46 % of the shared-memory wavefronts were still conflicts.
//   v4  (this file) the two lane roles are split by half-warp, so every 64-bit shared access phase sees one role
//       with a conflict-free 40-byte stride; the (eP+tP) rows and the logAdd coefficients live in shared memory in
//       8 replicas, lane l reading replica l & 7, which makes every LDS.128 conflict-free whatever rows the lanes
//       index.  Forward 24.0 -> 13.9 ms, backward 24.4 -> 15.1 ms per 5.6e8 cells.
//
//   * every cell of a diagonal is handled by TWO lanes (l and l + 16) that execute the SAME instruction stream of four
//     logAdds: the "match lane" runs the 5-term chain into the match state, the "gap lane" the four independent
//     2-term sums into the gap states (5-state cellCalculate has exactly 4 + 4 logAdds, stateMachine.c:399-429).
//     A 12-13 cell diagonal (expansion 10) fills 24-26 lanes and a warp issues 4 logAdds per diagonal instead of 8;
//   * logAdd is branch-free: the polynomial piece is picked by three compares;
//   * out-of-band neighbours are read from -inf guard cells at both ends of every ring slot (index clamp) instead of
//     per-state selects;
//   * the (eP + tP) rows are pre-arranged per lane role in shared memory so each lane fetches its 8 addends with
//     4 x LDS.128.
// Bit-exactness: the same doubles are added in the same order as the reference (see forward_cell<5> /
// backward_cell<5> in kernels.cu, which remain the readable statement and are used by the 3-state path).
#pragma once

// ---- branch-free logAdd ------------------------------------------------------------------------------------------
// lookup() coefficients of pairwiseAligner.c:239-245 ([piece][c3, c2, c1, c0]); float literals promoted to double.
__constant__ double c_la_coef[16] = {
    (double) -0.009350833524763f, (double) 0.130659527668286f, (double) 0.498799810682272f, (double) 0.693203116424741f,
    (double) -0.014532321752540f, (double) 0.139942324101744f, (double) 0.495635523139337f, (double) 0.692140569840976f,
    (double) -0.004605031767994f, (double) 0.063427417320019f, (double) 0.695956496475118f, (double) 0.514272634594009f,
    (double) -0.000458661602210f, (double) 0.009695946122598f, (double) 0.930734667215156f, (double) 0.168037164329057f};

enum CoefMode { COEF_SMEM = 1 };  // coefficients from a replicated shared-memory table (the constant bank and a mixed
                                  // fetch were measured slower in round 1 and removed)

#define W5_COPIES 8  // table replicas: lane l reads replica (l & 7), so the 8 lanes of a quarter-warp (one LDS.128
                     // phase) always hit 8 different 16-byte bank groups, whatever rows they index

// CM selects where the four cubic coefficients come from.  Measured (profiles/r1_ncu_summary.md): all four from the
// constant bank saturates the indexed-constant cache (the index differs per lane, every distinct address is a
// separate request: IDC 92 % busy); all four from shared memory loads the LSU instead.  COEF_MIXED splits them:
// (c3, c2) with one conflict-free LDS.128, (c1, c0) with two LDC.64.
// Shared-memory loads with an explicit 32-bit address and immediate offset: one address register serves both halves
// of a coefficient row (the compiler otherwise rebuilds a second base per load).
__device__ __forceinline__ double2 lds_v2(unsigned addr) {
    double2 v;
    asm("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(addr));
    return v;
}
__device__ __forceinline__ double2 lds_v2_128(unsigned addr) {
    double2 v;
    asm("ld.shared.v2.f64 {%0, %1}, [%2+128];" : "=d"(v.x), "=d"(v.y) : "r"(addr));
    return v;
}

// coef_addr: shared-space address of sh.coef[0][0][replica of this lane]; piece p lives 256 bytes further per step.
template <int CM>
__device__ __forceinline__ double la_bf(double x, double y, unsigned coef_addr) {
    const bool lt = x < y;
    const double hi = lt ? y : x;
    const double lo = lt ? x : y;
    const double d = __dsub_rn(hi, lo);  // +inf if lo = -inf, NaN if both are -inf: both fail (d < 7.5) below
    double c3, c2, c1, c0;
    {
        unsigned a = coef_addr;  // three predicated adds pick the row
        if (d > 1.00) a += 2 * W5_COPIES * 16;
        if (d > 2.50) a += 2 * W5_COPIES * 16;
        if (d > 4.50) a += 2 * W5_COPIES * 16;
        const double2 c32 = lds_v2(a), c10 = lds_v2_128(a);
        c3 = c32.x;
        c2 = c32.y;
        c1 = c10.x;
        c0 = c10.y;
    }
    double r = __dadd_rn(__dmul_rn(c3, d), c2);
    r = __dadd_rn(__dmul_rn(r, d), c1);
    r = __dadd_rn(__dmul_rn(r, d), c0);
    // result = d < 7.5 ? r + lo : hi, as one predicated add into a register preloaded with hi
    double res = hi;
    asm("{\n\t.reg .pred p;\n\tsetp.lt.f64 p, %1, 0d401E000000000000;\n\t@p add.rn.f64 %0, %2, %3;\n\t}"
        : "+d"(res)
        : "d"(d), "d"(r), "d"(lo));
    return res;
}

// Four logAdds, same instruction stream for both lane roles.  Slots t1..t8 hold the addends.
//   match lane (chain = true):  r = la(la(la(la(t1,t2),t4),t6),t8)      -> r4 is the match state
//   gap lane   (chain = false): r1 = la(t1,t2) r2 = la(t3,t4) r3 = la(t5,t6) r4 = la(t7,t8)
struct Quad {
    double r1, r2, r3, r4;
};
template <int CM>
__device__ __forceinline__ Quad four_logadds(bool chain, double t1, double t2, double t3, double t4, double t5, double t6,
                                             double t7, double t8, unsigned coef) {
    Quad q;
    q.r1 = la_bf<CM>(t1, t2, coef);
    q.r2 = la_bf<CM>(chain ? q.r1 : t3, t4, coef);
    q.r3 = la_bf<CM>(chain ? q.r2 : t5, t6, coef);
    q.r4 = la_bf<CM>(chain ? q.r3 : t7, t8, coef);
    return q;
}

struct Wave5Shared {
    double2 etp[2][25][4][W5_COPIES];  // [lane role][cX*5+cY][slot pair][replica]: (eP + tP) addends t1..t8
    double2 coef[4][2][W5_COPIES];     // [piece][(c3,c2) | (c1,c0)][replica]
    double boundary[4][5];             // start, ragged_start, end, ragged_end
};

// Forward tables.  Match lane: t1 = mi.M+eM0, t2 = mi.SX+eM1, t4 = mi.SY+eM2, t6 = mi.LX+eM3, t8 = mi.LY+eM4.
// Gap lane:   t1 = lo.M+eL0, t2 = lo.SX+eL1, t3 = lo.M+eL2, t4 = lo.LX+eL3, t5 = up.M+eU0, t6 = up.SY+eU1,
//             t7 = up.M+eU2, t8 = up.LY+eU3.
__device__ __forceinline__ double forward_table_entry(const DevModel &m, int role, int code, int slot) {
    const int cX = code / 5, cY = code % 5;
    if (role == 0) {
        switch (slot) {
            case 0: return m.etp_middle[code][0];
            case 1: return m.etp_middle[code][1];
            case 3: return m.etp_middle[code][2];
            case 5: return m.etp_middle[code][3];
            case 7: return m.etp_middle[code][4];
            default: return 0.0;
        }
    }
    return slot < 4 ? m.etp_lower[cX][slot] : m.etp_upper[cY][slot - 4];
}

// Backward tables (gather form, see backward_cell<5>).  code = X[x]*5 + Y[y], the symbols the neighbours consume.
// Match lane: t1 = nMi.M+eM0, t2 = nUp.SY+eU0, t4 = nUp.LY+eU2, t6 = nLo.SX+eL0, t8 = nLo.LX+eL2.
// Gap lane:   t1 = nMi.M+eM1, t2 = nLo.SX+eL1 (SX); t3 = nMi.M+eM2, t4 = nUp.SY+eU1 (SY);
//             t5 = nMi.M+eM3, t6 = nLo.LX+eL3 (LX); t7 = nMi.M+eM4, t8 = nUp.LY+eU3 (LY).
__device__ __forceinline__ double backward_table_entry(const DevModel &m, int role, int code, int slot) {
    const int cX = code / 5, cY = code % 5;
    const double *eM = m.etp_middle[code], *eL = m.etp_lower[cX], *eU = m.etp_upper[cY];
    if (role == 0) {
        switch (slot) {
            case 0: return eM[0];
            case 1: return eU[0];
            case 3: return eU[2];
            case 5: return eL[0];
            case 7: return eL[2];
            default: return 0.0;
        }
    }
    switch (slot) {
        case 0: return eM[1];
        case 1: return eL[1];
        case 2: return eM[2];
        case 3: return eU[1];
        case 4: return eM[3];
        case 5: return eL[3];
        case 6: return eM[4];
        default: return eU[3];
    }
}

template <bool FORWARD>
__device__ __forceinline__ void fill_wave5_shared(Wave5Shared &sh, const DevModel &m, int tid, int nthreads) {
    for (int i = tid; i < 4 * 2 * W5_COPIES; i += nthreads) {
        const int piece = i / (2 * W5_COPIES), half = (i / W5_COPIES) % 2, rep = i % W5_COPIES;
        sh.coef[piece][half][rep] = make_double2(c_la_coef[piece * 4 + half * 2], c_la_coef[piece * 4 + half * 2 + 1]);
    }
    for (int i = tid; i < 2 * 25 * 4 * W5_COPIES; i += nthreads) {
        const int role = i / (25 * 4 * W5_COPIES), code = (i / (4 * W5_COPIES)) % 25, sp = (i / W5_COPIES) % 4,
                  rep = i % W5_COPIES;
        const double a = FORWARD ? forward_table_entry(m, role, code, 2 * sp) : backward_table_entry(m, role, code, 2 * sp);
        const double b = FORWARD ? forward_table_entry(m, role, code, 2 * sp + 1)
                                 : backward_table_entry(m, role, code, 2 * sp + 1);
        sh.etp[role][code][sp][rep] = make_double2(a, b);
    }
    for (int i = tid; i < 20; i += nthreads) {
        const int which = i / 5, s = i % 5;
        const double *src = which == 0 ? m.start : which == 1 ? m.ragged_start : which == 2 ? m.end : m.ragged_end;
        sh.boundary[which][s] = src[s];
    }
}

// One ring slot = (ring_w + 2) cells: position 0 and position w+1 are -inf guards, cell i lives at position i+1.
// Position 0 is written once per warp (nothing else ever stores there); position w+1 after every diagonal.
__device__ __forceinline__ int clamp_pos(int i, int w) {  // cell index -> ring position, out-of-band -> a guard
    return min(max(i, -1), w) + 1;
}

__device__ __forceinline__ void write_high_guard(double *slot, int w, int lane) {
    if (lane < 5) slot[(w + 1) * 5 + lane] = neg_inf();
}

// ---- warp groups -------------------------------------------------------------------------------------------------
// G warps share one work item (region / traceback): warp gw of the group takes the 16-cell chunks gw, gw + G, ... of
// every diagonal, the ring lives once per group, and one named barrier per diagonal replaces the __syncwarp().  G = 1
// is the tRNA shape (12-13 cells per diagonal); wide bands (diagonalExpansion 50-500: 51-501 cells) use G = 2/4/8 so
// that a region with few, long, wide diagonals still fills the SM and a 60 KB ring is paid once per 8 warps.
template <int G>
__device__ __forceinline__ void group_sync(int group) {
    if (G == 1) __syncwarp();
    else asm volatile("bar.sync %0, %1;" ::"r"(group + 1), "n"(G * 32) : "memory");
}
template <int G>
__device__ __forceinline__ int next_item_group(int *counter, int lane, int gw, int group, volatile int *sh_item) {
    if (G == 1) return next_item(counter, lane);
    if (gw == 0 && lane == 0) sh_item[group] = atomicAdd(counter, 1);
    group_sync<G>(group);  // also separates the previous item's last reads of the ring from the next item's init
    const int item = sh_item[group];
    group_sync<G>(group);  // nobody re-writes sh_item before every warp of the group has read it
    return item;
}

// ------------------------------------------------------------------------------------------------ forward
// FULL = true : F holds all 5 states of every cell ([cell][5]); required by expectation mode.
// FULL = false: posterior mode -- F holds the match state of every cell, FT all states of the cells on the
//               diagonals where a total probability is evaluated (indexed through TotalDesc::full_off).
template <bool FULL, int CM, int G>
__global__ void __launch_bounds__(1024)
k_forward5(const __grid_constant__ DevModel model, Blob bl, int n_regions, double *__restrict__ F,
           double *__restrict__ FT, int *counter, int ring_w) {
    extern __shared__ __align__(16) unsigned char dyn_smem_raw[];
    __shared__ __align__(16) Wave5Shared sh;
    fill_wave5_shared<true>(sh, model, threadIdx.x, blockDim.x);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int role = lane >> 4;  // lanes 0-15 are match lanes, 16-31 gap lanes of cells 0..15 of a chunk: a 64-bit
                                 // shared access is served per half-warp, so each phase sees one role whose
                                 // 40-byte cell stride is conflict-free
    const bool chain = role == 0;
    const int jl = lane & 15;    // cell within a chunk of 16
    const int slot_doubles = (ring_w + 2) * 5;
    const int group = warp / G, gw = warp % G;  // G warps share a work item and its ring
    __shared__ int sh_item[16];
    double *ring = reinterpret_cast<double *>(dyn_smem_raw) + (size_t) group * 3 * slot_doubles;
    if (lane < 15) ring[(lane / 5) * slot_doubles + lane % 5] = neg_inf();  // low guards of the three slots
    __syncthreads();
    const double2 *etp_lane = &sh.etp[role][0][0][lane & (W5_COPIES - 1)];
    const unsigned coef_lane = (unsigned) __cvta_generic_to_shared(&sh.coef[0][0][lane & (W5_COPIES - 1)]);
    // per-lane state picks (see table comment): v4 = c1[role ? LX : SY], v6 = c2[role ? SY : LX]
    const int s4 = role ? LX : SY, s6 = role ? SY : LX;
    const int s_out4 = role ? LY : M;  // where r4 goes

    for (;;) {
        const int item = next_item_group<G>(counter, lane, gw, group, sh_item);
        if (item >= n_regions) break;
        const RegionDesc rg = bl.regions[bl.order_fwd[item]];
        const DiagDesc *dd = bl.diags + rg.diag_base;
        const int D = rg.lX + rg.lY;
        DiagDesc cur = dd[0], nxt = dd[1];
        int L1 = cur.xmyL, L2 = 0, w1 = (int) (nxt.cell_off - cur.cell_off), w2 = 0;
        double *slot = ring + slot_doubles, *slot1 = ring, *slot2 = ring + 2 * slot_doubles;  // d = 1: (d, d-1, d-2)
        // diagonal 0 (pairwiseAligner.c:658-659) goes to slot1; slot2 is the empty "d-2" of diagonal 1
        {
            const double *init = sh.boundary[(rg.flags & 1) ? 1 : 0];
            for (int i = gw * 32 + lane; i < w1 * 5; i += 32 * G) {
                const double v = init[i % 5];
                slot1[5 + i] = v;
                if (FULL) F[cur.cell_off * 5 + i] = v;
                else if (i % 5 == 0) F[cur.cell_off + i / 5] = v;
            }
            if (gw == 0) write_high_guard(slot1, w1, lane);
            if (gw == 0) write_high_guard(slot2, 0, lane);
        }
        // per-lane sequence pointers: cell (x, y) of diagonal d consumes X[x-1] and Y[y-1].  For x = 0 (y = 0) the
        // byte before the region is read and ignored: every term that uses it comes from an out-of-band neighbour.
        int x0 = (0 + L1) >> 1, y0 = (0 - L1) >> 1;  // coordinates of cell 0 of the previous diagonal
        // warp-uniform base pointers (shifted by the -1) + 32-bit per-lane offsets: one address add per load
        const uint8_t *sxb = bl.sym_x + rg.sym_x_off - 1, *syb = bl.sym_y + rg.sym_y_off - 1;
        unsigned xo = (unsigned) (x0 + jl), yo = (unsigned) (y0 - jl);
        group_sync<G>(group);
        for (int d = 1; d <= D; d++) {
            cur = nxt;
            nxt = dd[d + 1];
            const int L = cur.xmyL;
            const int w = (int) (nxt.cell_off - cur.cell_off);
            const int oLo = (L - 1 - L1) >> 1, oMi = (L - L2) >> 1;
            const int nx0 = (d + L) >> 1, ny0 = (d - L) >> 1;
            xo += (unsigned) (nx0 - x0);
            yo += (unsigned) (ny0 - y0);
            x0 = nx0;
            y0 = ny0;
            // per-lane neighbour geometry, branch-free: the match lane reads cell (d-2, xmy) twice, the gap lane the
            // lower (d-1, xmy-1) and upper (d-1, xmy+1) cells
            const double *nb_slot = chain ? slot2 : slot1;
            const int nb_off = chain ? oMi : oLo, nb_w = chain ? w2 : w1, nb_step = chain ? 0 : 1;
            double *f_match = F + cur.cell_off;  // posterior mode: the match plane
            for (int j0 = gw * 16; j0 < w; j0 += 16 * G) {
                const int j = j0 + jl;
                if (j < w) {
                    const int cX = __ldg(sxb + (xo + (unsigned) j0)), cY = __ldg(syb + (yo - (unsigned) j0));
                    const double2 *e = etp_lane + (cX * 5 + cY) * (4 * W5_COPIES);
                    const double *c1 = nb_slot + clamp_pos(j + nb_off, nb_w) * 5;
                    const double *c2 = nb_slot + clamp_pos(j + nb_off + nb_step, nb_w) * 5;
                    const double v1 = c1[M], v2 = c1[SX], v4 = c1[s4], v5 = c2[M], v6 = c2[s6], v8 = c2[LY];
                    const double2 e12 = e[0], e34 = e[W5_COPIES], e56 = e[2 * W5_COPIES], e78 = e[3 * W5_COPIES];
                    const Quad q = four_logadds<CM>(chain, __dadd_rn(v1, e12.x), __dadd_rn(v2, e12.y),
                                                    __dadd_rn(v1, e34.x), __dadd_rn(v4, e34.y), __dadd_rn(v5, e56.x),
                                                    __dadd_rn(v6, e56.y), __dadd_rn(v5, e78.x), __dadd_rn(v8, e78.y),
                                                    coef_lane);
                    double *out = slot + (j + 1) * 5;
                    out[s_out4] = q.r4;
                    if (!chain) {
                        out[SX] = q.r1;
                        out[LX] = q.r2;
                        out[SY] = q.r3;
                    }
                    if (!FULL && chain) f_match[j] = q.r4;  // 16 consecutive doubles per half-warp
                }
            }
            if (gw == 0) write_high_guard(slot, w, lane);
            group_sync<G>(group);
            // all five states leave through a coalesced copy of the finished ring slot: every diagonal in
            // expectation mode, only the diagonals that carry a total (1 in 10) in posterior mode
            if (FULL) {
                double *g = F + cur.cell_off * 5;
                for (int i = gw * 32 + lane; i < w * 5; i += 32 * G) g[i] = slot[5 + i];
            } else if (cur.tot_flags & MAB_DIAG_IS_TOTAL) {
                double *g = FT + (int64_t) bl.totals[cur.tot_flags & MAB_DIAG_TOT_MASK].full_off * 5;
                for (int i = gw * 32 + lane; i < w * 5; i += 32 * G) g[i] = slot[5 + i];
            }
            L2 = L1;
            L1 = L;
            w2 = w1;
            w1 = w;
            double *t = slot2;  // rotate (d, d-1, d-2) -> (d+1, d, d-1)
            slot2 = slot1;
            slot1 = slot;
            slot = t;
        }
    }
}

// ------------------------------------------------------------------------------------------------ backward
// One warp per traceback.  FULL as above (B / BT); `edge` receives the match state of diagonal from+1 of
// intermediate tracebacks (see TracebackDesc::edge_off).
template <bool FULL, int CM, int G>
__global__ void __launch_bounds__(1024)
k_backward5(const __grid_constant__ DevModel model, Blob bl, int n_tbs, double *__restrict__ B,
            double *__restrict__ BT, double *__restrict__ edge, int *counter, int ring_w) {
    extern __shared__ __align__(16) unsigned char dyn_smem_raw[];
    __shared__ __align__(16) Wave5Shared sh;
    fill_wave5_shared<false>(sh, model, threadIdx.x, blockDim.x);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int role = lane >> 4;
    const bool chain = role == 0;
    const int jl = lane & 15;
    const int slot_doubles = (ring_w + 2) * 5;
    const int group = warp / G, gw = warp % G;  // G warps share a work item and its ring
    __shared__ int sh_item[16];
    double *ring = reinterpret_cast<double *>(dyn_smem_raw) + (size_t) group * 3 * slot_doubles;
    if (lane < 15) ring[(lane / 5) * slot_doubles + lane % 5] = neg_inf();
    __syncthreads();
    const double2 *etp_lane = &sh.etp[role][0][0][lane & (W5_COPIES - 1)];
    const unsigned coef_lane = (unsigned) __cvta_generic_to_shared(&sh.coef[0][0][lane & (W5_COPIES - 1)]);
    const int s_out4 = role ? LY : M;
    // per-lane picks inside the neighbour cells, as byte-free state indices (see the table comment):
    //   match lane: v2 = nUp.SY v4 = nUp.LY v6 = nLo.SX v8 = nLo.LX;  gap lane: v2 = nLo.SX v4 = nUp.SY v6 = nLo.LX v8 = nUp.LY
    const int s2 = chain ? SY : SX, s4 = chain ? LY : SY, s6 = chain ? SX : LX, s8 = chain ? LX : LY;

    for (;;) {
        const int item = next_item_group<G>(counter, lane, gw, group, sh_item);
        if (item >= n_tbs) break;
        const TracebackDesc tb = bl.tbs[bl.order_bwd[item]];
        const RegionDesc rg = bl.regions[tb.region];
        const DiagDesc *dd = bl.diags + rg.diag_base;
        DiagDesc up1 = dd[tb.hi + 1], cur = dd[tb.hi];  // up1 = descriptor of diagonal d+1 while processing d
        int L1 = cur.xmyL, L2 = 0, w1 = (int) (up1.cell_off - cur.cell_off), w2 = 0;
        double *slot = ring + slot_doubles, *slot1 = ring, *slot2 = ring + 2 * slot_doubles;  // (d, d+1, d+2)
        // diagonal hi: (ragged)endStateProb on every cell (pairwiseAligner.c:680-681), into slot1
        {
            const double *init = sh.boundary[(tb.at_end && (rg.flags & 2)) ? 3 : 2];
            const bool keep = tb.hi <= tb.from;  // only the final traceback keeps its top diagonal
            double *bt = nullptr;
            if (!FULL && keep && (cur.tot_flags & MAB_DIAG_IS_TOTAL))
                bt = BT + (int64_t) bl.totals[cur.tot_flags & MAB_DIAG_TOT_MASK].full_off * 5;
            for (int i = gw * 32 + lane; i < w1 * 5; i += 32 * G) {
                const double v = init[i % 5];
                slot1[5 + i] = v;
                if (keep) {
                    if (FULL) B[cur.cell_off * 5 + i] = v;
                    else {
                        if (i % 5 == 0) B[cur.cell_off + i / 5] = v;
                        if (bt != nullptr) bt[i] = v;
                    }
                }
            }
            if (gw == 0) write_high_guard(slot1, w1, lane);
            if (gw == 0) write_high_guard(slot2, 0, lane);  // the empty "d+2" of diagonal hi-1
        }
        // the neighbours of target cell (x, y) consume X[x] and/or Y[y] (0-based); at x = lX (y = lY) the byte after
        // the region is read and ignored: the neighbours that would use it are out of band (-inf)
        int x0 = (tb.hi + L1) >> 1, y0 = (tb.hi - L1) >> 1;
        const uint8_t *sxb = bl.sym_x + rg.sym_x_off, *syb = bl.sym_y + rg.sym_y_off;  // warp-uniform bases
        unsigned xo = (unsigned) (x0 + jl), yo = (unsigned) (y0 - jl);                 // 32-bit per-lane offsets
        group_sync<G>(group);
        for (int d = tb.hi - 1; d > tb.lo; d--) {
            up1 = cur;
            cur = dd[d];
            const int L = cur.xmyL;
            const int w = (int) (up1.cell_off - cur.cell_off);
            const int oUp = (L - 1 - L1) >> 1;  // (d+1, xmy-1): the cell whose "upper" is the target
            const int oMi = (L - L2) >> 1;      // (d+2, xmy)
            const int nx0 = (d + L) >> 1, ny0 = (d - L) >> 1;
            xo += (unsigned) (nx0 - x0);
            yo += (unsigned) (ny0 - y0);
            x0 = nx0;
            y0 = ny0;
            const bool keep = d <= tb.from;
            // where the match lanes put their result: the match plane of B (posterior mode, kept diagonals), the
            // edge buffer (first discarded diagonal of an intermediate traceback, both modes), or nowhere
            double *b_match = d == tb.from + 1 ? edge + tb.edge_off : ((!FULL && keep) ? B + cur.cell_off : nullptr);
            for (int j0 = gw * 16; j0 < w; j0 += 16 * G) {
                const int j = j0 + jl;
                if (j < w) {
                    const int cX = __ldg(sxb + (xo + (unsigned) j0)), cY = __ldg(syb + (yo - (unsigned) j0));
                    const double2 *e = etp_lane + (cX * 5 + cY) * (4 * W5_COPIES);
                    const int iUp = j + oUp;
                    const double *nUp = slot1 + clamp_pos(iUp, w1) * 5;
                    const double *nLo = slot1 + clamp_pos(iUp + 1, w1) * 5;
                    const double *nMi = slot2 + clamp_pos(j + oMi, w2) * 5;
                    // match lane: (nMi.M, nUp.SY, -, nUp.LY, -, nLo.SX, -, nLo.LX)
                    // gap lane:   (nMi.M, nLo.SX, nMi.M, nUp.SY, nMi.M, nLo.LX, nMi.M, nUp.LY)
                    const double *p2 = chain ? nUp : nLo, *p4 = nUp, *p6 = nLo, *p8 = chain ? nLo : nUp;
                    const double vm = nMi[M], v2 = p2[s2], v4 = p4[s4], v6 = p6[s6], v8 = p8[s8];
                    const double2 e12 = e[0], e34 = e[W5_COPIES], e56 = e[2 * W5_COPIES], e78 = e[3 * W5_COPIES];
                    const Quad q = four_logadds<CM>(chain, __dadd_rn(vm, e12.x), __dadd_rn(v2, e12.y),
                                                    __dadd_rn(vm, e34.x), __dadd_rn(v4, e34.y), __dadd_rn(vm, e56.x),
                                                    __dadd_rn(v6, e56.y), __dadd_rn(vm, e78.x), __dadd_rn(v8, e78.y),
                                                    coef_lane);
                    // gap lane results: r1 = SX, r2 = SY, r3 = LX, r4 = LY
                    double *out = slot + (j + 1) * 5;
                    out[s_out4] = q.r4;
                    if (!chain) {
                        out[SX] = q.r1;
                        out[SY] = q.r2;
                        out[LX] = q.r3;
                    }
                    if (chain && b_match != nullptr) b_match[j] = q.r4;
                }
            }
            if (gw == 0) write_high_guard(slot, w, lane);
            group_sync<G>(group);
            if (keep) {  // all five states: coalesced copy of the finished ring slot (see k_forward5)
                if (FULL) {
                    double *g = B + cur.cell_off * 5;
                    for (int i = gw * 32 + lane; i < w * 5; i += 32 * G) g[i] = slot[5 + i];
                } else if (cur.tot_flags & MAB_DIAG_IS_TOTAL) {
                    double *g = BT + (int64_t) bl.totals[cur.tot_flags & MAB_DIAG_TOT_MASK].full_off * 5;
                    for (int i = gw * 32 + lane; i < w * 5; i += 32 * G) g[i] = slot[5 + i];
                }
            }
            L2 = L1;
            L1 = L;
            w2 = w1;
            w1 = w;
            double *t = slot2;  // rotate (d, d+1, d+2) -> (d-1, d, d+1)
            slot2 = slot1;
            slot1 = slot;
            slot = t;
        }
    }
}
