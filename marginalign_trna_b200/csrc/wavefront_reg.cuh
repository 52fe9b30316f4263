// wavefront_reg.cuh -- register-resident forward / backward wavefront kernels for narrow bands (band classes 0 and 1,
// mab_internal.h: every tRNA read, any read at diagonalExpansion 10).  Included by kernels.cu.
//
// Mapping.  A group of LPI lanes (a half-warp for class 0, a warp for class 1) owns one work item (a split region in
// the forward sweep, a traceback in the backward sweep).  The x-y coordinate of a cell picks its lane CIRCULARLY,
//     lane(xmy) = (xmy >> 1) mod LPI                      (arithmetic shift: 2k and 2k + 1 share lane k)
// so a lane alternates between x-y = 2k (even anti-diagonals) and 2k + 1 (odd ones), and the band may drift along the
// read without ever being re-based: it wraps around the lane ring.  The two previous anti-diagonals live in the
// registers of the lane that computed them:
//     middle neighbour (d-2, xmy)      own lane, two steps ago
//     even d: upper (d-1, xmy+1) own lane, lower (d-1, xmy-1) lane k-1        (one shuffle of the states it feeds)
//     odd  d: lower (d-1, xmy-1) own lane, upper (d-1, xmy+1) lane k+1
// The loop is unrolled over the two parities, so which neighbour is shuffled is static and the two register cells
// swap roles without moves.  A lane whose cell is outside the band looks up the "out of band" symbol 5, whose emission
// rows are -inf: every addend becomes -inf and the branch-free logAdd returns -inf, so the lane stores -inf for its
// neighbours without a single select (the reference skips transitions from cells outside the band,
// pairwiseAligner.c:419-425, 527-554, which is the same as adding log(0)).  The plan guarantees (band span <= 2 LPI - 2)
// that live cells never alias on the ring.
//
// (eP + tP) is formed here from the emission (one conflict-free LDS.64 per symbol pair) and the transition (an
// immediate constant-bank operand): the same double addition the host used to tabulate, so the sums are bit-identical
// to the reference's `eP + tP` (pairwiseAligner.c:314, 324), and 13 table loads per cell leave the shared-memory pipe,
// which bounds these kernels (profiles/r1_ncu_summary.md r1h).
#pragma once

#define RR_REPL 16  // replicas of the emission tables: lane l reads replica l & 15 -> LDS.64 is conflict-free

struct RegTables {
    double em[31][RR_REPL];  // match emission by cX * 5 + cY (symbols 0..4); 25..30 = -inf (cX = cY = 5: out of band)
    double ex[6][RR_REPL];   // gap-X emission by cX; [5] = -inf
    double ey[6][RR_REPL];   // gap-Y emission by cY; [5] = -inf
};

__device__ __forceinline__ void fill_reg_tables(RegTables &t, const DevModel &m, int tid, int nthreads) {
    const double NINF = neg_inf();
    for (int i = tid; i < 31 * RR_REPL; i += nthreads) t.em[i / RR_REPL][i % RR_REPL] = i / RR_REPL < 25 ? m.em_match[i / RR_REPL] : NINF;
    for (int i = tid; i < 6 * RR_REPL; i += nthreads) {
        t.ex[i / RR_REPL][i % RR_REPL] = i / RR_REPL < 5 ? m.em_gap_x[i / RR_REPL] : NINF;
        t.ey[i / RR_REPL][i % RR_REPL] = i / RR_REPL < 5 ? m.em_gap_y[i / RR_REPL] : NINF;
    }
}

__device__ __forceinline__ double lds_f64(unsigned addr) {
    double v;
    asm("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}

__device__ __forceinline__ double shfl_f64(unsigned mask, double v, int src) {
    return __hiloint2double(__shfl_sync(mask, __double2hiint(v), src), __shfl_sync(mask, __double2loint(v), src));
}

// what a lane needs to know about its group
struct RegLane {
    int lane;          // inside the group
    int src_dn, src_up;  // warp lanes of the ring neighbours k-1 / k+1
    unsigned gmask;
    CoefAddr cf;
    unsigned em_addr, ex_addr, ey_addr;  // shared addresses of this lane's replica of row 0
};

template <int LPI>
__device__ __forceinline__ RegLane reg_lane(const RegTables &t, const CoefTable &ct) {
    RegLane r;
    const int lane32 = threadIdx.x & 31;
    r.lane = lane32 & (LPI - 1);
    const int base = lane32 & ~(LPI - 1);
    r.src_dn = base | ((r.lane - 1) & (LPI - 1));
    r.src_up = base | ((r.lane + 1) & (LPI - 1));
    r.gmask = LPI == 32 ? FULL_MASK : (0xffffu << (lane32 & 16));
    r.cf = (unsigned) __cvta_generic_to_shared(&ct.coef[0][0][lane32 & (W5_COPIES - 1)]);
    r.em_addr = (unsigned) __cvta_generic_to_shared(&t.em[0][lane32 & (RR_REPL - 1)]);
    r.ex_addr = (unsigned) __cvta_generic_to_shared(&t.ex[0][lane32 & (RR_REPL - 1)]);
    r.ey_addr = (unsigned) __cvta_generic_to_shared(&t.ey[0][lane32 & (RR_REPL - 1)]);
    return r;
}

// One diagonal's geometry for a lane: its cell index j inside the diagonal (j < w: in band) and the cell coordinates.
struct RegGeom {
    int j, x, y;
    bool active;
};
template <int LPI>
__device__ __forceinline__ RegGeom reg_geom(int lane, int d, int L, int w) {
    RegGeom g;
    g.j = (lane - (L >> 1)) & (LPI - 1);
    g.active = g.j < w;
    const int xmy = L + 2 * g.j;
    g.x = (d + xmy) >> 1;
    g.y = d - g.x;
    return g;
}

// logAdd for the register kernels.  LA = 0: la_bf (compare, swap, subtract).  LA = 1: d = |x - y| straight from one
// subtraction (round-to-nearest is symmetric, so |fl(x - y)| == fl(max - min) bit for bit) and max / min picked by the
// sign bit of the difference: the compare-and-swap leaves the dependency chain in front of the polynomial and one
// DSETP leaves the FP64 pipe (tools/la_probe.cu: 9.0 instead of 9.6 SM cycles per warp-logAdd at saturation).
// x == y gives +0 (max = x as in the reference); -inf - -inf gives NaN: either operand is the right result (-inf).
template <int LA>
__device__ __forceinline__ double la_rr(double x, double y, unsigned coef_addr) {
    if (LA == 0) return la_bf<COEF_SMEM>(x, y, coef_addr);
    const double df = __dsub_rn(x, y);
    const bool lt = __double2hiint(df) < 0;
    const double hi = lt ? y : x, lo = lt ? x : y;
    const double d = fabs(df);
    unsigned a = coef_addr;
    if (d > 1.00) a += 2 * W5_COPIES * 16;
    if (d > 2.50) a += 2 * W5_COPIES * 16;
    if (d > 4.50) a += 2 * W5_COPIES * 16;
    const double2 c32 = lds_v2(a), c10 = lds_v2_128(a);
    double r = __dadd_rn(__dmul_rn(c32.x, d), c32.y);
    r = __dadd_rn(__dmul_rn(r, d), c10.x);
    r = __dadd_rn(__dmul_rn(r, d), c10.y);
    r = __dadd_rn(r, lo);
    return d < 7.5 ? r : hi;
}
#define RR_LG(x, y) la_rr<LA>((x), (y), ln.cf)

// ---- forward cell from register neighbours.  `p1` = this lane's cell on diagonal d-1, `p2` = on d-2 (overwritten with
// the new cell).  Q = d & 1.  Transition order as forward_cell<S> (stateMachine.c:399-429, 636-661).
template <int S, int Q, int LA>
__device__ __forceinline__ void reg_forward_cell(const DevModel &m, const RegLane &ln, const Cell<S> &p1, Cell<S> &p2,
                                                 double em, double ex, double ey) {
    const int src = Q ? ln.src_up : ln.src_dn;
    if constexpr (S == 5) {
        // the shuffled neighbour feeds M + two gap states; the own-lane neighbour the other two
        const double sM = shfl_f64(ln.gmask, p1.v[M], src);
        const double sA = shfl_f64(ln.gmask, p1.v[Q ? SY : SX], src);
        const double sB = shfl_f64(ln.gmask, p1.v[Q ? LY : LX], src);
        const double loM = Q ? p1.v[M] : sM, loSX = Q ? p1.v[SX] : sA, loLX = Q ? p1.v[LX] : sB;
        const double upM = Q ? sM : p1.v[M], upSY = Q ? sA : p1.v[SY], upLY = Q ? sB : p1.v[LY];
        Cell<5> c;
        c.v[SX] = RR_LG(__dadd_rn(loM, __dadd_rn(ex, m.tp[0][0])), __dadd_rn(loSX, __dadd_rn(ex, m.tp[0][1])));
        c.v[LX] = RR_LG(__dadd_rn(loM, __dadd_rn(ex, m.tp[0][2])), __dadd_rn(loLX, __dadd_rn(ex, m.tp[0][3])));
        double t = RR_LG(__dadd_rn(p2.v[M], __dadd_rn(em, m.tp[1][0])), __dadd_rn(p2.v[SX], __dadd_rn(em, m.tp[1][1])));
        t = RR_LG(t, __dadd_rn(p2.v[SY], __dadd_rn(em, m.tp[1][2])));
        t = RR_LG(t, __dadd_rn(p2.v[LX], __dadd_rn(em, m.tp[1][3])));
        c.v[M] = RR_LG(t, __dadd_rn(p2.v[LY], __dadd_rn(em, m.tp[1][4])));
        c.v[SY] = RR_LG(__dadd_rn(upM, __dadd_rn(ey, m.tp[2][0])), __dadd_rn(upSY, __dadd_rn(ey, m.tp[2][1])));
        c.v[LY] = RR_LG(__dadd_rn(upM, __dadd_rn(ey, m.tp[2][2])), __dadd_rn(upLY, __dadd_rn(ey, m.tp[2][3])));
        p2 = c;
    } else {
        const double sM = shfl_f64(ln.gmask, p1.v[M], src);
        const double sX = shfl_f64(ln.gmask, p1.v[SX], src);
        const double sY = shfl_f64(ln.gmask, p1.v[SY], src);
        const double loM = Q ? p1.v[M] : sM, loSX = Q ? p1.v[SX] : sX, loSY = Q ? p1.v[SY] : sY;
        const double upM = Q ? sM : p1.v[M], upSX = Q ? sX : p1.v[SX], upSY = Q ? sY : p1.v[SY];
        Cell<3> c;
        c.v[SX] = RR_LG(RR_LG(__dadd_rn(loM, __dadd_rn(ex, m.tp[0][0])), __dadd_rn(loSX, __dadd_rn(ex, m.tp[0][1]))),
                        __dadd_rn(loSY, __dadd_rn(ex, m.tp[0][2])));
        c.v[M] = RR_LG(RR_LG(__dadd_rn(p2.v[M], __dadd_rn(em, m.tp[1][0])), __dadd_rn(p2.v[SX], __dadd_rn(em, m.tp[1][1]))),
                       __dadd_rn(p2.v[SY], __dadd_rn(em, m.tp[1][2])));
        c.v[SY] = RR_LG(RR_LG(__dadd_rn(upM, __dadd_rn(ey, m.tp[2][0])), __dadd_rn(upSY, __dadd_rn(ey, m.tp[2][1]))),
                        __dadd_rn(upSX, __dadd_rn(ey, m.tp[2][2])));
        p2 = c;
    }
}

// ---- backward cell (gather form, backward_cell<S> in kernels.cu).  `p1` = this lane's cell on d+1, `p2` = on d+2
// (only its match state is read; overwritten with the new cell).  nUp = (d+1, xmy-1), nLo = (d+1, xmy+1).
template <int S, int Q, int LA>
__device__ __forceinline__ void reg_backward_cell(const DevModel &m, const RegLane &ln, const Cell<S> &p1, Cell<S> &p2,
                                                  double em, double ex, double ey) {
    const int src = Q ? ln.src_up : ln.src_dn;
    const double miM = p2.v[M];
    if constexpr (S == 5) {
        // even d: nUp lives in lane k-1 (shuffle its SY, LY), nLo in this lane; odd d: nLo in lane k+1 (SX, LX)
        const double sA = shfl_f64(ln.gmask, p1.v[Q ? SX : SY], src);
        const double sB = shfl_f64(ln.gmask, p1.v[Q ? LX : LY], src);
        const double upSY = Q ? p1.v[SY] : sA, upLY = Q ? p1.v[LY] : sB;
        const double loSX = Q ? sA : p1.v[SX], loLX = Q ? sB : p1.v[LX];
        const double mM = __dadd_rn(miM, __dadd_rn(em, m.tp[1][0]));
        Cell<5> c;
        double t = RR_LG(mM, __dadd_rn(upSY, __dadd_rn(ey, m.tp[2][0])));
        t = RR_LG(t, __dadd_rn(upLY, __dadd_rn(ey, m.tp[2][2])));
        t = RR_LG(t, __dadd_rn(loSX, __dadd_rn(ex, m.tp[0][0])));
        c.v[M] = RR_LG(t, __dadd_rn(loLX, __dadd_rn(ex, m.tp[0][2])));
        c.v[SX] = RR_LG(__dadd_rn(miM, __dadd_rn(em, m.tp[1][1])), __dadd_rn(loSX, __dadd_rn(ex, m.tp[0][1])));
        c.v[SY] = RR_LG(__dadd_rn(miM, __dadd_rn(em, m.tp[1][2])), __dadd_rn(upSY, __dadd_rn(ey, m.tp[2][1])));
        c.v[LX] = RR_LG(__dadd_rn(miM, __dadd_rn(em, m.tp[1][3])), __dadd_rn(loLX, __dadd_rn(ex, m.tp[0][3])));
        c.v[LY] = RR_LG(__dadd_rn(miM, __dadd_rn(em, m.tp[1][4])), __dadd_rn(upLY, __dadd_rn(ey, m.tp[2][3])));
        p2 = c;
    } else {
        const double s = shfl_f64(ln.gmask, p1.v[Q ? SX : SY], src);
        const double upSY = Q ? p1.v[SY] : s, loSX = Q ? s : p1.v[SX];
        // lower: M->X(0) X->X(1) Y->X(2); middle: M->M(0) X->M(1) Y->M(2); upper: M->Y(0) Y->Y(1) X->Y(2)
        Cell<3> c;
        c.v[M] = RR_LG(RR_LG(__dadd_rn(miM, __dadd_rn(em, m.tp[1][0])), __dadd_rn(upSY, __dadd_rn(ey, m.tp[2][0]))),
                       __dadd_rn(loSX, __dadd_rn(ex, m.tp[0][0])));
        c.v[SX] = RR_LG(RR_LG(__dadd_rn(miM, __dadd_rn(em, m.tp[1][1])), __dadd_rn(upSY, __dadd_rn(ey, m.tp[2][2]))),
                        __dadd_rn(loSX, __dadd_rn(ex, m.tp[0][1])));
        c.v[SY] = RR_LG(RR_LG(__dadd_rn(miM, __dadd_rn(em, m.tp[1][2])), __dadd_rn(upSY, __dadd_rn(ey, m.tp[2][1]))),
                        __dadd_rn(loSX, __dadd_rn(ex, m.tp[0][2])));
        p2 = c;
    }
}
#undef RR_LG

#define RR_THREADS 128

// ------------------------------------------------------------------------------------------------ staging (N1)
// Band descriptors: the DiagDesc entries of the item being swept are staged in shared memory by the TMA engine
// (cp.async.bulk, completion on an mbarrier) in windows of RR_DWIN diagonals, two windows per group: while the sweep
// reads one window the next one (in sweep direction) is in flight, so no global load sits on the sequential path.
// Symbols: the sequences live in HBM as 2-bit codes + an N mask (BlobHeader::off_pack_x); every RR_SYM_REFRESH
// diagonals the group unpacks the 128 bases its band can reach in that stretch (the band edges move by at most one
// base per diagonal) into a byte ring with coalesced 32-bit loads, and the per-cell symbol fetch is one LDS.U8.
#define RR_DWIN 64
#define RR_SYM_RING 128
#define RR_SYM_REFRESH 32

struct __align__(16) RegStage {
    DiagDesc ring[2 * RR_DWIN];
    unsigned long long bar[2];
    uint8_t sx[RR_SYM_RING], sy[RR_SYM_RING];
};

__device__ __forceinline__ void mbar_init(unsigned bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void *src, unsigned bytes, unsigned bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    unsigned done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                     : "=r"(done)
                     : "r"(bar), "r"(parity)
                     : "memory");
    } while (!done);
}

// The staged descriptor stream of one item.  Entries are read in one direction (DIR = +1 forward, -1 backward), one
// per diagonal; window k = entries [64 k, 64 k + 64) lives in slot k & 1.
template <int LPI, int DIR>
struct DescStream {
    const DiagDesc *g;      // region's first entry in global memory
    int n_ent;              // entries of the region that exist (lX + lY + 2)
    int k_last;             // last window the sweep will enter
    unsigned ring, bar;     // shared addresses of ring[0] / bar[0]
    unsigned gmask;
    unsigned phase;         // bit s = parity the next wait on slot s uses
    bool leader;

    __device__ __forceinline__ void issue(int k) {
        if (leader) {
            const int first = k * RR_DWIN, cnt = min(RR_DWIN, n_ent - first);
            const unsigned slot = k & 1, bytes = (unsigned) cnt * (unsigned) sizeof(DiagDesc);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // earlier generic reads of the slot vs the async write
            mbar_expect_tx(bar + 8 * slot, bytes);
            bulk_g2s(ring + slot * RR_DWIN * (unsigned) sizeof(DiagDesc), g + first, bytes, bar + 8 * slot);
        }
    }
    __device__ __forceinline__ void wait(int k) {
        const unsigned slot = k & 1;
        mbar_wait(bar + 8 * slot, (phase >> slot) & 1);
        phase ^= 1u << slot;
    }
    // first entry index `i0`; the sweep will read down / up to entry `i_end` inclusive
    __device__ __forceinline__ void open(const DiagDesc *g_, int n_ent_, int i0, int i_end) {
        g = g_;
        n_ent = n_ent_;
        k_last = i_end / RR_DWIN;
        const int k0 = i0 / RR_DWIN;
        __syncwarp(gmask);  // every lane is done with the previous item's windows
        issue(k0);
        if (k0 != k_last) issue(k0 + DIR);
        wait(k0);
    }
    // entry i; must be called with i moving by DIR each time (any entry may be read again while its window is current)
    __device__ __forceinline__ DiagDesc get(int i, bool entering) {
        if (entering) {  // first entry of a new window in sweep direction
            const int k = i / RR_DWIN;
            wait(k);
            __syncwarp(gmask);  // the other slot (window k - DIR) is fully consumed by every lane
            if (k != k_last) issue(k + DIR);
        }
        DiagDesc d;
        const unsigned a = ring + (unsigned) (i & (2 * RR_DWIN - 1)) * (unsigned) sizeof(DiagDesc);
        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(((uint32_t *) &d)[0]), "=r"(((uint32_t *) &d)[1]), "=r"(((uint32_t *) &d)[2]),
                       "=r"(((uint32_t *) &d)[3])
                     : "r"(a));
        return d;
    }
};

// Unpacks bases [b0, b0 + 128) of a region (b0 a multiple of 8; region-local, valid range [0, len)) into a byte ring
// (base b at ring[b & 127]; 4 = N or outside the sequence).  pack: the region's sequence array (BlobHeader::off_pack_x),
// off: global index of the region's base 0.
template <int LPI>
__device__ __forceinline__ void unpack_symbols(const uint32_t *__restrict__ pack, int64_t cw, int64_t off, int len, int b0,
                                               int lane, unsigned ring) {
    constexpr int K = RR_SYM_RING / LPI;  // bases per lane: 8 (half-warp) or 4 (warp)
    const int b = b0 + lane * K;
    const int64_t p = off + b + MAB_PACK_PAD_BASES;  // position in the padded arrays (>= 0: b0 >= -MAB_PACK_PAD_BASES)
    const uint32_t c0 = pack[p >> 4], c1 = pack[(p >> 4) + 1];
    const uint32_t m0 = pack[cw + (p >> 5)], m1 = pack[cw + (p >> 5) + 1];
    const uint32_t v = __funnelshift_r(c0, c1, (unsigned) (p & 15) * 2);
    uint32_t n = __funnelshift_r(m0, m1, (unsigned) (p & 31));
    const int lo = min(max(-b, 0), K), hi = min(max(len - b, 0), K);
    const uint32_t valid = hi > lo ? (((1u << hi) - 1u) & ~((1u << lo) - 1u)) : 0u;
    n = (n | ~valid) & ((1u << K) - 1u);
    uint32_t out[K / 4];
#pragma unroll
    for (int q = 0; q < K / 4; q++) {
        uint32_t t = (v >> (8 * q)) & 0xFFu;
        t = (t | (t << 12)) & 0x000F000Fu;
        t = (t | (t << 6)) & 0x03030303u;
        const uint32_t nb = (((n >> (4 * q)) & 0xFu) * 0x00204081u) & 0x01010101u;
        out[q] = (t & ~(nb * 3u)) | (nb << 2);
    }
    const unsigned a = ring + (unsigned) (b & (RR_SYM_RING - 1));
    if (K == 8) asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(a), "r"(out[0]), "r"(out[K / 4 - 1]) : "memory");
    else asm volatile("st.shared.b32 [%0], %1;" ::"r"(a), "r"(out[0]) : "memory");
}

__device__ __forceinline__ int lds_u8(unsigned addr) {
    unsigned v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
    return (int) v;
}

// per-block setup of the staging area: one mbarrier pair per group
template <int LPI>
__device__ __forceinline__ void stage_init(RegStage *stage) {
    constexpr int GROUPS = RR_THREADS / LPI;
    if (threadIdx.x < 2 * GROUPS)
        mbar_init((unsigned) __cvta_generic_to_shared(&stage[threadIdx.x >> 1].bar[threadIdx.x & 1]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

// ------------------------------------------------------------------------------------------------ forward
// Work items: order_fwd[item0 .. item1).  Outputs exactly what k_forward writes (F match plane + FT on total diagonals,
// or all states when FULL).
template <int S, bool FULL, int LPI, int LA, int MINB>
__global__ void __launch_bounds__(RR_THREADS, MINB)
k_forward_rr(const __grid_constant__ DevModel model, Blob bl, int item0, int item1, double *__restrict__ F,
             double *__restrict__ FT, int *counter) {
    __shared__ RegTables tb;
    __shared__ __align__(16) CoefTable ct;
    __shared__ RegStage stage[RR_THREADS / LPI];
    fill_reg_tables(tb, model, threadIdx.x, blockDim.x);
    fill_coef_table(ct, threadIdx.x, blockDim.x);
    stage_init<LPI>(stage);
    __syncthreads();
    const RegLane ln = reg_lane<LPI>(tb, ct);
    const int lane32 = threadIdx.x & 31;
    const double NINF = neg_inf();
    RegStage &sg = stage[threadIdx.x / LPI];
    DescStream<LPI, +1> ds;
    ds.ring = (unsigned) __cvta_generic_to_shared(&sg.ring[0]);
    ds.bar = (unsigned) __cvta_generic_to_shared(&sg.bar[0]);
    ds.gmask = ln.gmask;
    ds.phase = 0;
    ds.leader = ln.lane == 0;
    const unsigned sxr = (unsigned) __cvta_generic_to_shared(&sg.sx[0]), syr = (unsigned) __cvta_generic_to_shared(&sg.sy[0]);
    const int64_t cwx = mab_pack_code_words(bl.h_len_sym_x), cwy = mab_pack_code_words(bl.h_len_sym_y);

    for (;;) {
        const int item = item0 + next_item_sub<LPI>(counter, lane32, ln.gmask);
        if (item >= item1) break;
        const RegionDesc rg = bl.regions[bl.order_fwd[item]];
        const int D = rg.lX + rg.lY;
        ds.open(bl.diags + rg.diag_base, D + 2, 0, D + 1);
        DiagDesc cur = ds.get(0, false), nxt = ds.get(1, false);
        Cell<S> a, b = cell_fill<S>(NINF);
        {  // diagonal 0: (ragged)startStateProb on every cell, pairwiseAligner.c:658-659
            const double *init = (rg.flags & 1) ? model.ragged_start : model.start;
            const RegGeom g = reg_geom<LPI>(ln.lane, 0, cur.xmyL, (int) (nxt.cell_off - cur.cell_off));
#pragma unroll
            for (int s = 0; s < S; s++) a.v[s] = g.active ? init[s] : NINF;
            if (g.active) {
                if (FULL) cell_store<S>(F + (cur.cell_off + g.j) * S, a);
                else F[cur.cell_off + g.j] = a.v[M];
            }
        }
        // one diagonal: p1 = diagonal d-1, p2 = diagonal d-2, replaced by diagonal d
        auto step = [&](auto q_tag, int d, const Cell<S> &p1, Cell<S> &p2) {
            constexpr int Q = decltype(q_tag)::value;
            cur = nxt;
            nxt = ds.get(d + 1, ((d + 1) & (RR_DWIN - 1)) == 0);
            const int w = (int) (nxt.cell_off - cur.cell_off);
            if (((d - 1) & (RR_SYM_REFRESH - 1)) == 0) {
                // the cells of diagonals d .. d + 31 consume X[x - 1], Y[y - 1] with x >= x_lo(d), y >= y_lo(d) and
                // at most 31 + band width beyond them: refresh the rings from one base before those
                const int x_lo = (d + cur.xmyL) >> 1, y_lo = (d - (cur.xmyL + 2 * (w - 1))) >> 1;
                __syncwarp(ln.gmask);
                unpack_symbols<LPI>(bl.pack_x, cwx, rg.sym_x_off, rg.lX, (x_lo - 1) & ~7, ln.lane, sxr);
                unpack_symbols<LPI>(bl.pack_y, cwy, rg.sym_y_off, rg.lY, (y_lo - 1) & ~7, ln.lane, syr);
                __syncwarp(ln.gmask);
            }
            const RegGeom g = reg_geom<LPI>(ln.lane, d, cur.xmyL, w);
            int cX = 5, cY = 5;
            if (g.active) {  // position -1 (x = 0 or y = 0) unpacks as N, like the reference's `x > 0 ? sX[x-1] : n`
                cX = lds_u8(sxr + ((g.x - 1) & (RR_SYM_RING - 1)));
                cY = lds_u8(syr + ((g.y - 1) & (RR_SYM_RING - 1)));
            }
            const double em = lds_f64(ln.em_addr + (cX * 5 + cY) * (RR_REPL * 8));
            const double ex = lds_f64(ln.ex_addr + cX * (RR_REPL * 8));
            const double ey = lds_f64(ln.ey_addr + cY * (RR_REPL * 8));
            reg_forward_cell<S, Q, LA>(model, ln, p1, p2, em, ex, ey);
            if (g.active) {
                const int64_t c = cur.cell_off + g.j;
                if (FULL) cell_store<S>(F + c * S, p2);
                else {
                    F[c] = p2.v[M];
                    if (cur.tot_flags & MAB_DIAG_IS_TOTAL)
                        cell_store<S>(FT + ((int64_t) bl.totals[cur.tot_flags & MAB_DIAG_TOT_MASK].full_off + g.j) * S, p2);
                }
            }
        };
        int d = 1;
        for (; d + 1 <= D; d += 2) {
            step(std::integral_constant<int, 1>(), d, a, b);
            step(std::integral_constant<int, 0>(), d + 1, b, a);
        }
        if (d <= D) step(std::integral_constant<int, 1>(), d, a, b);
    }
}

// ------------------------------------------------------------------------------------------------ backward
// Work items: order_bwd[item0 .. item1).  Outputs exactly what k_backward writes (B, BT, edge).
template <int S, bool FULL, int LPI, int LA, int MINB>
__global__ void __launch_bounds__(RR_THREADS, MINB)
k_backward_rr(const __grid_constant__ DevModel model, Blob bl, int item0, int item1, double *__restrict__ B,
              double *__restrict__ BT, double *__restrict__ edge, int *counter) {
    __shared__ RegTables tb;
    __shared__ __align__(16) CoefTable ct;
    __shared__ RegStage stage[RR_THREADS / LPI];
    fill_reg_tables(tb, model, threadIdx.x, blockDim.x);
    fill_coef_table(ct, threadIdx.x, blockDim.x);
    stage_init<LPI>(stage);
    __syncthreads();
    const RegLane ln = reg_lane<LPI>(tb, ct);
    const int lane32 = threadIdx.x & 31;
    const double NINF = neg_inf();
    RegStage &sg = stage[threadIdx.x / LPI];
    DescStream<LPI, -1> ds;
    ds.ring = (unsigned) __cvta_generic_to_shared(&sg.ring[0]);
    ds.bar = (unsigned) __cvta_generic_to_shared(&sg.bar[0]);
    ds.gmask = ln.gmask;
    ds.phase = 0;
    ds.leader = ln.lane == 0;
    const unsigned sxr = (unsigned) __cvta_generic_to_shared(&sg.sx[0]), syr = (unsigned) __cvta_generic_to_shared(&sg.sy[0]);
    const int64_t cwx = mab_pack_code_words(bl.h_len_sym_x), cwy = mab_pack_code_words(bl.h_len_sym_y);

    for (;;) {
        const int item = item0 + next_item_sub<LPI>(counter, lane32, ln.gmask);
        if (item >= item1) break;
        const TracebackDesc t = bl.tbs[bl.order_bwd[item]];
        const RegionDesc rg = bl.regions[t.region];
        const int lX = rg.lX, lY = rg.lY;
        ds.open(bl.diags + rg.diag_base, lX + lY + 2, t.hi + 1, t.lo);
        int64_t c_above = ds.get(t.hi + 1, false).cell_off;  // first cell of the diagonal after `cur`
        DiagDesc cur = ds.get(t.hi, (t.hi & (RR_DWIN - 1)) == RR_DWIN - 1);
        DiagDesc nxt = ds.get(t.hi - 1, ((t.hi - 1) & (RR_DWIN - 1)) == RR_DWIN - 1);  // hi >= 1 always
        Cell<S> a, b = cell_fill<S>(NINF);
        auto store = [&](int d, const DiagDesc &dg, const RegGeom &g, const Cell<S> &c) {
            if (!g.active) return;
            if (d <= t.from) {
                const int64_t cell = dg.cell_off + g.j;
                if (FULL) cell_store<S>(B + cell * S, c);
                else {
                    B[cell] = c.v[M];
                    if (dg.tot_flags & MAB_DIAG_IS_TOTAL)
                        cell_store<S>(BT + ((int64_t) bl.totals[dg.tot_flags & MAB_DIAG_TOT_MASK].full_off + g.j) * S, c);
                }
            } else if (d == t.from + 1) {
                edge[t.edge_off + g.j] = c.v[M];
            }
        };
        {  // diagonal hi: (ragged)endStateProb on every cell, pairwiseAligner.c:680-681
            const double *init = (t.at_end && (rg.flags & 2)) ? model.ragged_end : model.end;
            const RegGeom g = reg_geom<LPI>(ln.lane, t.hi, cur.xmyL, (int) (c_above - cur.cell_off));
#pragma unroll
            for (int s = 0; s < S; s++) a.v[s] = g.active ? init[s] : NINF;
            store(t.hi, cur, g, a);
        }
        const int d_first = t.hi - 1;
        auto step = [&](auto q_tag, int d, const Cell<S> &p1, Cell<S> &p2) {
            constexpr int Q = decltype(q_tag)::value;
            c_above = cur.cell_off;
            cur = nxt;
            nxt = ds.get(d - 1, ((d - 1) & (RR_DWIN - 1)) == RR_DWIN - 1);  // d >= 1: entry 0 exists
            const int w = (int) (c_above - cur.cell_off);
            if (((d_first - d) & (RR_SYM_REFRESH - 1)) == 0) {
                // the neighbours of the cells of diagonals d .. d - 31 consume X[x], Y[y] with x <= x_hi(d), y <= y_hi(d)
                // and at most 31 + band width below them
                const int x_hi = (d + cur.xmyL + 2 * (w - 1)) >> 1, y_hi = (d - cur.xmyL) >> 1;
                __syncwarp(ln.gmask);
                unpack_symbols<LPI>(bl.pack_x, cwx, rg.sym_x_off, lX, (x_hi - 100) & ~7, ln.lane, sxr);
                unpack_symbols<LPI>(bl.pack_y, cwy, rg.sym_y_off, lY, (y_hi - 100) & ~7, ln.lane, syr);
                __syncwarp(ln.gmask);
            }
            const RegGeom g = reg_geom<LPI>(ln.lane, d, cur.xmyL, w);
            int cX = 5, cY = 5;
            if (g.active) {  // the neighbours consume X[x] and / or Y[y] (0-based); past the end they unpack as N
                cX = lds_u8(sxr + (g.x & (RR_SYM_RING - 1)));
                cY = lds_u8(syr + (g.y & (RR_SYM_RING - 1)));
            }
            const double em = lds_f64(ln.em_addr + (cX * 5 + cY) * (RR_REPL * 8));
            const double ex = lds_f64(ln.ex_addr + cX * (RR_REPL * 8));
            const double ey = lds_f64(ln.ey_addr + cY * (RR_REPL * 8));
            reg_backward_cell<S, Q, LA>(model, ln, p1, p2, em, ex, ey);
            store(d, cur, g, p2);
        };
        int d = t.hi - 1;
        if (d > t.lo && !(d & 1)) {  // peel an even diagonal so that the unrolled pair starts on an odd one
            step(std::integral_constant<int, 0>(), d, a, b);
            const Cell<S> tmp = a;
            a = b;
            b = tmp;
            d--;
        }
        for (; d - 1 > t.lo; d -= 2) {
            step(std::integral_constant<int, 1>(), d, a, b);
            step(std::integral_constant<int, 0>(), d - 1, b, a);
        }
        if (d > t.lo) step(std::integral_constant<int, 1>(), d, a, b);
    }
}

// ------------------------------------------------------------------------------------------------ fused backward (N2)
// Posterior mode: backward sweep + total probability + posterior match probabilities in ONE pass over a traceback, as
// the reference's loop does (pairwiseAligner.c:695-733 with the callbacks :566-615).  The backward cells never leave
// the registers: B, BT and the edge buffer are not written, k_totals and k_posterior_diag are not launched for these
// items.  Per kept diagonal the lane reads F (match state) of its cell, forms exp(F + B - total) and writes p_int;
// the per-diagonal pair count goes to counts[slot] for the ordered emission (k_emit_diag).
//
// Total probability (every 10th kept diagonal, diagonalCalculationTotalProbability :566-583): the per-cell dot products
// over the S states are formed by all lanes at once; the two logAdd chains that must run in cell order (logAdd is not
// associative) -- over the cells of diagonal d, and over the match states of diagonal d+1 that jump over d -- are
// carried by the even and the odd lanes side by side, every lane pulling the next value of ITS chain by shuffle, so a
// total costs max(w, w1) + 1 dependent logAdds for the whole warp (both half-warps share the instruction stream: all
// final tracebacks place their totals on the same steps).
template <int S, int LPI, int LA, int MINB>
__global__ void __launch_bounds__(RR_THREADS, MINB)
k_backward_post_rr(const __grid_constant__ DevModel model, Blob bl, int item0, int item1, const double *__restrict__ F,
                   const double *__restrict__ FT, double threshold, double log_reject, int32_t *__restrict__ pint,
                   int64_t *__restrict__ counts, int *counter) {
    __shared__ RegTables tb;
    __shared__ __align__(16) CoefTable ct;
    __shared__ RegStage stage[RR_THREADS / LPI];
    fill_reg_tables(tb, model, threadIdx.x, blockDim.x);
    fill_coef_table(ct, threadIdx.x, blockDim.x);
    stage_init<LPI>(stage);
    __syncthreads();
    const RegLane ln = reg_lane<LPI>(tb, ct);
    const int lane32 = threadIdx.x & 31, lane_base = lane32 & ~(LPI - 1);
    const double NINF = neg_inf();
    RegStage &sg = stage[threadIdx.x / LPI];
    DescStream<LPI, -1> ds;
    ds.ring = (unsigned) __cvta_generic_to_shared(&sg.ring[0]);
    ds.bar = (unsigned) __cvta_generic_to_shared(&sg.bar[0]);
    ds.gmask = ln.gmask;
    ds.phase = 0;
    ds.leader = ln.lane == 0;
    const unsigned sxr = (unsigned) __cvta_generic_to_shared(&sg.sx[0]), syr = (unsigned) __cvta_generic_to_shared(&sg.sy[0]);
    const int64_t cwx = mab_pack_code_words(bl.h_len_sym_x), cwy = mab_pack_code_words(bl.h_len_sym_y);
    const bool odd_lane = ln.lane & 1;

    for (;;) {
        const int item = item0 + next_item_sub<LPI>(counter, lane32, ln.gmask);
        if (item >= item1) break;
        const TracebackDesc t = bl.tbs[bl.order_bwd[item]];
        const RegionDesc rg = bl.regions[t.region];
        const int lX = rg.lX, lY = rg.lY;
        // emission slot of diagonal d (see the comment above diag_work in kernels.cu)
        const int64_t slot0 = rg.diag_base + 1 + (lX + lY - t.from) - t.lo - 1;
        ds.open(bl.diags + rg.diag_base, lX + lY + 2, t.hi + 1, t.lo);
        int64_t c_above = ds.get(t.hi + 1, false).cell_off;
        DiagDesc cur = ds.get(t.hi, (t.hi & (RR_DWIN - 1)) == RR_DWIN - 1);
        DiagDesc nxt = ds.get(t.hi - 1, ((t.hi - 1) & (RR_DWIN - 1)) == RR_DWIN - 1);
        Cell<S> a, b = cell_fill<S>(NINF);
        double total = NINF;
        // geometry of the previous step (diagonal d + 1): needed by the "jump over d" term of a total on d
        int64_t up_cell = 0;
        bool up_active = false;
        int up_first = 0, up_w = 0;

        // total probability on diagonal d (cells c, first lane lL, width w); through = d < hi
        auto total_on = [&](const DiagDesc &dg, const RegGeom &g, int w, const Cell<S> &c, const Cell<S> &c_up, bool through) {
            double cd = NINF, tv = NINF;
            if (g.active) {
                const double *ft = FT + ((int64_t) bl.totals[dg.tot_flags & MAB_DIAG_TOT_MASK].full_off + g.j) * S;
                cd = __dadd_rn(ft[0], c.v[0]);
#pragma unroll
                for (int s = 1; s < S; s++) cd = la_rr<LA>(cd, __dadd_rn(ft[s], c.v[s]), ln.cf);
            }
            if (through && up_active) tv = __dadd_rn(F[up_cell], c_up.v[M]);
            const int first = (ln.lane - g.j) & (LPI - 1);  // lane of the diagonal's first cell
            const int n_it = max(w, through ? up_w : 0);
            double acc = NINF;
            for (int j = 0; j < n_it; j++) {
                const double v1 = shfl_f64(ln.gmask, cd, lane_base | ((first + j) & (LPI - 1)));
                const double v2 = shfl_f64(ln.gmask, tv, lane_base | ((up_first + j) & (LPI - 1)));
                const double r = la_rr<LA>(acc, odd_lane ? v2 : v1, ln.cf);
                if (odd_lane ? (through && j < up_w) : (j < w)) acc = r;
            }
            const double t0 = shfl_f64(ln.gmask, acc, lane_base), t1 = shfl_f64(ln.gmask, acc, lane_base | 1);
            total = through ? la_rr<LA>(t0, t1, ln.cf) : t0;
        };
        // posteriors of kept diagonal d >= 1 (diagonalCalculationPosteriorMatchProbs :585-615).  exp() is needed only
        // where the pair can pass the threshold (~7 % of the cells), but in a warp some lane nearly always has such a
        // cell, and the library exp is ~40 dependent FP64 operations: a lane therefore parks its candidate (log value,
        // cell, emission slot) and the group evaluates the parked candidates together only when a lane that already
        // holds one finds another (every ~4 diagonals), off the recurrence's critical path.  Pair counts per diagonal
        // are accumulated with integer atomics (order-independent).
        double pend_s = 0.0;
        int64_t pend_cell = -1, pend_slot = 0;
        auto flush = [&]() {
            if (pend_cell >= 0) {
                int32_t v = -1;
                double p = exp(pend_s);
                if (p >= threshold) {
                    if (p > 1.0) p = 1.0;
                    v = (int32_t) floor(__dmul_rn(p, MAB_PROB_1));
                    atomicAdd((unsigned long long *) (counts + pend_slot), 1ull);
                }
                pint[pend_cell] = v;
                pend_cell = -1;
            }
        };
        auto posterior_on = [&](int d, const DiagDesc &dg, const RegGeom &g, const Cell<S> &c, double f_match) {
            const int64_t cell = dg.cell_off + g.j;
            const double s = __dsub_rn(__dadd_rn(f_match, c.v[M]), total);
            const bool cand = g.active && g.x > 0 && g.y > 0 && s >= log_reject;
            if (__any_sync(ln.gmask, cand && pend_cell >= 0)) flush();
            if (cand) {
                pend_s = s;
                pend_cell = cell;
                pend_slot = slot0 + d;
            } else if (g.active) {
                pint[cell] = -1;
            }
        };
        {  // diagonal hi: (ragged)endStateProb on every cell, pairwiseAligner.c:680-681
            const double *init = (t.at_end && (rg.flags & 2)) ? model.ragged_end : model.end;
            const int w = (int) (c_above - cur.cell_off);
            const RegGeom g = reg_geom<LPI>(ln.lane, t.hi, cur.xmyL, w);
#pragma unroll
            for (int s = 0; s < S; s++) a.v[s] = g.active ? init[s] : NINF;
            if (t.hi <= t.from) {  // a final traceback keeps its first diagonal, which then carries a total
                if (cur.tot_flags & MAB_DIAG_IS_TOTAL) total_on(cur, g, w, a, b, false);
                posterior_on(t.hi, cur, g, a, g.active ? F[cur.cell_off + g.j] : 0.0);
            }
            up_cell = cur.cell_off + g.j;
            up_active = g.active;
            up_first = (ln.lane - g.j) & (LPI - 1);
            up_w = w;
        }
        const int d_first = t.hi - 1;
        auto step = [&](auto q_tag, int d, const Cell<S> &p1, Cell<S> &p2) {
            constexpr int Q = decltype(q_tag)::value;
            c_above = cur.cell_off;
            cur = nxt;
            nxt = ds.get(d - 1, ((d - 1) & (RR_DWIN - 1)) == RR_DWIN - 1);
            const int w = (int) (c_above - cur.cell_off);
            if (((d_first - d) & (RR_SYM_REFRESH - 1)) == 0) {
                const int x_hi = (d + cur.xmyL + 2 * (w - 1)) >> 1, y_hi = (d - cur.xmyL) >> 1;
                __syncwarp(ln.gmask);
                unpack_symbols<LPI>(bl.pack_x, cwx, rg.sym_x_off, lX, (x_hi - 100) & ~7, ln.lane, sxr);
                unpack_symbols<LPI>(bl.pack_y, cwy, rg.sym_y_off, lY, (y_hi - 100) & ~7, ln.lane, syr);
                __syncwarp(ln.gmask);
            }
            const RegGeom g = reg_geom<LPI>(ln.lane, d, cur.xmyL, w);
            int cX = 5, cY = 5;
            double f_match = 0.0;
            if (g.active) {
                cX = lds_u8(sxr + (g.x & (RR_SYM_RING - 1)));
                cY = lds_u8(syr + (g.y & (RR_SYM_RING - 1)));
                if (d <= t.from) f_match = __ldg(F + cur.cell_off + g.j);  // in flight while the cell is computed
            }
            const double em = lds_f64(ln.em_addr + (cX * 5 + cY) * (RR_REPL * 8));
            const double ex = lds_f64(ln.ex_addr + cX * (RR_REPL * 8));
            const double ey = lds_f64(ln.ey_addr + cY * (RR_REPL * 8));
            reg_backward_cell<S, Q, LA>(model, ln, p1, p2, em, ex, ey);  // p1 (diagonal d + 1) is left untouched
            if (d <= t.from) {
                if (cur.tot_flags & MAB_DIAG_IS_TOTAL) total_on(cur, g, w, p2, p1, true);
                posterior_on(d, cur, g, p2, f_match);
            }
            up_cell = cur.cell_off + g.j;
            up_active = g.active;
            up_first = (ln.lane - g.j) & (LPI - 1);
            up_w = w;
        };
        int d = t.hi - 1;
        if (d > t.lo && !(d & 1)) {
            step(std::integral_constant<int, 0>(), d, a, b);
            const Cell<S> tmp = a;
            a = b;
            b = tmp;
            d--;
        }
        for (; d - 1 > t.lo; d -= 2) {
            step(std::integral_constant<int, 1>(), d, a, b);
            step(std::integral_constant<int, 0>(), d - 1, b, a);
        }
        if (d > t.lo) step(std::integral_constant<int, 1>(), d, a, b);
        flush();
    }
}
