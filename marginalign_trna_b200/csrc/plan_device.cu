// plan_device.cu -- the planner on the GPU ("sequence and band staging" of the north star).
//
// The host planner (plan.cpp) produces ~4.8 GB of band / schedule descriptors for 1 M tRNA reads and then pushes them
// over PCIe; planning and the copy cost 20x the DP kernels.  Here only the raw inputs (bases, cigars: ~0.4 GB) cross
// the bus and the same blob is built in HBM, by the same integer walks (plan_core.h), in count / scan / fill passes:
//
//   k_count_regions   one thread per alignment: validate, count split regions (D > 0) and their diagonals
//   scan              region counts -> first region of each alignment (the blob's aln_region array)
//   k_fill_regions    one thread per alignment: RegionDesc (geometry part), anchor cursor of each region, D + 1
//   scan              diagonals per region -> diag_base
//   k_bands           one thread per region: band_construct -> DiagDesc (region-local cell offsets), then the traceback
//                     schedule in count mode -> cells / tracebacks / totals / edge cells / full cells per region
//   scan              those five counts (one struct scan)
//   k_schedule        one thread per region: TracebackDesc / TotalDesc (the serial part of the schedule)
//   k_diag_fill       one thread per diagonal: cell offsets region-local -> global, per-diagonal total flags
//   sort x2           regions and tracebacks longest first (stable radix sort: same order as the host's stable_sort)
//   k_symbols         bases -> symbol codes
// Two tiny D2H reads (counts) size the later arrays.  The result is byte-identical to the host planner's blob except
// for where the arrays sit inside it (tests/test_gpu_parity.py::test_device_planner_matches_host_planner).
#include <cuda_runtime.h>

#include <cub/cub.cuh>

#include <algorithm>
#include <cstring>

#include "mab_internal.h"
#include "plan_core.h"

namespace {

using plan_core::AlnInput;
using plan_core::AnchorStream;
using plan_core::Split;

struct DevInput {
    int64_t n;
    const uint8_t *seq_x, *seq_y;
    const int64_t *off_x, *off_y;
    const uint8_t *ops;
    const int32_t *lens;
    const int64_t *off_ops;
    const int64_t *anchors, *off_anc;
};

__device__ __forceinline__ AlnInput aln_input(const DevInput &in, int64_t a) {
    AlnInput ai;
    ai.lX = in.off_x[a + 1] - in.off_x[a];
    ai.lY = in.off_y[a + 1] - in.off_y[a];
    ai.sx = in.seq_x + in.off_x[a];
    ai.sy = in.seq_y + in.off_y[a];
    if (in.ops != nullptr) {
        ai.ops = in.ops + in.off_ops[a];
        ai.lens = in.lens + in.off_ops[a];
        ai.n_ops = in.off_ops[a + 1] - in.off_ops[a];
        ai.anc = nullptr;
        ai.n_anc = 0;
    } else {
        ai.ops = nullptr;
        ai.lens = nullptr;
        ai.n_ops = 0;
        ai.anc = in.anchors + 2 * in.off_anc[a];
        ai.n_anc = in.off_anc[a + 1] - in.off_anc[a];
    }
    return ai;
}

struct RegionTmp {  // where a region's anchors start in its alignment's stream, and its rectangle
    int64_t x1, y1, x2, y2;
    int64_t o, l, j, k;
};

struct RegionCounts {
    int64_t cells, tbs, tots, edge, full;
};
struct SumCounts {
    __host__ __device__ RegionCounts operator()(const RegionCounts &a, const RegionCounts &b) const {
        return RegionCounts{a.cells + b.cells, a.tbs + b.tbs, a.tots + b.tots, a.edge + b.edge, a.full + b.full};
    }
};

struct Status {  // device-side status block (256 B), read back by the host
    unsigned long long n_diagonals;
    int error_code;
    int pad0;
    long long error_aln;
    long long max_width;
    unsigned long long n_class_regions[3], n_class_tbs[3];  // work items by band class (mab_band_class)
};

// ------------------------------------------------------------------------------------------------ kernels

struct RegionRec {  // one region as found by the walk
    int64_t x1, y1, x2, y2;
    AnchorStream cur;
    int flags;
};

__device__ __forceinline__ void store_region(const DevInput &in, int64_t a, int64_t r, const RegionRec &rec,
                                             RegionDesc *regions, RegionTmp *tmp, int64_t *diag_count) {
    const int64_t lX = rec.x2 - rec.x1, lY = rec.y2 - rec.y1;
    RegionDesc rg;
    rg.sym_x_off = in.off_x[a] + rec.x1;
    rg.sym_y_off = in.off_y[a] + rec.y1;
    rg.diag_base = 0;
    rg.lX = (int32_t) lX;
    rg.lY = (int32_t) lY;
    rg.x1 = (int32_t) rec.x1;
    rg.y1 = (int32_t) rec.y1;
    rg.aln = (int32_t) a;
    rg.flags = rec.flags;
    rg.tb_base = rg.n_tb = rg.tot_base = rg.n_tot = 0;
    regions[r] = rg;
    tmp[r] = RegionTmp{rec.x1, rec.y1, rec.x2, rec.y2, rec.cur.o, rec.cur.l, rec.cur.j, rec.cur.k};
    diag_count[r] = lX + lY + 1;
}

// WRITE = false: validate, count regions and diagonals.  WRITE = true: the same walk again, now that the scan has told
// every alignment where its regions go.  The first two regions of an alignment (all of them, for short reads) are
// held in registers and stored after the walk, where the lanes of the warp are together again; only further regions
// are stored from inside the walk.
template <bool WRITE>
__global__ void k_regions(DevInput in, mab_params p, int64_t *region_count /* [n+1] or aln_region in WRITE */,
                          Status *status, RegionDesc *regions, RegionTmp *tmp, int64_t *diag_count, int spread) {
    const int64_t gt = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t a_raw = gt >> spread;
    const bool live = a_raw < in.n && (gt & ((1 << spread) - 1)) == 0;
    const int64_t a = live ? a_raw : 0;  // idle lanes shadow alignment 0 so the whole warp stays in the stepping loop
    if (in.n == 0) return;
    const AlnInput ai = aln_input(in, a);
    bool more = live;
    if (!WRITE && live) {
        const int bad = plan_core::validate(ai);
        if (bad) {
            if (atomicCAS(&status->error_code, 0, bad) == 0) status->error_aln = a;
            region_count[a] = 0;
            more = false;
        }
    }
    const bool counted = more;
    int64_t n_regions = 0, n_diags = 0;
    const int64_t base = WRITE ? region_count[a] : 0;
    RegionRec r0{}, r1{};
    auto sink = [&](const Split &sp, const AnchorStream &cur, bool rl, bool rr) {
        const int64_t D = (sp.x2 - sp.x1) + (sp.y2 - sp.y1);
        if (D == 0) return;  // getPosteriorProbsWithBanding :650-652
        if (WRITE) {
            const RegionRec rec{sp.x1, sp.y1, sp.x2, sp.y2, cur, (rl ? 1 : 0) | (rr ? 2 : 0)};
            if (n_regions == 0) r0 = rec;
            else if (n_regions == 1) r1 = rec;
            else store_region(in, a, base + n_regions, rec, regions, tmp, diag_count);
        }
        n_regions++;
        n_diags += D + 1;
    };
    // one anchor per step, the warp re-joined at every step (lanes find their next anchor after different numbers
    // of bases; left to itself the compiler let them drift apart for the whole walk: 1.2 active lanes per instruction)
    plan_core::RegionWalker walker(ai, p);
    while (__any_sync(0xffffffffu, more)) {
        if (more) more = walker.step(sink);
    }
    if (!counted) return;
    if (WRITE) {
        if (n_regions > 0) store_region(in, a, base, r0, regions, tmp, diag_count);
        if (n_regions > 1) store_region(in, a, base + 1, r1, regions, tmp, diag_count);
    } else {
        region_count[a] = n_regions;
        if (n_diags) atomicAdd(&status->n_diagonals, (unsigned long long) n_diags);
    }
}

// The traceback schedule over a region whose DiagDesc entries hold region-local cell offsets.
struct WidthFromDiags {
    const DiagDesc *dd;
    int64_t D, n_cells;
    __device__ __forceinline__ int64_t operator()(int64_t d) const {
        return (d == D ? n_cells : dd[d + 1].cell_off) - dd[d].cell_off;
    }
};

struct CountSink {
    RegionCounts c{0, 0, 0, 0, 0};
    __device__ __forceinline__ void total(int64_t, int, int64_t w) {
        c.tots++;
        c.full += w;
    }
    __device__ __forceinline__ void diag_flags(int64_t, int64_t, bool, bool) {}
    __device__ __forceinline__ void traceback(int64_t, int64_t, int64_t, bool, int64_t, int64_t edge_cells) {
        c.tbs++;
        c.edge += edge_cells;
    }
};

__global__ void k_bands(DevInput in, mab_params p, int64_t n_regions, RegionDesc *regions, const RegionTmp *tmp,
                        const int64_t *diag_base, DiagDesc *diags, int32_t *diag_region, RegionCounts *counts,
                        Status *status, int spread) {
    const int64_t gt = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t r = gt >> spread;
    if (r >= n_regions || (gt & ((1 << spread) - 1)) != 0) return;
    const RegionDesc rg = regions[r];
    const RegionTmp t = tmp[r];
    const AlnInput ai = aln_input(in, rg.aln);
    AnchorStream cur;
    cur.o = t.o;
    cur.l = t.l;
    cur.j = t.j;
    cur.k = t.k;
    DiagDesc *dd = diags + diag_base[r];
    int32_t *dr = diag_region + diag_base[r];
    int64_t cell = 0, max_w = 0;
    plan_core::SpanTracker span;
    plan_core::band_walk(ai, p, Split{t.x1, t.y1, t.x2, t.y2}, cur, [&](int64_t xay, int64_t l, int64_t w) {
        DiagDesc d;
        d.cell_off = cell;
        d.xmyL = (int32_t) l;
        d.tot_flags = MAB_DIAG_TOT_MASK;
        dd[xay] = d;
        dr[xay] = (int32_t) r;
        cell += w;
        if (w > max_w) max_w = w;
        span.add(l, w);
    });
    regions[r].flags = rg.flags | span.flag_bits();
    const int64_t D = (int64_t) rg.lX + rg.lY;
    CountSink sink;
    plan_core::schedule_walk(p, D, WidthFromDiags{dd, D, cell}, sink);
    sink.c.cells = cell;
    counts[r] = sink.c;
    atomicMax(&status->max_width, (long long) max_w);
}

struct WriteSink {
    DiagDesc *dd;
    TracebackDesc *tbs;
    TotalDesc *totals;
    int32_t region;
    int64_t tot_base, edge_base, full_base;
    int64_t n_tb = 0, n_tot = 0, n_edge = 0, n_full = 0;
    __device__ __forceinline__ void total(int64_t q, int through, int64_t w) {
        TotalDesc td;
        td.region = region;
        td.d = (int32_t) q;
        td.edge_off = through == 0 ? MAB_THROUGH_NONE : (through == 1 ? (int32_t) (edge_base + n_edge) : MAB_THROUGH_MATRIX);
        td.full_off = (int32_t) (full_base + n_full);
        n_full += w;
        totals[n_tot++] = td;
    }
    // the per-diagonal flags are written by k_diag_fill, one thread per diagonal, from the tracebacks stored here
    __device__ __forceinline__ void diag_flags(int64_t, int64_t, bool, bool) {}
    __device__ __forceinline__ void traceback(int64_t lo, int64_t hi, int64_t from, bool at_end, int64_t n, int64_t edge_cells) {
        TracebackDesc tb;
        tb.region = region;
        tb.lo = (int32_t) lo;
        tb.hi = (int32_t) hi;
        tb.from = (int32_t) from;
        tb.at_end = at_end;
        tb.tot_base = (int32_t) (tot_base + n_tot - n);
        tb.n_tot = (int32_t) n;
        tb.edge_off = at_end ? -1 : (int32_t) (edge_base + n_edge);
        n_edge += edge_cells;
        tbs[n_tb++] = tb;
    }
};

__global__ void k_schedule(mab_params p, int64_t n_regions, RegionDesc *regions, const int64_t *diag_base,
                           const RegionCounts *base /* exclusive scan, [n_regions + 1] */, DiagDesc *diags,
                           TracebackDesc *tbs, TotalDesc *totals, int32_t *region_key, int32_t *tb_key,
                           int64_t n_diagonals, Status *status, int spread) {
    const int64_t gt = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if ((gt & ((1 << spread) - 1)) != 0) return;
    const int64_t r = gt >> spread;
    if (r == 0) {  // sentinel diagonal carrying the total cell count
        DiagDesc s;
        s.cell_off = base[n_regions].cells;
        s.xmyL = 0;
        s.tot_flags = MAB_DIAG_TOT_MASK;
        diags[n_diagonals] = s;
    }
    if (r >= n_regions) return;
    RegionDesc rg = regions[r];
    const RegionCounts b = base[r], nb = base[r + 1];
    DiagDesc *dd = diags + diag_base[r];
    const int64_t D = (int64_t) rg.lX + rg.lY, n_cells = nb.cells - b.cells;
    WriteSink sink{dd, tbs + b.tbs, totals + b.tots, (int32_t) r, b.tots, b.edge, b.full};
    plan_core::schedule_walk(p, D, WidthFromDiags{dd, D, n_cells}, sink);
    // sort keys: band class first (class 2 leads the order arrays), then length
    const int cls = mab_band_class(mab_region_span(rg.flags));
    const int32_t key_cap = (1 << 28) - 1;
    for (int64_t t = 0; t < sink.n_tb; t++)
        tb_key[b.tbs + t] = (cls << 28) | min(tbs[b.tbs + t].hi - tbs[b.tbs + t].lo, key_cap);
    atomicAdd(&status->n_class_regions[cls], 1ull);
    atomicAdd(&status->n_class_tbs[cls], (unsigned long long) sink.n_tb);
    rg.diag_base = diag_base[r];
    rg.tb_base = (int32_t) b.tbs;
    rg.n_tb = (int32_t) sink.n_tb;
    rg.tot_base = (int32_t) b.tots;
    rg.n_tot = (int32_t) sink.n_tot;
    regions[r] = rg;
    region_key[r] = (cls << 28) | (int32_t) min(D, (int64_t) key_cap);
}

// One thread per diagonal, after k_schedule: region-local cell offsets -> global, and the flags schedule_walk's
// diag_flags would have written (plan_core.h): diagonal q in (lo, from] of a traceback is the k = from - q-th posterior
// diagonal of that traceback; every tenth one carries a total.  Coalesced, where the per-region loops it replaces read
// and wrote every DiagDesc of a region from one thread (2 GB of sector traffic per 113 k tRNA regions).
__global__ void k_diag_fill(int64_t n_diagonals, const int32_t *__restrict__ diag_region, const int64_t *__restrict__ diag_base,
                            const RegionCounts *__restrict__ base, const RegionDesc *__restrict__ regions,
                            const TracebackDesc *__restrict__ tbs, DiagDesc *diags) {
    const int64_t g = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n_diagonals) return;
    const int32_t r = diag_region[g];
    const int32_t q = (int32_t) (g - diag_base[r]);
    DiagDesc d = diags[g];
    d.cell_off += base[r].cells;
    if (q >= 1) {
        const int32_t tb0 = regions[r].tb_base, n_tb = regions[r].n_tb;
        int32_t lo_i = 0, hi_i = n_tb - 1;  // first traceback with from >= q (their (lo, from] ranges tile 1 .. D)
        while (lo_i < hi_i) {
            const int32_t mid = (lo_i + hi_i) >> 1;
            if (tbs[tb0 + mid].from >= q) hi_i = mid;
            else lo_i = mid + 1;
        }
        const TracebackDesc tb = tbs[tb0 + lo_i];
        const int32_t k = tb.from - q;
        uint32_t tf = (uint32_t) (tb.tot_base + k / 10);
        if (k % 10 == 0) tf |= MAB_DIAG_IS_TOTAL;
        if (q - 2 < tb.lo) tf |= MAB_DIAG_NO_MIDDLE;
        d.tot_flags = tf;
    }
    diags[g] = d;
}

__global__ void k_symbols(const uint8_t *seq, int64_t n, uint8_t *sym /* 16 guard bytes on both sides */) {
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 16) {
        sym[i - 16] = 4;
        sym[n + i] = 4;
    }
    if (i < n) sym[i] = plan_core::symbol_of(seq[i]);
}

// 2-bit codes + N mask (BlobHeader::off_pack_x): one thread per mask word = 32 bases = two code words
__global__ void k_pack_symbols(const uint8_t *seq, int64_t n, uint32_t *pack) {
    const int64_t cw = mab_pack_code_words(n), mw = mab_pack_mask_words(n);
    const int64_t t = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= mw) return;
    uint32_t c0 = 0, c1 = 0, m = 0;
    for (int i = 0; i < 32; i++) {
        const int64_t b = t * 32 + i - MAB_PACK_PAD_BASES;  // position t * 32 + i holds base b
        if (b < 0) continue;
        if (b >= n) break;
        const uint32_t s = plan_core::symbol_of(seq[b]);
        if (s < 4) (i < 16 ? c0 : c1) |= s << (2 * (i & 15));
        else m |= 1u << i;
    }
    if (2 * t < cw) pack[2 * t] = c0;
    if (2 * t + 1 < cw) pack[2 * t + 1] = c1;
    pack[cw + t] = m;
}

__global__ void k_iota(int32_t *v, int64_t n) {
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) v[i] = (int32_t) i;
}

// ------------------------------------------------------------------------------------------------ host side

struct Arena {
    uint8_t *base;
    int64_t cap, off;
    bool ok = true;
    int64_t place(int64_t bytes) {  // 256-byte aligned sub-allocations; keeps counting past the end to report need
        const int64_t at = off;
        off = (off + bytes + 255) & ~int64_t(255);
        if (off > cap) ok = false;
        return at;
    }
    template <typename T>
    T *ptr(int64_t at) const {
        return reinterpret_cast<T *>(base + at);
    }
};

int cuda_fail(cudaError_t e, const char *what) {
    if (e == cudaSuccess) return MAB_OK;
    mab_set_error("CUDA error in %s: %s", what, cudaGetErrorString(e));
    return MAB_E_CUDA;
}

inline int grid(int64_t n, int block) { return (int) std::max<int64_t>(1, (n + block - 1) / block); }

// The walks are one thread per alignment / region.  A batch of few, long alignments (600 reads of 10 kb per chunk) would
// leave that to a handful of warps whose lanes, each in its own stretch of a different read, take turns: 12 ms per walk
// on 3 thread blocks.  Such batches are spread out, one item per 2^spread threads (only the first one works), so every
// item gets a warp -- and with it an instruction stream -- of its own; dense batches (tRNA) keep 32 items per warp.
inline int spread_for(int64_t n) {
    int s = 5;
    while (s > 0 && (n << s) > 300000) s--;
    return s;
}

}  // namespace

extern "C" int64_t mab_plan_device_arena_hint(const mab_params *p, int64_t n, int64_t len_x, int64_t len_y) {
    // generous first guess (no splits beyond one per eight alignments); the build call reports the exact need
    (void) p;
    const int64_t nR = n + n / 8 + 64, nD = len_x + len_y + 2 * nR;
    const int64_t nT = nR + nD / 900 + 64, nTot = nD / 10 + nT + 64;
    return 64 * 256 /* alignment slack of the sub-allocations */ + 4096 + 8 * (n + 2) * 2 + (len_x + len_y + 1024) +
           (mab_pack_words(len_x) + mab_pack_words(len_y)) * 4 + nR * (64 + 64 + 8 * 2 + 40 * 2 + 4 * 6) +
           (nD + 1) * (16 + 4) /* DiagDesc + diag_region */ + nT * (32 + 4 * 6) + nTot * 16 +
           (int64_t) 64 * 1024 * 1024 / 8 + (nR + nT) * 8;
}

extern "C" int mab_plan_build_device(const mab_params *p, int64_t n, const uint8_t *d_seq_x, const int64_t *d_off_x,
                                     int64_t len_x, const uint8_t *d_seq_y, const int64_t *d_off_y, int64_t len_y,
                                     const uint8_t *d_ops, const int32_t *d_lens, const int64_t *d_off_ops,
                                     const int64_t *d_anchors, const int64_t *d_off_anc, void *d_arena,
                                     int64_t arena_bytes, void *stream, mab_plan **out, int64_t *needed_bytes) {
    if (needed_bytes) *needed_bytes = 0;
    if (!p || !out || n < 0 || !d_off_x || !d_off_y || !d_arena || (!d_off_ops && !d_off_anc) ||
        ((uintptr_t) d_arena & 255)) {
        mab_set_error("mab_plan_build_device: bad argument (null pointer or arena not 256-byte aligned)");
        return MAB_E_ARG;
    }
    int rc = mab_params_validate(p);
    if (rc) return rc;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        mab_set_error("no CUDA device: libmargin_b200 has no CPU fallback");
        return MAB_E_NOGPU;
    }
    cudaStream_t st = (cudaStream_t) stream;
    const DevInput in{n, d_seq_x, d_seq_y, d_off_x, d_off_y, d_off_ops ? d_ops : nullptr, d_lens, d_off_ops,
                      d_off_ops ? nullptr : d_anchors, d_off_ops ? nullptr : d_off_anc};
    Arena ar{(uint8_t *) d_arena, arena_bytes, 0};
    BlobHeader h;
    std::memset(&h, 0, sizeof(h));
    h.magic = MAB_BLOB_MAGIC;
    h.n_alignments = n;
    h.len_sym_x = len_x;
    h.len_sym_y = len_y;

    // ---- phase 1: regions per alignment
    ar.place(sizeof(BlobHeader));
    const int64_t at_status = ar.place(256);
    h.off_aln_region = ar.place((n + 1) * 8);
    h.off_sym_x = ar.place(len_x + 32) + 16;
    h.off_sym_y = ar.place(len_y + 32) + 16;
    h.off_pack_x = ar.place(mab_pack_words(len_x) * 4);
    h.off_pack_y = ar.place(mab_pack_words(len_y) * 4);
    size_t scan_tmp1 = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, scan_tmp1, (int64_t *) nullptr, (int64_t *) nullptr, n + 1, st);
    const int64_t at_tmp1 = ar.place((int64_t) scan_tmp1);
    if (!ar.ok) {
        if (needed_bytes) *needed_bytes = mab_plan_device_arena_hint(p, n, len_x, len_y);
        mab_set_error("mab_plan_build_device: arena too small");
        return MAB_E_CAPACITY;
    }
    Status *d_status = ar.ptr<Status>(at_status);
    int64_t *d_aln_region = ar.ptr<int64_t>(h.off_aln_region);
    if ((rc = cuda_fail(cudaMemsetAsync(d_status, 0, 256, st), "memset"))) return rc;
    if ((rc = cuda_fail(cudaMemsetAsync(d_aln_region + n, 0, 8, st), "memset"))) return rc;
    const int spread_a = spread_for(n);
    if (n > 0)
        k_regions<false><<<grid(n << spread_a, 128), 128, 0, st>>>(in, *p, d_aln_region, d_status, nullptr, nullptr, nullptr,
                                                                 spread_a);
    {  // also for empty inputs: the guard bytes must be set
        k_symbols<<<grid(std::max<int64_t>(len_x, 16), 256), 256, 0, st>>>(d_seq_x, len_x, ar.ptr<uint8_t>(h.off_sym_x));
        k_symbols<<<grid(std::max<int64_t>(len_y, 16), 256), 256, 0, st>>>(d_seq_y, len_y, ar.ptr<uint8_t>(h.off_sym_y));
        k_pack_symbols<<<grid(mab_pack_mask_words(len_x), 256), 256, 0, st>>>(d_seq_x, len_x, ar.ptr<uint32_t>(h.off_pack_x));
        k_pack_symbols<<<grid(mab_pack_mask_words(len_y), 256), 256, 0, st>>>(d_seq_y, len_y, ar.ptr<uint32_t>(h.off_pack_y));
    }
    cub::DeviceScan::ExclusiveSum(ar.ptr<void>(at_tmp1), scan_tmp1, d_aln_region, d_aln_region, n + 1, st);
    Status hs;
    int64_t nR = 0;
    if ((rc = cuda_fail(cudaMemcpyAsync(&hs, d_status, sizeof(hs), cudaMemcpyDeviceToHost, st), "copy status"))) return rc;
    if ((rc = cuda_fail(cudaMemcpyAsync(&nR, d_aln_region + n, 8, cudaMemcpyDeviceToHost, st), "copy count"))) return rc;
    if ((rc = cuda_fail(cudaStreamSynchronize(st), "planner phase 1"))) return rc;
    if (hs.error_code) {
        static const char *const kWhy[] = {"", "a cigar op runs past the sequences", "the cigar does not span the sequences",
                                           "an anchor is out of order or out of range", "bad sequence lengths"};
        mab_set_error("alignment %lld: %s", (long long) hs.error_aln, kWhy[hs.error_code & 7]);
        return MAB_E_INPUT;
    }
    const int64_t nD = (int64_t) hs.n_diagonals;
    // capacity bounds for what phase 2 will count exactly: a traceback needs >= min_diags - traceback_diags - 1 new
    // diagonals (or the region's end); a traceback of k posterior diagonals evaluates ceil(k / 10) totals
    const int64_t tb_gap = std::max<int64_t>(1, p->min_diags_between_traceback - p->traceback_diagonals - 1);
    const int64_t capT = nR + nD / tb_gap + 1, capTot = nD / 10 + capT + 1;
    if (nR > INT32_MAX || capT > INT32_MAX || capTot >= (int64_t) MAB_DIAG_TOT_MASK) {
        mab_set_error("mab_plan: batch too large for 32-bit region/traceback indices; split it");
        return MAB_E_ARG;
    }
    h.n_regions = nR;
    h.n_diagonals = nD;

    // ---- phase 2 layout
    h.off_regions = ar.place(nR * (int64_t) sizeof(RegionDesc));
    h.off_diags = ar.place((nD + 1) * (int64_t) sizeof(DiagDesc));
    h.off_tracebacks = ar.place(capT * (int64_t) sizeof(TracebackDesc));
    h.off_totals = ar.place(capTot * (int64_t) sizeof(TotalDesc));
    h.off_diag_region = ar.place((nD + 1) * 4);
    h.off_order_fwd = ar.place(nR * 4);
    h.off_order_bwd = ar.place(capT * 4);
    const int64_t at_tmp = ar.place(nR * (int64_t) sizeof(RegionTmp));
    const int64_t at_diag_base = ar.place((nR + 1) * 8);
    const int64_t at_counts = ar.place((nR + 1) * (int64_t) sizeof(RegionCounts));
    const int64_t at_bases = ar.place((nR + 1) * (int64_t) sizeof(RegionCounts));
    const int64_t at_rkey = ar.place(nR * 4), at_rkey2 = ar.place(nR * 4), at_riota = ar.place(nR * 4);
    const int64_t at_tkey = ar.place(capT * 4), at_tkey2 = ar.place(capT * 4), at_tiota = ar.place(capT * 4);
    size_t scan_tmp2 = 0, scan_tmp3 = 0, sort_tmp_r = 0, sort_tmp_t = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, scan_tmp2, (int64_t *) nullptr, (int64_t *) nullptr, nR + 1, st);
    cub::DeviceScan::ExclusiveScan(nullptr, scan_tmp3, (RegionCounts *) nullptr, (RegionCounts *) nullptr, SumCounts(),
                                   RegionCounts{0, 0, 0, 0, 0}, nR + 1, st);
    cub::DeviceRadixSort::SortPairsDescending(nullptr, sort_tmp_r, (int32_t *) nullptr, (int32_t *) nullptr,
                                              (int32_t *) nullptr, (int32_t *) nullptr, nR, 0, 32, st);
    cub::DeviceRadixSort::SortPairsDescending(nullptr, sort_tmp_t, (int32_t *) nullptr, (int32_t *) nullptr,
                                              (int32_t *) nullptr, (int32_t *) nullptr, capT, 0, 32, st);
    const int64_t at_cub = ar.place((int64_t) std::max(std::max(scan_tmp2, scan_tmp3), std::max(sort_tmp_r, sort_tmp_t)));
    h.blob_bytes = ar.off;
    if (needed_bytes) *needed_bytes = ar.off;
    if (!ar.ok) {
        mab_set_error("mab_plan_build_device: arena too small: need %lld bytes, got %lld", (long long) ar.off,
                      (long long) arena_bytes);
        return MAB_E_CAPACITY;
    }
    RegionDesc *d_regions = ar.ptr<RegionDesc>(h.off_regions);
    DiagDesc *d_diags = ar.ptr<DiagDesc>(h.off_diags);
    RegionTmp *d_tmp = ar.ptr<RegionTmp>(at_tmp);
    int64_t *d_diag_base = ar.ptr<int64_t>(at_diag_base);
    RegionCounts *d_counts = ar.ptr<RegionCounts>(at_counts), *d_bases = ar.ptr<RegionCounts>(at_bases);
    void *d_cub = ar.ptr<void>(at_cub);
    RegionCounts totals{0, 0, 0, 0, 0};
    if ((rc = cuda_fail(cudaMemsetAsync(ar.ptr<int32_t>(h.off_diag_region) + nD, 0xff, 4, st), "memset"))) return rc;
    if (nR > 0) {
        if ((rc = cuda_fail(cudaMemsetAsync(d_diag_base + nR, 0, 8, st), "memset"))) return rc;
        if ((rc = cuda_fail(cudaMemsetAsync(d_counts + nR, 0, sizeof(RegionCounts), st), "memset"))) return rc;
        const int spread_r = spread_for(nR);
        k_regions<true><<<grid(n << spread_a, 128), 128, 0, st>>>(in, *p, d_aln_region, d_status, d_regions, d_tmp, d_diag_base,
                                                                 spread_a);
        cub::DeviceScan::ExclusiveSum(d_cub, scan_tmp2, d_diag_base, d_diag_base, nR + 1, st);
        k_bands<<<grid(nR << spread_r, 128), 128, 0, st>>>(in, *p, nR, d_regions, d_tmp, d_diag_base, d_diags,
                                                           ar.ptr<int32_t>(h.off_diag_region), d_counts, d_status, spread_r);
        cub::DeviceScan::ExclusiveScan(d_cub, scan_tmp3, d_counts, d_bases, SumCounts(), RegionCounts{0, 0, 0, 0, 0}, nR + 1,
                                       st);
        int32_t *rkey = ar.ptr<int32_t>(at_rkey), *rkey2 = ar.ptr<int32_t>(at_rkey2), *riota = ar.ptr<int32_t>(at_riota);
        int32_t *tkey = ar.ptr<int32_t>(at_tkey), *tkey2 = ar.ptr<int32_t>(at_tkey2), *tiota = ar.ptr<int32_t>(at_tiota);
        k_schedule<<<grid(nR << spread_r, 128), 128, 0, st>>>(*p, nR, d_regions, d_diag_base, d_bases, d_diags,
                                                              ar.ptr<TracebackDesc>(h.off_tracebacks),
                                                              ar.ptr<TotalDesc>(h.off_totals), rkey, tkey, nD, d_status,
                                                              spread_r);
        k_diag_fill<<<grid(nD, 256), 256, 0, st>>>(nD, ar.ptr<int32_t>(h.off_diag_region), d_diag_base, d_bases, d_regions,
                                                   ar.ptr<TracebackDesc>(h.off_tracebacks), d_diags);
        if ((rc = cuda_fail(cudaMemcpyAsync(&totals, d_bases + nR, sizeof(totals), cudaMemcpyDeviceToHost, st), "copy")))
            return rc;
        if ((rc = cuda_fail(cudaMemcpyAsync(&hs, d_status, sizeof(hs), cudaMemcpyDeviceToHost, st), "copy"))) return rc;
        if ((rc = cuda_fail(cudaStreamSynchronize(st), "planner phase 2"))) return rc;
        const int64_t nT = totals.tbs;
        k_iota<<<grid(nR, 256), 256, 0, st>>>(riota, nR);
        cub::DeviceRadixSort::SortPairsDescending(d_cub, sort_tmp_r, rkey, rkey2, riota, ar.ptr<int32_t>(h.off_order_fwd),
                                                  nR, 0, 32, st);
        k_iota<<<grid(nT, 256), 256, 0, st>>>(tiota, nT);
        size_t bytes_t = sort_tmp_t;
        cub::DeviceRadixSort::SortPairsDescending(d_cub, bytes_t, tkey, tkey2, tiota, ar.ptr<int32_t>(h.off_order_bwd), nT,
                                                  0, 32, st);
    } else {
        DiagDesc s;
        s.cell_off = 0;
        s.xmyL = 0;
        s.tot_flags = MAB_DIAG_TOT_MASK;
        if ((rc = cuda_fail(cudaMemcpyAsync(d_diags, &s, sizeof(s), cudaMemcpyHostToDevice, st), "sentinel"))) return rc;
    }
    if (totals.edge > INT32_MAX || totals.full > INT32_MAX) {
        mab_set_error("mab_plan: batch too large for 32-bit side-buffer offsets; split it");
        return MAB_E_ARG;
    }
    h.n_cells = totals.cells;
    h.n_tracebacks = totals.tbs;
    h.n_totals = totals.tots;
    h.n_edge_cells = totals.edge;
    h.n_full_cells = totals.full;
    h.max_width = hs.max_width;
    for (int c = 0; c < 3; c++) {
        h.n_class_regions[c] = nR > 0 ? (int64_t) hs.n_class_regions[c] : 0;
        h.n_class_tbs[c] = nR > 0 ? (int64_t) hs.n_class_tbs[c] : 0;
    }
    h.n_anchors = -1;  // not materialised on the device

    mab_plan *plan = new mab_plan();
    plan->params = *p;
    plan->hdr = h;
    plan->d_blob = d_arena;
    // the header itself (kernels read the arrays through offsets held in the host copy; the device copy is for tools)
    if ((rc = cuda_fail(cudaMemcpyAsync(d_arena, &plan->hdr, sizeof(BlobHeader), cudaMemcpyHostToDevice, st), "header"))) {
        delete plan;
        return rc;
    }
    if ((rc = cuda_fail(cudaGetLastError(), "planner launch"))) {
        delete plan;
        return rc;
    }
    *out = plan;
    return MAB_OK;
}

extern "C" const void *mab_plan_device_blob(const mab_plan *plan) { return plan ? plan->d_blob : nullptr; }
