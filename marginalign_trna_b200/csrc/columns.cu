// columns.cu -- marginCaller's consumer of the posterior pairs on the GPU.
//
// The reference (src/margin/marginCallerLib.py:39-74) runs cPecanRealign --outputAllPosteriorProbs once per read, parses
// the "refPos \t readPos \t %f" lines back and adds every probability to a dict keyed by (reference, position) with one
// slot per read base; variantCallSamFileTargetFn (:171-217) then merges the dicts and evaluates, per reference column,
// the posterior of each possible true base (calcBasePosteriorProbs, :81-92).  Here the pairs never leave HBM:
//
//   k_col_accumulate  one warp per alignment: every pair adds its probability to table[column][read base].  The
//       probability is what the reference re-reads from the text file: p_int / 1e7 printed with "%f" (cPecanRealign.c:
//       346-351), i.e. rounded to 6 decimals.  The rounding is reproduced exactly in integers (units of 1e-6; a p_int
//       ending in 5 is rounded the way printf rounds the binary double, decided by an exact FMA residual) and the sums
//       are 64-bit integer atomics, so the table is deterministic whatever the order of the reads.
//   k_col_call        one thread per column: calcBasePosteriorProbs in FP64 (log / exp), columns without observations
//       or with a non-ACGT reference base give zeros.
#include <cuda_runtime.h>

#include <cstdint>

#include "mab_internal.h"

namespace {

__device__ __forceinline__ int base_index(uint8_t c) {  // BASES = "ACGT", upper-cased (marginCallerLib.py:13, 64)
    switch (c) {
        case 'A': case 'a': return 0;
        case 'C': case 'c': return 1;
        case 'G': case 'g': return 2;
        case 'T': case 't': return 3;
        default: return -1;
    }
}

// round(p_int / 1e7 to 6 decimals) in units of 1e-6, as printf("%f") rounds the double (double) p_int / 1e7
__device__ __forceinline__ long long micro_units(int32_t p_int) {
    const long long q = p_int / 10;
    const int r = p_int - (int) (q * 10);
    if (r != 5) return q + (r > 5 ? 1 : 0);
    // decimal tie: the binary quotient lies just above or just below it; the FMA residual is exact
    const double v = __ddiv_rn((double) p_int, MAB_PROB_1);
    const double res = __fma_rn(v, MAB_PROB_1, -(double) p_int);
    if (res > 0.0) return q + 1;
    if (res < 0.0) return q;
    return q + (q & 1);  // exactly representable tie: round half to even
}

__global__ void __launch_bounds__(256)
k_col_accumulate(int64_t n, const int64_t *__restrict__ pair_off, const int32_t *__restrict__ x,
                 const int32_t *__restrict__ y, const int32_t *__restrict__ p, const uint8_t *__restrict__ seq_y,
                 const int64_t *__restrict__ off_y, const int64_t *__restrict__ col_base, long long *table,
                 int64_t n_columns) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = ((int64_t) gridDim.x * blockDim.x) >> 5;
    for (int64_t al = warp; al < n; al += n_warps) {
        const int64_t p0 = pair_off[al], p1 = pair_off[al + 1], base = col_base[al];
        const uint8_t *read = seq_y + off_y[al];
        for (int64_t i = p0 + lane; i < p1; i += 32) {
            const int b = base_index(read[y[i]]);  // N and other wildcards are ignored (:65)
            const int64_t col = base + x[i];
            if (b >= 0 && col >= 0 && col < n_columns)
                atomicAdd((unsigned long long *) &table[col * 4 + b], (unsigned long long) micro_units(p[i]));
        }
    }
}

// calcBasePosteriorProbs (marginCallerLib.py:81-92) with observations normalised by their sum (:199-203)
__global__ void __launch_bounds__(256)
k_col_call(const long long *__restrict__ table, int64_t n_columns, const uint8_t *__restrict__ ref,
           const double *__restrict__ log_err /* [true base][observed base] */,
           const double *__restrict__ log_evo /* [reference base][true base] */, double *__restrict__ post) {
    const int64_t c = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_columns) return;
    double e[4], total = 0.0;
#pragma unroll
    for (int b = 0; b < 4; b++) {
        e[b] = __dmul_rn((double) table[c * 4 + b], 1e-6);
        total = __dadd_rn(total, e[b]);
    }
    const int rb = base_index(ref[c]);
    if (!(total > 0.0) || rb < 0) {
#pragma unroll
        for (int b = 0; b < 4; b++) post[c * 4 + b] = 0.0;
        return;
    }
    double lp[4];
#pragma unroll
    for (int m = 0; m < 4; m++) {
        double s = 0.0;  // reduce(lambda x, y: x + y, ...) over the observed bases, left to right
#pragma unroll
        for (int o = 0; o < 4; o++) {
            const double term = __dmul_rn(log_err[m * 4 + o], __ddiv_rn(e[o], total));
            s = o == 0 ? term : __dadd_rn(s, term);
        }
        lp[m] = __dadd_rn(log_evo[rb * 4 + m], s);
    }
    double t = lp[0];  // reduce(lambda x, y: x + log(1 + exp(y - x)), logBaseProbs)
#pragma unroll
    for (int m = 1; m < 4; m++) t = __dadd_rn(t, log(__dadd_rn(1.0, exp(__dsub_rn(lp[m], t)))));
#pragma unroll
    for (int m = 0; m < 4; m++) post[c * 4 + m] = exp(__dsub_rn(lp[m], t));
}

int col_check(cudaError_t e, const char *what) {
    if (e == cudaSuccess) return MAB_OK;
    mab_set_error("%s: %s", what, cudaGetErrorString(e));
    return MAB_E_CUDA;
}

}  // namespace

extern "C" int mab_columns_accumulate(int64_t n, const int64_t *d_pair_off, const int32_t *d_x, const int32_t *d_y,
                                      const int32_t *d_p, const uint8_t *d_seq_y, const int64_t *d_off_y,
                                      const int64_t *d_col_base, int64_t *d_table, int64_t n_columns, void *stream) {
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev < 1) {
        mab_set_error("no CUDA device: the column accumulation has no CPU fallback");
        return MAB_E_NOGPU;
    }
    if (n < 0 || n_columns < 0 || (n > 0 && (!d_pair_off || !d_seq_y || !d_off_y || !d_col_base || !d_table))) {
        mab_set_error("mab_columns_accumulate: bad arguments");
        return MAB_E_ARG;
    }
    if (n == 0) return MAB_OK;
    int sms = 148, dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t blocks = (n + 7) / 8 < (int64_t) sms * 8 ? (n + 7) / 8 : (int64_t) sms * 8;
    void *t = mab_launch_timer_begin(MAB_K_COL_ACC, stream);
    k_col_accumulate<<<(unsigned) blocks, 256, 0, (cudaStream_t) stream>>>(n, d_pair_off, d_x, d_y, d_p, d_seq_y, d_off_y,
                                                                          d_col_base, (long long *) d_table, n_columns);
    mab_launch_timer_end(t);
    return col_check(cudaGetLastError(), "column accumulate launch");
}

extern "C" int mab_columns_call(const int64_t *d_table, int64_t n_columns, const uint8_t *d_ref, const double *d_log_err,
                                const double *d_log_evo, double *d_post, void *stream) {
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev < 1) {
        mab_set_error("no CUDA device: the column caller has no CPU fallback");
        return MAB_E_NOGPU;
    }
    if (n_columns < 0 || (n_columns > 0 && (!d_table || !d_ref || !d_log_err || !d_log_evo || !d_post))) {
        mab_set_error("mab_columns_call: bad arguments");
        return MAB_E_ARG;
    }
    if (n_columns == 0) return MAB_OK;
    void *t = mab_launch_timer_begin(MAB_K_COL_CALL, stream);
    k_col_call<<<(unsigned) ((n_columns + 255) / 256), 256, 0, (cudaStream_t) stream>>>(
        (const long long *) d_table, n_columns, d_ref, d_log_err, d_log_evo, d_post);
    mab_launch_timer_end(t);
    return col_check(cudaGetLastError(), "column call launch");
}
