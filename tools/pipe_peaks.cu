// pipe_peaks.cu -- measures the non-tensor issue peaks the pair-HMM roofline is quoted against (SURVEY 8d asks the
// builder to measure them: MEASURED_PEAKS.json only has HBM and bf16 tensor numbers).
//
//   dadd / dmul   : independent chains of non-contracted FP64 add / multiply (what the bit-exact DP is made of)
//   dfma          : FP64 FMA (nominal "FP64 TFLOP/s" figure, for context; the DP cannot use it)
//   dsetp         : FP64 compare feeding an integer select (does DSETP share the FP64 pipe?)
//   fadd / ffma   : FP32 add / FMA
//   iadd          : INT32 add (ALU pipe)
//   dadd_lat      : one dependent DADD chain per warp, one warp per SM sub-partition -> latency in cycles
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/pipe_peaks tools/pipe_peaks.cu
// Prints one JSON object on stdout.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>

#define ITERS 4096
#define CHAINS 8

template <int OP>
__global__ void k_pipe(double *out, double a, double b, int iters) {
    double v[CHAINS];
    float f[CHAINS];
    int n[CHAINS];
#pragma unroll
    for (int c = 0; c < CHAINS; c++) {
        v[c] = a + threadIdx.x * 1e-9 + c;
        f[c] = (float) v[c];
        n[c] = threadIdx.x + c;
    }
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int c = 0; c < CHAINS; c++) {
            if (OP == 0) v[c] = __dadd_rn(v[c], b);
            if (OP == 1) v[c] = __dmul_rn(v[c], b);
            if (OP == 2) v[c] = __fma_rn(v[c], b, a);
            if (OP == 3) n[c] += (v[c] > b) ? 3 : 1;                     // DSETP + SEL + IADD per step
            if (OP == 7) v[c] = __dadd_rn(v[c], (double) f[c]);          // F2F.F64.F32 + DADD per step
            if (OP == 8) f[c] = __fadd_rn(f[c], 1.0f), v[c] = (double) f[c];  // FADD + F2F.F64.F32 per step
            if (OP == 4) f[c] = __fadd_rn(f[c], (float) b);
            if (OP == 5) f[c] = __fmaf_rn(f[c], (float) b, (float) a);
            if (OP == 6) n[c] = n[c] * 3 + i;
        }
    }
    double s = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; c++) s += v[c] + f[c] + n[c];
    if (s == 12345.678) out[0] = s;
}

__global__ void k_dadd_latency(double *out, double b, int iters, long long *cycles) {
    double v = threadIdx.x;
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
        v = __dadd_rn(v, b);
        v = __dadd_rn(v, b);
        v = __dadd_rn(v, b);
        v = __dadd_rn(v, b);
    }
    long long t1 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) cycles[0] = t1 - t0;
    if (v == 12345.678) out[0] = v;
}

template <int OP>
double run(int sms, double *d_out, int ops_per_step) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int blocks = sms * 8, threads = 256;
    k_pipe<OP><<<blocks, threads>>>(d_out, 1.0000001, 0.9999999, ITERS);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0);
        k_pipe<OP><<<blocks, threads>>>(d_out, 1.0000001, 0.9999999, ITERS);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double ops = (double) blocks * threads * ITERS * CHAINS * ops_per_step;
    return ops / (best * 1e-3) / 1e12;  // T lane-ops / s
}

int main() {
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, 0) != cudaSuccess) {
        fprintf(stderr, "no device\n");
        return 1;
    }
    double *d_out;
    long long *d_cyc, h_cyc = 0;
    cudaMalloc(&d_out, 8);
    cudaMalloc(&d_cyc, 8);
    const int sms = p.multiProcessorCount;
    const double dadd = run<0>(sms, d_out, 1), dmul = run<1>(sms, d_out, 1), dfma = run<2>(sms, d_out, 1),
                 dsetp = run<3>(sms, d_out, 1), fadd = run<4>(sms, d_out, 1), ffma = run<5>(sms, d_out, 1),
                 imad = run<6>(sms, d_out, 1), f2f_dadd = run<7>(sms, d_out, 1), f2f_fadd = run<8>(sms, d_out, 1);
    k_dadd_latency<<<1, 32>>>(d_out, 0.5, 4096, d_cyc);
    cudaDeviceSynchronize();
    k_dadd_latency<<<1, 32>>>(d_out, 0.5, 4096, d_cyc);
    cudaMemcpy(&h_cyc, d_cyc, 8, cudaMemcpyDeviceToHost);
    int clk = 0;
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"clock_khz_attr\": %d, \"dadd_tops\": %.3f, \"dmul_tops\": %.3f, "
           "\"dfma_tops\": %.3f, \"dsetp_sel_tops\": %.3f, \"fadd_tops\": %.3f, \"ffma_tops\": %.3f, \"imad_tops\": %.3f, "
           "\"f2f_f64_f32_plus_dadd_tops\": %.3f, \"f2f_f64_f32_plus_fadd_tops\": %.3f, \"dadd_latency_cycles\": %.2f}\n",
           p.name, sms, clk, dadd, dmul, dfma, dsetp, fadd, ffma, imad, f2f_dadd, f2f_fadd, (double) h_cyc / (4096.0 * 4));
    return cudaGetLastError() == cudaSuccess ? 0 : 2;
}
