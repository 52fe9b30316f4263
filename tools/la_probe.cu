// la_probe.cu -- microbenchmarks behind the wavefront kernels' design (run on the GPU box: nvcc -arch=sm_100a -O3
// -fmad=false -o tools/la_probe tools/la_probe.cu && tools/la_probe):
//   1. dependent-issue latency of DADD / DMUL / DSETP+FSEL chains (one warp per SM sub-partition);
//   2. logAdd throughput per SM as a function of resident warps x independent logAdds per thread (ILP),
//      i.e. how much parallelism the FP64 pipe needs before it saturates.
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ double2 lds_v2(unsigned addr) { double2 v; asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(addr)); return v; }
__device__ __forceinline__ double2 lds_v2_128(unsigned addr) { double2 v; asm volatile("ld.shared.v2.f64 {%0, %1}, [%2+128];" : "=d"(v.x), "=d"(v.y) : "r"(addr)); return v; }

__constant__ double c_coef[16] = {
    (double) -0.009350833524763f, (double) 0.130659527668286f, (double) 0.498799810682272f, (double) 0.693203116424741f,
    (double) -0.014532321752540f, (double) 0.139942324101744f, (double) 0.495635523139337f, (double) 0.692140569840976f,
    (double) -0.004605031767994f, (double) 0.063427417320019f, (double) 0.695956496475118f, (double) 0.514272634594009f,
    (double) -0.000458661602210f, (double) 0.009695946122598f, (double) 0.930734667215156f, (double) 0.168037164329057f};

// LA: bit 0 = integer compares; bits 1.. = coefficient source: 0 doubles (2 x LDS.128), 1 floats (LDS.128 + 4 cvt),
// 2 high words (LDS.128) + low bytes (LDS.32 + 4 PRMT), 3 floats in registers picked by FSEL (no shared memory),
// 4 = (c3, c2) floats by LDS.64 + (c1, c0) doubles by LDS.128; bit 4 (16) = |x - y| and the sign bit instead of the
// compare-and-swap in front of the subtraction
struct Tabs {
    double2 d[4][2][8];
    float4 f[4][8];
    uint4 hi[4][8];
    unsigned lo[4][32];
    float2 f32[4][16];
};
struct Addr { unsigned d, f, hi, lo, f32; };
template <int LA, int N>
__device__ __forceinline__ void la_n(const Addr &a, double (&x)[N], const double (&y)[N]) {
    double hi[N], lo[N], d[N], q[N], c3[N], c2[N], c1[N], c0[N];
    unsigned pc[N];
    bool use[N];
    constexpr int SRC = (LA >> 1) & 7;
#pragma unroll
    for (int i = 0; i < N; i++) {
        if (LA & 16) {
            const double df = __dsub_rn(x[i], y[i]);
            const bool lt = __double2hiint(df) < 0;  // x < y (x == y gives +0; NaN: the result is `hi` = either)
            hi[i] = lt ? y[i] : x[i]; lo[i] = lt ? x[i] : y[i];
            d[i] = fabs(df);
        } else {
            const bool lt = x[i] < y[i]; hi[i] = lt ? y[i] : x[i]; lo[i] = lt ? x[i] : y[i];
            d[i] = __dsub_rn(hi[i], lo[i]);
        }
    }
#pragma unroll
    for (int i = 0; i < N; i++) {
        pc[i] = 0;
        if (LA & 1) {
            const unsigned h = (unsigned) __double2hiint(d[i]), l = (unsigned) __double2loint(d[i]);
            unsigned hp;
            asm("{\n\t.reg .u32 t;\n\tadd.cc.u32 t, %1, 0xffffffff;\n\taddc.u32 %0, %2, 0;\n\t}" : "=r"(hp) : "r"(l), "r"(h));
            if (SRC == 3) {
                pc[i] = (hp > 0x3FF00000u ? 1 : 0) | (hp > 0x40040000u ? 2 : 0) | (hp > 0x40120000u ? 4 : 0);
            } else {
                if (hp > 0x3FF00000u) pc[i] += 128;
                if (hp > 0x40040000u) pc[i] += 128;
                if (hp > 0x40120000u) pc[i] += 128;
            }
            use[i] = h < 0x401E0000u;
        } else {
            if (SRC == 3) {
                pc[i] = (d[i] > 1.0 ? 1 : 0) | (d[i] > 2.5 ? 2 : 0) | (d[i] > 4.5 ? 4 : 0);
            } else {
                if (d[i] > 1.0) pc[i] += 128;
                if (d[i] > 2.5) pc[i] += 128;
                if (d[i] > 4.5) pc[i] += 128;
            }
            use[i] = d[i] < 7.5;
        }
    }
#pragma unroll
    for (int i = 0; i < N; i++) {
        if (SRC == 0) {
            const double2 u = lds_v2(a.d + 2 * pc[i]), v = lds_v2_128(a.d + 2 * pc[i]);
            c3[i] = u.x, c2[i] = u.y, c1[i] = v.x, c0[i] = v.y;
        } else if (SRC == 1) {
            float f3, f2, f1, f0;
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(f3), "=f"(f2), "=f"(f1), "=f"(f0) : "r"(a.f + pc[i]));
            c3[i] = (double) f3, c2[i] = (double) f2, c1[i] = (double) f1, c0[i] = (double) f0;
        } else if (SRC == 2) {
            unsigned h3, h2, h1, h0, l;
            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(h3), "=r"(h2), "=r"(h1), "=r"(h0) : "r"(a.hi + pc[i]));
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(l) : "r"(a.lo + pc[i]));
            c3[i] = __hiloint2double(h3, __byte_perm(l, 0, 0x0444));
            c2[i] = __hiloint2double(h2, __byte_perm(l, 0, 0x1444));
            c1[i] = __hiloint2double(h1, __byte_perm(l, 0, 0x2444));
            c0[i] = __hiloint2double(h0, __byte_perm(l, 0, 0x3444));
        } else if (SRC == 3) {
            const bool p1 = pc[i] & 1, p2 = pc[i] & 2, p3 = pc[i] & 4;
#define PICK(k) (p2 ? (p3 ? (float) c_coef[12 + k] : (float) c_coef[8 + k]) : (p1 ? (float) c_coef[4 + k] : (float) c_coef[k]))
            c3[i] = (double) PICK(0), c2[i] = (double) PICK(1), c1[i] = (double) PICK(2), c0[i] = (double) PICK(3);
#undef PICK
        } else {
            float f3, f2;
            asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(f3), "=f"(f2) : "r"(a.f32 + pc[i]));
            const double2 v = lds_v2_128(a.d + 2 * pc[i]);
            c3[i] = (double) f3, c2[i] = (double) f2, c1[i] = v.x, c0[i] = v.y;
        }
    }
#pragma unroll
    for (int i = 0; i < N; i++) q[i] = __dmul_rn(c3[i], d[i]);
#pragma unroll
    for (int i = 0; i < N; i++) q[i] = __dadd_rn(q[i], c2[i]);
#pragma unroll
    for (int i = 0; i < N; i++) q[i] = __dmul_rn(q[i], d[i]);
#pragma unroll
    for (int i = 0; i < N; i++) q[i] = __dadd_rn(q[i], c1[i]);
#pragma unroll
    for (int i = 0; i < N; i++) q[i] = __dmul_rn(q[i], d[i]);
#pragma unroll
    for (int i = 0; i < N; i++) q[i] = __dadd_rn(q[i], c0[i]);
#pragma unroll
    for (int i = 0; i < N; i++) q[i] = __dadd_rn(q[i], lo[i]);
#pragma unroll
    for (int i = 0; i < N; i++) x[i] = use[i] ? q[i] : hi[i];
}

template <int LA, int N>
__global__ void k_la(const double *in, double *out, int iters, long long *cycles) {
    __shared__ __align__(16) Tabs tab;
    for (int i = threadIdx.x; i < 128; i += blockDim.x) {
        const int piece = i / 32, rep = i % 32;
        const double *c = c_coef + piece * 4;
        unsigned lo = 0;
        for (int k = 0; k < 4; k++) lo |= ((unsigned) __double2loint(c[k]) >> 24) << (8 * k);
        tab.lo[piece][rep] = lo;
        if (rep < 16) tab.f32[piece][rep] = make_float2((float) c[0], (float) c[1]);
        if (rep < 8) {
            tab.d[piece][0][rep] = make_double2(c[0], c[1]);
            tab.d[piece][1][rep] = make_double2(c[2], c[3]);
            tab.f[piece][rep] = make_float4((float) c[0], (float) c[1], (float) c[2], (float) c[3]);
            tab.hi[piece][rep] = make_uint4(__double2hiint(c[0]), __double2hiint(c[1]), __double2hiint(c[2]), __double2hiint(c[3]));
        }
    }
    __syncthreads();
    Addr a;
    a.d = (unsigned) __cvta_generic_to_shared(&tab.d[0][0][threadIdx.x & 7]);
    a.f = (unsigned) __cvta_generic_to_shared(&tab.f[0][threadIdx.x & 7]);
    a.hi = (unsigned) __cvta_generic_to_shared(&tab.hi[0][threadIdx.x & 7]);
    a.lo = (unsigned) __cvta_generic_to_shared(&tab.lo[0][threadIdx.x & 31]);
    a.f32 = (unsigned) __cvta_generic_to_shared(&tab.f32[0][threadIdx.x & 15]);
    double x[N], y[N];
#pragma unroll
    for (int i = 0; i < N; i++) { x[i] = in[(threadIdx.x + 37 * i) & 255]; y[i] = in[(threadIdx.x + 91 * i + 13) & 255]; }
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
        la_n<LA, N>(a, x, y);
#pragma unroll
        for (int i = 0; i < N; i++) y[i] = __dadd_rn(y[i], -0.37);  // keep |x - y| moving through the pieces
    }
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < N; i++) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

template <int OP>
__global__ void k_lat(const double *in, double *out, int iters, long long *cycles) {
    double x = in[threadIdx.x], y = in[threadIdx.x + 32];
    const long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 16; u++) {
            if (OP == 0) x = __dadd_rn(x, y);
            if (OP == 1) x = __dmul_rn(x, y);
            if (OP == 2) { const bool lt = x < y; const double t = lt ? y : x; y = lt ? x : y; x = t; }
            if (OP == 3) x = __fma_rn(x, y, y);
        }
    }
    const long long t1 = clock64();
    out[threadIdx.x] = x + y;
    if (threadIdx.x == 0) *cycles = t1 - t0;
}

int main() {
    double h[512];
    for (int i = 0; i < 512; i++) h[i] = -1.0 - 0.173 * (i % 97);
    double *d_in, *d_out;
    long long *d_cyc, cyc;
    cudaMalloc(&d_in, sizeof(h));
    cudaMalloc(&d_out, 8 * 2048 * 1024);
    cudaMalloc(&d_cyc, 8);
    cudaMemcpy(d_in, h, sizeof(h), cudaMemcpyHostToDevice);
    const int iters = 2000;
    const char *names[4] = {"DADD", "DMUL", "DSETP+2xFSEL swap", "DFMA"};
    for (int op = 0; op < 4; op++) {
        if (op == 0) k_lat<0><<<1, 32>>>(d_in, d_out, iters, d_cyc);
        if (op == 1) k_lat<1><<<1, 32>>>(d_in, d_out, iters, d_cyc);
        if (op == 2) k_lat<2><<<1, 32>>>(d_in, d_out, iters, d_cyc);
        if (op == 3) k_lat<3><<<1, 32>>>(d_in, d_out, iters, d_cyc);
        cudaMemcpy(&cyc, d_cyc, 8, cudaMemcpyDeviceToHost);
        printf("{\"probe\": \"latency\", \"op\": \"%s\", \"cycles_per_op\": %.2f}\n", names[op], (double) cyc / (iters * 16.0));
    }
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    const int sms = prop.multiProcessorCount;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int las[] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 16, 17, 18, 19, 20, 21, 24, 25};
    for (int la : las) {
        for (int warps = 8; warps <= 32; warps += 4) {  // resident warps per SM (one block per SM)
            for (int ilp = 1; ilp <= 2; ilp++) {
                for (int rep = 0; rep < 2; rep++) {
                    cudaEventRecord(e0);
#define RUN(LA) if (la == LA) { if (ilp == 1) k_la<LA, 1><<<sms, warps * 32>>>(d_in, d_out, iters, d_cyc); else k_la<LA, 2><<<sms, warps * 32>>>(d_in, d_out, iters, d_cyc); }
                    RUN(0) RUN(1) RUN(2) RUN(3) RUN(4) RUN(5) RUN(6) RUN(7) RUN(8) RUN(9) RUN(16) RUN(17) RUN(18) RUN(19) RUN(20) RUN(21) RUN(24) RUN(25)
                    cudaEventRecord(e1);
                    cudaEventSynchronize(e1);
                }
                float ms;
                cudaEventElapsedTime(&ms, e0, e1);
                cudaMemcpy(&cyc, d_cyc, 8, cudaMemcpyDeviceToHost);
                if (warps == 8 && ilp == 1) {  // every variant must give the same bits
                    double v;
                    cudaMemcpy(&v, d_out + 5, 8, cudaMemcpyDeviceToHost);
                    printf("{\"probe\": \"checksum\", \"la\": %d, \"v\": \"%a\"}\n", la, v);
                }
                const double n_la = (double) sms * warps * 32 * ilp * iters;
                printf("{\"probe\": \"logadd\", \"la\": %d, \"warps_per_sm\": %d, \"ilp\": %d, \"ms\": %.3f, \"G_logadd_per_s\": %.1f, "
                       "\"sm_cycles_per_warp_logadd\": %.2f}\n", la, warps, ilp, ms, n_la / ms / 1e6, (double) cyc / (iters * ilp * warps));
            }
        }
    }
    double hc[256];
    cudaMemcpy(hc, d_out, sizeof(hc), cudaMemcpyDeviceToHost);
    printf("{\"probe\": \"checksum\", \"v\": %.17g}\n", hc[5]);
    return 0;
}
