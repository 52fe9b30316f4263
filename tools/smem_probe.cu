// smem_probe.cu -- how many LSU cycles does a shared-memory table lookup cost on B200, by access width and by how
// many DISTINCT addresses a warp touches?  The pair-HMM wavefront fetches four logAdd coefficients (32 B) per lane
// from a 4-row table; this probe decides between replicated tables (conflict-free, every lane its own 16 B) and a
// single copy (few distinct rows, served by broadcast).
//
// Every kernel runs `iters` x 8 independent loads per thread; the row index is derived from a per-thread value that
// changes every iteration (so nothing is hoisted).  Reported: SM-cycles per warp-level load instruction.
//
//   pattern 0: all lanes the same 16-byte row                     (pure broadcast)
//   pattern 1: row = pseudo-random in 0..3, single table copy     (<= 4 distinct addresses per warp)
//   pattern 2: row = pseudo-random in 0..3, 8 replicas (lane & 7) (what wavefront v4 does)
//   pattern 3: every lane its own row, consecutive 16 B           (conflict-free, 512 B per warp)
//   pattern 4: as 1 but only 26 of 32 lanes active
//   pattern 5: as 2 but lanes 8-15 and 24-31 predicated off (two empty quarter-warps)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/smem_probe tools/smem_probe.cu
#include <cuda_runtime.h>

#include <cstdio>

#define UNROLL 8

template <int WIDTH, int PATTERN>
__global__ void __launch_bounds__(256) k_probe(double *out, int iters, unsigned seed) {
    __shared__ __align__(16) double tab[4 * 2 * 8 * 2 + 32 * 2];  // rows x halves x replicas x 2 doubles (+ pattern 3)
    for (int i = threadIdx.x; i < (int) (sizeof(tab) / 8); i += blockDim.x) tab[i] = 1.0 + i * 1e-3;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    unsigned state = seed + threadIdx.x * 2654435761u + blockIdx.x * 40503u;
    const unsigned base = (unsigned) __cvta_generic_to_shared(tab);
    unsigned acc = 0;
    bool active = true;
    if (PATTERN == 4) active = lane < 13 || (lane >= 16 && lane < 29);
    if (PATTERN == 5) active = (lane & 8) == 0;
    for (int it = 0; it < iters; it++) {
        state = state * 1664525u + 1013904223u;  // one LCG step per 8 loads; load u uses bits 2u+8, 2u+9
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            const unsigned r = (state >> (2 * u + 8)) & 3;
            unsigned addr;
            if (PATTERN == 0) addr = base + ((it + u) & 3) * 16;
            else if (PATTERN == 1 || PATTERN == 4) addr = base + r * 32;
            else if (PATTERN == 2 || PATTERN == 5) addr = base + r * 256 + (lane & 7) * 16;
            else addr = base + lane * WIDTH + (r & 1) * 512;
            if (active) {
                if (WIDTH == 16) {
                    unsigned a, b, c, d;
                    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(addr));
                    acc ^= a ^ b;
                    acc ^= c ^ d;
                } else if (WIDTH == 8) {
                    unsigned a, b;
                    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(a), "=r"(b) : "r"(addr));
                    acc ^= a ^ b;
                } else {
                    unsigned a;
                    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(a) : "r"(addr));
                    acc ^= a;
                }
            }
        }
    }
    if (acc == 0x12345678u) out[0] = acc;
}

// The same loop without the load: its time is subtracted (address arithmetic + accumulate).
template <int WIDTH>
__global__ void __launch_bounds__(256) k_empty(double *out, int iters, unsigned seed) {
    unsigned state = seed + threadIdx.x * 2654435761u + blockIdx.x * 40503u;
    unsigned acc = 0;
    for (int it = 0; it < iters; it++) {
        state = state * 1664525u + 1013904223u;
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            const unsigned r = (state >> (2 * u + 8)) & 3;
            acc ^= r * 32 + (WIDTH == 16 ? u : 0);
        }
    }
    if (acc == 0x12345678u) out[0] = acc;
}

template <typename K>
float time_kernel(K launch) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    launch();
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 3; r++) {
        cudaEventRecord(e0);
        launch();
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return best;
}

int main() {
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, 0) != cudaSuccess) return 1;
    double *d_out;
    cudaMalloc(&d_out, 8);
    const int sms = p.multiProcessorCount, blocks = sms * 8, threads = 256, iters = 2048;
    int clk_khz = 0;
    cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const double warp_loads_per_sm = 8.0 * (threads / 32) * (double) iters * UNROLL;
    printf("{\"gpu\": \"%s\", \"sms\": %d", p.name, sms);
#define RUN(W, P)                                                                                                   \
    {                                                                                                               \
        float ms = time_kernel([&] { k_probe<W, P><<<blocks, threads>>>(d_out, iters, 12345u); });                  \
        float ms0 = time_kernel([&] { k_empty<W><<<blocks, threads>>>(d_out, iters, 12345u); });                    \
        printf(", \"w%d_p%d_cycles_per_warp_load\": %.3f, \"w%d_p%d_cycles_net\": %.3f", W, P,                       \
               ms * 1e-3 * clk_khz * 1e3 / warp_loads_per_sm, W, P, (ms - ms0) * 1e-3 * clk_khz * 1e3 / warp_loads_per_sm); \
    }
    RUN(16, 0) RUN(16, 1) RUN(16, 2) RUN(16, 3) RUN(16, 4) RUN(16, 5)
    RUN(8, 0) RUN(8, 1) RUN(8, 2) RUN(8, 3)
    RUN(4, 0) RUN(4, 1) RUN(4, 3)
    printf("}\n");
    return cudaGetLastError() == cudaSuccess ? 0 : 2;
}
