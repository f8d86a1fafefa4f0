// MUFU.EX2 issue rate per SM sub-partition as a function of resident warps (1 CTA per SM, W warps).
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(float* out, long long* cyc, int iters) {
    float v[16];
    for (int i = 0; i < 16; ++i) v[i] = threadIdx.x * 1e-3f + i;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(v[i]));
    }
    long long t1 = clock64();
    float s = 0;
    for (int i = 0; i < 16; ++i) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
__global__ void kf(float* out, long long* cyc, int iters) {   // FFMA2-heavy polynomial stand-in, for comparison
    float v[16];
    for (int i = 0; i < 16; ++i) v[i] = threadIdx.x * 1e-3f + i;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = fmaf(v[i], 0.999f, 0.001f);
    }
    long long t1 = clock64();
    float s = 0;
    for (int i = 0; i < 16; ++i) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
int main() {
    float* out; long long* cyc; cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
    const int iters = 2000;
    for (int warps : {4, 8, 12, 16, 32}) {
        for (int which = 0; which < 2; ++which) {
            if (which == 0) k<<<148, warps * 32>>>(out, cyc, iters); else kf<<<148, warps * 32>>>(out, cyc, iters);
            cudaDeviceSynchronize();
            long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
            double per_warp_instr = (double)c / (iters * 16.0);          // cycles between this warp's instructions
            double per_smsp = per_warp_instr / (warps / 4.0);            // cycles per warp-instruction per sub-partition
            printf("%s warps/SM=%2d (per SMSP %d): %.2f clk per warp-instr per warp, %.2f clk per warp-instr per SMSP -> %.1f lanes/clk/SM\n",
                   which == 0 ? "MUFU.EX2" : "FFMA    ", warps, warps / 4, per_warp_instr, per_smsp, 32.0 * 4 / per_smsp);
        }
    }
    return 0;
}
