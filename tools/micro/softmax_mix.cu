// The softmax instruction mix of the attention kernel in isolation (no TMEM, no MMA, no barriers): one thread owns a
// 128-element score row per "key tile", kept in registers (empty asm statements stand in for the TMEM load / store). Cycles per tile per warp as a function of
// warps per scheduler and of which parts of the mix are present -> what the FP pipes alone allow.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I comfy_gen_server_b200/csrc -o tools/micro/softmax_mix tools/micro/softmax_mix.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "common.cuh"
using namespace cgs;

template <int WPS, int POLY_PAIRS, int POLY_PERIOD, bool DO_MAX, bool DO_SUM, bool DO_PACK, bool DO_SCALE>
__global__ void __launch_bounds__(WPS * 128, 1) k(float* out, long long* cyc, int iters, float scale) {
    uint32_t sr[128];
#pragma unroll
    for (int i = 0; i < 128; ++i) sr[i] = __float_as_uint(0.01f * i + threadIdx.x * 1e-3f - 0.5f);
    float m_run = -INFINITY, l_run = 0.f;
    const uint64_t scale2 = pack_f32x2(scale, scale);
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 128; ++i) asm volatile("" : "+r"(sr[i]));      // "freshly loaded" scores
        float m_new = m_run;
        if (DO_MAX) {
            float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
            for (int i = 0; i < 128; i += 8) {
                mx0 = fmax3(mx0, __uint_as_float(sr[i + 0]), __uint_as_float(sr[i + 1]));
                mx1 = fmax3(mx1, __uint_as_float(sr[i + 2]), __uint_as_float(sr[i + 3]));
                mx2 = fmax3(mx2, __uint_as_float(sr[i + 4]), __uint_as_float(sr[i + 5]));
                mx3 = fmax3(mx3, __uint_as_float(sr[i + 6]), __uint_as_float(sr[i + 7]));
            }
            m_new = fmaxf(fmax3(mx0, mx1, mx2), mx3) * scale;
        }
        if (it == 0) m_run = DO_MAX ? m_new : 1.0f;
        else if (m_new > m_run + 8.f) { l_run *= ex2_approx(m_run - m_new); m_run = m_new; }
        const uint64_t negm2 = pack_f32x2(-m_run, -m_run);
        uint32_t pr[64];
        uint64_t acc0 = 0ull, acc1 = 0ull;
#pragma unroll
        for (int i = 0; i < 64; ++i) {
            uint64_t x2 = pack_u32x2(sr[2 * i], sr[2 * i + 1]);
            if (DO_SCALE) x2 = ffma2(x2, scale2, negm2);
            uint64_t e2;
            if ((i % POLY_PERIOD) < POLY_PAIRS) {
                e2 = ex2_poly2(x2);
            } else {
                float lo, hi;
                unpack_f32x2(x2, lo, hi);
                e2 = pack_f32x2(ex2_approx(lo), ex2_approx(hi));
            }
            if (DO_SUM) { if (i & 1) acc1 = fadd2(acc1, e2); else acc0 = fadd2(acc0, e2); }
            float elo, ehi;
            unpack_f32x2(e2, elo, ehi);
            if (DO_PACK) pr[i] = pack_bf16x2(elo, ehi);
            else pr[i] = __float_as_uint(elo) ^ __float_as_uint(ehi);
        }
        float s0, s1, s2, s3;
        unpack_f32x2(acc0, s0, s1);
        unpack_f32x2(acc1, s2, s3);
        l_run += (s0 + s1) + (s2 + s3);
#pragma unroll
        for (int i = 0; i < 64; ++i) asm volatile("" ::"r"(pr[i]));         // "stored" probabilities
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = l_run + m_run;
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}

template <int WPS, int PP, int PER, bool MX, bool SM, bool PK, bool SC>
void run1(float* out, long long* cyc) {
    const int iters = 400;
    k<WPS, PP, PER, MX, SM, PK, SC><<<148, WPS * 128>>>(out, cyc, iters, 0.18f);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf(" [%s]", cudaGetErrorString(e)); return; }
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    printf("  %dw/sched: %5.0f clk/round (%4.0f per warp-tile)", WPS, (double)c / iters, (double)c / iters / WPS);
}
template <int PP, int PER, bool MX, bool SM, bool PK, bool SC>
void run(const char* name, float* out, long long* cyc) {
    printf("%-38s", name);
    run1<1, PP, PER, MX, SM, PK, SC>(out, cyc);
    run1<2, PP, PER, MX, SM, PK, SC>(out, cyc);
    run1<3, PP, PER, MX, SM, PK, SC>(out, cyc);
    run1<4, PP, PER, MX, SM, PK, SC>(out, cyc);
    printf("\n");
}

int main() {
    float* out; long long* cyc; cudaMalloc(&out, 148 * 512 * 4); cudaMalloc(&cyc, 8);
    run<1, 4, true, true, true, true>("full mix, 25% poly", out, cyc);
    run<0, 4, true, true, true, true>("full mix, MUFU only", out, cyc);
    run<1, 2, true, true, true, true>("full mix, 50% poly", out, cyc);
    run<3, 8, true, true, true, true>("full mix, 37.5% poly", out, cyc);
    run<1, 4, false, true, true, true>("no max, 25% poly", out, cyc);
    run<1, 4, true, false, true, true>("no row sum, 25% poly", out, cyc);
    run<1, 4, true, true, false, true>("no bf16 pack (xor instead), 25% poly", out, cyc);
    run<1, 4, true, true, true, false>("no scale FFMA2, 25% poly", out, cyc);
    run<1, 4, false, false, false, false>("exp only, 25% poly", out, cyc);
    run<0, 4, false, false, false, false>("exp only, MUFU only", out, cyc);
    return 0;
}
