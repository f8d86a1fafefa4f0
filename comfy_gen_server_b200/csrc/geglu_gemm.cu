// G1: GEGLU projection as one tcgen05 GEMM with the gate fused in the epilogue.
//
//   reference:  h = Linear(C -> 8C)(x);  value, gate = h.chunk(2, -1);  y = value * gelu(gate)
//               (comfy/ldm/modules/attention.py:56-63; exact-erf GELU)
//   here:       y[m, n] = (x[m,:] . W[n,:] + b[n]) * gelu(x[m,:] . W[N+n,:] + b[N+n])
//               with the [M, 2N] intermediate never leaving TMEM.
//
// The same kernel (EPI_LINEAR) serves plain Linear layers  y = x W^T (+ b) (+ residual)
// (reference comfy/ops.py:40-52), used for to_q/to_k/to_v/to_out/ff.net.2/proj_in/proj_out.
//
// Structure (one persistent CTA per SM, 192 threads):
//   warp 0      TMA producer: A tile [128 x 64] and B tile [256 x 64] per k-block, 4-stage ring
//   warp 1      TMEM allocator + single-thread tcgen05.mma issuer (M=128, N=256, K=16 x 4 per block)
//   warps 2..5  epilogue: tcgen05.ld accumulator -> bias/GELU/gate -> bf16 -> global
// Accumulators are double-buffered in TMEM (2 x 256 fp32 columns = all 512 columns), so the
// epilogue of tile i overlaps the MMAs of tile i+1.
//
// For GEGLU the B tile's rows [0,128) come from W rows [n0, n0+128) (value half) and rows
// [128,256) from W rows [N+n0, N+n0+128) (gate half): one N=256 MMA produces both halves of
// the same 128 output columns side by side in TMEM.
//
// Tensor-core bound. Algorithmic FLOPs = 2 * M * K * (2N) for GEGLU, 2 * M * K * N for linear.
#include "common.cuh"

namespace cgs {

namespace gemm {
constexpr int BM = 128;
constexpr int BK = 64;
constexpr int BN_HALF = 128;   // rows per B TMA box
constexpr int UMMA_N = 256;
constexpr int STAGES = 4;
constexpr int A_BYTES = BM * BK * 2;          // 16 KiB
constexpr int B_BYTES = UMMA_N * BK * 2;      // 32 KiB
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 256 + 1024;  // + barriers + alignment slack
constexpr int THREADS = 192;
enum { EPI_LINEAR = 0, EPI_GEGLU = 1 };
}  // namespace gemm

__device__ __forceinline__ float gelu_erf(float x) {
    return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}

struct alignas(16) BF8 {
    __nv_bfloat162 v[4];
};

template <int EPI>
__global__ void __launch_bounds__(gemm::THREADS, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __nv_bfloat16* __restrict__ bias, const __nv_bfloat16* residual,
            __nv_bfloat16* out, int M, int N, int K, int ldo, int ldr) {
    using namespace gemm;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
    uint64_t* empty = full + STAGES;
    uint64_t* tfull = empty + STAGES;
    uint64_t* tempty = tfull + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int tiles_m = (M + BM - 1) / BM;
    const int tiles_n = (EPI == EPI_GEGLU) ? (N / BN_HALF) : ((N + UMMA_N - 1) / UMMA_N);
    const int num_tiles = tiles_m * tiles_n;
    const int kblocks = (K + BK - 1) / BK;

    if (warp == 0 && elect_one()) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
    }
    if (warp == 1) {
        if (elect_one()) {
            for (int i = 0; i < STAGES; ++i) {
                mbar_init(&full[i], 1);
                mbar_init(&empty[i], 1);
            }
            for (int i = 0; i < 2; ++i) {
                mbar_init(&tfull[i], 1);
                mbar_init(&tempty[i], 128);
            }
            fence_mbar_init();
        }
        __syncwarp();
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ------------------------------ TMA producer ------------------------------
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                const int m0 = (t % tiles_m) * BM;
                const int nb = t / tiles_m;
                const int row0 = (EPI == EPI_GEGLU) ? nb * BN_HALF : nb * UMMA_N;
                const int row1 = (EPI == EPI_GEGLU) ? N + nb * BN_HALF : nb * UMMA_N + BN_HALF;
                for (int kb = 0; kb < kblocks; ++kb) {
                    mbar_wait(&empty[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * STAGE_BYTES;
                    uint8_t* sb = sa + A_BYTES;
                    mbar_expect_tx(&full[stage], STAGE_BYTES);
                    tma_load_2d(sa, &tmA, &full[stage], kb * BK, m0);
                    tma_load_2d(sb, &tmB, &full[stage], kb * BK, row0);
                    tma_load_2d(sb + BN_HALF * BK * 2, &tmB, &full[stage], kb * BK, row1);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ------------------------------ MMA issuer --------------------------------
        if (elect_one()) {
            const uint32_t idesc = umma_idesc_bf16(BM, UMMA_N, 0);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                mbar_wait(&tempty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * UMMA_N;
                for (int kb = 0; kb < kblocks; ++kb) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
                    const uint64_t adesc = umma_desc_sw128(sa, 16, 1024);
                    const uint64_t bdesc = umma_desc_sw128(sa + A_BYTES, 16, 1024);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) {
                        // +32 bytes per k-step inside the 128-byte swizzle row (>> 4 == 2)
                        umma_ss(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
                    }
                    tc_commit(&empty[stage]);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                tc_commit(&tfull[acc]);
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
        __syncwarp();
    } else {
        // ------------------------------ epilogue ----------------------------------
        const int quad = warp & 3;  // TMEM lane quadrant this warp may access
        const int row_in_tile = quad * 32 + lane;
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
            const int m0 = (t % tiles_m) * BM;
            const int nb = t / tiles_m;
            const int row = m0 + row_in_tile;
            mbar_wait(&tfull[acc], acc_phase);
            tc_fence_after();
            const uint32_t taddr = tmem_base + acc * UMMA_N + (static_cast<uint32_t>(quad * 32) << 16);
            if (EPI == EPI_GEGLU) {
                const int n0 = nb * BN_HALF;
#pragma unroll 1
                for (int c = 0; c < 4; ++c) {
                    uint32_t v[32], g[32];
                    tmem_ld_x32(taddr + c * 32, v);
                    tmem_ld_x32(taddr + BN_HALF + c * 32, g);
                    tmem_wait_ld();
                    const int col = n0 + c * 32;
                    BF8 o[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        BF8 bv, bg;
                        if (bias != nullptr) {
                            bv = *reinterpret_cast<const BF8*>(bias + col + q * 8);
                            bg = *reinterpret_cast<const BF8*>(bias + N + col + q * 8);
                        }
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            float2 fv = make_float2(__uint_as_float(v[q * 8 + 2 * j]),
                                                    __uint_as_float(v[q * 8 + 2 * j + 1]));
                            float2 fg = make_float2(__uint_as_float(g[q * 8 + 2 * j]),
                                                    __uint_as_float(g[q * 8 + 2 * j + 1]));
                            if (bias != nullptr) {
                                const float2 b0 = __bfloat1622float2(bv.v[j]);
                                const float2 b1 = __bfloat1622float2(bg.v[j]);
                                fv.x += b0.x; fv.y += b0.y;
                                fg.x += b1.x; fg.y += b1.y;
                            }
                            o[q].v[j] = __floats2bfloat162_rn(fv.x * gelu_erf(fg.x),
                                                              fv.y * gelu_erf(fg.y));
                        }
                    }
                    if (row < M) {
                        BF8* dst = reinterpret_cast<BF8*>(out + (size_t)row * ldo + col);
#pragma unroll
                        for (int q = 0; q < 4; ++q) dst[q] = o[q];
                    }
                }
            } else {
                const int n0 = nb * UMMA_N;
#pragma unroll 1
                for (int c = 0; c < 8; ++c) {
                    const int col = n0 + c * 32;
                    if (col >= N) break;  // warp-uniform
                    uint32_t v[32];
                    tmem_ld_x32(taddr + c * 32, v);
                    tmem_wait_ld();
                    BF8 o[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        BF8 bv, rv;
                        if (bias != nullptr) bv = *reinterpret_cast<const BF8*>(bias + col + q * 8);
                        if (residual != nullptr && row < M)
                            rv = *reinterpret_cast<const BF8*>(residual + (size_t)row * ldr + col + q * 8);
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            float2 fv = make_float2(__uint_as_float(v[q * 8 + 2 * j]),
                                                    __uint_as_float(v[q * 8 + 2 * j + 1]));
                            if (bias != nullptr) {
                                const float2 b0 = __bfloat1622float2(bv.v[j]);
                                fv.x += b0.x; fv.y += b0.y;
                            }
                            if (residual != nullptr && row < M) {
                                const float2 r0 = __bfloat1622float2(rv.v[j]);
                                fv.x += r0.x; fv.y += r0.y;
                            }
                            o[q].v[j] = __floats2bfloat162_rn(fv.x, fv.y);
                        }
                    }
                    if (row < M) {
                        BF8* dst = reinterpret_cast<BF8*>(out + (size_t)row * ldo + col);
#pragma unroll
                        for (int q = 0; q < 4; ++q) dst[q] = o[q];
                    }
                }
            }
            tc_fence_before();
            mbar_arrive(&tempty[acc]);
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, 512);
}

static int launch_gemm(int epi, const void* x, const void* w, const void* bias,
                       const void* residual, void* out, long long M, int N, int K, long long lda,
                       long long ldw, long long ldo, long long ldr, cudaStream_t stream) {
    using namespace gemm;
    const char* who = epi == EPI_GEGLU ? "geglu_gemm" : "linear";
    if (M < 0 || N <= 0 || K <= 0) {
        set_last_error("%s: bad sizes M=%lld N=%d K=%d", who, M, N, K);
        return kBadArgument;
    }
    if (M == 0) return kOk;
    if (M > 0x7fffffffLL) {
        set_last_error("%s: M=%lld too large", who, M);
        return kUnsupportedShape;
    }
    if (K % 8 != 0 || lda % 8 != 0 || ldw % 8 != 0 || ldo % 8 != 0 || (residual && ldr % 8 != 0)) {
        set_last_error("%s: K and leading dimensions must be multiples of 8 (K=%d lda=%lld ldw=%lld ldo=%lld)",
                       who, K, lda, ldw, ldo);
        return kUnsupportedShape;
    }
    if (epi == EPI_GEGLU && N % BN_HALF != 0) {
        set_last_error("geglu_gemm: N=%d must be a multiple of %d", N, BN_HALF);
        return kUnsupportedShape;
    }
    if (epi == EPI_LINEAR && N % 32 != 0) {
        set_last_error("linear: N=%d must be a multiple of 32", N);
        return kUnsupportedShape;
    }
    if ((reinterpret_cast<uintptr_t>(out) & 0xF) || (reinterpret_cast<uintptr_t>(bias) & 0xF) ||
        (reinterpret_cast<uintptr_t>(residual) & 0xF)) {
        set_last_error("%s: out/bias/residual pointers must be 16-byte aligned", who);
        return kBadArgument;
    }
    CUtensorMap tmA, tmB;
    {
        const uint64_t dims[2] = {(uint64_t)K, (uint64_t)M};
        const uint64_t strides[1] = {(uint64_t)lda * 2};
        const uint32_t box[2] = {BK, BM};
        int rc = make_tmap_bf16(&tmA, x, 2, dims, strides, box);
        if (rc != kOk) return rc;
    }
    {
        const uint64_t rows = (epi == EPI_GEGLU) ? 2ull * N : (uint64_t)N;
        const uint64_t dims[2] = {(uint64_t)K, rows};
        const uint64_t strides[1] = {(uint64_t)ldw * 2};
        const uint32_t box[2] = {BK, BN_HALF};
        int rc = make_tmap_bf16(&tmB, w, 2, dims, strides, box);
        if (rc != kOk) return rc;
    }
    const int tiles_m = (int)((M + BM - 1) / BM);
    const int tiles_n = (epi == EPI_GEGLU) ? N / BN_HALF : (N + UMMA_N - 1) / UMMA_N;
    const long long tiles = (long long)tiles_m * tiles_n;
    const int grid = (int)(tiles < sm_count() ? tiles : sm_count());
    auto bp = static_cast<const __nv_bfloat16*>(bias);
    auto rp = static_cast<const __nv_bfloat16*>(residual);
    auto op = static_cast<__nv_bfloat16*>(out);
    if (epi == EPI_GEGLU) {
        CGS_CUDA_CHECK(cudaFuncSetAttribute(gemm_kernel<EPI_GEGLU>,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
        gemm_kernel<EPI_GEGLU><<<grid, THREADS, SMEM_BYTES, stream>>>(tmA, tmB, bp, rp, op, (int)M, N, K,
                                                                      (int)ldo, (int)ldr);
    } else {
        CGS_CUDA_CHECK(cudaFuncSetAttribute(gemm_kernel<EPI_LINEAR>,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
        gemm_kernel<EPI_LINEAR><<<grid, THREADS, SMEM_BYTES, stream>>>(tmA, tmB, bp, rp, op, (int)M, N, K,
                                                                       (int)ldo, (int)ldr);
    }
    CGS_CUDA_CHECK(cudaGetLastError());
    return kOk;
}

}  // namespace cgs

extern "C" int cgs_geglu_gemm_fwd(const void* x, const void* w, const void* bias, void* y,
                                  long long M, int N, int K, long long lda, long long ldw,
                                  long long ldy, void* stream) {
    return cgs::launch_gemm(cgs::gemm::EPI_GEGLU, x, w, bias, nullptr, y, M, N, K, lda, ldw, ldy, 0,
                            static_cast<cudaStream_t>(stream));
}

extern "C" int cgs_linear_fwd(const void* x, const void* w, const void* bias, const void* residual,
                              void* y, long long M, int N, int K, long long lda, long long ldw,
                              long long ldy, long long ldr, void* stream) {
    return cgs::launch_gemm(cgs::gemm::EPI_LINEAR, x, w, bias, residual, y, M, N, K, lda, ldw, ldy,
                            ldr, static_cast<cudaStream_t>(stream));
}
