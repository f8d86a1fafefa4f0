// L1: LayerNorm over the channel dimension of token-major activations, bf16 in / bf16 out,
// statistics in fp32 (two-pass in registers). Replaces the reference's ops.LayerNorm ->
// F.layer_norm call (reference comfy/ops.py:97-113) used as norm1/norm2/norm3 of every
// BasicTransformerBlock (reference comfy/ldm/modules/attention.py:450-453).
//
// Also the fused residual form  x <- x + r ; y <- LN(x)  which is the reference's
// `x += n` (attention.py:527,567) followed by the next norm (attention.py:534,570) in one pass.
//
// HBM-bound: one warp per row, 16-byte loads kept in registers between the two passes, so each
// element is read once and written once. Algorithmic bytes = 2 * rows * C * 2.
#include "common.cuh"

namespace cgs {

struct alignas(16) Vec8 {
    __nv_bfloat162 v[4];
};

__device__ __forceinline__ Vec8 ld_stream(const Vec8* p) {
    Vec8 r;
    uint32_t* u = reinterpret_cast<uint32_t*>(&r);
    asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3])
                 : "l"(p));
    return r;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// MAXV = max 16-byte vectors per lane  (C <= 256 * MAXV). The row is kept in registers as packed bf16 (4 registers
// per 8 elements) and unpacked in each of the three passes: fewer registers -> 4 blocks (32 rows) per SM in flight.
template <int MAXV, bool ADD_RESIDUAL>
__global__ void __launch_bounds__(256, (MAXV <= 5 ? 4 : 2))
layernorm_kernel(const __nv_bfloat16* x, const __nv_bfloat16* res,
                 const __nv_bfloat16* __restrict__ gamma, const __nv_bfloat16* __restrict__ beta,
                 __nv_bfloat16* x_out, __nv_bfloat16* __restrict__ y, long long rows,
                 int C, float eps) {
    const int lane = threadIdx.x & 31;
    const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int nvec = C >> 3;
    const Vec8* xr = reinterpret_cast<const Vec8*>(x + row * C);
    const Vec8* rr = ADD_RESIDUAL ? reinterpret_cast<const Vec8*>(res + row * C) : nullptr;

    Vec8 a[MAXV];
    Vec8 r[ADD_RESIDUAL ? MAXV : 1];
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int v = lane + i * 32;
        if (v < nvec) {
            // x_out may alias x (in-place residual update): plain loads on that path.
            a[i] = ADD_RESIDUAL ? xr[v] : ld_stream(xr + v);
            if (ADD_RESIDUAL) r[i] = ld_stream(rr + v);
        }
    }
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int v = lane + i * 32;
        if (v < nvec) {
            if (ADD_RESIDUAL) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    // the reference adds in bf16 (x += n on bf16 tensors): round the sum to bf16
                    // before normalising, and store it as the new residual stream.
                    a[i].v[j] = __hadd2_rn(a[i].v[j], r[i].v[j]);
                }
                reinterpret_cast<Vec8*>(x_out + row * C)[v] = a[i];
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 t = __bfloat1622float2(a[i].v[j]);
                sum += t.x + t.y;
            }
        }
    }
    const float mean = warp_sum(sum) / (float)C;
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        if (lane + i * 32 < nvec) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 t = __bfloat1622float2(a[i].v[j]);
                const float d0 = t.x - mean, d1 = t.y - mean;
                sq += d0 * d0 + d1 * d1;
            }
        }
    }
    const float rstd = rsqrtf(warp_sum(sq) / (float)C + eps);

    const Vec8* g8 = reinterpret_cast<const Vec8*>(gamma);
    const Vec8* b8 = reinterpret_cast<const Vec8*>(beta);
    Vec8* yr = reinterpret_cast<Vec8*>(y + row * C);
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int v = lane + i * 32;
        if (v < nvec) {
            const Vec8 g = g8[v];
            const Vec8 b = b8[v];
            Vec8 o;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 t = __bfloat1622float2(a[i].v[j]);
                const float2 gg = __bfloat1622float2(g.v[j]);
                const float2 bb = __bfloat1622float2(b.v[j]);
                const float o0 = (t.x - mean) * rstd * gg.x + bb.x;
                const float o1 = (t.y - mean) * rstd * gg.y + bb.y;
                o.v[j] = __floats2bfloat162_rn(o0, o1);
            }
            yr[v] = o;
        }
    }
}

// Plain LayerNorm, the form every block of the large-batch path uses now that the residual adds ride in the GEMM
// epilogues. The one-row-per-warp kernel above reached only 0.40-0.45 of the measured copy bandwidth at the UNet's
// sizes -- and a persistent, prefetching version of it gained nothing: it is ISSUE-bound, not memory-bound
// (~15 scalar instructions per element: the bf16 pairs are unpacked three times, the statistics and the affine
// transform run one lane-element at a time). This kernel
//   * unpacks a row's bf16 pairs ONCE into packed fp32x2 registers,
//   * runs every pass on the packed FADD2 / FFMA2 pipe: sum, (x - mean)^2, and y = x * (rstd * gamma) + (beta -
//     mean * rstd * gamma) -- 4.5 instructions per element,
//   * keeps gamma / beta as fp32 in shared memory (conflict-free float4 arrays), filled once per CTA,
//   * is persistent (3 CTAs per SM) with the NEXT row's 16-byte loads issued before the current row's reductions.
// MAXV 16-byte vectors per lane; LASTV lanes hold a valid last vector (32 = the row fills every lane: C = 256 * MAXV).
// Both are compile-time so that the row loop carries no per-vector bounds tests: with run-time bounds every unrolled
// vector became its own divergence region (BSSY / BSYNC, BRA.DIV around the shuffles) and a row cost 570 warp
// instructions, half of them bookkeeping (ncu: issue slots 50 % busy, DRAM 24 %).
template <int MAXV, int LASTV>
__global__ void __launch_bounds__(256, 3)
layernorm_stream_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ gamma,
                        const __nv_bfloat16* __restrict__ beta, __nv_bfloat16* __restrict__ y, int rows, float eps) {
    constexpr int NVEC = (MAXV - 1) * 32 + LASTV;
    constexpr int C = NVEC * 8;
    extern __shared__ float4 gb_smem[];                 // [4][NVEC]: gamma lo / hi halves, beta lo / hi halves
    const int lane = threadIdx.x & 31;
    float4* g_lo = gb_smem;
    float4* g_hi = g_lo + NVEC;
    float4* b_lo = g_hi + NVEC;
    float4* b_hi = b_lo + NVEC;
    for (int v = threadIdx.x; v < NVEC; v += blockDim.x) {
        const Vec8 g = reinterpret_cast<const Vec8*>(gamma)[v];
        const Vec8 b = reinterpret_cast<const Vec8*>(beta)[v];
        const float2 g0 = __bfloat1622float2(g.v[0]), g1 = __bfloat1622float2(g.v[1]);
        const float2 g2 = __bfloat1622float2(g.v[2]), g3 = __bfloat1622float2(g.v[3]);
        const float2 b0 = __bfloat1622float2(b.v[0]), b1 = __bfloat1622float2(b.v[1]);
        const float2 b2 = __bfloat1622float2(b.v[2]), b3 = __bfloat1622float2(b.v[3]);
        g_lo[v] = make_float4(g0.x, g0.y, g1.x, g1.y);
        g_hi[v] = make_float4(g2.x, g2.y, g3.x, g3.y);
        b_lo[v] = make_float4(b0.x, b0.y, b1.x, b1.y);
        b_hi[v] = make_float4(b2.x, b2.y, b3.x, b3.y);
    }
    __syncthreads();
    const int stride = gridDim.x * (blockDim.x >> 5);
    int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const bool last_ok = LASTV == 32 || lane < LASTV;      // this lane holds a vector in the last position
    const Vec8* xv = reinterpret_cast<const Vec8*>(x) + lane;
    Vec8* yv = reinterpret_cast<Vec8*>(y) + lane;
    Vec8 nxt[MAXV];
    const Vec8 zero = {};
    if (row < rows) {
        const Vec8* p = xv + (size_t)row * NVEC;
#pragma unroll
        for (int i = 0; i < MAXV; ++i) nxt[i] = (i < MAXV - 1 || last_ok) ? ld_stream(p + i * 32) : zero;
    }
    constexpr float inv_c = 1.0f / (float)C;
    for (; row < rows; row += stride) {
        // bf16 pair -> packed fp32 pair: the low half shifted up, the high half masked
        uint64_t a[MAXV][4];
#pragma unroll
        for (int i = 0; i < MAXV; ++i) {
            const uint32_t* w = reinterpret_cast<const uint32_t*>(&nxt[i]);
#pragma unroll
            for (int j = 0; j < 4; ++j) a[i][j] = pack_u32x2(w[j] << 16, w[j] & 0xffff0000u);
        }
        if (row + stride < rows) {
            const Vec8* p = xv + (size_t)(row + stride) * NVEC;
#pragma unroll
            for (int i = 0; i < MAXV; ++i) nxt[i] = (i < MAXV - 1 || last_ok) ? ld_stream(p + i * 32) : zero;
        }
        uint64_t s2 = 0ull;
#pragma unroll
        for (int i = 0; i < MAXV; ++i) {
#pragma unroll
            for (int j = 0; j < 4; ++j) s2 = fadd2(s2, a[i][j]);        // (absent vectors are zeros)
        }
        float s_lo, s_hi;
        unpack_f32x2(s2, s_lo, s_hi);
        const float mean = warp_sum(s_lo + s_hi) * inv_c;
        const uint64_t negmean2 = pack_f32x2(-mean, -mean);
        uint64_t q2 = 0ull;
#pragma unroll
        for (int i = 0; i < MAXV; ++i) {
            uint64_t qi = 0ull;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint64_t d2 = fadd2(a[i][j], negmean2);
                qi = ffma2(d2, d2, qi);
            }
            if (i < MAXV - 1 || last_ok) q2 = fadd2(q2, qi);           // (a select, not a branch)
        }
        float q_lo, q_hi;
        unpack_f32x2(q2, q_lo, q_hi);
        const float rstd = rsqrtf(warp_sum(q_lo + q_hi) * inv_c + eps);
        const uint64_t rstd2 = pack_f32x2(rstd, rstd);
        Vec8* yr = yv + (size_t)row * NVEC;
#pragma unroll
        for (int i = 0; i < MAXV; ++i) {
            const int v = lane + i * 32;
            if (i < MAXV - 1 || last_ok) {
                const float4 gl = g_lo[v], gh = g_hi[v], bl = b_lo[v], bh = b_hi[v];
                const uint64_t g2[4] = {pack_f32x2(gl.x, gl.y), pack_f32x2(gl.z, gl.w), pack_f32x2(gh.x, gh.y),
                                        pack_f32x2(gh.z, gh.w)};
                const uint64_t b2[4] = {pack_f32x2(bl.x, bl.y), pack_f32x2(bl.z, bl.w), pack_f32x2(bh.x, bh.y),
                                        pack_f32x2(bh.z, bh.w)};
                Vec8 ov;
                uint32_t* ow = reinterpret_cast<uint32_t*>(&ov);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const uint64_t gs = ffma2(g2[j], rstd2, 0ull);             // rstd * gamma
                    const uint64_t bs = ffma2(negmean2, gs, b2[j]);            // beta - mean * rstd * gamma
                    const uint64_t y2 = ffma2(a[i][j], gs, bs);
                    float lo, hi;
                    unpack_f32x2(y2, lo, hi);
                    ow[j] = pack_bf16x2(lo, hi);
                }
                yr[i * 32] = ov;
            }
        }
    }
}

template <bool ADD>
static int launch_ln(const void* x, const void* res, const void* gamma, const void* beta,
                     void* x_out, void* y, long long rows, int C, float eps, cudaStream_t stream) {
    if (rows <= 0) return kOk;
    if (C <= 0 || (C & 7) != 0) {
        set_last_error("layernorm: C=%d must be a positive multiple of 8", C);
        return kUnsupportedShape;
    }
    if (C > 2560) {
        set_last_error("layernorm: C=%d exceeds the supported maximum of 2560", C);
        return kUnsupportedShape;
    }
    const int warps = 8;
    const unsigned grid = (unsigned)((rows + warps - 1) / warps);
    auto xp = static_cast<const __nv_bfloat16*>(x);
    auto rp = static_cast<const __nv_bfloat16*>(res);
    auto gp = static_cast<const __nv_bfloat16*>(gamma);
    auto bp = static_cast<const __nv_bfloat16*>(beta);
    auto xo = static_cast<__nv_bfloat16*>(x_out);
    auto yp = static_cast<__nv_bfloat16*>(y);
    if (!ADD && rows < 0x7fffffffLL / 160) {
        // packed-math persistent variant (see layernorm_stream_kernel) for the UNets' channel counts
        const long long resident = (long long)sm_count() * 3;
        const unsigned pgrid = (unsigned)(grid < resident ? grid : resident);
        const size_t smem = (size_t)(C >> 3) * 4 * sizeof(float4);
        const int r = (int)rows;
        bool done = true;
        switch (C) {
            case 1280: layernorm_stream_kernel<5, 32><<<pgrid, warps * 32, smem, stream>>>(xp, gp, bp, yp, r, eps); break;
            case 1024: layernorm_stream_kernel<4, 32><<<pgrid, warps * 32, smem, stream>>>(xp, gp, bp, yp, r, eps); break;
            case 768: layernorm_stream_kernel<3, 32><<<pgrid, warps * 32, smem, stream>>>(xp, gp, bp, yp, r, eps); break;
            case 640: layernorm_stream_kernel<3, 16><<<pgrid, warps * 32, smem, stream>>>(xp, gp, bp, yp, r, eps); break;
            case 512: layernorm_stream_kernel<2, 32><<<pgrid, warps * 32, smem, stream>>>(xp, gp, bp, yp, r, eps); break;
            case 320: layernorm_stream_kernel<2, 8><<<pgrid, warps * 32, smem, stream>>>(xp, gp, bp, yp, r, eps); break;
            case 256: layernorm_stream_kernel<1, 32><<<pgrid, warps * 32, smem, stream>>>(xp, gp, bp, yp, r, eps); break;
            default: done = false;
        }
        if (done) {
            CGS_CUDA_CHECK(cudaGetLastError());
            return kOk;
        }
    }
    if (C <= 512)
        layernorm_kernel<2, ADD><<<grid, warps * 32, 0, stream>>>(xp, rp, gp, bp, xo, yp, rows, C, eps);
    else if (C <= 768)
        layernorm_kernel<3, ADD><<<grid, warps * 32, 0, stream>>>(xp, rp, gp, bp, xo, yp, rows, C, eps);
    else if (C <= 1280)
        layernorm_kernel<5, ADD><<<grid, warps * 32, 0, stream>>>(xp, rp, gp, bp, xo, yp, rows, C, eps);
    else
        layernorm_kernel<10, ADD><<<grid, warps * 32, 0, stream>>>(xp, rp, gp, bp, xo, yp, rows, C, eps);
    CGS_CUDA_CHECK(cudaGetLastError());
    return kOk;
}

}  // namespace cgs

extern "C" int cgs_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y,
                                 long long rows, int C, float eps, void* stream) {
    return cgs::launch_ln<false>(x, nullptr, gamma, beta, nullptr, y, rows, C, eps,
                                 static_cast<cudaStream_t>(stream));
}

extern "C" int cgs_add_layernorm_fwd(const void* x, const void* residual, const void* gamma,
                                     const void* beta, void* x_out, void* y, long long rows, int C,
                                     float eps, void* stream) {
    return cgs::launch_ln<true>(x, residual, gamma, beta, x_out, y, rows, C, eps,
                                static_cast<cudaStream_t>(stream));
}
