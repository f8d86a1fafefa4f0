// N1 (first "next" row): GroupNorm(32) with the SiLU that always follows it in a ResBlock fused in,
// NCHW bf16 in / bf16 out, statistics in fp32. Replaces ops.GroupNorm -> F.group_norm followed by nn.SiLU
// (reference comfy/ldm/modules/diffusionmodules/openaimodel.py:177-181, 213-217, 805-807; three ATen kernels:
// RowwiseMoments, the affine apply, and SiLU = 5 passes over the tensor) and SpatialTransformer.norm
// (comfy/ldm/modules/attention.py:595, eps 1e-6, no activation).
//
// One CTA per (sample, group): the group's channels are contiguous in NCHW (cpg * H*W elements, 80-330 KB), so
// pass 1 (sum, sum of squares) streams it from HBM and pass 2 (normalise, affine, SiLU, store) re-reads it from
// L2: one HBM read + one HBM write. Algorithmic bytes = 2 * N*C*H*W * 2.
#include "common.cuh"

namespace cgs {

struct alignas(16) GN8 {
    __nv_bfloat162 v[4];
};

template <bool SILU>
__global__ void __launch_bounds__(512)
groupnorm_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ add_nc,
                 const __nv_bfloat16* __restrict__ gamma, const __nv_bfloat16* __restrict__ beta,
                 __nv_bfloat16* __restrict__ y, int C, int HW, int groups, float eps) {
    const int g = blockIdx.x, n = blockIdx.y;
    const int cpg = C / groups;
    const long long len = (long long)cpg * HW;
    const long long base = ((long long)n * C + (long long)g * cpg) * HW;
    const GN8* xv = reinterpret_cast<const GN8*>(x + base);
    const int nvec = (int)(len >> 3);

    const int vec_per_ch = HW >> 3;
    // optional per-(sample, channel) addend: the ResBlock's `h + emb_out[..., None, None]` (openaimodel.py:259),
    // added in bf16 like the reference, without a separate pass over h
    const __nv_bfloat16* add_n = add_nc != nullptr ? add_nc + (size_t)n * C + g * cpg : nullptr;
    float s = 0.f, ss = 0.f;
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
        GN8 a = xv[i];
        if (add_n != nullptr) {
            const __nv_bfloat162 e2 = __bfloat162bfloat162(add_n[i / vec_per_ch]);
#pragma unroll
            for (int j = 0; j < 4; ++j) a.v[j] = __hadd2_rn(a.v[j], e2);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 f = __bfloat1622float2(a.v[j]);
            s += f.x + f.y;
            ss += f.x * f.x + f.y * f.y;
        }
    }
    __shared__ float red[2][16];
    __shared__ float stats[2];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, o);
        ss += __shfl_xor_sync(0xffffffffu, ss, o);
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { red[0][warp] = s; red[1][warp] = ss; }
    __syncthreads();
    if (warp == 0) {
        s = lane < (blockDim.x >> 5) ? red[0][lane] : 0.f;
        ss = lane < (blockDim.x >> 5) ? red[1][lane] : 0.f;
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) {
            s += __shfl_xor_sync(0xffffffffu, s, o);
            ss += __shfl_xor_sync(0xffffffffu, ss, o);
        }
        if (lane == 0) {
            const float mean = s / (float)len;
            const float var = fmaxf(ss / (float)len - mean * mean, 0.f);
            stats[0] = mean;
            stats[1] = rsqrtf(var + eps);
        }
    }
    __syncthreads();
    const float mean = stats[0], rstd = stats[1];

    GN8* yv = reinterpret_cast<GN8*>(y + base);
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
        const int c = g * cpg + i / vec_per_ch;
        const float ga = __bfloat162float(gamma[c]) * rstd;
        const float be = __bfloat162float(beta[c]) - mean * ga;
        GN8 a = xv[i];          // second read: L2-resident
        if (add_n != nullptr) {
            const __nv_bfloat162 e2 = __bfloat162bfloat162(add_n[i / vec_per_ch]);
#pragma unroll
            for (int j = 0; j < 4; ++j) a.v[j] = __hadd2_rn(a.v[j], e2);
        }
        GN8 o;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 f = __bfloat1622float2(a.v[j]);
            float v0 = fmaf(f.x, ga, be), v1 = fmaf(f.y, ga, be);
            if (SILU) {
                v0 = v0 / (1.f + __expf(-v0));
                v1 = v1 / (1.f + __expf(-v1));
            }
            o.v[j] = __floats2bfloat162_rn(v0, v1);
        }
        yv[i] = o;
    }
}

}  // namespace cgs

extern "C" int cgs_groupnorm_fwd(const void* x, const void* add_nc, const void* gamma, const void* beta, void* y,
                                 int N, int C, int HW, int groups, float eps, int silu, void* stream) {
    using namespace cgs;
    if (N < 0 || C <= 0 || HW <= 0 || groups <= 0 || C % groups != 0) {
        set_last_error("groupnorm: bad sizes N=%d C=%d HW=%d groups=%d", N, C, HW, groups);
        return kBadArgument;
    }
    if (N == 0) return kOk;
    if (HW % 8 != 0) {
        set_last_error("groupnorm: H*W=%d must be a multiple of 8", HW);
        return kUnsupportedShape;
    }
    if (N > 65535) {
        set_last_error("groupnorm: N=%d exceeds the grid limit", N);
        return kUnsupportedShape;
    }
    dim3 grid(groups, N);
    auto xp = static_cast<const __nv_bfloat16*>(x);
    auto ap = static_cast<const __nv_bfloat16*>(add_nc);
    auto gp = static_cast<const __nv_bfloat16*>(gamma);
    auto bp = static_cast<const __nv_bfloat16*>(beta);
    auto yp = static_cast<__nv_bfloat16*>(y);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (silu)
        groupnorm_kernel<true><<<grid, 512, 0, st>>>(xp, ap, gp, bp, yp, C, HW, groups, eps);
    else
        groupnorm_kernel<false><<<grid, 512, 0, st>>>(xp, ap, gp, bp, yp, C, HW, groups, eps);
    CGS_CUDA_CHECK(cudaGetLastError());
    return kOk;
}
