// Row softmax in place, bf16: x[r, :] <- softmax(scale * x[r, :]).
//
// Used by the VAE decoder's mid-block attention (one head, d = 512, h*w tokens), which runs the reference's
// `normal_attention` / `slice_attention` algorithm (comfy/ldm/modules/diffusionmodules/model.py:160-209:
// s = bmm(q, k) * scale; s = softmax(s); out = bmm(v, s^T), all in the compute dtype) on this library's GEMM: the
// [tokens, tokens] score matrix is written by the first GEMM, normalised here and read by the second. One CTA per row,
// the row stays in registers between the passes (max, exp + sum, store), fp32 arithmetic, exp2 with the scale folded in.
// HBM-bound: 2 * rows * cols * 2 bytes.
#include "common.cuh"

namespace cgs {

constexpr int kSoftmaxThreads = 512;
constexpr int kSoftmaxMaxVec = 8;      // 16-byte vectors per thread: cols <= 512 * 8 * 8 = 32768

struct alignas(16) SM8 {
    __nv_bfloat162 v[4];
};

__device__ __forceinline__ float block_reduce(float v, float* red, bool is_max) {
    for (int o = 16; o > 0; o >>= 1) {
        const float t = __shfl_xor_sync(0xffffffffu, v, o);
        v = is_max ? fmaxf(v, t) : v + t;
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();                       // `red` may still be read from the previous reduction
    if (lane == 0) red[warp] = v;
    __syncthreads();
    v = red[0];
    for (int w = 1; w < kSoftmaxThreads / 32; ++w) v = is_max ? fmaxf(v, red[w]) : v + red[w];   // fixed order
    return v;
}

// VEC = 16-byte vectors per thread (compile time: the row lives in registers; up to 4 vectors two CTAs fit on an SM)
template <int VEC>
__global__ void __launch_bounds__(kSoftmaxThreads, VEC <= 4 ? 2 : 1)
softmax_rows_kernel(__nv_bfloat16* __restrict__ x, int cols, long long ld, float scale_log2) {
    __shared__ float red[kSoftmaxThreads / 32];
    __nv_bfloat16* row = x + (size_t)blockIdx.x * ld;
    const int nvec = cols / 8;
    SM8 raw[VEC];
    float mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < VEC; ++i) {
        const int v = threadIdx.x + i * kSoftmaxThreads;
        if (v < nvec) {
            raw[i] = *reinterpret_cast<const SM8*>(row + (size_t)v * 8);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 t = __bfloat1622float2(raw[i].v[j]);
                mx = fmaxf(mx, fmaxf(t.x, t.y));
            }
        }
    }
    mx = block_reduce(mx, red, true);
    // scale > 0: max(scale * x) = scale * max(x). A row of -inf only (cannot come out of a GEMM) would give NaN, as torch does.
    const float off = mx * scale_log2;
    float e[VEC][8];            // the exponentials stay in registers between the sum and the store
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < VEC; ++i) {
        const int v = threadIdx.x + i * kSoftmaxThreads;
        if (v < nvec) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 t = __bfloat1622float2(raw[i].v[j]);
                e[i][2 * j] = ex2_approx(fmaf(t.x, scale_log2, -off));
                e[i][2 * j + 1] = ex2_approx(fmaf(t.y, scale_log2, -off));
                sum += e[i][2 * j] + e[i][2 * j + 1];
            }
        }
    }
    sum = block_reduce(sum, red, false);
    const float inv = 1.0f / sum;
#pragma unroll
    for (int i = 0; i < VEC; ++i) {
        const int v = threadIdx.x + i * kSoftmaxThreads;
        if (v < nvec) {
            SM8 o;
#pragma unroll
            for (int j = 0; j < 4; ++j) o.v[j] = __floats2bfloat162_rn(e[i][2 * j] * inv, e[i][2 * j + 1] * inv);
            *reinterpret_cast<SM8*>(row + (size_t)v * 8) = o;
        }
    }
}

}  // namespace cgs

extern "C" int cgs_softmax_rows_fwd(void* x, long long rows, int cols, long long ld, float scale, void* stream) {
    using namespace cgs;
    if (rows < 0 || cols <= 0 || ld < cols || !(scale > 0.f)) {
        set_last_error("softmax_rows: bad arguments rows=%lld cols=%d ld=%lld scale=%f", rows, cols, ld, (double)scale);
        return kBadArgument;
    }
    if (rows == 0) return kOk;
    if (cols % 8 != 0 || ld % 8 != 0 || cols > kSoftmaxThreads * kSoftmaxMaxVec * 8 || rows > 0x7fffffffLL ||
        (reinterpret_cast<uintptr_t>(x) & 0xF)) {
        set_last_error("softmax_rows: cols=%d must be a multiple of 8 and at most %d, rows 16-byte aligned", cols,
                       kSoftmaxThreads * kSoftmaxMaxVec * 8);
        return kUnsupportedShape;
    }
    const int vec = (cols / 8 + kSoftmaxThreads - 1) / kSoftmaxThreads;
    auto xp = static_cast<__nv_bfloat16*>(x);
    const float sl = scale * 1.4426950408889634f;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (vec <= 2) softmax_rows_kernel<2><<<(unsigned)rows, kSoftmaxThreads, 0, st>>>(xp, cols, ld, sl);
    else if (vec <= 4) softmax_rows_kernel<4><<<(unsigned)rows, kSoftmaxThreads, 0, st>>>(xp, cols, ld, sl);
    else softmax_rows_kernel<8><<<(unsigned)rows, kSoftmaxThreads, 0, st>>>(xp, cols, ld, sl);
    CGS_CUDA_CHECK(cudaGetLastError());
    return kOk;
}
