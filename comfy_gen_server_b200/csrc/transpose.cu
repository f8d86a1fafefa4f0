// Layout change at the edges of a SpatialTransformer: NCHW <-> token-major, optionally fused with
// the residual add that follows it.
//     out[n, c, r] = in[n, r, c] (+ residual[n, c, r])          in [N, R, C], out / residual [N, C, R], bf16
// Replaces `rearrange(x, 'b c h w -> b (h w) c').contiguous()` (in = NCHW viewed as [N, C, HW], no residual) and
// `rearrange(x, 'b (h w) c -> b c h w').contiguous()` + `x + x_in`
// (reference comfy/ldm/modules/attention.py:628, 636, 639). ATen runs the permuted copy as a non-vectorised
// strided kernel (0.13 ms for 42 MB, 0.66 TB/s); this is a 64 x 64 shared-memory tile transpose with 16-byte global
// accesses on both sides. HBM-bound: algorithmic bytes = 2 * N*R*C * 2 (+ residual read).
#include "common.cuh"

namespace cgs {

constexpr int TT = 64;

__global__ void __launch_bounds__(256)
transpose_add_kernel(const __nv_bfloat16* __restrict__ in, const __nv_bfloat16* __restrict__ residual,
                     __nv_bfloat16* __restrict__ out, int R, int C) {
    __shared__ __align__(16) unsigned short tile[TT][TT + 2];   // +2 halves: row stride 33 words, conflict-free columns
    const int n = blockIdx.z;
    const int r0 = blockIdx.y * TT, c0 = blockIdx.x * TT;
    const unsigned short* src = reinterpret_cast<const unsigned short*>(in) + (size_t)n * R * C;
    const int chunk = (threadIdx.x & 7) * 8, lrow = threadIdx.x >> 3;   // 8 x 16-byte chunks per 64-wide row, 32 rows per pass
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int r = r0 + lrow + 32 * i, c = c0 + chunk;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (r < R && c < C) v = *reinterpret_cast<const uint4*>(src + (size_t)r * C + c);
        unsigned short* t = &tile[lrow + 32 * i][chunk];
        *reinterpret_cast<uint32_t*>(t + 0) = v.x;
        *reinterpret_cast<uint32_t*>(t + 2) = v.y;
        *reinterpret_cast<uint32_t*>(t + 4) = v.z;
        *reinterpret_cast<uint32_t*>(t + 6) = v.w;
    }
    __syncthreads();
    unsigned short* dst = reinterpret_cast<unsigned short*>(out) + (size_t)n * R * C;
    const unsigned short* res = reinterpret_cast<const unsigned short*>(residual);
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int cc = lrow + 32 * i;           // output row inside the tile = input column
        const int c = c0 + cc, r = r0 + chunk;  // 8 consecutive output columns = input rows r .. r+7
        if (c < C && r < R) {
            unsigned short e[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) e[k] = tile[chunk + k][cc];
            uint4 v;
            v.x = e[0] | (uint32_t(e[1]) << 16);
            v.y = e[2] | (uint32_t(e[3]) << 16);
            v.z = e[4] | (uint32_t(e[5]) << 16);
            v.w = e[6] | (uint32_t(e[7]) << 16);
            const size_t o = (size_t)n * R * C + (size_t)c * R + r;
            if (res != nullptr) {
                const uint4 rv = *reinterpret_cast<const uint4*>(res + o);
                __nv_bfloat162* a = reinterpret_cast<__nv_bfloat162*>(&v);
                const __nv_bfloat162* b = reinterpret_cast<const __nv_bfloat162*>(&rv);
#pragma unroll
                for (int k = 0; k < 4; ++k) a[k] = __hadd2_rn(a[k], b[k]);   // bf16 add, as the reference's x + x_in
            }
            *reinterpret_cast<uint4*>(dst + (size_t)c * R + r) = v;
        }
    }
}

// out = a + b (+ bias[channel]) for two bf16 tensors with identical dense layouts (the ResBlock's
// `skip_connection(x) + h`, openaimodel.py:264, and the SpatialTransformer's `x + x_in` in the channels-last
// pipeline). ATen sends these adds down its strided, non-vectorised path when the operands are channels_last views
// (2.7 TB/s measured); this is the plain 16-byte-per-thread streaming form. The optional per-channel bias is the
// bias of the convolution that produced b (and of a 1x1 skip convolution): cuDNN's fprop has no bias epilogue
// through ATen, which adds it in a separate full pass; here it rides along. channel = (i / inner) % channels with
// inner = 1 for channels_last (8 consecutive channels per thread) and H*W for NCHW (one channel per thread).
template <bool BIAS>
__global__ void __launch_bounds__(256)
add_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, const __nv_bfloat16* __restrict__ bias,
           uint4* __restrict__ out, long long n8, int channels, int inner) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n8) return;
    const uint4 x = a[i];
    const uint4 y = b[i];
    const __nv_bfloat162* xv = reinterpret_cast<const __nv_bfloat162*>(&x);
    const __nv_bfloat162* yv = reinterpret_cast<const __nv_bfloat162*>(&y);
    uint4 r;
    __nv_bfloat162* rv = reinterpret_cast<__nv_bfloat162*>(&r);
    if constexpr (BIAS) {
        float bf[8];
        if (inner == 1) {
            const uint4 bb = *reinterpret_cast<const uint4*>(bias + (i * 8) % channels);
            const __nv_bfloat162* bv = reinterpret_cast<const __nv_bfloat162*>(&bb);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float2 f = __bfloat1622float2(bv[k]);
                bf[2 * k] = f.x;
                bf[2 * k + 1] = f.y;
            }
        } else {
            const float v = __bfloat162float(bias[((i * 8) / inner) % channels]);
#pragma unroll
            for (int k = 0; k < 8; ++k) bf[k] = v;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float2 fx = __bfloat1622float2(xv[k]), fy = __bfloat1622float2(yv[k]);
            rv[k] = __floats2bfloat162_rn(fx.x + fy.x + bf[2 * k], fx.y + fy.y + bf[2 * k + 1]);
        }
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) rv[k] = __hadd2_rn(xv[k], yv[k]);
    }
    out[i] = r;
}

// Channel concatenation of two channels-last images (the UNet's skip connections, `th.cat([h, hs.pop()], dim=1)`,
// openaimodel.py:869): out[p, 0:Ca] = a[p, :], out[p, Ca:Ca+Cb] = b[p, :] for every pixel p, 16 bytes per thread.
// ATen's CatArrayBatchedCopy moved these at 1.3 TB/s. bias_a (optional, [Ca]) is added to the `a` part: the bias of
// the Upsample convolution that produced `a`, which ATen would add in a separate full pass over the 2x-larger image.
__global__ void __launch_bounds__(256)
cat_channels_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, const __nv_bfloat16* __restrict__ bias_a,
                    uint4* __restrict__ out, long long pixels, int ca8, int cb8) {
    const int c8 = ca8 + cb8;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= pixels * c8) return;
    const long long pix = i / c8;
    const int c = (int)(i - pix * c8);
    if (c >= ca8) {
        out[i] = b[pix * cb8 + (c - ca8)];
        return;
    }
    uint4 v = a[pix * ca8 + c];
    if (bias_a != nullptr) {
        const uint4 bb = *reinterpret_cast<const uint4*>(bias_a + c * 8);
        __nv_bfloat162* vv = reinterpret_cast<__nv_bfloat162*>(&v);
        const __nv_bfloat162* bv = reinterpret_cast<const __nv_bfloat162*>(&bb);
#pragma unroll
        for (int k = 0; k < 4; ++k) vv[k] = __hadd2_rn(vv[k], bv[k]);
    }
    out[i] = v;
}

// Nearest-neighbour 2x up-sampling of a channels-last image (Upsample.forward, openaimodel.py:86-95, for the exact
// doubling every SD / SDXL level uses): out[n, 2h+dy, 2w+dx, :] = x[n, h, w, :]. One thread = 16 bytes of one input
// pixel, written to its four output pixels. ATen's upsample_nearest2d_nhwc ran at 0.44 TB/s.
__global__ void __launch_bounds__(256)
upsample2x_nhwc_kernel(const uint4* __restrict__ x, uint4* __restrict__ out, long long n_in8, int H, int W, int c8) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_in8) return;
    const int c = (int)(i % c8);
    const long long pix = i / c8;
    const int w = (int)(pix % W);
    const long long r = pix / W;
    const int h = (int)(r % H);
    const long long n = r / H;
    const uint4 v = x[i];
    const long long row0 = ((n * (2 * H) + 2 * h) * (2LL * W) + 2 * w) * c8 + c;
    const long long row1 = row0 + 2LL * W * c8;
    out[row0] = v;
    out[row0 + c8] = v;
    out[row1] = v;
    out[row1 + c8] = v;
}

}  // namespace cgs

extern "C" int cgs_cat_channels_nhwc_fwd(const void* a, const void* b, const void* bias_a, void* out, long long pixels,
                                         int ca, int cb, void* stream) {
    using namespace cgs;
    if (pixels < 0 || ca <= 0 || cb <= 0) {
        set_last_error("cat_channels: bad sizes pixels=%lld ca=%d cb=%d", pixels, ca, cb);
        return kBadArgument;
    }
    if (pixels == 0) return kOk;
    if (ca % 8 != 0 || cb % 8 != 0 || ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) |
                                         reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(bias_a)) & 0xF)) {
        set_last_error("cat_channels: channel counts must be multiples of 8 and pointers 16-byte aligned");
        return kUnsupportedShape;
    }
    const long long n8 = pixels * ((ca + cb) / 8);
    const long long blocks = (n8 + 255) / 256;
    if (blocks > 0x7fffffffLL) {
        set_last_error("cat_channels: tensor too large");
        return kUnsupportedShape;
    }
    cat_channels_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const uint4*>(a), static_cast<const uint4*>(b), static_cast<const __nv_bfloat16*>(bias_a),
        static_cast<uint4*>(out), pixels, ca / 8, cb / 8);
    CGS_CUDA_CHECK(cudaGetLastError());
    return kOk;
}

extern "C" int cgs_upsample_nearest2x_nhwc_fwd(const void* x, void* out, int N, int H, int W, int C, void* stream) {
    using namespace cgs;
    if (N < 0 || H <= 0 || W <= 0 || C <= 0) {
        set_last_error("upsample2x: bad sizes N=%d H=%d W=%d C=%d", N, H, W, C);
        return kBadArgument;
    }
    if (N == 0) return kOk;
    if (C % 8 != 0 || ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 0xF)) {
        set_last_error("upsample2x: C must be a multiple of 8 and pointers 16-byte aligned");
        return kUnsupportedShape;
    }
    const long long n8 = (long long)N * H * W * (C / 8);
    const long long blocks = (n8 + 255) / 256;
    if (blocks > 0x7fffffffLL) {
        set_last_error("upsample2x: tensor too large");
        return kUnsupportedShape;
    }
    upsample2x_nhwc_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const uint4*>(x), static_cast<uint4*>(out), n8, H, W, C / 8);
    CGS_CUDA_CHECK(cudaGetLastError());
    return kOk;
}

extern "C" int cgs_add_fwd(const void* a, const void* b, const void* bias, void* out, long long n, int channels,
                           long long inner, void* stream) {
    using namespace cgs;
    if (n < 0) {
        set_last_error("add: negative size");
        return kBadArgument;
    }
    if (n == 0) return kOk;
    if (n % 8 != 0 || ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) |
                        reinterpret_cast<uintptr_t>(out)) & 0xF)) {
        set_last_error("add: element count must be a multiple of 8 and pointers 16-byte aligned");
        return kUnsupportedShape;
    }
    if (bias != nullptr) {
        const bool nhwc_ok = inner == 1 && channels > 0 && channels % 8 == 0 &&
                             (reinterpret_cast<uintptr_t>(bias) & 0xF) == 0;
        const bool nchw_ok = inner > 1 && inner % 8 == 0 && inner <= 0x7fffffffLL && channels > 0;
        if (!nhwc_ok && !nchw_ok) {
            set_last_error("add: bias needs channels %% 8 == 0 (channels_last, inner 1) or H*W %% 8 == 0 (NCHW); "
                           "got channels=%d inner=%lld", channels, inner);
            return kUnsupportedShape;
        }
    }
    const long long n8 = n / 8;
    const long long blocks = (n8 + 255) / 256;
    if (blocks > 0x7fffffffLL) {
        set_last_error("add: tensor too large");
        return kUnsupportedShape;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (bias != nullptr)
        add_kernel<true><<<(unsigned)blocks, 256, 0, st>>>(
            static_cast<const uint4*>(a), static_cast<const uint4*>(b), static_cast<const __nv_bfloat16*>(bias),
            static_cast<uint4*>(out), n8, channels, (int)inner);
    else
        add_kernel<false><<<(unsigned)blocks, 256, 0, st>>>(
            static_cast<const uint4*>(a), static_cast<const uint4*>(b), nullptr, static_cast<uint4*>(out), n8, 0, 1);
    CGS_CUDA_CHECK(cudaGetLastError());
    return kOk;
}

extern "C" int cgs_transpose_add_fwd(const void* in, const void* residual, void* out, int N, int R, int C,
                                     void* stream) {
    using namespace cgs;
    if (N < 0 || R <= 0 || C <= 0) {
        set_last_error("transpose_add: bad sizes N=%d R=%d C=%d", N, R, C);
        return kBadArgument;
    }
    if (N == 0) return kOk;
    if (R % 8 != 0 || C % 8 != 0) {
        set_last_error("transpose_add: R=%d and C=%d must be multiples of 8", R, C);
        return kUnsupportedShape;
    }
    if (N > 65535 || (R + TT - 1) / TT > 65535) {
        set_last_error("transpose_add: N=%d / R=%d exceed the grid limits", N, R);
        return kUnsupportedShape;
    }
    dim3 grid((C + TT - 1) / TT, (R + TT - 1) / TT, N);
    transpose_add_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(in), static_cast<const __nv_bfloat16*>(residual),
        static_cast<__nv_bfloat16*>(out), R, C);
    CGS_CUDA_CHECK(cudaGetLastError());
    return kOk;
}
