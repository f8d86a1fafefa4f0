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

#include <atomic>

namespace cgs {

// x * sigmoid(x) = h + h * tanh(h), h = x / 2: one MUFU (tanh.approx, rel. error ~2^-11, below the bf16 rounding of the
// result) and two FMA-pipe instructions. The IEEE division of x / (1 + exp(-x)) made the normalise kernels issue-bound
// (~20 instructions per channel), and exp + reciprocal are two MUFUs per channel.
__device__ __forceinline__ float silu_fast(float x) {
    const float h = 0.5f * x;
    float t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
    return fmaf(h, t, h);
}


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
                v0 = silu_fast(v0);
                v1 = silu_fast(v1);
            }
            o.v[j] = __floats2bfloat162_rn(v0, v1);
        }
        yv[i] = o;
    }
}


// ---------------------------------------------------------------------------------------------------------------
// Channels-last (NHWC) form: x, y [N, H*W, C]. Keeping the UNet's activations NHWC end to end removes
// cuDNN's per-convolution NCHW<->NHWC conversion kernels and makes the SpatialTransformer's token layout free.
// A group's channels are now only cpg (10..80) contiguous elements per pixel, so the work is split by pixel rows,
// not by group:
//   kernel 1  every CTA reduces a chunk of pixel rows for ALL 32 groups -> partial (sum, sum of squares) per
//             (sample, chunk, group) in a small global scratch buffer;
//   kernel 2  every CTA first adds up its sample's partials (L2-resident, a few KB), then normalises its chunk.
// Thread <-> channel mapping is fixed (a thread always owns the same 8 channels), so gamma/beta/group ids live in
// registers; a vector of 8 channels straddles at most two groups (cpg >= 8, or cpg == 4: exactly two).
// ---------------------------------------------------------------------------------------------------------------
constexpr int kGnMaxGroups = 32;

struct GnNhwcGeom {
    int nvec;           // C / 8
    int rows_per_iter;  // pixel rows handled per block iteration
    int threads;        // nvec * rows_per_iter
    int chunks;         // CTAs per sample
    int rows_per_chunk;
};

static GnNhwcGeom gn_nhwc_geom(int N, int C, int HW) {
    GnNhwcGeom g;
    g.nvec = C / 8;
    g.rows_per_iter = g.nvec >= 256 ? 1 : 256 / g.nvec;
    g.threads = g.nvec * g.rows_per_iter;
    // enough CTAs to fill the GPU a few times over. Up to 128 partials per sample the normalise kernel adds them up
    // itself; beyond that (one or two large images: the VAE decoder) a small kernel in between does it once --
    // 128 CTAs for a 268 MB image ran the pair of kernels at a third of the HBM bandwidth.
    int want = (sm_count() * 4 + N - 1) / (N > 0 ? N : 1);
    const int cap = (long long)HW * C * 2 >= (64LL << 20) ? 1024 : 128;      // (per sample: 64 MB and more)
    if (want > cap) want = cap;
    if (want < 1) want = 1;
    int rows = (HW + want - 1) / want;
    rows = ((rows + g.rows_per_iter - 1) / g.rows_per_iter) * g.rows_per_iter;
    g.rows_per_chunk = rows;
    g.chunks = (HW + rows - 1) / rows;
    return g;
}

__device__ __forceinline__ void gn_load_row_vec(const __nv_bfloat16* xp, const __nv_bfloat162 (&e)[4], bool has_add,
                                                float (&f)[8]) {
    GN8 a = *reinterpret_cast<const GN8*>(xp);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        if (has_add) a.v[j] = __hadd2_rn(a.v[j], e[j]);      // bf16 add, as the reference's h + emb
        const float2 t = __bfloat1622float2(a.v[j]);
        f[2 * j] = t.x;
        f[2 * j + 1] = t.y;
    }
}

__global__ void __launch_bounds__(320)
gn_nhwc_stats_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ add_nc,
                     float* __restrict__ partial, int C, int HW, int groups, int nvec, int rows_per_iter,
                     int rows_per_chunk) {
    // per-thread partials, combined in a fixed order (no atomics: results are reproducible run to run)
    __shared__ float part[320][4];
    const int n = blockIdx.y, chunk = blockIdx.x;
    const int vec = threadIdx.x % nvec, rsub = threadIdx.x / nvec;
    const int cpg = C / groups;
    const int c0 = vec * 8;
    const int g0 = c0 / cpg;
    const int split = min(8, (g0 + 1) * cpg - c0);     // channels [0, split) of this vector belong to g0
    __nv_bfloat162 e[4];
    const bool has_add = add_nc != nullptr;
    if (has_add) {
        const GN8 ev = *reinterpret_cast<const GN8*>(add_nc + (size_t)n * C + c0);
#pragma unroll
        for (int j = 0; j < 4; ++j) e[j] = ev.v[j];
    }
    // one (sum, sum of squares) pair per channel of the vector: four shared accumulators were four dependent FADD
    // chains of length 8 per 16 bytes, which capped the kernel at half of HBM speed
    float sj[8], qj[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) sj[j] = qj[j] = 0.f;
    const int r_begin = chunk * rows_per_chunk;
    const int r_end = min(HW, r_begin + rows_per_chunk);
    const __nv_bfloat16* xn = x + (size_t)n * HW * C + c0;
    auto accumulate = [&](GN8 a) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (has_add) a.v[j] = __hadd2_rn(a.v[j], e[j]);      // bf16 add, as the reference's h + emb
            const float2 t = __bfloat1622float2(a.v[j]);
            sj[2 * j] += t.x;
            qj[2 * j] = fmaf(t.x, t.x, qj[2 * j]);
            sj[2 * j + 1] += t.y;
            qj[2 * j + 1] = fmaf(t.y, t.y, qj[2 * j + 1]);
        }
    };
    int r = r_begin + rsub;
    for (; r + 3 * rows_per_iter < r_end; r += 4 * rows_per_iter) {      // four rows per trip, loads first
        GN8 raw[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) raw[u] = *reinterpret_cast<const GN8*>(xn + (size_t)(r + u * rows_per_iter) * C);
#pragma unroll
        for (int u = 0; u < 4; ++u) accumulate(raw[u]);
    }
    for (; r < r_end; r += rows_per_iter) accumulate(*reinterpret_cast<const GN8*>(xn + (size_t)r * C));
    float sa = 0.f, qa = 0.f, sb = 0.f, qb = 0.f;       // channels [0, split) belong to group g0, the rest to g0 + 1
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        if (j < split) { sa += sj[j]; qa += qj[j]; }
        else           { sb += sj[j]; qb += qj[j]; }
    }
    part[threadIdx.x][0] = sa;
    part[threadIdx.x][1] = qa;
    part[threadIdx.x][2] = sb;
    part[threadIdx.x][3] = qb;
    __syncthreads();
    if (threadIdx.x < groups) {
        // group g owns channels [g*cpg, (g+1)*cpg): the vectors that touch it, each contributing its "a" half when
        // the vector starts in g and its "b" half when it started in g-1 and spilled over
        const int g = threadIdx.x;
        const int v_first = (g * cpg) / 8, v_last = ((g + 1) * cpg - 1) / 8;
        float s = 0.f, q = 0.f;
        for (int v = v_first; v <= v_last; ++v) {
            const int vg0 = (v * 8) / cpg;                 // group of the vector's first channel
            for (int rs = 0; rs < rows_per_iter; ++rs) {
                const float* pp = part[rs * nvec + v];
                if (vg0 == g) { s += pp[0]; q += pp[1]; }
                else          { s += pp[2]; q += pp[3]; }  // vg0 == g - 1
            }
        }
        float* dst = partial + (((size_t)n * gridDim.x + chunk) * groups + g) * 2;
        dst[0] = s;
        dst[1] = q;
    }
}

// (more than 128 chunks per sample) adds up a sample's partials once, in a fixed order: final[n][g] = (sum, sum sq)
__global__ void __launch_bounds__(256)
gn_nhwc_finalize_kernel(const float* __restrict__ partial, float* __restrict__ final_, int chunks, int groups) {
    __shared__ float red[256][2];
    const int n = blockIdx.x;
    const int g = threadIdx.x % groups, stripe = threadIdx.x / groups, stripes = blockDim.x / groups;
    float s = 0.f, q = 0.f;
    if (stripe < stripes) {
        const float* src = partial + ((size_t)n * chunks * groups + g) * 2;
        for (int k = stripe; k < chunks; k += stripes) {
            s += src[(size_t)k * groups * 2];
            q += src[(size_t)k * groups * 2 + 1];
        }
    }
    red[threadIdx.x][0] = s;
    red[threadIdx.x][1] = q;
    __syncthreads();
    if (threadIdx.x < groups) {
        s = q = 0.f;
        for (int k = 0; k < stripes; ++k) {
            s += red[k * groups + g][0];
            q += red[k * groups + g][1];
        }
        final_[((size_t)n * groups + g) * 2] = s;
        final_[((size_t)n * groups + g) * 2 + 1] = q;
    }
}

template <bool SILU>
__global__ void __launch_bounds__(320)
gn_nhwc_apply_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ add_nc,
                     const __nv_bfloat16* __restrict__ gamma, const __nv_bfloat16* __restrict__ beta,
                     const float* __restrict__ partial, int chunks, __nv_bfloat16* __restrict__ y, int C, int HW,
                     int groups, int nvec, int rows_per_iter, int rows_per_chunk, float eps) {
    __shared__ float stat[kGnMaxGroups][2];    // mean, rstd
    const int n = blockIdx.y, chunk = blockIdx.x;    // `chunks` partials per sample (1 after the finalize kernel)
    const int cpg = C / groups;
    if (threadIdx.x < groups) {
        float s = 0.f, q = 0.f;
        const float* src = partial + ((size_t)n * chunks * groups + threadIdx.x) * 2;
        for (int k = 0; k < chunks; ++k) {
            s += src[(size_t)k * groups * 2];
            q += src[(size_t)k * groups * 2 + 1];
        }
        const float cnt = (float)cpg * (float)HW;
        const float mean = s / cnt;
        const float var = fmaxf(q / cnt - mean * mean, 0.f);
        stat[threadIdx.x][0] = mean;
        stat[threadIdx.x][1] = rsqrtf(var + eps);
    }
    __syncthreads();
    const int vec = threadIdx.x % nvec, rsub = threadIdx.x / nvec;
    const int c0 = vec * 8;
    const int g0 = c0 / cpg;
    const int split = min(8, (g0 + 1) * cpg - c0);
    float ga[8], be[8];
    {
        const GN8 gv = *reinterpret_cast<const GN8*>(gamma + c0);
        const GN8 bv = *reinterpret_cast<const GN8*>(beta + c0);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int g = j < split ? g0 : g0 + 1;
            const float mean = stat[g][0], rstd = stat[g][1];
            const float gam = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(&gv)[j]);
            const float bet = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(&bv)[j]);
            ga[j] = gam * rstd;
            be[j] = bet - mean * ga[j];
        }
    }
    __nv_bfloat162 e[4];
    const bool has_add = add_nc != nullptr;
    if (has_add) {
        const GN8 ev = *reinterpret_cast<const GN8*>(add_nc + (size_t)n * C + c0);
#pragma unroll
        for (int j = 0; j < 4; ++j) e[j] = ev.v[j];
    }
    const int r_begin = chunk * rows_per_chunk;
    const int r_end = min(HW, r_begin + rows_per_chunk);
    auto finish = [&](GN8 a, size_t off) {
        GN8 o;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (has_add) a.v[j] = __hadd2_rn(a.v[j], e[j]);      // bf16 add, as the reference's h + emb
            const float2 t = __bfloat1622float2(a.v[j]);
            float v0 = fmaf(t.x, ga[2 * j], be[2 * j]);
            float v1 = fmaf(t.y, ga[2 * j + 1], be[2 * j + 1]);
            if (SILU) {
                v0 = silu_fast(v0);
                v1 = silu_fast(v1);
            }
            o.v[j] = __floats2bfloat162_rn(v0, v1);
        }
        *reinterpret_cast<GN8*>(y + off) = o;
    };
    // four rows per trip, loads first (memory-level parallelism: see the statistics kernel)
    const size_t base = (size_t)n * HW * C + c0;
    int r = r_begin + rsub;
    for (; r + 3 * rows_per_iter < r_end; r += 4 * rows_per_iter) {
        GN8 raw[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            raw[u] = *reinterpret_cast<const GN8*>(x + base + (size_t)(r + u * rows_per_iter) * C);
#pragma unroll
        for (int u = 0; u < 4; ++u) finish(raw[u], base + (size_t)(r + u * rows_per_iter) * C);
    }
    for (; r < r_end; r += rows_per_iter)
        finish(*reinterpret_cast<const GN8*>(x + base + (size_t)r * C), base + (size_t)r * C);
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

extern "C" long long cgs_groupnorm_nhwc_scratch_bytes(int N, int C, int HW, int groups) {
    if (N <= 0 || C <= 0 || HW <= 0 || groups <= 0) return 0;
    const cgs::GnNhwcGeom g = cgs::gn_nhwc_geom(N, C, HW);
    return (long long)N * (g.chunks + 1) * groups * 2 * (long long)sizeof(float);      // partials + their sums
}

extern "C" int cgs_groupnorm_nhwc_fwd(const void* x, const void* add_nc, const void* gamma, const void* beta, void* y,
                                      void* scratch, long long scratch_bytes, int N, int C, int HW, int groups,
                                      float eps, int silu, void* stream) {
    using namespace cgs;
    if (N < 0 || C <= 0 || HW <= 0 || groups <= 0 || C % groups != 0) {
        set_last_error("groupnorm_nhwc: bad sizes N=%d C=%d HW=%d groups=%d", N, C, HW, groups);
        return kBadArgument;
    }
    if (N == 0) return kOk;
    // a 16-byte vector of 8 channels may touch at most two groups: 8 or more channels per group, or exactly 4
    const int cpg_ = groups > 0 ? C / groups : 0;
    if (C % 8 != 0 || (cpg_ < 8 && cpg_ != 4) || groups > kGnMaxGroups || C / 8 > 320) {
        set_last_error("groupnorm_nhwc: C=%d groups=%d unsupported (C %% 8 == 0, C/groups >= 8 or == 4, groups <= 32, C <= 2560)",
                       C, groups);
        return kUnsupportedShape;
    }
    if (N > 65535) {
        set_last_error("groupnorm_nhwc: N=%d exceeds the grid limit", N);
        return kUnsupportedShape;
    }
    const GnNhwcGeom g = gn_nhwc_geom(N, C, HW);
    const long long need = (long long)N * (g.chunks + 1) * groups * 2 * (long long)sizeof(float);
    if (scratch == nullptr || scratch_bytes < need) {
        set_last_error("groupnorm_nhwc: scratch buffer of %lld bytes needed (got %lld)", need, scratch_bytes);
        return kScratchTooSmall;
    }
    dim3 grid(g.chunks, N);
    auto xp = static_cast<const __nv_bfloat16*>(x);
    auto ap = static_cast<const __nv_bfloat16*>(add_nc);
    auto gp = static_cast<const __nv_bfloat16*>(gamma);
    auto bp = static_cast<const __nv_bfloat16*>(beta);
    auto yp = static_cast<__nv_bfloat16*>(y);
    auto sp = static_cast<float*>(scratch);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    gn_nhwc_stats_kernel<<<grid, g.threads, 0, st>>>(xp, ap, sp, C, HW, groups, g.nvec, g.rows_per_iter,
                                                     g.rows_per_chunk);
    CGS_CUDA_CHECK(cudaGetLastError());
    const float* stats = sp;
    int stat_chunks = g.chunks;
    if (g.chunks > 128) {
        float* fin = sp + (size_t)N * g.chunks * groups * 2;
        gn_nhwc_finalize_kernel<<<N, 256, 0, st>>>(sp, fin, g.chunks, groups);
        CGS_CUDA_CHECK(cudaGetLastError());
        stats = fin;
        stat_chunks = 1;
    }
    if (silu)
        gn_nhwc_apply_kernel<true><<<grid, g.threads, 0, st>>>(xp, ap, gp, bp, stats, stat_chunks, yp, C, HW, groups,
                                                               g.nvec, g.rows_per_iter, g.rows_per_chunk, eps);
    else
        gn_nhwc_apply_kernel<false><<<grid, g.threads, 0, st>>>(xp, ap, gp, bp, stats, stat_chunks, yp, C, HW, groups,
                                                                g.nvec, g.rows_per_iter, g.rows_per_chunk, eps);
    CGS_CUDA_CHECK(cudaGetLastError());
    return kOk;
}

// ---------------------------------------------------------------------------------------------------------------
// Allocation-free variant of the channels-last GroupNorm.
//
// cgs_groupnorm_nhwc_fwd needs a scratch buffer that somebody has to cudaMalloc. Some callers cannot: a custom node
// or serving worker that lives inside the host framework's memory budget (the reference sizes its batches from
// model_management.get_free_memory(), which cannot see memory taken behind the framework allocator's back), a plain C
// caller that owns no device allocator at all, or code on a latency path where the device-wide synchronisation of
// cudaMalloc / cudaFree is not acceptable. For them the partial-statistics buffer lives in the library image itself:
// a fixed 1 MiB __device__ array that every GPU loading this library carries until the process exits.
//
// What the caller has to know:
//   * the first call on each device looks up the array's device address (cudaGetSymbolAddress) with that device
//     current -- make that first call during set-up, before any stream capture is started;
//   * problems whose partial statistics do not fit (N * (chunks + 1) * groups * 8 bytes > 1 MiB; roughly N > 32 at the
//     largest feature maps) are refused with CGS_SCRATCH_TOO_SMALL, nothing is launched;
//   * all calls on a device share the one array and it keeps whatever the previous call left in it. Nothing has to
//     be zeroed between calls (the statistics kernel overwrites every slot the normalise kernel reads), but two calls
//     whose kernels overlap on the GPU -- different streams, or different host threads on the same device -- race on
//     it. Keep them on one stream, or order the streams with events.
//   * neither this function nor anything it calls allocates or frees memory.
// ---------------------------------------------------------------------------------------------------------------
namespace cgs {
constexpr long long kGnStaticScratchBytes = 1 << 20;
static __device__ __align__(256) float g_gn_static_scratch[kGnStaticScratchBytes / sizeof(float)];
// In host code the name above is only a handle for the runtime, not a device address: the address each device
// assigned to its copy has to be asked for, per device, and is cached here.
// (atomic slots: two host threads may make a device's first call at the same time; both look the address up and
// store the same value)
static std::atomic<float*> g_gn_static_scratch_addr[64];
}  // namespace cgs

extern "C" int cgs_groupnorm_nhwc_noalloc_fwd(const void* x, const void* add_nc, const void* gamma, const void* beta,
                                              void* y, int N, int C, int HW, int groups, float eps, int silu,
                                              void* stream) {
    using namespace cgs;
    int dev = -1;
    CGS_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) {
        set_last_error("groupnorm_nhwc_noalloc: device ordinal %d not supported (0..63)", dev);
        return kBadArgument;
    }
    if (N > 0 && C > 0 && HW > 0 && groups > 0) {
        const long long need = cgs_groupnorm_nhwc_scratch_bytes(N, C, HW, groups);
        if (need > kGnStaticScratchBytes) {
            set_last_error("groupnorm_nhwc_noalloc: this problem needs %lld bytes of partial statistics but the built-in "
                           "buffer holds %lld; use cgs_groupnorm_nhwc_fwd with a caller-provided scratch buffer",
                           need, kGnStaticScratchBytes);
            return kScratchTooSmall;
        }
    }
    // The scratch array used is the CURRENT device's copy: the tensors have to live there too, or the kernels
    // would be handed a pointer into another GPU's memory.
    if (x != nullptr) {
        cudaPointerAttributes attr;
        CGS_CUDA_CHECK(cudaPointerGetAttributes(&attr, x));
        if (attr.type == cudaMemoryTypeDevice && attr.device != dev) {
            set_last_error("groupnorm_nhwc_noalloc: x lives on device %d but device %d is current; make the tensors' "
                           "device current before the call", attr.device, dev);
            return kBadArgument;
        }
    }
    float* scratch = g_gn_static_scratch_addr[dev].load(std::memory_order_acquire);
    if (scratch == nullptr) {
        // once per device, with that device current (it is: cudaGetDevice above)
        CGS_CUDA_CHECK(cudaGetSymbolAddress(reinterpret_cast<void**>(&scratch), g_gn_static_scratch));
        if (scratch == nullptr) {
            set_last_error("groupnorm_nhwc_noalloc: cudaGetSymbolAddress returned a null address on device %d", dev);
            return kCudaError;
        }
        g_gn_static_scratch_addr[dev].store(scratch, std::memory_order_release);
    }
    return cgs_groupnorm_nhwc_fwd(x, add_nc, gamma, beta, y, scratch, kGnStaticScratchBytes, N, C, HW, groups, eps,
                                  silu, stream);
}
