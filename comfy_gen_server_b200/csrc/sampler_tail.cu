// C1 + S1: the classifier-free-guidance combine and the sampler update, fused.
//
// What the reference runs as ~15 + ~20 ATen launches per step on <= 4 MiB tensors:
//   calculate_input      x / sqrt(sigma^2 + sigma_data^2)         comfy/model_sampling.py:6-8
//   calculate_denoised   EPS:  x - out * sigma                      comfy/model_sampling.py:10-12
//                        V:    x*sd^2/(s^2+sd^2) - out*s*sd/sqrt(s^2+sd^2)   :27-29
//   cfg_function         uncond + (cond - uncond) * cond_scale      comfy/samplers.py:240
//   sample_euler         d = (x - D)/sigma ; x += d * dt            comfy/k_diffusion/sampling.py:137-143
//   sample_dpmpp_2m      x = c1*x - c2*(c3*D - c4*D_old)            comfy/k_diffusion/sampling.py:583-592
//
// All arithmetic is fp32 with explicit round-to-nearest intrinsics in the reference's operation
// order (no FMA contraction), so results are bit-identical to the reference's eager composition
// for the same scalar coefficients. The per-step scalar coefficients are computed once per
// schedule on the host (no device->host sync inside the sampling loop).
//
// HBM-bound, float4-vectorised, grid-stride. Algorithmic bytes per latent element:
//   cfg_denoise: 4 (x) + 2*esz (model out, uncond+cond) + 4 (D)
//   euler step:  4 (x) + 4 (D) + 4 (x')        dpmpp2m step: 4 + 4 + 4 (D_old) + 4
//   fused tail:  4 (x) + 2*esz + [4 (D_old)] + 4 (D) + 4 (x')
#include "common.cuh"

namespace cgs {

struct Out4 {
    float v[4];
};

template <bool BF16>
__device__ __forceinline__ Out4 load_model_out(const void* p, long long i4) {
    Out4 r;
    if (BF16) {
        const uint2 u = __ldg(reinterpret_cast<const uint2*>(p) + i4);
        const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&u.x);
        const __nv_bfloat162 b = *reinterpret_cast<const __nv_bfloat162*>(&u.y);
        const float2 fa = __bfloat1622float2(a), fb = __bfloat1622float2(b);
        r.v[0] = fa.x; r.v[1] = fa.y; r.v[2] = fb.x; r.v[3] = fb.y;
    } else {
        const float4 f = __ldg(reinterpret_cast<const float4*>(p) + i4);
        r.v[0] = f.x; r.v[1] = f.y; r.v[2] = f.z; r.v[3] = f.w;
    }
    return r;
}

// pred_type: 0 = EPS, 1 = V_PREDICTION. sigma_data == 1 for every ModelSamplingDiscrete model.
__device__ __forceinline__ float denoise_one(float x, float out, float sigma, float sigma_data,
                                             int pred_type) {
    if (pred_type == 0) {
        return __fsub_rn(x, __fmul_rn(out, sigma));
    }
    const float sd2 = __fmul_rn(sigma_data, sigma_data);
    const float den = __fadd_rn(__fmul_rn(sigma, sigma), sd2);
    const float a = __fdiv_rn(__fmul_rn(x, sd2), den);
    const float b = __fdiv_rn(__fmul_rn(__fmul_rn(out, sigma), sigma_data), __fsqrt_rn(den));
    return __fsub_rn(a, b);
}

__device__ __forceinline__ float cfg_one(float dc, float du, float scale) {
    return __fadd_rn(du, __fmul_rn(__fsub_rn(dc, du), scale));
}

enum TailMode : int { kTailNone = 0, kTailEuler = 1, kTailDpmpp2mFirst = 2, kTailDpmpp2m = 3 };

struct TailCoef {
    // euler: c0 = sigma_hat, c1 = dt.   dpmpp2m: x' = c0*x - c1*(c2*D - c3*D_old)
    float c0, c1, c2, c3;
};

__device__ __forceinline__ float step_one(int mode, float x, float d, float d_old, TailCoef k) {
    if (mode == kTailEuler) {
        const float dd = __fdiv_rn(__fsub_rn(x, d), k.c0);
        return __fadd_rn(x, __fmul_rn(dd, k.c1));
    }
    float den = d;
    if (mode == kTailDpmpp2m) den = __fsub_rn(__fmul_rn(k.c2, d), __fmul_rn(k.c3, d_old));
    return __fsub_rn(__fmul_rn(k.c0, x), __fmul_rn(k.c1, den));
}

// One kernel for the whole tail. If out_cond == nullptr the model output is skipped and
// `denoised` is an input (pure sampler step). If mode == kTailNone only D is produced.
template <bool BF16>
__global__ void __launch_bounds__(256)
tail_kernel(const float* __restrict__ x, const void* __restrict__ out_uncond,
            const void* __restrict__ out_cond, const float* __restrict__ sigma,
            const float* __restrict__ d_old, float* __restrict__ denoised,
            float* __restrict__ x_next, long long n4, long long per_sample4, float cond_scale,
            float sigma_data, int pred_type, int mode, TailCoef k, const float* __restrict__ dev_step) {
    if (dev_step != nullptr) {
        // sampler mode and coefficients of this step live in device memory: the launch (and a CUDA graph holding it)
        // is the same for every step of the schedule
        mode = static_cast<int>(__ldg(dev_step));
        k.c0 = __ldg(dev_step + 1);
        k.c1 = __ldg(dev_step + 2);
        k.c2 = __ldg(dev_step + 3);
        k.c3 = __ldg(dev_step + 4);
    }
    // (large tensors get half as many threads with two float4 groups each: the second group's loads are in flight
    // while the first is being computed -- cfg_denoise alone went from 0.79 to above 0.80 of the copy bandwidth)
#pragma unroll 2
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
         i += (long long)gridDim.x * blockDim.x) {
        const float4 xv = __ldg(reinterpret_cast<const float4*>(x) + i);
        const float xs[4] = {xv.x, xv.y, xv.z, xv.w};
        float d[4];
        if (out_cond != nullptr) {
            const float s = __ldg(sigma + ((n4 >> 31) == 0 ? (long long)((unsigned)i / (unsigned)per_sample4)
                                                           : i / per_sample4));
            const Out4 oc = load_model_out<BF16>(out_cond, i);
            if (out_uncond != nullptr) {
                const Out4 ou = load_model_out<BF16>(out_uncond, i);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float dc = denoise_one(xs[j], oc.v[j], s, sigma_data, pred_type);
                    const float du = denoise_one(xs[j], ou.v[j], s, sigma_data, pred_type);
                    d[j] = cfg_one(dc, du, cond_scale);
                }
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    d[j] = denoise_one(xs[j], oc.v[j], s, sigma_data, pred_type);
            }
            reinterpret_cast<float4*>(denoised)[i] = make_float4(d[0], d[1], d[2], d[3]);
        } else {
            const float4 dv = __ldg(reinterpret_cast<const float4*>(denoised) + i);
            d[0] = dv.x; d[1] = dv.y; d[2] = dv.z; d[3] = dv.w;
        }
        if (mode != kTailNone) {
            float o[4] = {0.f, 0.f, 0.f, 0.f};
            if (mode == kTailDpmpp2m) {
                const float4 ov = __ldg(reinterpret_cast<const float4*>(d_old) + i);
                o[0] = ov.x; o[1] = ov.y; o[2] = ov.z; o[3] = ov.w;
            }
            float4 r;
            r.x = step_one(mode, xs[0], d[0], o[0], k);
            r.y = step_one(mode, xs[1], d[1], o[1], k);
            r.z = step_one(mode, xs[2], d[2], o[2], k);
            r.w = step_one(mode, xs[3], d[3], o[3], k);
            reinterpret_cast<float4*>(x_next)[i] = r;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// CFG pair split across two GPUs: the exchange of the two halves of the model output and the tail, in ONE kernel.
//
// Each rank of a pair evaluated one half of the UNet batch (cond or uncond). Both need both halves for
// cfg_function (comfy/samplers.py:234-247). Instead of a collective followed by the tail kernel, every CTA
//   1. pushes its chunk of this rank's half into the partner's receive buffer (ordinary 8-byte global stores
//      through a peer mapping: they cross NVLink once, as posted writes), fences, and raises the partner's flag
//      for that chunk (release store, system scope);
//   2. waits for its own flag (acquire load, system scope): the partner's chunk has landed in local memory;
//   3. runs the tail on (x, its half, the partner's half).
// Chunk i of one rank pairs with chunk i of the other (same shapes, same grid), so the exchange is pipelined chunk
// by chunk and needs no grid-wide or cross-GPU barrier. The receive buffer has two slots (step parity): a rank can
// only be one step ahead of its partner, because its next step needs the partner's push of that step. Flags are the
// step sequence number and only ever grow. The step's sequence number, sampler mode and coefficients are read from
// device memory, so the launch is the same for every step and sits inside the replayed step graph.
// A partner that never arrives trips a watchdog (trap) instead of hanging the GPU.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void st_release_sys(int* p, int v) {
    asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(256)
tail_peer_kernel(const float* __restrict__ x, const uint2* __restrict__ out_local, const uint2* recv_data,
                 uint2* peer_data, const int* recv_flags, int* peer_flags, const float* __restrict__ sigma,
                 const float* __restrict__ d_old, float* __restrict__ denoised, float* __restrict__ x_next,
                 long long n4, long long per_sample4, float cond_scale, float sigma_data, int pred_type,
                 int local_is_cond, const float* __restrict__ dev_step) {
    const int mode = static_cast<int>(__ldg(dev_step));
    TailCoef k;
    k.c0 = __ldg(dev_step + 1);
    k.c1 = __ldg(dev_step + 2);
    k.c2 = __ldg(dev_step + 3);
    k.c3 = __ldg(dev_step + 4);
    const int seq = __float_as_int(__ldg(dev_step + 5));
    const long long slot_off = (seq & 1) ? n4 : 0;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i < n4;

    uint2 mine = make_uint2(0u, 0u);
    if (live) {
        mine = __ldg(out_local + i);
        peer_data[slot_off + i] = mine;                  // remote store into the partner's receive slot
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        st_release_sys(peer_flags + blockIdx.x, seq + 1);
        long long t0 = clock64();
        while (ld_acquire_sys(recv_flags + blockIdx.x) < seq + 1) {
            if (clock64() - t0 > 60000000000LL) __trap();    // ~30 s: the partner rank is gone
        }
    }
    __syncthreads();
    if (!live) return;
    // written by the partner GPU: read past the (non-coherent) L1
    const uint2 theirs = __ldcv(recv_data + slot_off + i);
    const uint2 uc = local_is_cond ? theirs : mine;      // uncond half
    const uint2 cc = local_is_cond ? mine : theirs;      // cond half
    const float2 u0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&uc.x));
    const float2 u1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&uc.y));
    const float2 c0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&cc.x));
    const float2 c1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&cc.y));
    const float ou[4] = {u0.x, u0.y, u1.x, u1.y};
    const float oc[4] = {c0.x, c0.y, c1.x, c1.y};
    const float4 xv = __ldg(reinterpret_cast<const float4*>(x) + i);
    const float xs[4] = {xv.x, xv.y, xv.z, xv.w};
    const float s = __ldg(sigma + i / per_sample4);
    float d[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float dc = denoise_one(xs[j], oc[j], s, sigma_data, pred_type);
        const float du = denoise_one(xs[j], ou[j], s, sigma_data, pred_type);
        d[j] = cfg_one(dc, du, cond_scale);
    }
    reinterpret_cast<float4*>(denoised)[i] = make_float4(d[0], d[1], d[2], d[3]);
    float o[4] = {0.f, 0.f, 0.f, 0.f};
    if (mode == kTailDpmpp2m) {
        const float4 ov = __ldg(reinterpret_cast<const float4*>(d_old) + i);
        o[0] = ov.x; o[1] = ov.y; o[2] = ov.z; o[3] = ov.w;
    }
    float4 r;
    r.x = step_one(mode, xs[0], d[0], o[0], k);
    r.y = step_one(mode, xs[1], d[1], o[1], k);
    r.z = step_one(mode, xs[2], d[2], o[2], k);
    r.w = step_one(mode, xs[3], d[3], o[3], k);
    reinterpret_cast<float4*>(x_next)[i] = r;
}

// x (fp32) -> UNet input (bf16), duplicated `copies` times along the batch (uncond, cond halves):
//   out[c, b, ...] = bf16( x[b, ...] / sqrt(sigma[b]^2 + sigma_data^2) )
__global__ void __launch_bounds__(256)
unet_input_kernel(const float* __restrict__ x, const float* __restrict__ sigma,
                  __nv_bfloat16* __restrict__ out, long long n4, long long per_sample4, int copies,
                  float sigma_data) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
         i += (long long)gridDim.x * blockDim.x) {
        const float s = __ldg(sigma + i / per_sample4);
        const float den =
            __fsqrt_rn(__fadd_rn(__fmul_rn(s, s), __fmul_rn(sigma_data, sigma_data)));
        const float4 xv = __ldg(reinterpret_cast<const float4*>(x) + i);
        uint2 packed;
        packed.x = pack_bf16x2(__fdiv_rn(xv.x, den), __fdiv_rn(xv.y, den));
        packed.y = pack_bf16x2(__fdiv_rn(xv.z, den), __fdiv_rn(xv.w, den));
        for (int c = 0; c < copies; ++c) reinterpret_cast<uint2*>(out)[(long long)c * n4 + i] = packed;
    }
}

// One float4 group per thread when the tensor allows it (the hardware block scheduler balances a streaming
// kernel better than a capped grid-stride loop: 74 % -> see profiles/); the loop only matters beyond 2^31 threads.
static unsigned tail_grid(long long n4) {
    long long want = (n4 + 255) / 256;
    if (n4 >= (1LL << 22)) want = (want + 1) / 2;          // >= 64 MiB of latents: two groups per thread
    const long long cap = 0x7fffffffLL / 256;
    return (unsigned)(want < cap ? (want < 1 ? 1 : want) : cap);
}

static int check_tail_shape(const char* who, long long batch, long long per_sample) {
    if (batch < 0 || per_sample < 0) {
        set_last_error("%s: negative size", who);
        return kBadArgument;
    }
    if (per_sample % 4 != 0) {
        set_last_error("%s: per-sample element count %lld must be a multiple of 4", who, per_sample);
        return kUnsupportedShape;
    }
    return kOk;
}

static int launch_tail(const float* x, const void* ou, const void* oc, int out_is_bf16,
                       const float* sigma, const float* d_old, float* denoised, float* x_next,
                       long long batch, long long per_sample, float cond_scale, float sigma_data,
                       int pred_type, int mode, TailCoef k, cudaStream_t stream,
                       const float* dev_step = nullptr) {
    int rc = check_tail_shape("sampler tail", batch, per_sample);
    if (rc != kOk) return rc;
    const long long n4 = batch * per_sample / 4;
    if (n4 == 0) return kOk;
    if (pred_type != 0 && pred_type != 1) {
        set_last_error("sampler tail: pred_type %d (0 = EPS, 1 = V_PREDICTION)", pred_type);
        return kBadArgument;
    }
    const unsigned grid = tail_grid(n4);
    if (out_is_bf16)
        tail_kernel<true><<<grid, 256, 0, stream>>>(x, ou, oc, sigma, d_old, denoised, x_next, n4,
                                                    per_sample / 4, cond_scale, sigma_data,
                                                    pred_type, mode, k, dev_step);
    else
        tail_kernel<false><<<grid, 256, 0, stream>>>(x, ou, oc, sigma, d_old, denoised, x_next, n4,
                                                     per_sample / 4, cond_scale, sigma_data,
                                                     pred_type, mode, k, dev_step);
    CGS_CUDA_CHECK(cudaGetLastError());
    return kOk;
}

}  // namespace cgs

using cgs::TailCoef;

extern "C" int cgs_unet_input_fwd(const float* x, const float* sigma, void* out_bf16,
                                  long long batch, long long per_sample, int copies,
                                  float sigma_data, void* stream) {
    int rc = cgs::check_tail_shape("unet_input", batch, per_sample);
    if (rc != cgs::kOk) return rc;
    const long long n4 = batch * per_sample / 4;
    if (n4 == 0 || copies <= 0) return cgs::kOk;
    cgs::unet_input_kernel<<<cgs::tail_grid(n4), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        x, sigma, static_cast<__nv_bfloat16*>(out_bf16), n4, per_sample / 4, copies, sigma_data);
    CGS_CUDA_CHECK(cudaGetLastError());
    return cgs::kOk;
}

extern "C" int cgs_cfg_denoise_fwd(const float* x, const void* out_uncond, const void* out_cond,
                                   int out_is_bf16, const float* sigma, float cond_scale,
                                   int pred_type, float sigma_data, float* denoised,
                                   long long batch, long long per_sample, void* stream) {
    if (out_cond == nullptr) {
        cgs::set_last_error("cfg_denoise: out_cond must not be null");
        return cgs::kBadArgument;
    }
    return cgs::launch_tail(x, out_uncond, out_cond, out_is_bf16, sigma, nullptr, denoised, nullptr,
                            batch, per_sample, cond_scale, sigma_data, pred_type, cgs::kTailNone,
                            TailCoef{0, 0, 0, 0}, static_cast<cudaStream_t>(stream));
}

extern "C" int cgs_sampler_step_euler(const float* x, const float* denoised, float sigma_hat,
                                      float dt, float* x_next, long long n, void* stream) {
    return cgs::launch_tail(x, nullptr, nullptr, 0, nullptr, nullptr, const_cast<float*>(denoised),
                            x_next, 1, n, 0.f, 1.f, 0, cgs::kTailEuler, TailCoef{sigma_hat, dt, 0, 0},
                            static_cast<cudaStream_t>(stream));
}

extern "C" int cgs_sampler_step_dpmpp2m(const float* x, const float* denoised,
                                        const float* old_denoised, float c_x, float c_d,
                                        float c_cur, float c_old, float* x_next, long long n,
                                        void* stream) {
    const int mode = old_denoised ? cgs::kTailDpmpp2m : cgs::kTailDpmpp2mFirst;
    return cgs::launch_tail(x, nullptr, nullptr, 0, nullptr, old_denoised,
                            const_cast<float*>(denoised), x_next, 1, n, 0.f, 1.f, 0, mode,
                            TailCoef{c_x, c_d, c_cur, c_old}, static_cast<cudaStream_t>(stream));
}

// Whole tail in one launch: model output (uncond, cond) -> denoised (CFG) -> sampler update.
// sampler: 1 = euler (c0 = sigma_hat, c1 = dt); 2 = dpmpp_2m first/last-step form (c0, c1);
//          3 = dpmpp_2m multistep form (c0..c3, old_denoised required).
extern "C" int cgs_cfg_sampler_tail_fwd(const float* x, const void* out_uncond,
                                        const void* out_cond, int out_is_bf16, const float* sigma,
                                        float cond_scale, int pred_type, float sigma_data,
                                        const float* old_denoised, int sampler, float c0, float c1,
                                        float c2, float c3, float* denoised, float* x_next,
                                        long long batch, long long per_sample, void* stream) {
    if (out_cond == nullptr) {
        cgs::set_last_error("cfg_sampler_tail: out_cond must not be null");
        return cgs::kBadArgument;
    }
    if (sampler < 1 || sampler > 3 || (sampler == 3 && old_denoised == nullptr)) {
        cgs::set_last_error("cfg_sampler_tail: sampler mode %d invalid (or old_denoised missing)",
                            sampler);
        return cgs::kBadArgument;
    }
    return cgs::launch_tail(x, out_uncond, out_cond, out_is_bf16, sigma, old_denoised, denoised,
                            x_next, batch, per_sample, cond_scale, sigma_data, pred_type, sampler,
                            TailCoef{c0, c1, c2, c3}, static_cast<cudaStream_t>(stream));
}

// The same tail with the step's sampler mode and coefficients read from device memory: step[0] = mode (1, 2 or 3 as
// a float), step[1..4] = c0..c3. old_denoised must always be a valid buffer (it is read only in mode 3).
extern "C" int cgs_cfg_sampler_tail_dev_fwd(const float* x, const void* out_uncond,
                                            const void* out_cond, int out_is_bf16, const float* sigma,
                                            float cond_scale, int pred_type, float sigma_data,
                                            const float* old_denoised, const float* step,
                                            float* denoised, float* x_next, long long batch,
                                            long long per_sample, void* stream) {
    if (out_cond == nullptr || old_denoised == nullptr || step == nullptr) {
        cgs::set_last_error("cfg_sampler_tail_dev: out_cond, old_denoised and step must not be null");
        return cgs::kBadArgument;
    }
    return cgs::launch_tail(x, out_uncond, out_cond, out_is_bf16, sigma, old_denoised, denoised,
                            x_next, batch, per_sample, cond_scale, sigma_data, pred_type, cgs::kTailDpmpp2m,
                            TailCoef{0, 0, 0, 0}, static_cast<cudaStream_t>(stream), step);
}

// CFG-pair split over two GPUs: exchange + tail in one launch (see tail_peer_kernel). All `recv_*` pointers are this
// GPU's own memory, `peer_*` the partner's receive buffer / flags through a peer mapping. Model outputs are bf16.
// step[0..4] as in cgs_cfg_sampler_tail_dev_fwd, step[5] = the bit pattern of the int32 sequence number of this
// exchange (0, 1, 2, ... identical on both ranks). flags: one int32 per 1024 latent elements, zero-initialised once.
extern "C" int cgs_cfg_sampler_tail_peer_fwd(const float* x, const void* out_local, int local_is_cond,
                                             const void* recv_data, void* peer_data, const int* recv_flags,
                                             int* peer_flags, long long flag_capacity, const float* sigma,
                                             float cond_scale, int pred_type, float sigma_data,
                                             const float* old_denoised, const float* step, float* denoised,
                                             float* x_next, long long batch, long long per_sample, void* stream) {
    using namespace cgs;
    if (!x || !out_local || !recv_data || !peer_data || !recv_flags || !peer_flags || !sigma || !old_denoised ||
        !step || !denoised || !x_next) {
        set_last_error("cfg_sampler_tail_peer: null pointer argument");
        return kBadArgument;
    }
    int rc = check_tail_shape("cfg_sampler_tail_peer", batch, per_sample);
    if (rc != kOk) return rc;
    if (pred_type != 0 && pred_type != 1) {
        set_last_error("cfg_sampler_tail_peer: pred_type %d (0 = EPS, 1 = V_PREDICTION)", pred_type);
        return kBadArgument;
    }
    const long long n4 = batch * per_sample / 4;
    if (n4 == 0) return kOk;
    const long long grid = (n4 + 255) / 256;
    // every CTA waits for its partner CTA while holding its SM slot: all of them must be resident at once
    if (grid > flag_capacity || grid > (long long)sm_count() * 8) {
        set_last_error("cfg_sampler_tail_peer: %lld chunks exceed the flag capacity %lld or the resident-CTA limit %lld",
                       grid, flag_capacity, (long long)sm_count() * 8);
        return kUnsupportedShape;
    }
    tail_peer_kernel<<<(unsigned)grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        x, static_cast<const uint2*>(out_local), static_cast<const uint2*>(recv_data), static_cast<uint2*>(peer_data),
        recv_flags, peer_flags, sigma, old_denoised, denoised, x_next, n4, per_sample / 4, cond_scale, sigma_data,
        pred_type, local_is_cond, step);
    CGS_CUDA_CHECK(cudaGetLastError());
    return kOk;
}
