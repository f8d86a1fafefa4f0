// A1/A2: multi-head attention forward on tcgen05 / TMEM with TMA-staged Q/K/V tiles and an
// online softmax. Replaces the reference's optimized_attention(q, k, v, heads) =
// attention_pytorch -> F.scaled_dot_product_attention (comfy/ldm/modules/attention.py:337-349),
// non-causal, no dropout, scale 1/sqrt(d), mask None (the UNet never passes one,
// attention.py:411-414, 520, 560).
//
// Layout: q [B, Nq, H, d], k/v [B, Nk, H, d] bf16 with d contiguous and arbitrary (16-byte
// aligned) strides for B, N and H -- i.e. the reference's [B, N, H*d] tensors as they are, or
// column slices of a fused QKV projection. Output [B, Nq, H, d] bf16. No transposes, no copies.
//
// One CTA handles NQT query tiles of 128 rows for one (batch, head):
//   warps [0, 4*NQT)   softmax warpgroups, one per query tile; thread == query row
//   warp  4*NQT        TMA producer (Q once; K/V ring of KST stages, 128 keys per tile)
//   warp  4*NQT+1      TMEM allocator + single-thread tcgen05.mma issuer
// Per key tile j and query tile t:
//   S_t = Q_t K_j^T        SS MMA, M=128, N<=128, K=d            -> TMEM  (128 fp32 columns)
//   softmax warpgroup t:   tcgen05.ld S, running max (lazy: O is only rescaled when the max
//                          grows by > 2^8), P = exp2(S*scale - m) as bf16 -> tcgen05.st over S
//   O_t += P_t V_j         TS MMA (A = P from TMEM, B = V MN-major from SMEM), fp32 in TMEM
// With NQT = 2 the two warpgroups ping-pong: the tensor core works on one tile's MMAs while
// the other tile is in its softmax. Ragged Nq / Nk and head dims that are not multiples of 64
// (40, 80, 160) are handled by TMA out-of-bounds zero fill plus column masking; the MMA K/N
// extents are trimmed to multiples of 16.
//
// Tensor-core bound for self-attention. Algorithmic FLOPs = 4 * B * H * Nq * Nk * d.
#include "common.cuh"

namespace cgs {

namespace attn {
constexpr int BLOCK_M = 128;
constexpr int BLOCK_N = 128;
constexpr int TILE_BYTES = 128 * 128;  // one [128 rows x 64 bf16] swizzled box
constexpr float kRescaleThreshold = 8.0f;  // log2 units
}  // namespace attn

struct AttnParams {
    __nv_bfloat16* out;
    long long o_stride_b, o_stride_n, o_stride_h;  // elements
    int B, H, Nq, Nk, d;
    float scale_log2;
};

template <int NKB, int NQT, int KST>
struct AttnSmem {
    static constexpr int Q_OFF = 0;
    static constexpr int K_OFF = Q_OFF + NQT * NKB * attn::TILE_BYTES;
    static constexpr int V_OFF = K_OFF + KST * NKB * attn::TILE_BYTES;
    static constexpr int BAR_OFF = V_OFF + KST * NKB * attn::TILE_BYTES;
    static constexpr int NBARS = 3 * NQT + 4 * KST + NQT;
    static constexpr int BYTES = BAR_OFF + NBARS * 8 + 16 + 1024;
};

template <int NKB, int NQT, int KST>
__global__ void __launch_bounds__((NQT * 4 + 2) * 32, 1)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const AttnParams p) {
    using namespace attn;
    using L = AttnSmem<NKB, NQT, KST>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
    uint64_t* q_full = bars;                 // [NQT]
    uint64_t* s_full = q_full + NQT;         // [NQT]
    uint64_t* p_full = s_full + NQT;         // [NQT]
    uint64_t* o_full = p_full + NQT;         // [NQT]
    uint64_t* k_full = o_full + NQT;         // [KST]
    uint64_t* k_empty = k_full + KST;        // [KST]
    uint64_t* v_full = k_empty + KST;        // [KST]
    uint64_t* v_empty = v_full + KST;        // [KST]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(v_empty + KST);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    constexpr int PRODUCER_WARP = NQT * 4;
    constexpr int MMA_WARP = NQT * 4 + 1;

    const int q0 = blockIdx.x * (BLOCK_M * NQT);
    const int head = blockIdx.y;
    const int batch = blockIdx.z;
    const int n_tiles = (p.Nk + BLOCK_N - 1) / BLOCK_N;
    const int ksteps_d = (p.d + 15) >> 4;     // K steps of the QK^T MMA
    const int dv_n = ksteps_d << 4;           // N of the PV MMA (d rounded up to 16)

    constexpr uint32_t S_COL = 0;                 // S_t at S_COL + t*128 ; P_t aliases S_t
    constexpr uint32_t O_COL = NQT * BLOCK_N;     // O_t at O_COL + t*NKB*64

    if (warp == PRODUCER_WARP && elect_one()) {
        tma_prefetch_desc(&tmQ);
        tma_prefetch_desc(&tmK);
        tma_prefetch_desc(&tmV);
    }
    if (warp == MMA_WARP) {
        if (elect_one()) {
            for (int t = 0; t < NQT; ++t) {
                mbar_init(&q_full[t], 1);
                mbar_init(&s_full[t], 1);
                mbar_init(&p_full[t], 128);
                mbar_init(&o_full[t], 1);
            }
            for (int s = 0; s < KST; ++s) {
                mbar_init(&k_full[s], 1);
                mbar_init(&k_empty[s], 1);
                mbar_init(&v_full[s], 1);
                mbar_init(&v_empty[s], 1);
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

    if (warp == PRODUCER_WARP) {
        // ------------------------------------------------------------------ TMA producer
        if (elect_one()) {
            for (int t = 0; t < NQT; ++t) {
                mbar_expect_tx(&q_full[t], NKB * TILE_BYTES);
                for (int kb = 0; kb < NKB; ++kb)
                    tma_load_4d(smem + L::Q_OFF + (t * NKB + kb) * TILE_BYTES, &tmQ, &q_full[t],
                                kb * 64, head, q0 + t * BLOCK_M, batch);
            }
            int stage = 0;
            uint32_t phase = 0;
            for (int j = 0; j < n_tiles; ++j) {
                mbar_wait(&k_empty[stage], phase ^ 1);
                mbar_expect_tx(&k_full[stage], NKB * TILE_BYTES);
                for (int kb = 0; kb < NKB; ++kb)
                    tma_load_4d(smem + L::K_OFF + (stage * NKB + kb) * TILE_BYTES, &tmK,
                                &k_full[stage], kb * 64, head, j * BLOCK_N, batch);
                mbar_wait(&v_empty[stage], phase ^ 1);
                mbar_expect_tx(&v_full[stage], NKB * TILE_BYTES);
                for (int kb = 0; kb < NKB; ++kb)
                    tma_load_4d(smem + L::V_OFF + (stage * NKB + kb) * TILE_BYTES, &tmV,
                                &v_full[stage], kb * 64, head, j * BLOCK_N, batch);
                if (++stage == KST) { stage = 0; phase ^= 1; }
            }
        }
        __syncwarp();
    } else if (warp == MMA_WARP) {
        // ------------------------------------------------------------------ MMA issuer
        if (elect_one()) {
            const uint32_t idesc_o = umma_idesc_bf16(BLOCK_M, dv_n, 1);
            auto n_mma_of = [&](int j) {
                const int valid = min(BLOCK_N, p.Nk - j * BLOCK_N);
                return (valid + 15) & ~15;
            };
            auto issue_s = [&](int t, int stage, int n_mma) {
                const uint32_t idesc_s = umma_idesc_bf16(BLOCK_M, n_mma, 0);
                const uint32_t d_tmem = tmem_base + S_COL + t * BLOCK_N;
                for (int s = 0; s < ksteps_d; ++s) {
                    const int kb = s >> 2, k = s & 3;
                    const uint64_t adesc = umma_desc_sw128(
                        smem_u32(smem + L::Q_OFF + (t * NKB + kb) * TILE_BYTES) + 32 * k, 16, 1024);
                    const uint64_t bdesc = umma_desc_sw128(
                        smem_u32(smem + L::K_OFF + (stage * NKB + kb) * TILE_BYTES) + 32 * k, 16, 1024);
                    umma_ss(d_tmem, adesc, bdesc, idesc_s, s != 0);
                }
            };
            auto issue_pv = [&](int t, int stage, int n_mma, bool accumulate) {
                const uint32_t d_tmem = tmem_base + O_COL + t * (NKB * 64);
                const uint32_t a_tmem = tmem_base + S_COL + t * BLOCK_N;
                const uint32_t vbase = smem_u32(smem + L::V_OFF + stage * NKB * TILE_BYTES);
                for (int ks = 0; ks < (n_mma >> 4); ++ks) {
                    // 16 key rows per step = 2 swizzle groups of 8 rows = 2048 bytes;
                    // LBO = distance between 64-wide head-dim blocks (one TMA box).
                    const uint64_t bdesc = umma_desc_sw128(vbase + ks * 2048, TILE_BYTES, 1024);
                    umma_ts(d_tmem, a_tmem + ks * 8, bdesc, idesc_o, (accumulate || ks != 0) ? 1u : 0u);
                }
            };

            // prologue: S_t(0) for every query tile
            mbar_wait(&k_full[0], 0);
            for (int t = 0; t < NQT; ++t) {
                mbar_wait(&q_full[t], 0);
                tc_fence_after();
                issue_s(t, 0, n_mma_of(0));
                tc_commit(&s_full[t]);
            }
            tc_commit(&k_empty[0]);

            int stage = 0;
            uint32_t phase = 0;
            for (int j = 0; j < n_tiles; ++j) {
                int nstage = stage + 1;
                uint32_t nphase = phase;
                if (nstage == KST) { nstage = 0; nphase ^= 1; }
                const bool has_next = (j + 1 < n_tiles);
                if (has_next) mbar_wait(&k_full[nstage], nphase);
                mbar_wait(&v_full[stage], phase);
                const int n_mma = n_mma_of(j);
                for (int t = 0; t < NQT; ++t) {
                    mbar_wait(&p_full[t], j & 1);
                    tc_fence_after();
                    issue_pv(t, stage, n_mma, j > 0);
                    tc_commit(&o_full[t]);
                    if (t == NQT - 1) tc_commit(&v_empty[stage]);
                    if (has_next) {
                        issue_s(t, nstage, n_mma_of(j + 1));
                        tc_commit(&s_full[t]);
                        if (t == NQT - 1) tc_commit(&k_empty[nstage]);
                    }
                }
                stage = nstage;
                phase = nphase;
            }
        }
        __syncwarp();
    } else {
        // ------------------------------------------------------------------ softmax warpgroups
        const int t = warp >> 2;
        const int quad = warp & 3;
        const int row_in_tile = quad * 32 + lane;
        const uint32_t lane_field = static_cast<uint32_t>(quad * 32) << 16;
        const uint32_t s_addr = tmem_base + lane_field + S_COL + t * BLOCK_N;
        const uint32_t o_addr = tmem_base + lane_field + O_COL + t * (NKB * 64);

        float m_run = -INFINITY;  // running max, already multiplied by scale_log2
        float l_run = 0.f;
        const float scale = p.scale_log2;

        for (int j = 0; j < n_tiles; ++j) {
            mbar_wait(&s_full[t], j & 1);
            tc_fence_after();
            uint32_t sr[128];
            tmem_ld_x32(s_addr + 0, sr + 0);
            tmem_ld_x32(s_addr + 32, sr + 32);
            tmem_ld_x32(s_addr + 64, sr + 64);
            tmem_ld_x32(s_addr + 96, sr + 96);
            tmem_wait_ld();

            const int n_valid = p.Nk - j * BLOCK_N;
            if (n_valid < BLOCK_N) {
#pragma unroll
                for (int i = 0; i < 128; ++i)
                    if (i >= n_valid) sr[i] = 0xff800000u;  // -inf
            }
            float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
            for (int i = 0; i < 128; i += 4) {
                mx0 = fmaxf(mx0, __uint_as_float(sr[i + 0]));
                mx1 = fmaxf(mx1, __uint_as_float(sr[i + 1]));
                mx2 = fmaxf(mx2, __uint_as_float(sr[i + 2]));
                mx3 = fmaxf(mx3, __uint_as_float(sr[i + 3]));
            }
            const float m_new = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * scale;

            float alpha = 1.f;
            bool need_rescale = false;
            if (j == 0) {
                m_run = m_new;
            } else if (m_new > m_run + kRescaleThreshold) {
                alpha = ex2_approx(m_run - m_new);
                l_run *= alpha;
                m_run = m_new;
                need_rescale = true;
            }
            const float neg_m = -m_run;
            float sum0 = 0.f, sum1 = 0.f, sum2 = 0.f, sum3 = 0.f;
            uint32_t pr[64];
#pragma unroll
            for (int i = 0; i < 128; i += 4) {
                const float e0 = ex2_approx(fmaf(__uint_as_float(sr[i + 0]), scale, neg_m));
                const float e1 = ex2_approx(fmaf(__uint_as_float(sr[i + 1]), scale, neg_m));
                const float e2 = ex2_approx(fmaf(__uint_as_float(sr[i + 2]), scale, neg_m));
                const float e3 = ex2_approx(fmaf(__uint_as_float(sr[i + 3]), scale, neg_m));
                sum0 += e0; sum1 += e1; sum2 += e2; sum3 += e3;
                pr[(i >> 1) + 0] = pack_bf16x2(e0, e1);
                pr[(i >> 1) + 1] = pack_bf16x2(e2, e3);
            }
            l_run += (sum0 + sum1) + (sum2 + sum3);

            // P over the first 64 columns of S (all of S is already in registers)
            tmem_st_x32(s_addr + 0, pr + 0);
            tmem_st_x32(s_addr + 32, pr + 32);

            if (__any_sync(0xffffffffu, need_rescale)) {
                // PV(j-1) has completed by the time S(j) is visible (issue order), so this does
                // not stall; the wait keeps the ordering explicit.
                mbar_wait(&o_full[t], (j - 1) & 1);
                tc_fence_after();
                for (int c = 0; c < (dv_n >> 4); ++c) {
                    uint32_t orr[16];
                    tmem_ld_x16(o_addr + c * 16, orr);
                    tmem_wait_ld();
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        orr[i] = __float_as_uint(__uint_as_float(orr[i]) * alpha);
                    tmem_st_x16(o_addr + c * 16, orr);
                }
            }
            tmem_wait_st();
            tc_fence_before();
            mbar_arrive(&p_full[t]);
        }

        // epilogue: O / l -> bf16 -> global
        mbar_wait(&o_full[t], (n_tiles - 1) & 1);
        tc_fence_after();
        const float inv_l = 1.f / l_run;
        const int q_row = q0 + t * BLOCK_M + row_in_tile;
        __nv_bfloat16* orow = p.out + (long long)batch * p.o_stride_b +
                              (long long)q_row * p.o_stride_n + (long long)head * p.o_stride_h;
        for (int c = 0; c < (dv_n >> 4); ++c) {
            uint32_t orr[16];
            tmem_ld_x16(o_addr + c * 16, orr);
            tmem_wait_ld();
            if (q_row < p.Nq) {
                uint4 w0, w1;
                w0.x = pack_bf16x2(__uint_as_float(orr[0]) * inv_l, __uint_as_float(orr[1]) * inv_l);
                w0.y = pack_bf16x2(__uint_as_float(orr[2]) * inv_l, __uint_as_float(orr[3]) * inv_l);
                w0.z = pack_bf16x2(__uint_as_float(orr[4]) * inv_l, __uint_as_float(orr[5]) * inv_l);
                w0.w = pack_bf16x2(__uint_as_float(orr[6]) * inv_l, __uint_as_float(orr[7]) * inv_l);
                w1.x = pack_bf16x2(__uint_as_float(orr[8]) * inv_l, __uint_as_float(orr[9]) * inv_l);
                w1.y = pack_bf16x2(__uint_as_float(orr[10]) * inv_l, __uint_as_float(orr[11]) * inv_l);
                w1.z = pack_bf16x2(__uint_as_float(orr[12]) * inv_l, __uint_as_float(orr[13]) * inv_l);
                w1.w = pack_bf16x2(__uint_as_float(orr[14]) * inv_l, __uint_as_float(orr[15]) * inv_l);
                uint4* dst = reinterpret_cast<uint4*>(orow + c * 16);
                dst[0] = w0;
                if (c * 16 + 8 < p.d) dst[1] = w1;
            }
        }
        tc_fence_before();
    }

    __syncthreads();
    if (warp == MMA_WARP) tmem_dealloc(tmem_base, 512);
}

template <int NKB, int NQT, int KST>
static int launch_attn(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv,
                       const AttnParams& p, cudaStream_t stream) {
    using L = AttnSmem<NKB, NQT, KST>;
    auto kern = attn_fwd_kernel<NKB, NQT, KST>;
    CGS_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::BYTES));
    dim3 grid((p.Nq + attn::BLOCK_M * NQT - 1) / (attn::BLOCK_M * NQT), p.H, p.B);
    kern<<<grid, (NQT * 4 + 2) * 32, L::BYTES, stream>>>(tq, tk, tv, p);
    CGS_CUDA_CHECK(cudaGetLastError());
    return kOk;
}

static int make_qkv_map(CUtensorMap* m, const void* ptr, int B, int N, int H, int d,
                        long long sb, long long sn, long long sh) {
    const uint64_t dims[4] = {(uint64_t)d, (uint64_t)H, (uint64_t)N, (uint64_t)B};
    const uint64_t strides[3] = {(uint64_t)sh * 2, (uint64_t)sn * 2, (uint64_t)sb * 2};
    const uint32_t box[4] = {64, 1, 128, 1};
    return make_tmap_bf16(m, ptr, 4, dims, strides, box);
}

}  // namespace cgs

extern "C" int cgs_attention_fwd(const void* q, const void* k, const void* v, void* out, int B,
                                 int H, int Nq, int Nk, int d, long long q_stride_b,
                                 long long q_stride_n, long long q_stride_h, long long k_stride_b,
                                 long long k_stride_n, long long k_stride_h, long long v_stride_b,
                                 long long v_stride_n, long long v_stride_h, long long o_stride_b,
                                 long long o_stride_n, long long o_stride_h, float softmax_scale,
                                 void* stream) {
    using namespace cgs;
    if (B < 0 || H <= 0 || Nq < 0 || Nk < 0 || d <= 0) {
        set_last_error("attention: bad sizes B=%d H=%d Nq=%d Nk=%d d=%d", B, H, Nq, Nk, d);
        return kBadArgument;
    }
    if (B == 0 || Nq == 0) return kOk;
    if (Nk == 0) {
        set_last_error("attention: Nk == 0 (softmax over an empty key set is undefined)");
        return kBadArgument;
    }
    if (d % 8 != 0 || d > 192) {
        set_last_error("attention: head dim %d unsupported (multiple of 8, <= 192)", d);
        return kUnsupportedShape;
    }
    if (H > 65535 || B > 65535) {
        set_last_error("attention: H=%d / B=%d exceed the grid limits", H, B);
        return kUnsupportedShape;
    }
    if ((o_stride_b | o_stride_n | o_stride_h) % 8 != 0 || (reinterpret_cast<uintptr_t>(out) & 0xF)) {
        set_last_error("attention: output strides/pointer must be 16-byte aligned");
        return kBadArgument;
    }
    CUtensorMap tq, tk, tv;
    int rc;
    if ((rc = make_qkv_map(&tq, q, B, Nq, H, d, q_stride_b, q_stride_n, q_stride_h)) != kOk) return rc;
    if ((rc = make_qkv_map(&tk, k, B, Nk, H, d, k_stride_b, k_stride_n, k_stride_h)) != kOk) return rc;
    if ((rc = make_qkv_map(&tv, v, B, Nk, H, d, v_stride_b, v_stride_n, v_stride_h)) != kOk) return rc;

    AttnParams p;
    p.out = static_cast<__nv_bfloat16*>(out);
    p.o_stride_b = o_stride_b;
    p.o_stride_n = o_stride_n;
    p.o_stride_h = o_stride_h;
    p.B = B; p.H = H; p.Nq = Nq; p.Nk = Nk; p.d = d;
    p.scale_log2 = softmax_scale * 1.4426950408889634f;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool two_tiles = Nq > attn::BLOCK_M;
    if (d <= 64) return two_tiles ? launch_attn<1, 2, 3>(tq, tk, tv, p, st) : launch_attn<1, 1, 3>(tq, tk, tv, p, st);
    if (d <= 128) return two_tiles ? launch_attn<2, 2, 2>(tq, tk, tv, p, st) : launch_attn<2, 1, 2>(tq, tk, tv, p, st);
    return launch_attn<3, 1, 1>(tq, tk, tv, p, st);
}
