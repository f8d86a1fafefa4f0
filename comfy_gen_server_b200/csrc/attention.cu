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
//   warps [0, 4*NQT)          softmax warpgroups, one per query tile; thread == query row
//   warp  4*NQT               TMA producer (Q once; K/V ring of KST stages, 128 keys per tile)
//   warps 4*NQT+1 .. +NQT     one single-thread tcgen05.mma issuer per query tile (the first also owns TMEM);
//                             the service warpgroup is filled up to four warps and hands its registers to the
//                             softmax warpgroups with setmaxnreg (40 vs 224 / 200 registers per thread).
//                             One issuer for both tiles was the bottleneck of the first versions (~3100 cycles
//                             of serial issue work per key tile, head-of-line blocking between the tiles).
// Per key tile j and query tile t:
//   S_t = Q_t K_j^T        SS MMA, M=128, N<=128, K=d            -> TMEM  (128 fp32 columns)
//   softmax warpgroup t:   tcgen05.ld S into registers and immediately hand the S buffer back (s_free), so the
//                          tensor core computes S_t(j+1) while this warpgroup is still in softmax(j);
//                          row max with 3-input FMNMX; lazy running max (O is rescaled only when the max grows
//                          by > 2^8); x = S*scale - m with packed FFMA2; P = 2^x (MUFU.EX2), row sum with
//                          packed FADD2, P packed to bf16 -> tcgen05.st into its own TMEM region
//   O_t += P_t V_j         TS MMA (A = P from TMEM, B = V MN-major from SMEM), fp32 in TMEM
// TMEM columns: S_t at t*128, P_t at NQT*128 + t*64, O_t at NQT*192 + t*dv_n.
// (Measured on the way here, profiles/: P aliased over S serialises S(j+1) behind PV(j) and left each
// warpgroup waiting ~1200 of 3400 cycles per tile; a row-sum MMA against a ones matrix costs a full MMA slot
// per k-step; ex2.approx.f16x2 is two MUFUs plus a PRMT; kind::f16 rejects fp16 A with bf16 B.)
// With NQT = 2 the two warpgroups ping-pong: the tensor core works on one tile's MMAs while
// the other tile is in its softmax. Ragged Nq / Nk and head dims that are not multiples of 64
// (40, 80, 160) are handled by TMA out-of-bounds zero fill plus column masking; the MMA K/N
// extents are trimmed to multiples of 16.
//
// Tensor-core bound for self-attention. Algorithmic FLOPs = 4 * B * H * Nq * Nk * d.
#include "common.cuh"

#include <stdlib.h>

#include <type_traits>

namespace cgs {

namespace attn {
constexpr int BLOCK_M = 128;
constexpr int TILE_BYTES = 128 * 128;  // one [128 rows x 64 bf16] swizzled box (a Q tile; a 128-key K or V tile)
constexpr float kRescaleThreshold = 8.0f;  // log2 units
#ifndef CGS_ATTN_SETMAXNREG
#define CGS_ATTN_SETMAXNREG 1
#endif
// 4 softmax warps per query tile + one service warpgroup (TMA producer, one MMA issuer per tile, filler)
template <int NQT> constexpr int kThreads = (NQT * 4 + 4) * 32;
constexpr int kRegsService = 40;
// registers per softmax thread after the hand-over: what the launch gives every thread (64 K / threads per SM,
// rounded down to 8) times all threads, minus the service warpgroup's share, over the softmax threads
template <int NQT> constexpr int kRegsLaunch = (65536 / (kThreads<NQT> * (NQT == 1 ? 2 : 1))) / 8 * 8;
template <int NQT> constexpr int kRegsSoftmax =
    ((kRegsLaunch<NQT> * kThreads<NQT> - 128 * kRegsService) / (kThreads<NQT> - 128)) / 8 * 8;
#ifndef CGS_ATTN_POLY_PAIRS          // tuning overrides for tools/attn_variants.py
#define CGS_ATTN_POLY_PAIRS 1
#endif
#ifndef CGS_ATTN_POLY_PERIOD
#define CGS_ATTN_POLY_PERIOD 4
#endif
#ifndef CGS_ATTN_WAITMODE            // how the service warps wait: 0 spin, 1 try_wait with suspend hint, 2 nanosleep
#define CGS_ATTN_WAITMODE 1
#endif
#ifndef CGS_ATTN_EXPT                // timing experiments with wrong results (tools/attn_variants.py): 1 no exp2,
#define CGS_ATTN_EXPT 0              // 2 no softmax arithmetic at all, 3 no row max, 4 no output stores
#endif
constexpr int kPolyPairs = CGS_ATTN_POLY_PAIRS;    // of every kPolyPeriod element pairs, this many use the FMA-pipe exp2
constexpr int kPolyPeriod = CGS_ATTN_POLY_PERIOD;
}  // namespace attn

struct AttnParams {
    long long* dbg;   // optional timeline buffer (CGS_ATTN_TRACE), nullptr in production
    __nv_bfloat16* out;
    int tmem_cols;   // power of two >= NQT * (192 + dv_n)
    long long o_stride_b, o_stride_n, o_stride_h;  // elements
    int B, H, Nq, Nk, d;
    float scale_log2;
    // optional additive mask (fp32, added to the scaled scores; key stride 1, 0-strides broadcast)
    const float* mask;
    long long mask_sb, mask_sh, mask_sq;   // elements
    float inv_scale;                        // 1 / softmax_scale: S + mask * inv_scale == (S * scale + mask) / scale
};

// The timeline stamps split the softmax loop into basic blocks and keep ptxas from overlapping its phases, so they
// are compiled in only by `CGS_BUILD_TRACE=1 python comfy_gen_server_b200/build.py` (-DCGS_ATTN_TRACING=1).
#ifndef CGS_ATTN_TRACING
#define CGS_ATTN_TRACING 0
#endif
#if !CGS_ATTN_TRACING
#define CGS_TRACE(role, j, e) do { } while (0)
#define CGS_TRACE_CTA(e) do { } while (0)
#else
#define CGS_TRACE(role, j, e)                                                                      \
    do {                                                                                           \
        if (p.dbg != nullptr && trace_me && (j) < 64) p.dbg[((role) * 64 + (j)) * 8 + (e)] = clock64(); \
    } while (0)

// whole-CTA events (entry, set-up done, epilogue begin/end, exit) of CTA (0,0,0), stamped by thread 0
#define CGS_TRACE_CTA(e)                                                                            \
    do {                                                                                            \
        if (p.dbg != nullptr && (blockIdx.x | blockIdx.y | blockIdx.z | threadIdx.x) == 0)          \
            p.dbg[(3 * 64 + 63) * 8 + (e)] = clock64();                                             \
    } while (0)
#endif

template <int NKB, int NQT, int KST, int BN>
struct AttnSmem {
    static constexpr int KV_TILE = BN * 128;          // one [BN keys x 64 bf16] swizzled box
    // one- or two-key-tile items (text cross-attention) are as long as one HBM round trip: two Q buffers, so that the
    // query tile of the item after next is already on its way
    static constexpr int QBUF = (NQT == 1 && NKB == 1 && KST == 1) ? 2 : 1;
    static constexpr int Q_OFF = 0;
    static constexpr int K_OFF = Q_OFF + QBUF * NQT * NKB * attn::TILE_BYTES;
    static constexpr int V_OFF = K_OFF + KST * NKB * KV_TILE;
    static constexpr int O_OFF = V_OFF + KST * NKB * KV_TILE;      // one [128 rows x 64 bf16] output staging box per query tile
    static constexpr int BAR_OFF = O_OFF + (NQT == 1 ? attn::TILE_BYTES : 0);   // (only the one-tile variants stage their output)
    static constexpr int NBARS = (4 + 2 * QBUF) * NQT + 4 * KST;
    static constexpr int BYTES = BAR_OFF + NBARS * 8 + 16 + 1024;
};

template <int NKB, int NQT, int KST, int BN, bool MASK>
__global__ void __launch_bounds__(attn::kThreads<NQT>, (NQT == 1 ? 2 : 1))
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO, const AttnParams p) {
    using namespace attn;
    using L = AttnSmem<NKB, NQT, KST, BN>;
    constexpr int BLOCK_N = BN;
    constexpr int KV_TILE = L::KV_TILE;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
    constexpr int QBUF = L::QBUF;
    uint64_t* q_full = bars;                 // [QBUF * NQT]
    uint64_t* s_full = q_full + QBUF * NQT;  // [NQT]
    uint64_t* p_full = s_full + NQT;         // [NQT]
    uint64_t* o_full = p_full + NQT;         // [NQT]
    uint64_t* s_free = o_full + NQT;         // [NQT]
    uint64_t* q_empty = s_free + NQT;        // [QBUF * NQT]
    uint64_t* k_full = q_empty + QBUF * NQT; // [KST]
    uint64_t* k_empty = k_full + KST;        // [KST]
    uint64_t* v_full = k_empty + KST;        // [KST]
    uint64_t* v_empty = v_full + KST;        // [KST]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(v_empty + KST);

    CGS_TRACE_CTA(0);
    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    constexpr int PRODUCER_WARP = NQT * 4;
    constexpr int MMA_WARP = PRODUCER_WARP + 1;      // first of NQT issuer warps

    // Persistent CTAs: work item w = (batch, head, block of NQT query tiles), query block fastest so that the CTAs
    // working on one (batch, head) at the same time share its K/V in L2. Barriers, the K/V ring and the TMEM
    // buffers carry over from one item to the next: Q(w+1), K/V(w+1) load and S(w+1) is computed while the
    // softmax warps are still in the epilogue of w, and TMEM allocation / barrier set-up happen once per CTA.
    // (Text cross-attention is one key tile per item: set-up + first-load latency + epilogue used to be 3/4 of a
    // CTA's life.)
    const int n_qblk = (p.Nq + BLOCK_M * NQT - 1) / (BLOCK_M * NQT);
    const int n_items = n_qblk * p.H * p.B;
    auto item_coords = [&](int w, int& q0, int& head, int& batch) {
        const int qb = w % n_qblk;
        const int r = w / n_qblk;
        q0 = qb * (BLOCK_M * NQT);
        head = r % p.H;
        batch = r / p.H;
    };
    const int n_tiles = (p.Nk + BLOCK_N - 1) / BLOCK_N;
    const int ksteps_d = (p.d + 15) >> 4;     // K steps of the QK^T MMA
    const int dv_n = ksteps_d << 4;           // N of the PV MMA (d rounded up to 16)

    constexpr uint32_t S_COL = 0;                 // S_t at t*128
    constexpr uint32_t P_COL = NQT * BLOCK_N;                     // P_t (bf16, 2 per column) at P_COL + t * BLOCK_N / 2
    constexpr uint32_t O_COL = NQT * (BLOCK_N + BLOCK_N / 2);     // O_t at O_COL + t * dv_n
    const uint32_t o_stride = dv_n;

    if (warp == PRODUCER_WARP && elect_one()) {
        tma_prefetch_desc(&tmQ);
        tma_prefetch_desc(&tmK);
        tma_prefetch_desc(&tmV);
        tma_prefetch_desc(&tmO);
    }
    if (warp == MMA_WARP) {
        if (elect_one()) {
            for (int t = 0; t < QBUF * NQT; ++t) {
                mbar_init(&q_full[t], 1);
                mbar_init(&q_empty[t], 1);
            }
            for (int t = 0; t < NQT; ++t) {
                mbar_init(&s_full[t], 1);
                mbar_init(&p_full[t], 128);
                mbar_init(&o_full[t], 1);
                mbar_init(&s_free[t], 128);
            }
            for (int s = 0; s < KST; ++s) {
                mbar_init(&k_full[s], 1);
                mbar_init(&k_empty[s], NQT);      // one tcgen05.commit per issuer warp
                mbar_init(&v_full[s], 1);
                mbar_init(&v_empty[s], NQT);
            }
            fence_mbar_init();
        }
        __syncwarp();
        tmem_alloc(tmem_slot, p.tmem_cols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    CGS_TRACE_CTA(1);

    // Register hand-over between warpgroups (as in every warp-specialised Blackwell kernel): the TMA / MMA
    // warpgroup needs a few dozen registers, the softmax threads hold a whole 128-column score row plus the packed
    // probabilities. The launch gives every thread the same share (168 with 384 threads, 128 with 2 x 256).

    if (warp >= PRODUCER_WARP) {
#if CGS_ATTN_SETMAXNREG
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(attn::kRegsService));
#endif
    if (warp == PRODUCER_WARP) {
        // ------------------------------------------------------------------ TMA producer
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            uint32_t it = 0;                          // items done by this CTA
            for (int w = blockIdx.x; w < n_items; w += gridDim.x, ++it) {
            int q0, head, batch;
            item_coords(w, q0, head, batch);
            const uint32_t qslot = it % QBUF, quse = it / QBUF;      // Q buffer of this item and how often it was used
            for (int t = 0; t < NQT; ++t) {
                // the S MMAs of the item that used this buffer last have read Q
                if (quse > 0) mbar_wait_parked<CGS_ATTN_WAITMODE>(&q_empty[qslot * NQT + t], (quse - 1) & 1);
                mbar_expect_tx(&q_full[qslot * NQT + t], NKB * TILE_BYTES);
                for (int kb = 0; kb < NKB; ++kb)
                    tma_load_4d(smem + L::Q_OFF + ((qslot * NQT + t) * NKB + kb) * TILE_BYTES, &tmQ,
                                &q_full[qslot * NQT + t], kb * 64, head, q0 + t * BLOCK_M, batch);
            }
            for (int j = 0; j < n_tiles; ++j) {
                mbar_wait_parked<CGS_ATTN_WAITMODE>(&k_empty[stage], phase ^ 1);
                mbar_expect_tx(&k_full[stage], NKB * KV_TILE);
                for (int kb = 0; kb < NKB; ++kb)
                    tma_load_4d(smem + L::K_OFF + (stage * NKB + kb) * KV_TILE, &tmK,
                                &k_full[stage], kb * 64, head, j * BLOCK_N, batch);
                mbar_wait_parked<CGS_ATTN_WAITMODE>(&v_empty[stage], phase ^ 1);
                mbar_expect_tx(&v_full[stage], NKB * KV_TILE);
                for (int kb = 0; kb < NKB; ++kb)
                    tma_load_4d(smem + L::V_OFF + (stage * NKB + kb) * KV_TILE, &tmV,
                                &v_full[stage], kb * 64, head, j * BLOCK_N, batch);
                if (++stage == KST) { stage = 0; phase ^= 1; }
            }
            }
        }
        __syncwarp();
    } else if (warp < MMA_WARP + NQT) {     // (the remaining warps only fill the service warpgroup up to four)
        // ------------------------------------------------------------------ MMA issuer for query tile t
        if (elect_one()) {
            const int t = warp - MMA_WARP;
            const bool trace_me = blockIdx.x == 0;
            const uint32_t idesc_o = umma_idesc_bf16(BLOCK_M, dv_n, 1);
            // descriptor halves: K-major operands (Q, K): LBO 16 B, SBO 1024 B; MN-major V: LBO = one TMA box
            const uint64_t dk = umma_desc_sw128(0, 16, 1024);
            const uint64_t dvv = umma_desc_sw128(0, KV_TILE, 1024);
            const uint32_t hi_k = static_cast<uint32_t>(dk >> 32), hi_v = static_cast<uint32_t>(dvv >> 32);
            // low word = start address >> 4 in bits [0,14) and the leading-dimension byte offset >> 4 in bits [16,30)
            const uint32_t q_lo0 = static_cast<uint32_t>(dk) | ((smem_u32(smem + L::Q_OFF + t * NKB * TILE_BYTES) & 0x3FFFF) >> 4);
            uint32_t q_lo = q_lo0;
            const uint32_t k_lo0 = static_cast<uint32_t>(dk) | ((smem_u32(smem + L::K_OFF) & 0x3FFFF) >> 4);
            const uint32_t v_lo0 = static_cast<uint32_t>(dvv) | ((smem_u32(smem + L::V_OFF) & 0x3FFFF) >> 4);
            constexpr uint32_t STAGE_LO = (NKB * KV_TILE) >> 4;
            const uint32_t s_tmem = tmem_base + S_COL + t * BLOCK_N;
            const uint32_t p_tmem = tmem_base + P_COL + t * (BLOCK_N / 2);
            const uint32_t o_tmem = tmem_base + O_COL + t * o_stride;
            auto n_mma_of = [&](int j) {
                const int valid = min(BLOCK_N, p.Nk - j * BLOCK_N);
                return (valid + 15) & ~15;
            };
            auto issue_s = [&](int stage, int n_mma) {
                const uint32_t idesc_s = umma_idesc_bf16(BLOCK_M, n_mma, 0);
                const uint32_t k_lo = k_lo0 + stage * STAGE_LO;
#pragma unroll
                for (int s = 0; s < NKB * 4; ++s) {
                    if (s < ksteps_d) {
                        // 64-wide block kb = s / 4 (one TMA box further on), 32 bytes per k-step inside it
                        const uint32_t off_q = (s >> 2) * (TILE_BYTES >> 4) + (s & 3) * 2;
                        const uint32_t off_k = (s >> 2) * (KV_TILE >> 4) + (s & 3) * 2;
                        umma_ss_lh(s_tmem, q_lo + off_q, hi_k, k_lo + off_k, hi_k, idesc_s, s != 0);
                    }
                }
            };
            auto issue_pv = [&](int stage, int n_mma, uint32_t accumulate) {
                const uint32_t v_lo = v_lo0 + stage * STAGE_LO;
                const int ksteps = n_mma >> 4;
#pragma unroll
                for (int ks = 0; ks < BLOCK_N / 16; ++ks) {
                    if (ks < ksteps) {
                        // 16 key rows per step = 2 swizzle groups of 8 rows = 2048 bytes
                        umma_ts_lh(o_tmem, p_tmem + ks * 8, v_lo + ks * (2048 >> 4), hi_v, idesc_o,
                                   (accumulate | ks) != 0 ? 1u : 0u);
                    }
                }
            };

            int stage = 0;
            uint32_t phase = 0;
            uint32_t g = 0;                           // key tiles done by this issuer (all items): barrier phases
            uint32_t it = 0;
            for (int w = blockIdx.x; w < n_items; w += gridDim.x, ++it) {
            // S of the item's first key tile: needs Q, K and (after the first item) the S buffer handed back
            const uint32_t qslot = it % QBUF, quse = it / QBUF;
            q_lo = q_lo0 + qslot * ((NQT * NKB * TILE_BYTES) >> 4);
            mbar_wait_parked<CGS_ATTN_WAITMODE>(&q_full[qslot * NQT + t], quse & 1);
            mbar_wait_parked<CGS_ATTN_WAITMODE>(&k_full[stage], phase);
            if (g > 0) mbar_wait_parked<CGS_ATTN_WAITMODE>(&s_free[t], (g - 1) & 1);
            tc_fence_after();
            issue_s(stage, n_mma_of(0));
            tc_commit(&s_full[t]);
            tc_commit(&k_empty[stage]);
            for (int j = 0; j < n_tiles; ++j, ++g) {
                int nstage = stage + 1;
                uint32_t nphase = phase;
                if (nstage == KST) { nstage = 0; nphase ^= 1; }
                if (j + 1 < n_tiles) {
                    // S(j+1): needs K(j+1) and the S buffer handed back (S(j) copied into registers)
                    mbar_wait_parked<CGS_ATTN_WAITMODE>(&k_full[nstage], nphase);
                    CGS_TRACE(2 + t, g, 0);
                    mbar_wait_parked<CGS_ATTN_WAITMODE>(&s_free[t], g & 1);
                    tc_fence_after();
                    issue_s(nstage, n_mma_of(j + 1));
                    tc_commit(&s_full[t]);
                    tc_commit(&k_empty[nstage]);
                    CGS_TRACE(2 + t, g, 1);
                } else {
                    tc_commit(&q_empty[(it % QBUF) * NQT + t]);   // every S MMA of this item has been issued: Q may be replaced
                }
                mbar_wait_parked<CGS_ATTN_WAITMODE>(&v_full[stage], phase);
                CGS_TRACE(2 + t, g, 2);
                mbar_wait_parked<CGS_ATTN_WAITMODE>(&p_full[t], g & 1);
                CGS_TRACE(2 + t, g, 3);
                tc_fence_after();
                issue_pv(stage, n_mma_of(j), j > 0 ? 1u : 0u);
                tc_commit(&o_full[t]);
                tc_commit(&v_empty[stage]);
                CGS_TRACE(2 + t, g, 4);
                stage = nstage;
                phase = nphase;
            }
            }
        }
        __syncwarp();
    }
    } else {
        // ------------------------------------------------------------------ softmax warps
        // thread == query row. ncu with two softmax warps per scheduler (NQT = 2, 128 keys): issue active 51 %, XU
        // 54 %, FMA 36 %, ALU 31 %, tensor 35 %: nothing saturated, the TMEM / barrier latencies of a tile are only
        // covered by the one other warp. (Half a row per thread -- 4 warps per scheduler but still two phases --
        // measured 4-5 % slower.)
#if CGS_ATTN_SETMAXNREG
        asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(attn::kRegsSoftmax<NQT>));
#endif
        constexpr int COLS = BLOCK_N;
        const int t = warp >> 2;
        const int quad = warp & 3;
        const int row_in_tile = quad * 32 + lane;
        const uint32_t lane_field = static_cast<uint32_t>(quad * 32) << 16;
        const uint32_t s_addr = tmem_base + lane_field + S_COL + t * BLOCK_N;
        const uint32_t p_addr = tmem_base + lane_field + P_COL + t * (BLOCK_N / 2);
        const uint32_t o_addr = tmem_base + lane_field + O_COL + t * o_stride;

        const bool trace_me = blockIdx.x == 0 && quad == 0 && lane == 0;
        const float scale = p.scale_log2;
        const uint64_t scale2 = pack_f32x2(scale, scale);
        uint32_t g = 0;           // key tiles done by this thread (all items): barrier phases

        for (int w = blockIdx.x; w < n_items; w += gridDim.x) {
        int q0, head, batch;
        item_coords(w, q0, head, batch);
        float m_run = -INFINITY;  // running max, already multiplied by scale_log2
        float m_raw = -INFINITY;  // the same before scaling (used by the masked variant only)
        float l_run = 0.f;        // row sum
        // One key tile. The last tile of a key sequence that is not a multiple of the tile width is a second instance
        // of this body with the masking code in it: the loop over full tiles carries none of it.
        auto key_tile = [&](auto ragged_tag, int j) {
            constexpr bool RAGGED = decltype(ragged_tag)::value;
            CGS_TRACE(t, g, 0);
            mbar_wait(&s_full[t], g & 1);
            CGS_TRACE(t, g, 1);
            tc_fence_after();
            uint32_t sr[COLS];
#pragma unroll
            for (int c = 0; c + 32 <= COLS; c += 32) tmem_ld_x32(s_addr + c, sr + c);
            if constexpr (COLS % 32 != 0) tmem_ld_x16(s_addr + (COLS / 32) * 32, sr + (COLS / 32) * 32);
            tmem_wait_ld();
            // S(j) is in registers: give the buffer back so S(j+1) is computed during this softmax
            tc_fence_before();
            mbar_arrive(&s_free[t]);
            CGS_TRACE(t, g, 2);

            const int n_valid = p.Nk - j * BLOCK_N;
            if constexpr (MASK) {
                // optimized_attention_masked (attention.py:111-123 / SDPA attn_mask): scores * scale + mask. Not on
                // the UNet's path (its mask is always None); one scalar load per score, no tuning.
                const int q_row_m = min(q0 + t * BLOCK_M + row_in_tile, p.Nq - 1);
                const float* mrow = p.mask + (long long)batch * p.mask_sb + (long long)head * p.mask_sh +
                                    (long long)q_row_m * p.mask_sq + (long long)j * BLOCK_N;
#pragma unroll 8
                for (int i = 0; i < COLS; ++i)
                    if (i < n_valid)
                        // (-inf entries are taken as -1e30, like boolean masks: a key tile whose entries are all
                        // -inf would otherwise make the running max -inf and alpha = 2^(-inf + inf) a NaN)
                        sr[i] = __float_as_uint(fmaf(fmaxf(__ldg(mrow + i), -1.0e30f), p.inv_scale, __uint_as_float(sr[i])));
            }
            if constexpr (RAGGED) {
#pragma unroll
                for (int i = 0; i < COLS; ++i)
                    if (i >= n_valid) sr[i] = 0xff800000u;  // -inf
            }
            float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
            for (int i = 0; i < ((CGS_ATTN_EXPT == 2 || CGS_ATTN_EXPT == 3) ? 8 : COLS); i += 8) {
                mx0 = fmax3(mx0, __uint_as_float(sr[i + 0]), __uint_as_float(sr[i + 1]));
                mx1 = fmax3(mx1, __uint_as_float(sr[i + 2]), __uint_as_float(sr[i + 3]));
                mx2 = fmax3(mx2, __uint_as_float(sr[i + 4]), __uint_as_float(sr[i + 5]));
                mx3 = fmax3(mx3, __uint_as_float(sr[i + 6]), __uint_as_float(sr[i + 7]));
            }
            float mx = fmaxf(fmax3(mx0, mx1, mx2), mx3);
            const float m_new = mx * scale;

            // lazy running max, branch-free (a per-thread branch here put a reconvergence point between the max and
            // the exponentials): the reference point moves only when the tile's max exceeds it by > 2^8, or on the
            // item's first tile (m_run = -inf, alpha = 2^-inf = 0, l_run = 0)
            const bool grow = (j == 0) || (m_new > m_run + kRescaleThreshold);
            const float m_next = grow ? m_new : m_run;
            const float alpha = ex2_approx(m_run - m_next);         // 1 when the reference point stays
            const bool need_rescale = grow && j > 0;
            l_run *= alpha;
            m_run = m_next;
            m_raw = grow ? mx : m_raw;
            const uint64_t negm2 = pack_f32x2(-m_run, -m_run);
            // masked variant: x = (s - m) * scale, subtraction first. With s * scale - m * scale in one FMA the
            // rounding of m * scale (half an ulp of m) is left in x; harmless for |m| ~ 10, an overflow for a row
            // whose keys are all masked (m ~ -1e30), which must come out as the uniform average of attention.py:113-117
            const uint64_t negmr2 = pack_f32x2(-m_raw, -m_raw);
            uint32_t pr[COLS / 2];
            uint64_t acc0 = 0ull, acc1 = 0ull;   // two packed (lo, hi) partial row sums
            CGS_TRACE(t, g, 3);
            // x = s * scale - m (packed FFMA2); 2^x: the MUFU gives 16 lanes/clk/SM, so kPolyPairs of every
            // kPolyPeriod pairs go to a degree-3 polynomial on the FMA pipe; row sum (packed FADD2); bf16 pack
#pragma unroll
            for (int i = 0; i < COLS / 2; ++i) {
                const uint64_t s2 = pack_u32x2(sr[2 * i], sr[2 * i + 1]);
                const uint64_t x2 = MASK ? ffma2(fadd2(s2, negmr2), scale2, 0ull) : ffma2(s2, scale2, negm2);
                uint64_t e2;
                if (CGS_ATTN_EXPT == 1 || CGS_ATTN_EXPT == 2) {
                    e2 = x2;            // timing experiment: no exponentials
                } else if ((i % kPolyPeriod) < kPolyPairs) {
                    e2 = ex2_poly2(x2);
                } else {
                    float lo, hi;
                    unpack_f32x2(x2, lo, hi);
                    e2 = pack_f32x2(ex2_approx(lo), ex2_approx(hi));
                }
                if (i & 1) acc1 = fadd2(acc1, e2); else acc0 = fadd2(acc0, e2);
                float elo, ehi;
                unpack_f32x2(e2, elo, ehi);
                pr[i] = CGS_ATTN_EXPT == 2 ? sr[2 * i] : pack_bf16x2(elo, ehi);
            }
            CGS_TRACE(t, g, 4);
            {
                float s0, s1, s2, s3;
                unpack_f32x2(acc0, s0, s1);
                unpack_f32x2(acc1, s2, s3);
                l_run += (s0 + s1) + (s2 + s3);
            }
            // PV of the previous key tile must have finished reading P (and O must be quiescent for a rescale). It was
            // issued a whole softmax ago, so this wait is normally already satisfied. (For g == 0 it asks for the phase
            // before the first one, which an initialised barrier reports as complete: no special case.)
            mbar_wait(&o_full[t], (g - 1) & 1);
            tc_fence_after();
#pragma unroll
            for (int c = 0; c + 32 <= COLS / 2; c += 32) tmem_st_x32(p_addr + c, pr + c);
            if constexpr ((COLS / 2) % 32 != 0) tmem_st_x16(p_addr + (COLS / 64) * 32, pr + (COLS / 64) * 32);
            if (__any_sync(0xffffffffu, need_rescale)) {
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
            CGS_TRACE(t, g, 5);
        };
        const int n_full = p.Nk / BLOCK_N;
        for (int j = 0; j < n_full; ++j, ++g) key_tile(std::false_type{}, j);
        if (n_full < n_tiles) {
            key_tile(std::true_type{}, n_full);
            ++g;
        }

        // epilogue: O / l -> bf16 -> global. (The next item's first PV, which overwrites O, is issued only after
        // every thread of this warpgroup has arrived on p_full for it, i.e. after this epilogue.)
        mbar_wait(&o_full[t], (g - 1) & 1);
        tc_fence_after();
        CGS_TRACE_CTA(2);
        const float inv_l = 1.f / l_run;
        if constexpr (NQT == 1) {
        // registers -> 128-byte-swizzled staging box -> one TMA store per 64-column block (row-strided 16-byte global
        // stores from 128 threads cost a third of a cross-attention launch; the store clips ragged Nq and d < 64)
        uint8_t* stage_row = smem + L::O_OFF + t * TILE_BYTES + row_in_tile * 128;
        const int sw = row_in_tile & 7;
        const bool store_leader = (quad == 0 && lane == 0);
        for (int kb = 0; kb < NKB; ++kb) {
            if (kb * 64 >= dv_n) break;
            uint4 w[8];
            uint32_t oall[64];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                if (kb * 64 + c * 16 < dv_n) {
                    tmem_ld_x16(o_addr + kb * 64 + c * 16, oall + c * 16);
                } else {
#pragma unroll
                    for (int i = 0; i < 16; ++i) oall[c * 16 + i] = 0u;
                }
            }
            tmem_wait_ld();
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const uint32_t* orr = oall + c * 16;
                w[2 * c].x = pack_bf16x2(__uint_as_float(orr[0]) * inv_l, __uint_as_float(orr[1]) * inv_l);
                w[2 * c].y = pack_bf16x2(__uint_as_float(orr[2]) * inv_l, __uint_as_float(orr[3]) * inv_l);
                w[2 * c].z = pack_bf16x2(__uint_as_float(orr[4]) * inv_l, __uint_as_float(orr[5]) * inv_l);
                w[2 * c].w = pack_bf16x2(__uint_as_float(orr[6]) * inv_l, __uint_as_float(orr[7]) * inv_l);
                w[2 * c + 1].x = pack_bf16x2(__uint_as_float(orr[8]) * inv_l, __uint_as_float(orr[9]) * inv_l);
                w[2 * c + 1].y = pack_bf16x2(__uint_as_float(orr[10]) * inv_l, __uint_as_float(orr[11]) * inv_l);
                w[2 * c + 1].z = pack_bf16x2(__uint_as_float(orr[12]) * inv_l, __uint_as_float(orr[13]) * inv_l);
                w[2 * c + 1].w = pack_bf16x2(__uint_as_float(orr[14]) * inv_l, __uint_as_float(orr[15]) * inv_l);
            }
            if (store_leader) tma_store_wait_read();          // the previous store has finished reading the box
            named_bar_sync(1, 128);
#pragma unroll
            for (int c = 0; c < 8; ++c)
                *reinterpret_cast<uint4*>(stage_row + ((c ^ sw) << 4)) = w[c];
            fence_proxy_async_smem();
            named_bar_sync(1, 128);
            if (store_leader && CGS_ATTN_EXPT != 4) {
                tma_store_4d(&tmO, smem + L::O_OFF + t * TILE_BYTES, kb * 64, head, q0 + t * BLOCK_M, batch);
                tma_store_commit();
            }
        }
        } else {
        // long key sequences: the epilogue is a per-mille of the item; plain 16-byte stores, no staging traffic
        const int q_row = q0 + t * BLOCK_M + row_in_tile;
        __nv_bfloat16* orow = p.out + (long long)batch * p.o_stride_b +
                              (long long)q_row * p.o_stride_n + (long long)head * p.o_stride_h;
        for (int c = 0; c < (dv_n >> 4); ++c) {
            uint32_t orr[16];
            tmem_ld_x16(o_addr + c * 16, orr);
            tmem_wait_ld();
            if (q_row < p.Nq && CGS_ATTN_EXPT != 4) {
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
        }
        tc_fence_before();
        CGS_TRACE_CTA(3);
        }
        if (NQT == 1 && quad == 0 && lane == 0) tma_store_wait_all();     // shared memory must outlive the last store's reads
    }

    __syncthreads();
    CGS_TRACE_CTA(4);
    if (warp == MMA_WARP) tmem_dealloc(tmem_base, p.tmem_cols);
}

template <int NKB, int NQT, int KST, int BN = 128, bool MASK = false>
static int launch_attn(const CUtensorMap& tq, const CUtensorMap& tk, const CUtensorMap& tv, const CUtensorMap& to,
                       const AttnParams& p, cudaStream_t stream) {
    using L = AttnSmem<NKB, NQT, KST, BN>;
    auto kern = attn_fwd_kernel<NKB, NQT, KST, BN, MASK>;
    CGS_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::BYTES));
    const long long items = (long long)((p.Nq + attn::BLOCK_M * NQT - 1) / (attn::BLOCK_M * NQT)) * p.H * p.B;
    // one-tile variants: persistent CTAs (two per SM) walk the items; multi-tile variants: one CTA per item, handed
    // out by the hardware scheduler (an item is tens of key tiles long, and static round-robin measured 3 % slower)
    const long long resident = (long long)sm_count() * 2;
    const unsigned grid = (unsigned)((NQT > 1 || items < resident) ? items : resident);
    kern<<<grid, attn::kThreads<NQT>, L::BYTES, stream>>>(tq, tk, tv, to, p);
    CGS_CUDA_CHECK(cudaGetLastError());
    return kOk;
}

static int make_qkv_map(CUtensorMap* m, const void* ptr, int B, int N, int H, int d,
                        long long sb, long long sn, long long sh, int box_rows = 128) {
    const uint64_t dims[4] = {(uint64_t)d, (uint64_t)H, (uint64_t)N, (uint64_t)B};
    const uint64_t strides[3] = {(uint64_t)sh * 2, (uint64_t)sn * 2, (uint64_t)sb * 2};
    const uint32_t box[4] = {64, 1, (uint32_t)box_rows, 1};
    return make_tmap_bf16(m, ptr, 4, dims, strides, box);
}

static int attention_entry(const void* q, const void* k, const void* v, void* out, int B,
                           int H, int Nq, int Nk, int d, long long q_stride_b,
                           long long q_stride_n, long long q_stride_h, long long k_stride_b,
                           long long k_stride_n, long long k_stride_h, long long v_stride_b,
                           long long v_stride_n, long long v_stride_h, long long o_stride_b,
                           long long o_stride_n, long long o_stride_h, float softmax_scale,
                           const float* mask, long long mask_sb, long long mask_sh, long long mask_sq,
                           void* stream) {
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
    if ((long long)((Nq + 127) / 128) * H * B > 0x7fffffffLL) {
        set_last_error("attention: too many (batch, head, query tile) work items");
        return kUnsupportedShape;
    }
    if ((o_stride_b | o_stride_n | o_stride_h) % 8 != 0 || (reinterpret_cast<uintptr_t>(out) & 0xF)) {
        set_last_error("attention: output strides/pointer must be 16-byte aligned");
        return kBadArgument;
    }
    AttnParams p;
    p.dbg = nullptr;
    // developer knobs, read once per process (tools/attn_trace*.py, tools/kernel_bench.py set them before the first call)
    static const struct Knobs {
        long long* trace;
        int max_1cta;
        bool no_3tile, two_by_64, no_bn96;
        Knobs() {
            const char* e = getenv("CGS_ATTN_TRACE");
            trace = e ? reinterpret_cast<long long*>(strtoull(e, nullptr, 0)) : nullptr;
            e = getenv("CGS_ATTN_MAX_TILES_1CTA");
            max_1cta = e ? atoi(e) : 8;
            no_3tile = getenv("CGS_ATTN_NO_3TILE") != nullptr;
            two_by_64 = getenv("CGS_ATTN_2X64") != nullptr;
            no_bn96 = getenv("CGS_ATTN_NO_BN96") != nullptr;
        }
    } knobs;
    p.dbg = knobs.trace;
    p.out = static_cast<__nv_bfloat16*>(out);
    p.o_stride_b = o_stride_b;
    p.o_stride_n = o_stride_n;
    p.o_stride_h = o_stride_h;
    p.B = B; p.H = H; p.Nq = Nq; p.Nk = Nk; p.d = d;
    p.scale_log2 = softmax_scale * 1.4426950408889634f;
    p.mask = mask;
    p.mask_sb = mask_sb; p.mask_sh = mask_sh; p.mask_sq = mask_sq;
    p.inv_scale = 1.0f / softmax_scale;
    if (mask != nullptr && !(softmax_scale > 0.f)) {
        set_last_error("attention: a mask needs a positive softmax scale");
        return kBadArgument;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int dv_n = ((d + 15) / 16) * 16;
    const int n_tiles = (Nk + 127) / 128;
    // Long key sequences at d <= 64: three query tiles per CTA against 64-key tiles (TMEM 3 x (64 S + 32 P + 64 O)
    // = 480 columns), so every scheduler has three softmax warps in three different phases to cover the TMEM and
    // barrier latencies of one another. d <= 128: two query tiles against 128-key tiles. Short key sequences (text
    // cross-attention, up to 8 key tiles): one tile per CTA, 256 TMEM columns, two CTAs per SM, so the set-up and
    // epilogue of one CTA overlap the loop of the other.
    const int max_1cta = knobs.max_1cta;
    const bool multi = Nq > attn::BLOCK_M && n_tiles > max_1cta && mask == nullptr;   // masked: one-tile variants only
    int nqt = 1, bn = 128;
    if (multi && d <= 64 && !knobs.no_3tile) { nqt = 3; bn = 64; }
    else if (multi && d <= 64 && knobs.two_by_64) { nqt = 2; bn = 64; }
    else if (multi && 2 * (192 + dv_n) <= 512) { nqt = 2; }
    // text cross-attention (77 keys, 154 when the prompts' lengths differ): 96-key tiles -- the softmax runs over
    // every column of a tile, so a 128-column tile spends 40 % of its exponentials on padding
    const bool short_keys = !multi && mask == nullptr && d <= 128 && Nk <= 192 && !knobs.no_bn96;
    if (short_keys) bn = 96;
    int cols = 32;
    while (cols < nqt * (bn + bn / 2 + dv_n)) cols *= 2;
    p.tmem_cols = cols;
    CUtensorMap tq, tk, tv, to;
    int rc;
    if ((rc = make_qkv_map(&to, out, B, Nq, H, d, o_stride_b, o_stride_n, o_stride_h)) != kOk) return rc;
    if ((rc = make_qkv_map(&tq, q, B, Nq, H, d, q_stride_b, q_stride_n, q_stride_h)) != kOk) return rc;
    if ((rc = make_qkv_map(&tk, k, B, Nk, H, d, k_stride_b, k_stride_n, k_stride_h, bn)) != kOk) return rc;
    if ((rc = make_qkv_map(&tv, v, B, Nk, H, d, v_stride_b, v_stride_n, v_stride_h, bn)) != kOk) return rc;
    if (mask != nullptr) {
        if (d <= 64) return n_tiles > 2 ? launch_attn<1, 1, 2, 128, true>(tq, tk, tv, to, p, st)
                                        : launch_attn<1, 1, 1, 128, true>(tq, tk, tv, to, p, st);
        if (d <= 128) return n_tiles > 2 ? launch_attn<2, 1, 2, 128, true>(tq, tk, tv, to, p, st)
                                         : launch_attn<2, 1, 1, 128, true>(tq, tk, tv, to, p, st);
        return launch_attn<3, 1, 1, 128, true>(tq, tk, tv, to, p, st);
    }
    if (d <= 64) {
        if (nqt == 3) return launch_attn<1, 3, 4, 64>(tq, tk, tv, to, p, st);
        if (nqt == 2 && bn == 64) return launch_attn<1, 2, 4, 64>(tq, tk, tv, to, p, st);
        if (nqt == 2) return launch_attn<1, 2, 3>(tq, tk, tv, to, p, st);
        if (short_keys) return launch_attn<1, 1, 1, 96>(tq, tk, tv, to, p, st);
        return n_tiles > 2 ? launch_attn<1, 1, 2>(tq, tk, tv, to, p, st) : launch_attn<1, 1, 1>(tq, tk, tv, to, p, st);
    }
    if (d <= 128) {
        if (short_keys) return launch_attn<2, 1, 1, 96>(tq, tk, tv, to, p, st);
        if (nqt == 2) return launch_attn<2, 2, 2>(tq, tk, tv, to, p, st);
        return n_tiles > 2 ? launch_attn<2, 1, 2>(tq, tk, tv, to, p, st) : launch_attn<2, 1, 1>(tq, tk, tv, to, p, st);
    }
    return launch_attn<3, 1, 1>(tq, tk, tv, to, p, st);
}

}  // namespace cgs

extern "C" int cgs_attention_fwd(const void* q, const void* k, const void* v, void* out, int B,
                                 int H, int Nq, int Nk, int d, long long q_stride_b,
                                 long long q_stride_n, long long q_stride_h, long long k_stride_b,
                                 long long k_stride_n, long long k_stride_h, long long v_stride_b,
                                 long long v_stride_n, long long v_stride_h, long long o_stride_b,
                                 long long o_stride_n, long long o_stride_h, float softmax_scale,
                                 void* stream) {
    return cgs::attention_entry(q, k, v, out, B, H, Nq, Nk, d, q_stride_b, q_stride_n, q_stride_h, k_stride_b,
                                k_stride_n, k_stride_h, v_stride_b, v_stride_n, v_stride_h, o_stride_b, o_stride_n,
                                o_stride_h, softmax_scale, nullptr, 0, 0, 0, stream);
}

extern "C" int cgs_attention_masked_fwd(const void* q, const void* k, const void* v, void* out, int B,
                                        int H, int Nq, int Nk, int d, long long q_stride_b,
                                        long long q_stride_n, long long q_stride_h, long long k_stride_b,
                                        long long k_stride_n, long long k_stride_h, long long v_stride_b,
                                        long long v_stride_n, long long v_stride_h, long long o_stride_b,
                                        long long o_stride_n, long long o_stride_h, float softmax_scale,
                                        const float* mask, long long mask_stride_b, long long mask_stride_h,
                                        long long mask_stride_q, void* stream) {
    if (mask == nullptr) {
        cgs::set_last_error("attention_masked: mask must not be null (use cgs_attention_fwd)");
        return cgs::kBadArgument;
    }
    return cgs::attention_entry(q, k, v, out, B, H, Nq, Nk, d, q_stride_b, q_stride_n, q_stride_h, k_stride_b,
                                k_stride_n, k_stride_h, v_stride_b, v_stride_n, v_stride_h, o_stride_b, o_stride_n,
                                o_stride_h, softmax_scale, mask, mask_stride_b, mask_stride_h, mask_stride_q, stream);
}
