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
// Structure (persistent, 576 threads per CTA):
//   warp 0       TMA producer
//   warp 1       TMEM allocator + single-thread tcgen05.mma issuer
//   warps 2..17  epilogue (4 warps per TMEM lane quadrant, each owning a 32-column slice of a 128-column pass):
//                tcgen05.ld accumulator -> bias / GELU / gate / residual -> bf16 -> swizzled SMEM staging -> TMA store.
// Two instantiations of the mainloop:
//   PAIR = false  one CTA per SM, tile M128 x N256: A [128 x 64] + B [256 x 64] per k-block, 4-stage ring.
//   PAIR = true   CTA pairs (cluster of 2, cta_group::2), tile M256 x N256 per pair: each CTA stages its own
//                 A [128 x 64] and HALF of B [128 x 64] (32 KiB per k-block instead of 48, 6-stage ring) and owns the
//                 accumulators of its 128 rows. The first round's mainloop was bound by the L2 -> SMEM operand feed
//                 (narrower tiles lost exactly in proportion to their arithmetic intensity, profiles/r01_*): halving
//                 the B traffic per CTA is the fix. The leader CTA issues the MMAs; its commits arrive on the
//                 barriers of both CTAs; the peer's TMA credits its bytes to the leader's `full` barrier and its
//                 epilogue warps arrive on the leader's `tempty` barrier through the cluster's shared window.
//                 Used when the problem has enough M256 x N256 tiles to fill the 74 pairs.
// 16 epilogue warps, not 4: the first version (4 warps, erff) ran one warp per scheduler at IPC 0.24 and
// needed 23.8 k cycles per tile against 5.1 k cycles of MMA (ncu, profiles/r01_*): latency-bound. Tiles are walked
// in groups of 8 M-blocks x all N-blocks so the A panel of a group stays in L2 while its N-blocks are produced.
// Accumulators are double-buffered in TMEM (2 x 256 fp32 columns = all 512 columns), so the
// epilogue of tile i overlaps the MMAs of tile i+1.
//
// For GEGLU the B tile's rows [0,128) come from W rows [n0, n0+128) (value half) and rows
// [128,256) from W rows [N+n0, N+n0+128) (gate half): one N=256 MMA produces both halves of
// the same 128 output columns side by side in TMEM. (With CTA pairs the leader stages the value rows and the peer
// the gate rows.)
//
// Wave quantisation (EPI_LINEAR): the SDXL linears (M = 16384 / 65536, N = 1280 / 640) give 4.3 rounds of tiles, and
// one image per GPU gives 80 tiles for 148 SMs. The tiles of the last, partly filled round are therefore cut along N
// into halves or quarters (a narrower MMA on the same A rows: an independent set of output columns -- no reduction,
// no workspace, and every output element is still the same sum in the same order, so results do not change), but
// only as far as every worker still gets at most ONE piece: a piece is not proportionally cheaper than a tile (same
// A traffic, same k-loop, lower arithmetic intensity), so a worker with two of them would be the slowest of the
// round. On CTA pairs a narrow piece stages just the B rows it needs (32-row boxes). Measured against the first
// scheme of this round (64-column groups dealt out evenly, some workers with two): +6 % on to_out, +11 % on
// ff.net.2, +28 % at M = 2048. (Splitting those tiles along K instead -- partial accumulators through global
// memory -- was built and measured first: 19 MB of fp32 partials per launch cost as much as the idle SMs.)
// The same mechanism serves the last column of tiles when N is not a multiple of 256 (N = 640: an N = 128 MMA).
//
// CONV (EPI_LINEAR, CTA pairs): the same GEMM as a 3x3 / stride 1 / pad 1 convolution over a channels-last image
// (reference: ResBlock / Upsample convolutions, comfy/ldm/modules/diffusionmodules/openaimodel.py:83-96, 177-232;
// operator comfy/ops.py:67-80), implicit -- no im2col matrix exists anywhere. A tile's 128 rows are a TH x TW
// rectangle of pixels of one image; k-block (tap r,s ; 64 channels) of its A operand is ONE 4-D TMA box
// {64 ch, TW, TH, 1} starting at pixel (h0 + r - 1, w0 + s - 1): the tensor map's out-of-bounds zero fill IS the
// padding, and the box lands in shared memory in exactly the K-major, 128-byte-swizzled [128 x 64] layout the MMA
// reads. B is the weight as stored in channels-last memory, [Cout][3][3][Cin] = [Cout, 9 Cin] K-major. Output and
// residual tiles use 4-D maps of the same box, so bias + residual (ResBlock: `skip(x) + h`) ride in the epilogue.
//
// WIDE (N a multiple of 320: the 320-channel level of every SD / SDXL UNet): 256 + 64 columns as two tiles would load
// A twice and run the 64-column tile at a quarter of the arithmetic intensity (bound by the operand feed: 77 %).
// Instead one tile is 320 columns: two MMAs per k-step (N = 256 and N = 64) on the same A tile, accumulator
// 320 TMEM columns, single-buffered (640 would not fit): the epilogue's TMEM reads are exposed (~5 % of a K = 2880
// mainloop), the A traffic halves.
//
// Tensor-core bound. Algorithmic FLOPs = 2 * M * K * (2N) for GEGLU, 2 * M * K * N for linear, 2 * pixels * 9 Cin * Cout
// for the convolution.
#include "common.cuh"

#ifndef CGS_GEMM_EXPT
#define CGS_GEMM_EXPT 0
#endif

namespace cgs {

// -DCGS_GEMM_TRACE: clock stamps of CTA 0's MMA thread and first epilogue warp, per work item (tools/gemm_trace.py)
#ifdef CGS_GEMM_TRACE
__device__ long long* g_gemm_trace = nullptr;
#define GEMM_STAMP(item, slot)                                                                        \
    do {                                                                                              \
        if (g_gemm_trace != nullptr && blockIdx.x == 0 && (item) < 32) g_gemm_trace[(item) * 16 + (slot)] = clock64(); \
    } while (0)
#else
#define GEMM_STAMP(item, slot) do { } while (0)
#endif

namespace gemm {
constexpr int BM = 128;                       // rows per CTA
constexpr int BK = 64;
constexpr int BN_HALF = 128;                  // rows per B TMA box
constexpr int UMMA_N = 256;
constexpr int A_BYTES = BM * BK * 2;          // 16 KiB
constexpr int OUT_BOX_BYTES = BM * 128;                  // [128 rows x 64 bf16], 128-byte swizzle
constexpr int OUT_STAGE_BYTES = 2 * OUT_BOX_BYTES;       // 128 output columns per pass
constexpr int EPI_WARPS = 16;
constexpr int THREADS = (2 + EPI_WARPS) * 32;
constexpr int GROUP_M = 8;
enum { EPI_LINEAR = 0, EPI_GEGLU = 1 };

constexpr int WIDE_N = 320;                   // WIDE tiles: 256 + 64 columns
constexpr int B2_ROWS = (WIDE_N - UMMA_N) / 2; // rows of the second B box per CTA
constexpr int B2_BYTES = B2_ROWS * BK * 2;    // 4 KiB

// a tile's rows as a rectangle of pixels (CONV)
struct ConvGeom {
    int H, W, TH, TW, tiles_h, tiles_w, n_img, cin, cblocks;
    int taps, taps_w;     // filter taps (9 / 3 for 3x3; 4 / 2 for one phase of the 2x-upsample convolution)
    int off_h, off_w;     // input pixel of tap (0, 0) relative to the output pixel (-1, -1 for 3x3 / pad 1)
    int whole_tiles;      // last, partly filled round of tiles: 0 = 64-column groups dealt out evenly, 1 = whole tiles,
                          // 2 = halves or quarters of a tile, at most one per worker (Schedule)
};

template <bool PAIR, int EPI, bool WIDE = false>
struct Cfg {
    static_assert(!WIDE || (PAIR && EPI == EPI_LINEAR), "WIDE tiles exist for linear epilogues on CTA pairs");
    // EPI_LINEAR stages the residual tile in shared memory too (32 KiB): one pipeline stage fewer
    // (experiment bit 16: no residual staging, one more stage -- residual must be NULL)
    static constexpr bool RES = (EPI == EPI_LINEAR) && !(CGS_GEMM_EXPT & 16);
    static constexpr int RES_STAGE_BYTES = RES ? OUT_STAGE_BYTES : 0;
    static constexpr int STAGES = (WIDE ? 4 : PAIR ? (EPI == EPI_LINEAR ? 5 : 6) : (EPI == EPI_LINEAR ? 3 : 4)) +
                                  ((EPI == EPI_LINEAR && !RES) ? 1 : 0);
    static constexpr int B_BYTES = (PAIR ? BN_HALF : UMMA_N) * BK * 2 + (WIDE ? B2_BYTES : 0);   // per CTA: 16 / 20 / 32 KiB
    static constexpr int TILE_N = WIDE ? WIDE_N : UMMA_N;                     // output columns per tile (EPI_LINEAR)
    static constexpr int GROUPS = TILE_N / 64;                                // 64-column groups per tile
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int TILE_M = PAIR ? 2 * BM : BM;                         // rows per tile (per pair / per CTA)
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + OUT_STAGE_BYTES + RES_STAGE_BYTES + 256 + 1024;  // + barriers + slack
};
}  // namespace gemm

// Exact-erf GELU, 0.5 x (1 + erf(x / sqrt 2)), with erf from Abramowitz & Stegun 7.1.26
// (|error| <= 1.5e-7, far below bf16 resolution), branch-free: 1 MUFU.RCP + 1 MUFU.EX2 + ~10 FMA-pipe ops.
__device__ __forceinline__ float gelu_erf(float x) {
    const float z = fabsf(x) * 0.70710678118654752440f;
    const float t = __fdividef(1.0f, fmaf(0.3275911f, z, 1.0f));   // MUFU.RCP, denominator in [1, inf)
    float p = fmaf(1.061405429f, t, -1.453152027f);
    p = fmaf(p, t, 1.421413741f);
    p = fmaf(p, t, -0.284496736f);
    p = fmaf(p, t, 0.254829592f);
    p *= t;
    const float e = ex2_approx(-1.4426950408889634f * z * z);
    const float erf_abs = fmaf(-p, e, 1.0f);                 // erf(|x| / sqrt 2)
    const float half_erf = copysignf(0.5f * erf_abs, x);
    return x * (0.5f + half_erf);
}

// tile index -> (m block, n block): groups of GROUP_M M-blocks, N-major inside a group
__device__ __forceinline__ void tile_coords(int t, int tiles_m, int tiles_n, int& mb, int& nb) {
    const int per_group = gemm::GROUP_M * tiles_n;
    const int g = t / per_group;
    const int r = t - g * per_group;
    const int gm = min(gemm::GROUP_M, tiles_m - g * gemm::GROUP_M);
    nb = r / gm;
    mb = g * gemm::GROUP_M + (r - nb * gm);
}

// One unit of work of a persistent worker: `width` output columns of one tile, starting `n_off` columns into it
// (a whole tile: n_off 0, width = what the tile has of N).
struct Piece {
    int tile;
    int n_off;
    int width;      // multiple of 32; 0 = nothing to do (all roles skip it)
};

// Full rounds: worker w takes tiles w, w + W, ... whole. Last, partly filled round (rem < W tiles), EPI_LINEAR only:
// its rem * G column groups (64 columns each; G = 4, or 5 for WIDE tiles) are dealt out contiguously and evenly to all workers, so a worker gets
// the end of one tile and / or the beginning of the next.
template <int EPI, int TILE_N = gemm::UMMA_N>
struct Schedule {
    static constexpr int G = TILE_N / 64;
    int workers, num_tiles, tiles_m, tiles_n, N, full_rounds, rem, active, groups;
    bool split;
    int gpt, gw;          // (mode 2) pieces per tile of the last round (1, 2 or 4) and their width
    // mode 2: a narrow piece is not proportionally cheaper than a tile (same A traffic, same k-loop), so a worker
    // with two of them is the slowest of the round. Halve or quarter the tiles only as far as every worker still gets
    // at most one piece; otherwise keep them whole.
    __device__ Schedule(int workers_, int tiles_m_, int tiles_n_, int N_, int mode = 0)
        : workers(workers_), num_tiles(tiles_m_ * tiles_n_), tiles_m(tiles_m_), tiles_n(tiles_n_), N(N_),
          split(mode != 1 && EPI == gemm::EPI_LINEAR), gpt(0), gw(TILE_N) {
        full_rounds = num_tiles / workers;
        rem = num_tiles - full_rounds * workers;
        groups = rem * G;
        active = split ? (groups < workers ? groups : workers) : rem;
        if (split && mode == 2) {
            gpt = (TILE_N % 256 == 0 && rem * 4 <= workers) ? 4 : (TILE_N % 128 == 0 && rem * 2 <= workers) ? 2 : 1;
            gw = TILE_N / gpt;
            active = rem * gpt;
        }
    }
    __device__ int tile_width(int tile) const {
        if (EPI == gemm::EPI_GEGLU) return gemm::UMMA_N;
        int mb, nb;
        tile_coords(tile, tiles_m, tiles_n, mb, nb);
        const int left = N - nb * TILE_N;
        return left < TILE_N ? left : TILE_N;
    }
    __device__ int items(int w) const {
        if (w >= active) return full_rounds;
        if (!split || gpt > 0) return full_rounds + 1;
        const int a = (int)((long long)groups * w / active), b = (int)((long long)groups * (w + 1) / active);
        return full_rounds + ((b - 1) / G - a / G) + 1;
    }
    __device__ Piece item(int w, int i) const {
        Piece p;
        p.n_off = 0;
        if (i < full_rounds) {
            p.tile = w + i * workers;
            p.width = tile_width(p.tile);
            return p;
        }
        if (!split) {
            p.tile = full_rounds * workers + w;
            p.width = tile_width(p.tile);
            return p;
        }
        if (gpt > 0) {
            p.tile = full_rounds * workers + w / gpt;
            p.n_off = (w % gpt) * gw;
            const int have = tile_width(p.tile) - p.n_off;
            p.width = have < gw ? have : gw;
            if (p.width < 0) p.width = 0;
            return p;
        }
        const int a = (int)((long long)groups * w / active), b = (int)((long long)groups * (w + 1) / active);
        const int slot = a / G + (i - full_rounds);
        const int g0 = (a > slot * G ? a : slot * G) - slot * G;
        const int g1 = (b < slot * G + G ? b : slot * G + G) - slot * G;
        p.tile = full_rounds * workers + slot;
        p.n_off = g0 * 64;
        const int have = tile_width(p.tile) - p.n_off;
        p.width = (g1 - g0) * 64 < have ? (g1 - g0) * 64 : have;
        if (p.width < 0) p.width = 0;
        return p;
    }
};

struct alignas(16) BF8 {
    __nv_bfloat162 v[4];
};

// 128-row tile index -> image, first pixel row / column of its TH x TW rectangle (n >= n_img: a tile past the end,
// its loads come back as zeros and its stores are clipped)
__device__ __forceinline__ void conv_tile(const gemm::ConvGeom& g, int mt, int& n, int& h0, int& w0) {
    const int per_img = g.tiles_h * g.tiles_w;
    n = mt / per_img;
    const int r = mt - n * per_img;
    const int ht = r / g.tiles_w;
    h0 = ht * g.TH;
    w0 = (r - ht * g.tiles_w) * g.TW;
}

template <int EPI, bool PAIR, bool CONV = false, bool WIDE = false>
__global__ void __launch_bounds__(gemm::THREADS, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ CUtensorMap tmB2,
            const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmR,
            const __nv_bfloat16* __restrict__ bias, const __nv_bfloat16* residual,
            __nv_bfloat16* out, int M, int N, int K, int ldo, int ldr, const gemm::ConvGeom geom) {
    using namespace gemm;
    static_assert(!CONV || (PAIR && EPI == EPI_LINEAR), "the convolution runs on CTA pairs with the linear epilogue");
    using C = Cfg<PAIR, EPI, WIDE>;
    constexpr int TILE_N = C::TILE_N;
    constexpr int STAGES = C::STAGES;
    constexpr int STAGE_BYTES = C::STAGE_BYTES;
    // timing experiments with wrong results are compile-time only (-DCGS_GEMM_EXPT=<bits>, tools/gemm_dbg.sh builds a
    // private copy of the library): 1 no GELU, 2 no stores, 4 reload A instead of B, 8 one MMA per k-block
    constexpr int dbg = CGS_GEMM_EXPT;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    uint8_t* out_stage = smem + STAGES * STAGE_BYTES;
    uint8_t* res_stage = out_stage + OUT_STAGE_BYTES;                 // (EPI_LINEAR) residual tile, same swizzled layout
    uint64_t* full = reinterpret_cast<uint64_t*>(res_stage + C::RES_STAGE_BYTES);
    uint64_t* empty = full + STAGES;
    uint64_t* tfull = empty + STAGES;
    uint64_t* tempty = tfull + 2;
    uint64_t* rfull = tempty + 2;                                     // residual tile has landed
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(rfull + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = PAIR ? cluster_ctarank() : 0u;        // 0 = leader of the pair
    const int worker = PAIR ? (blockIdx.x >> 1) : blockIdx.x;   // pair (or CTA) index
    const int workers = PAIR ? (gridDim.x >> 1) : gridDim.x;
    const int tiles_m = (M + C::TILE_M - 1) / C::TILE_M;
    const int tiles_n = (EPI == EPI_GEGLU) ? (N / BN_HALF) : ((N + TILE_N - 1) / TILE_N);
    const int kblocks = CONV ? geom.taps * geom.cblocks : (K + BK - 1) / BK;
    const Schedule<EPI, TILE_N> sched(workers, tiles_m, tiles_n, N, geom.whole_tiles);
    const int n_items = sched.items(worker);

    if (warp == 0 && elect_one()) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        if (WIDE) tma_prefetch_desc(&tmB2);
        tma_prefetch_desc(&tmO);
        if (EPI == EPI_LINEAR && residual != nullptr) tma_prefetch_desc(&tmR);
    }
    if (warp == 1) {
        if (elect_one()) {
            for (int i = 0; i < STAGES; ++i) {
                mbar_init(&full[i], 1);
                mbar_init(&empty[i], 1);
            }
            for (int i = 0; i < 2; ++i) {
                mbar_init(&tfull[i], 1);
                mbar_init(&tempty[i], PAIR ? 2 * EPI_WARPS : EPI_WARPS);   // pair: both CTAs' epilogue warps
            }
            mbar_init(rfull, 1);
            fence_mbar_init();
        }
        __syncwarp();
        if (PAIR) {
            tmem_alloc_pair(tmem_slot, 512);
            tmem_relinquish_pair();
        } else {
            tmem_alloc(tmem_slot, 512);
            tmem_relinquish();
        }
    }
    tc_fence_before();
    if (PAIR) cluster_sync_all(); else __syncthreads();      // barriers of both CTAs initialised before any remote use
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ------------------------------ TMA producer ------------------------------
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (int i = 0; i < n_items; ++i) {
                const Piece pc = sched.item(worker, i);
                if (pc.width <= 0) continue;
                int mb, nb;
                tile_coords(pc.tile, tiles_m, tiles_n, mb, nb);
                const int m0 = mb * C::TILE_M + (int)rank * BM;
                int img = 0, h0 = 0, w0 = 0;
                if (CONV) conv_tile(geom, mb * 2 + (int)rank, img, h0, w0);
                // B rows of this piece. GEGLU: value rows + gate rows of the tile's 128 output columns. Linear: `width`
                // consecutive rows of W; with CTA pairs each CTA stages half of them (a full 128-row box is loaded,
                // the MMA reads the first width / 2 rows of it).
                // WIDE, a whole 320-column tile: 128 + 32 rows per CTA (columns [0, 256) and [256, 320) of the tile)
                const bool wide_piece = WIDE && pc.width > UMMA_N;
                const int w1 = wide_piece ? UMMA_N : pc.width;
                const int row0 = (EPI == EPI_GEGLU) ? nb * BN_HALF : nb * TILE_N + pc.n_off;
                const int row1 = (EPI == EPI_GEGLU) ? N + nb * BN_HALF : row0 + (PAIR ? w1 / 2 : BN_HALF);
                const bool two_boxes = PAIR || pc.width > BN_HALF;
                // (pairs, EPI_LINEAR) a piece narrower than 256 columns stages only the B rows it needs
                const int b_boxes = (w1 / 2 + B2_ROWS - 1) / B2_ROWS;
                const bool narrow_b = PAIR && EPI == EPI_LINEAR && w1 < UMMA_N;
                int tap = 0, cb = 0;                       // (CONV) filter tap 0..8 and 64-channel block inside it
                for (int kb = 0; kb < kblocks; ++kb) {
                    mbar_wait(&empty[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * STAGE_BYTES;
                    uint8_t* sb = sa + A_BYTES;
                    if (PAIR) {
                        // both CTAs' bytes are credited to the leader's barrier; the leader announces them
                        if (rank == 0)
                            mbar_expect_tx(&full[stage], 2 * (A_BYTES + (narrow_b ? b_boxes * B2_BYTES : BN_HALF * BK * 2) +
                                                              (wide_piece ? B2_BYTES : 0)));
                        int kcoord = kb * BK;
                        if (CONV) {
                            const int r = tap / geom.taps_w, sft = tap - geom.taps_w * r;
                            tma_load_4d_pair(sa, &tmA, &full[stage], cb * BK, w0 + sft + geom.off_w, h0 + r + geom.off_h,
                                             img);
                            kcoord = tap * geom.cin + cb * BK;
                            if (++cb == geom.cblocks) { cb = 0; ++tap; }
                        } else {
                            tma_load_2d_pair(sa, &tmA, &full[stage], kcoord, m0);
                        }
                        if (narrow_b) {              // w1 / 2 rows as 32-row boxes (the 128-row box costs a full tile's feed)
                            for (int bx = 0; bx < b_boxes; ++bx)
                                tma_load_2d_pair(sb + bx * B2_BYTES, &tmB2, &full[stage], kcoord,
                                                 (rank == 0 ? row0 : row1) + bx * B2_ROWS);
                        } else {
                            tma_load_2d_pair(sb, &tmB, &full[stage], kcoord, rank == 0 ? row0 : row1);
                        }
                        if (wide_piece)
                            tma_load_2d_pair(sb + BN_HALF * BK * 2, &tmB2, &full[stage], kcoord,
                                             row0 + UMMA_N + (int)rank * B2_ROWS);
                    } else {
                        mbar_expect_tx(&full[stage], two_boxes ? STAGE_BYTES : A_BYTES + BN_HALF * BK * 2);
                        tma_load_2d(sa, &tmA, &full[stage], kb * BK, m0);
                        if (dbg & 4) {   // experiment: reload the A box instead of B (1/3 of the distinct L2 lines)
                            tma_load_2d(sb, &tmA, &full[stage], kb * BK, m0);
                            tma_load_2d(sb + BN_HALF * BK * 2, &tmA, &full[stage], kb * BK, m0);
                        } else {
                            tma_load_2d(sb, &tmB, &full[stage], kb * BK, row0);
                            if (two_boxes) tma_load_2d(sb + BN_HALF * BK * 2, &tmB, &full[stage], kb * BK, row1);
                        }
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ------------------------------ MMA issuer (leader CTA only for pairs) ------
        if (rank == 0 && elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int i = 0; i < n_items; ++i) {
                const Piece pc = sched.item(worker, i);
                if (pc.width <= 0) continue;
                const bool wide_piece = WIDE && pc.width > UMMA_N;
                const uint32_t idesc = umma_idesc_bf16(C::TILE_M, wide_piece ? UMMA_N : pc.width, 0);
                const uint32_t idesc2 = umma_idesc_bf16(C::TILE_M, WIDE_N - UMMA_N, 0);
                GEMM_STAMP(i, 0);
                mbar_wait(&tempty[acc], acc_phase ^ 1);
                tc_fence_after();
                GEMM_STAMP(i, 1);
                const uint32_t d_tmem = tmem_base + acc * UMMA_N;
                for (int kb = 0; kb < kblocks; ++kb) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    if (kb == 0) GEMM_STAMP(i, 2);
                    const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
                    const uint64_t adesc = umma_desc_sw128(sa, 16, 1024);
                    const uint64_t bdesc = umma_desc_sw128(sa + A_BYTES, 16, 1024);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) {
                        // +32 bytes per k-step inside the 128-byte swizzle row (>> 4 == 2)
                        if (!(dbg & 8) || k == 0) {
                            if (PAIR) umma_ss_pair(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
                            else umma_ss(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
                        }
                    }
                    if (wide_piece) {          // columns [256, 320) of the tile: same A, the CTA's second B box
                        const uint64_t b2desc = umma_desc_sw128(sa + A_BYTES + BN_HALF * BK * 2, 16, 1024);
#pragma unroll
                        for (int k = 0; k < BK / 16; ++k)
                            umma_ss_pair(d_tmem + UMMA_N, adesc + 2 * k, b2desc + 2 * k, idesc2, (kb | k) != 0);
                    }
                    if (PAIR) tc_commit_pair(&empty[stage]); else tc_commit(&empty[stage]);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                if (PAIR) tc_commit_pair(&tfull[acc]); else tc_commit(&tfull[acc]);
                GEMM_STAMP(i, 3);
                if (WIDE) acc_phase ^= 1;                       // one accumulator buffer
                else if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
        __syncwarp();
    } else {
        // ------------------------------ epilogue ----------------------------------
        const int quad = warp & 3;               // TMEM lane quadrant this warp may access
        const int cg = (warp - 2) >> 2;          // 32-column group 0..3 of a 128-column pass
        const int row_in_tile = quad * 32 + lane;
        const bool store_leader = (warp == 2 && lane == 0);
        constexpr int EPI_THREADS = EPI_WARPS * 32;
        // this thread's 64 bytes of the staging tile: box = cg / 2, 16-byte chunks ((cg & 1) * 4 + j) ^ (row & 7)
        uint8_t* my_stage = out_stage + (cg >> 1) * OUT_BOX_BYTES + row_in_tile * 128;
        const int chunk0 = (cg & 1) * 4;
        const int sw = row_in_tile & 7;
        // the accumulator buffer goes back to the MMA issuer: its `tempty` barrier lives in the leader CTA
        const uint32_t tempty_addr0 = PAIR ? mapa_shared(smem_u32(&tempty[0]), 0) : 0u;
        auto release_acc = [&](int acc) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (PAIR) mbar_arrive_cluster(tempty_addr0 + acc * 8);
                else mbar_arrive(&tempty[acc]);
            }
        };
        // The residual tile comes in through TMA into a staging tile of the output's swizzled layout: read straight
        // from global memory it was 64 bytes per thread at a 2.5 KB row stride, which made the epilogue of the K = 1280
        // projections slower than their mainloop (to_out + residual ran at half the speed of the same GEMM without).
        const bool use_res = C::RES && residual != nullptr;
        const uint8_t* my_res = res_stage + (cg >> 1) * OUT_BOX_BYTES + row_in_tile * 128;
        uint32_t res_phase = 0;
        int img = 0, h0 = 0, w0 = 0;                                          // (CONV) this CTA's pixel rectangle
        auto load_residual = [&](int c0, int left, int row_base) {          // store_leader only
            mbar_expect_tx(rfull, (left > 64 ? 2 : 1) * OUT_BOX_BYTES);
            if (CONV) {
                tma_load_4d(res_stage, &tmR, rfull, c0, w0, h0, img);
                if (left > 64) tma_load_4d(res_stage + OUT_BOX_BYTES, &tmR, rfull, c0 + 64, w0, h0, img);
            } else {
                tma_load_2d(res_stage, &tmR, rfull, c0, row_base);
                if (left > 64) tma_load_2d(res_stage + OUT_BOX_BYTES, &tmR, rfull, c0 + 64, row_base);
            }
        };
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int i = 0; i < n_items; ++i) {
            const Piece pc = sched.item(worker, i);
            if (pc.width <= 0) continue;
            int mb, nb;
            tile_coords(pc.tile, tiles_m, tiles_n, mb, nb);
            const int row_base = mb * C::TILE_M + (int)rank * BM;
            const int row = row_base + row_in_tile;
            const int col_base = nb * TILE_N + pc.n_off;        // (EPI_LINEAR) first output column of this piece
            if (CONV) conv_tile(geom, mb * 2 + (int)rank, img, h0, w0);
            // (the previous item's last residual reads are behind its second named barrier: the tile is free)
            if (use_res && store_leader) load_residual(col_base, pc.width, row_base);
            if (store_leader) GEMM_STAMP(i, 4);
            mbar_wait(&tfull[acc], acc_phase);
            tc_fence_after();
            if (store_leader) GEMM_STAMP(i, 5);
            const uint32_t taddr = tmem_base + acc * UMMA_N + (static_cast<uint32_t>(quad * 32) << 16);
            const int PASSES = (EPI == EPI_GEGLU) ? 1 : (pc.width + 127) / 128;
#pragma unroll 1
            for (int pass = 0; pass < PASSES; ++pass) {
                BF8 o[4];
                int col;
                if (EPI == EPI_GEGLU) {
                    col = nb * BN_HALF + cg * 32;
                    uint32_t v[32], g[32];
                    tmem_ld_x32(taddr + cg * 32, v);
                    tmem_ld_x32(taddr + BN_HALF + cg * 32, g);
                    tmem_wait_ld();
                    // accumulator is in registers: hand the TMEM buffer back to the MMA warp now
                    release_acc(acc);
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
                            o[q].v[j] = (dbg & 1) ? __floats2bfloat162_rn(fv.x * fg.x, fv.y * fg.y)
                                                  : __floats2bfloat162_rn(fv.x * gelu_erf(fg.x), fv.y * gelu_erf(fg.y));
                        }
                    }
                } else {
                    col = col_base + pass * 128 + cg * 32;
                    uint32_t v[32];
                    tmem_ld_x32(taddr + pass * 128 + cg * 32, v);
                    tmem_wait_ld();
                    if (store_leader) GEMM_STAMP(i, 6 + pass);
                    if (pass == PASSES - 1) release_acc(acc);
                    // widths are multiples of 32: a 32-column group is all in or all out of the piece
                    const bool col_ok = pass * 128 + cg * 32 < pc.width;
                    if (use_res) {
                        mbar_wait(rfull, res_phase);
                        res_phase ^= 1;
                    }
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        BF8 bv, rv;
                        const bool use_b = bias != nullptr && col_ok;
                        const bool use_r = use_res && col_ok && (CONV || row < M);   // (CONV: rows outside come back 0)
                        if (use_b) bv = *reinterpret_cast<const BF8*>(bias + col + q * 8);
                        if (use_r) rv = *reinterpret_cast<const BF8*>(my_res + (((chunk0 + q) ^ sw) << 4));
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            float2 fv = make_float2(__uint_as_float(v[q * 8 + 2 * j]),
                                                    __uint_as_float(v[q * 8 + 2 * j + 1]));
                            if (use_b) {
                                const float2 b0 = __bfloat1622float2(bv.v[j]);
                                fv.x += b0.x; fv.y += b0.y;
                            }
                            if (use_r) {
                                const float2 r0 = __bfloat1622float2(rv.v[j]);
                                fv.x += r0.x; fv.y += r0.y;
                            }
                            o[q].v[j] = __floats2bfloat162_rn(fv.x, fv.y);
                        }
                    }
                }
                // staging tile must no longer be read by the previous TMA store
                if (store_leader) tma_store_wait_read();
                named_bar_sync(1, EPI_THREADS);
                // every thread has its residual values in registers: the second pass's tile may come in
                if (use_res && store_leader && pass + 1 < PASSES)
                    load_residual(col_base + (pass + 1) * 128, pc.width - (pass + 1) * 128, row_base);
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    *reinterpret_cast<BF8*>(my_stage + (((chunk0 + q) ^ sw) << 4)) = o[q];
                fence_proxy_async_smem();
                named_bar_sync(1, EPI_THREADS);
                if (store_leader && !(dbg & 2) && (CONV ? img < geom.n_img : row_base < M)) {
                    // 64-column boxes; pieces are multiples of 64 columns except at the edge of N, where the tensor map clips
                    const int c0 = (EPI == EPI_GEGLU) ? nb * BN_HALF : col_base + pass * 128;
                    const int left = (EPI == EPI_GEGLU) ? BN_HALF : pc.width - pass * 128;
                    if (CONV) {
                        if (left > 0) tma_store_4d(&tmO, out_stage, c0, w0, h0, img);
                        if (left > 64) tma_store_4d(&tmO, out_stage + OUT_BOX_BYTES, c0 + 64, w0, h0, img);
                    } else {
                        if (left > 0) tma_store_2d(&tmO, out_stage, c0, row_base);
                        if (left > 64) tma_store_2d(&tmO, out_stage + OUT_BOX_BYTES, c0 + 64, row_base);
                    }
                    tma_store_commit();
                }
            }
            if (store_leader) GEMM_STAMP(i, 10);
            if (WIDE) acc_phase ^= 1;
            else if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if (store_leader) tma_store_wait_all();
    }

    tc_fence_before();
    if (PAIR) cluster_sync_all(); else __syncthreads();       // nobody leaves while its partner still needs it
    if (warp == 1) {
        if (PAIR) tmem_dealloc_pair(tmem_base, 512);
        else tmem_dealloc(tmem_base, 512);
    }
}

template <int EPI, bool PAIR, bool CONV = false, bool WIDE = false>
static int launch_variant(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmB2, const CUtensorMap& tmO,
                          const CUtensorMap& tmR, const __nv_bfloat16* bp, const __nv_bfloat16* rp, __nv_bfloat16* op,
                          int M, int N, int K, int ldo, int ldr, const gemm::ConvGeom& geom, long long tiles,
                          cudaStream_t stream) {
    using namespace gemm;
    using C = Cfg<PAIR, EPI, WIDE>;
    auto kern = gemm_kernel<EPI, PAIR, CONV, WIDE>;
    CGS_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
    const int sms = sm_count();
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute attr[1];
    // EPI_LINEAR deals the last round out in 64-column pieces: up to 4 workers per tile can be put to use
    const long long useful = (EPI == EPI_LINEAR && geom.whole_tiles != 1) ? tiles * C::GROUPS : tiles;
    if (PAIR) {
        const long long pairs = useful < sms / 2 ? useful : sms / 2;
        cfg.gridDim = dim3((unsigned)(2 * pairs));
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 2;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
    } else {
        cfg.gridDim = dim3((unsigned)(useful < sms ? useful : sms));
    }
    cfg.blockDim = dim3(THREADS);
    cfg.dynamicSmemBytes = C::SMEM_BYTES;
    cfg.stream = stream;
    CGS_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, tmA, tmB, tmB2, tmO, tmR, bp, rp, op, M, N, K, ldo, ldr, geom));
    return kOk;
}

// which mainloop: 0 = automatic, 1 = one CTA per tile (M128), 2 = CTA pairs (M256). (tools/kernel_bench.py compares)
static int g_gemm_force_variant = 0;
// convolution tiles: 0 = automatic, 1 = never 320-column tiles, 2 = 320-column tiles whenever Cout allows
static int g_conv_force_wide = 0;
// (measured, profiles/r02_conv_bench.json: with K = 9 Cin >= 2880 a narrow piece costs nearly a whole tile's time --
// the k-loop is as long and the operand feed almost the same -- so the convolution keeps its last round whole)
static int g_conv_whole_tiles = 1;
// the linear layers' last round: 0 = 64-column groups dealt out evenly, 1 = whole tiles, 2 = halves / quarters, one per worker
// (measured, profiles/r02_kernel_bench_gemm.json `split_policy`: halves / quarters +6 % on to_out, +11 % on ff.net.2 at
// M = 16384, +28 % at M = 2048 over the even dealing of 64-column groups)
static int g_linear_split_mode = 2;

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
    CUtensorMap tmO;
    {
        const uint64_t dims[2] = {(uint64_t)N, (uint64_t)M};
        const uint64_t strides[1] = {(uint64_t)ldo * 2};
        const uint32_t box[2] = {64, BM};
        int rc = make_tmap_bf16(&tmO, out, 2, dims, strides, box);
        if (rc != kOk) return rc;
    }
    CUtensorMap tmR = tmO;
    if (residual != nullptr) {
        const uint64_t dims[2] = {(uint64_t)N, (uint64_t)M};
        const uint64_t strides[1] = {(uint64_t)ldr * 2};
        const uint32_t box[2] = {64, BM};
        int rc = make_tmap_bf16(&tmR, residual, 2, dims, strides, box);
        if (rc != kOk) return rc;
    }
    const int tiles_n = (epi == EPI_GEGLU) ? N / BN_HALF : (N + UMMA_N - 1) / UMMA_N;
    const long long tiles1 = ((M + BM - 1) / BM) * tiles_n;
    const long long tiles2 = ((M + 2 * BM - 1) / (2 * BM)) * tiles_n;
    // CTA pairs pay off when the mainloop is long enough to be bound by the operand feed: measured on the SDXL shapes
    // (profiles/r02_kernel_bench.json), K >= 1024 wins 7-15 % with pairs at every M, K = 640 loses 3-14 % (ten k-blocks
    // per tile: the epilogue hand-over between two SMs is a visible part of every tile)
    (void)tiles2;
    bool pair = K >= 1024;
    if (g_gemm_force_variant == 1) pair = false;
    if (g_gemm_force_variant == 2) pair = true;
    auto bp = static_cast<const __nv_bfloat16*>(bias);
    auto rp = static_cast<const __nv_bfloat16*>(residual);
    auto op = static_cast<__nv_bfloat16*>(out);
    ConvGeom none = {};
    none.whole_tiles = g_linear_split_mode;
    CUtensorMap tmB2 = tmB;
    if (epi == EPI_LINEAR) {
        const uint64_t dims[2] = {(uint64_t)K, (uint64_t)N};
        const uint64_t strides[1] = {(uint64_t)ldw * 2};
        const uint32_t box[2] = {BK, B2_ROWS};
        int rc = make_tmap_bf16(&tmB2, w, 2, dims, strides, box);
        if (rc != kOk) return rc;
    }
    if (epi == EPI_GEGLU) {
        return pair ? launch_variant<EPI_GEGLU, true>(tmA, tmB, tmB, tmO, tmR, bp, rp, op, (int)M, N, K, (int)ldo, (int)ldr, none, tiles2, stream)
                    : launch_variant<EPI_GEGLU, false>(tmA, tmB, tmB, tmO, tmR, bp, rp, op, (int)M, N, K, (int)ldo, (int)ldr, none, tiles1, stream);
    }
    return pair ? launch_variant<EPI_LINEAR, true>(tmA, tmB, tmB2, tmO, tmR, bp, rp, op, (int)M, N, K, (int)ldo, (int)ldr, none, tiles2, stream)
                : launch_variant<EPI_LINEAR, false>(tmA, tmB, tmB2, tmO, tmR, bp, rp, op, (int)M, N, K, (int)ldo, (int)ldr, none, tiles1, stream);
}

// Convolution over channels-last images as an implicit GEMM (see the CONV notes at the top): taps_h x taps_w filter
// taps, tap (0, 0) reading the input pixel (off_h, off_w) away from the output pixel, zero outside the image.
// x: [n_img, H, W, Cin] with `ldx` elements between consecutive pixels (>= Cin: a channel slice of a wider buffer is
// fine), w: [Cout, taps_h, taps_w, Cin] contiguous. Output / residual pixel (n, h, w) lives at
// base + n * s[2] + h * s[1] + w * s[0] elements -- dense images, or one phase of a 2x finer lattice.
struct ImgStrides {
    long long s[3];
};

static int launch_conv(const char* who, const void* x, const void* w, const void* bias, const void* residual, void* out,
                       int n_img, int H, int W, int Cin, int Cout, int taps_h, int taps_w, int off_h, int off_w,
                       long long ldx, ImgStrides so, ImgStrides sr, cudaStream_t stream) {
    using namespace gemm;
    if (n_img < 0 || H <= 0 || W <= 0 || Cin <= 0 || Cout <= 0) {
        set_last_error("%s: bad sizes n=%d H=%d W=%d Cin=%d Cout=%d", who, n_img, H, W, Cin, Cout);
        return kBadArgument;
    }
    if (n_img == 0) return kOk;
    bool ok = Cin % 8 == 0 && Cout % 32 == 0 && ldx % 8 == 0 && ldx >= Cin;
    for (int i = 0; i < 3; ++i) ok = ok && so.s[i] % 8 == 0 && (!residual || sr.s[i] % 8 == 0);
    if (!ok) {
        set_last_error("%s: Cin (%d) and the pixel strides must be multiples of 8, Cout (%d) of 32", who, Cin, Cout);
        return kUnsupportedShape;
    }
    if ((reinterpret_cast<uintptr_t>(x) & 0xF) || (reinterpret_cast<uintptr_t>(w) & 0xF) ||
        (reinterpret_cast<uintptr_t>(out) & 0xF) || (reinterpret_cast<uintptr_t>(bias) & 0xF) ||
        (reinterpret_cast<uintptr_t>(residual) & 0xF)) {
        set_last_error("%s: pointers must be 16-byte aligned", who);
        return kBadArgument;
    }
    // a tile's 128 rows: TH x TW pixels, TW the largest power of two dividing W (at most 128)
    ConvGeom g;
    g.H = H;
    g.W = W;
    g.TW = 1;
    while (g.TW < BM && W % (g.TW * 2) == 0) g.TW *= 2;
    g.TH = BM / g.TW;
    g.tiles_w = W / g.TW;
    g.tiles_h = (H + g.TH - 1) / g.TH;
    g.n_img = n_img;
    g.cin = Cin;
    g.cblocks = (Cin + BK - 1) / BK;
    g.taps = taps_h * taps_w;
    g.taps_w = taps_w;
    g.off_h = off_h;
    g.off_w = off_w;
    g.whole_tiles = g_conv_whole_tiles;
    const long long mtiles = (long long)n_img * g.tiles_h * g.tiles_w;       // 128-row tiles
    if (mtiles * BM > 0x7fffffffLL) {
        set_last_error("%s: %lld pixel tiles are too many", who, mtiles);
        return kUnsupportedShape;
    }
    auto img_map = [&](CUtensorMap* tm, const void* base, int channels, const long long* st) {
        const uint64_t dims[4] = {(uint64_t)channels, (uint64_t)W, (uint64_t)H, (uint64_t)n_img};
        const uint64_t strides[3] = {(uint64_t)st[0] * 2, (uint64_t)st[1] * 2, (uint64_t)st[2] * 2};
        const uint32_t box[4] = {64, (uint32_t)g.TW, (uint32_t)g.TH, 1};
        return make_tmap_bf16(tm, base, 4, dims, strides, box);
    };
    CUtensorMap tmA, tmB, tmB2, tmO, tmR;
    const long long sx[3] = {ldx, (long long)W * ldx, (long long)H * W * ldx};
    int rc = img_map(&tmA, x, Cin, sx);
    if (rc != kOk) return rc;
    rc = img_map(&tmO, out, Cout, so.s);
    if (rc != kOk) return rc;
    tmR = tmO;
    if (residual != nullptr) {
        rc = img_map(&tmR, residual, Cout, sr.s);
        if (rc != kOk) return rc;
    }
    const int K = g.taps * Cin;
    const long long ldy = so.s[0], ldr = sr.s[0];
    {
        const uint64_t dims[2] = {(uint64_t)K, (uint64_t)Cout};
        const uint64_t strides[1] = {(uint64_t)K * 2};
        const uint32_t box[2] = {BK, BN_HALF};
        rc = make_tmap_bf16(&tmB, w, 2, dims, strides, box);
        if (rc != kOk) return rc;
        const uint32_t box2[2] = {BK, B2_ROWS};
        rc = make_tmap_bf16(&tmB2, w, 2, dims, strides, box2);
        if (rc != kOk) return rc;
    }
    const long long tiles_m = (mtiles + 1) / 2;
    const int M = (int)(mtiles * BM);
    auto bp = static_cast<const __nv_bfloat16*>(bias);
    auto rp = static_cast<const __nv_bfloat16*>(residual);
    auto op = static_cast<__nv_bfloat16*>(out);
    bool wide = Cout % WIDE_N == 0 && Cout % UMMA_N != 0;      // 320, 960: tiles of 256 would end in a 64-column one
    if (g_conv_force_wide == 1) wide = false;
    if (g_conv_force_wide == 2) wide = Cout % WIDE_N == 0;
    if (wide)
        return launch_variant<EPI_LINEAR, true, true, true>(tmA, tmB, tmB2, tmO, tmR, bp, rp, op, M, Cout, K, (int)ldy,
                                                            (int)ldr, g, tiles_m * (Cout / WIDE_N), stream);
    return launch_variant<EPI_LINEAR, true, true, false>(tmA, tmB, tmB2, tmO, tmR, bp, rp, op, M, Cout, K, (int)ldy,
                                                         (int)ldr, g, tiles_m * ((Cout + UMMA_N - 1) / UMMA_N), stream);
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

extern "C" int cgs_conv3x3_nhwc_fwd(const void* x, const void* w, const void* bias, const void* residual, void* y,
                                    int n_img, int H, int W, int Cin, int Cout, long long ldx, long long ldy,
                                    long long ldr, void* stream) {
    if (ldy < Cout || (residual && ldr < Cout)) {
        cgs::set_last_error("conv3x3: pixel strides of y / residual are smaller than Cout");
        return cgs::kBadArgument;
    }
    const cgs::ImgStrides so = {{ldy, (long long)W * ldy, (long long)H * W * ldy}};
    const cgs::ImgStrides sr = {{ldr, (long long)W * ldr, (long long)H * W * ldr}};
    return cgs::launch_conv("conv3x3", x, w, bias, residual, y, n_img, H, W, Cin, Cout, 3, 3, -1, -1, ldx, so, sr,
                            static_cast<cudaStream_t>(stream));
}

// conv3x3(nearest-neighbour 2x up-sampling of x) without the up-sampled image: output pixel (2i + a, 2j + b) sees
// only 2 x 2 distinct input pixels, so each of the four phases (a, b) is a 2 x 2 convolution of x with the 3x3 taps
// that fall on the same input pixel added up -- 16 instead of 36 multiply-adds per input pixel, channel pair and
// output, and no 4x larger intermediate. w4: [4 phases (a * 2 + b)][Cout][2][2][Cin], see cgs_b200.h.
extern "C" int cgs_upsample2x_conv3x3_nhwc_fwd(const void* x, const void* w4, const void* bias, void* y, int n_img,
                                               int H, int W, int Cin, int Cout, long long ldx, long long ldy,
                                               void* stream) {
    if (ldy < Cout) {
        cgs::set_last_error("upsample2x_conv3x3: pixel stride of y is smaller than Cout");
        return cgs::kBadArgument;
    }
    const long long W2 = 2LL * W, H2 = 2LL * H;
    const cgs::ImgStrides so = {{2 * ldy, 2 * W2 * ldy, H2 * W2 * ldy}};
    for (int a = 0; a < 2; ++a) {
        for (int b = 0; b < 2; ++b) {
            const auto* wp = static_cast<const __nv_bfloat16*>(w4) + (size_t)(a * 2 + b) * Cout * 4 * Cin;
            auto* yp = static_cast<__nv_bfloat16*>(y) + (a * W2 + b) * ldy;
            const int rc = cgs::launch_conv("upsample2x_conv3x3", x, wp, bias, nullptr, yp, n_img, H, W, Cin, Cout, 2, 2,
                                            a - 1, b - 1, ldx, so, so, static_cast<cudaStream_t>(stream));
            if (rc != cgs::kOk) return rc;
        }
    }
    return cgs::kOk;
}

#ifdef CGS_GEMM_TRACE
extern "C" int cgs_gemm_set_trace(void* buf) {
    long long* p = static_cast<long long*>(buf);
    return cudaMemcpyToSymbol(cgs::g_gemm_trace, &p, sizeof(p)) == cudaSuccess ? 0 : 1;
}
#endif

// Developer switch for the micro-benchmarks: 0 automatic, 1 always one CTA per tile, 2 always CTA pairs. Both
// mainloops compute the same products in the same order per output element; results are identical.
extern "C" int cgs_gemm_set_variant(int variant) {
    // 10 + v: the convolution's tile shape (0 automatic, 1 never 320-column tiles, 2 always when Cout % 320 == 0)
    if (variant >= 10 && variant <= 12) {
        cgs::g_conv_force_wide = variant - 10;
        return cgs::kOk;
    }
    if (variant >= 20 && variant <= 22) {      // the convolution's last round: 20 whole tiles (default), 21 64-column
        cgs::g_conv_whole_tiles = variant == 20 ? 1 : variant == 21 ? 0 : 2;      // groups, 22 halves / quarters
        return cgs::kOk;
    }
    if (variant >= 30 && variant <= 32) {      // the linear layers' last round: 30 64-column groups, 31 whole, 32 halves / quarters
        cgs::g_linear_split_mode = variant - 30;
        return cgs::kOk;
    }
    if (variant < 0 || variant > 2) {
        cgs::set_last_error("gemm_set_variant: %d (0 automatic, 1 single CTA, 2 CTA pairs)", variant);
        return cgs::kBadArgument;
    }
    cgs::g_gemm_force_variant = variant;
    return cgs::kOk;
}
