/*
 * cgs_b200.h -- C-ABI of libcgs_b200.so: the B200 (sm_100a) kernels for the per-step UNet denoise
 * hot path of the ComfyUI-derived Comfy-Gen-Server.
 *
 * The reference is pure Python/PyTorch and has no FFI of its own; every entry point below
 * replaces a PyTorch library call made by the reference at the cited file:line (paths relative
 * to the reference tree). The Python boundary (comfy_gen_server_b200/_native.py) binds these
 * with ctypes and registers them as torch.ops.cgs_b200.*; INTEGRATION.md shows the binding a
 * maintainer of the reference would add.
 *
 * Conventions
 *   - plain pointers and sizes only; all pointers are DEVICE pointers on the current device;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - return value: 0 = ok, non-zero = error (enum below); cgs_last_error() returns a
 *     thread-local human-readable message for the last non-zero return;
 *   - no entry point allocates device memory, synchronises the device, or copies to the host;
 *   - bf16 tensors are CUDA __nv_bfloat16; "elements" means bf16/float elements, not bytes;
 *   - inputs are borrowed and never modified, outputs are fully overwritten (rows/columns the
 *     sizes exclude are never touched).
 */
#ifndef CGS_B200_H
#define CGS_B200_H

#ifdef __cplusplus
extern "C" {
#endif

enum cgs_status {
    CGS_OK = 0,
    CGS_BAD_ARGUMENT = 1,
    CGS_UNSUPPORTED_SHAPE = 2,
    CGS_CUDA_ERROR = 3,
    CGS_DRIVER_ENTRY_MISSING = 4,
    CGS_SCRATCH_TOO_SMALL = 5
};

/* Thread-local message describing the last non-zero status returned on this thread. */
const char* cgs_last_error(void);
int cgs_abi_version(void);

/*
 * Attention forward (A1 self-attention, A2 77/154-key cross-attention; same entry point).
 * Replaces optimized_attention(q, k, v, heads, mask=None) == attention_pytorch ->
 * F.scaled_dot_product_attention   (comfy/ldm/modules/attention.py:337-349, 352-368, 412).
 *   q [B, Nq, H, d], k/v [B, Nk, H, d], out [B, Nq, H, d]; bf16; d contiguous, every stride a
 *   multiple of 8 elements; d a multiple of 8, <= 192. softmax_scale is 1/sqrt(d) for the
 *   reference's call. Non-causal, no mask, no dropout. fp32 accumulation and softmax.
 */
int cgs_attention_fwd(const void* q, const void* k, const void* v, void* out,
                      int B, int H, int Nq, int Nk, int d,
                      long long q_stride_b, long long q_stride_n, long long q_stride_h,
                      long long k_stride_b, long long k_stride_n, long long k_stride_h,
                      long long v_stride_b, long long v_stride_n, long long v_stride_h,
                      long long o_stride_b, long long o_stride_n, long long o_stride_h,
                      float softmax_scale, void* stream);

/*
 * The same with an additive mask: softmax(q k^T * scale + mask) v. Replaces optimized_attention_masked(q, k, v, heads,
 * mask) == attention_pytorch with attn_mask / attention_basic's mask branch (comfy/ldm/modules/attention.py:111-123,
 * 337-349, 353-368). The UNet never passes a mask (attention.py:411-414, 520, 560); this entry exists so that the
 * rebinding of optimized_attention_masked keeps the reference's other callers (nodes_sag.py:130, Stable Cascade) working.
 *   mask: fp32, logically [B, H, Nq, Nk] with key stride 1 and element strides mask_stride_{b,h,q}; a stride of 0
 *   broadcasts that dimension. Boolean "keep" masks are passed as 0 / -1e30 by the host (attention.py:113-117 fills with
 *   -finfo.max, so a fully masked row is a uniform average there and here).
 */
int cgs_attention_masked_fwd(const void* q, const void* k, const void* v, void* out,
                             int B, int H, int Nq, int Nk, int d,
                             long long q_stride_b, long long q_stride_n, long long q_stride_h,
                             long long k_stride_b, long long k_stride_n, long long k_stride_h,
                             long long v_stride_b, long long v_stride_n, long long v_stride_h,
                             long long o_stride_b, long long o_stride_n, long long o_stride_h,
                             float softmax_scale, const float* mask, long long mask_stride_b,
                             long long mask_stride_h, long long mask_stride_q, void* stream);

/*
 * LayerNorm forward (L1). Replaces ops.LayerNorm.forward -> F.layer_norm
 * (comfy/ops.py:97-113) as norm1/norm2/norm3 (comfy/ldm/modules/attention.py:450-453).
 *   x, y [rows, C] bf16 contiguous; gamma, beta [C] bf16; C multiple of 8, <= 2560;
 *   fp32 statistics, eps as given (the reference uses torch's default 1e-5).
 */
int cgs_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y,
                      long long rows, int C, float eps, void* stream);

/*
 * Residual add + LayerNorm in one pass: x_out = bf16(x + residual); y = LN(x_out).
 * Replaces `x += n` followed by the next norm (comfy/ldm/modules/attention.py:527+534,
 * 567+570). x_out may alias x (in-place update, as the reference does).
 */
int cgs_add_layernorm_fwd(const void* x, const void* residual, const void* gamma,
                          const void* beta, void* x_out, void* y,
                          long long rows, int C, float eps, void* stream);

/*
 * GroupNorm (+ optional fused SiLU), NCHW. Replaces ops.GroupNorm -> F.group_norm and the nn.SiLU that follows
 * it in ResBlock.in_layers / out_layers and UNetModel.out
 * (comfy/ldm/modules/diffusionmodules/openaimodel.py:177-181, 213-217, 805-807), and SpatialTransformer.norm
 * (comfy/ldm/modules/attention.py:595; silu = 0, eps = 1e-6).
 *   x, y [N, C, H*W] bf16 contiguous; gamma, beta [C] bf16; C % groups == 0; H*W % 8 == 0; fp32 statistics.
 *   add_nc [N, C] bf16 or NULL: a per-(sample, channel) addend applied (in bf16) before the norm -- the ResBlock's
 *   `h + emb_out[..., None, None]` (openaimodel.py:255-259) without a separate pass over h.
 */
int cgs_groupnorm_fwd(const void* x, const void* add_nc, const void* gamma, const void* beta, void* y,
                      int N, int C, int HW, int groups, float eps, int silu, void* stream);

/*
 * The same GroupNorm (+ SiLU, + addend) for channels-last activations: x, y [N, H*W, C] bf16 (torch
 * memory_format=channels_last). Two launches (partial statistics per pixel-row chunk, then normalise; three for
 * images of 64 MB and more per sample, whose up to 1024 partials a small kernel adds up in between) that
 * communicate through `scratch`, a caller-provided device buffer of at least
 * cgs_groupnorm_nhwc_scratch_bytes(N, C, HW, groups) bytes; it need not be initialised and may be reused by the next
 * call on the same stream. C % 8 == 0, C / groups >= 8 or == 4 (the VAE's 128-channel level), groups <= 32, C <= 2560.
 */
long long cgs_groupnorm_nhwc_scratch_bytes(int N, int C, int HW, int groups);
int cgs_groupnorm_nhwc_fwd(const void* x, const void* add_nc, const void* gamma, const void* beta, void* y,
                           void* scratch, long long scratch_bytes,
                           int N, int C, int HW, int groups, float eps, int silu, void* stream);

/*
 * cgs_groupnorm_nhwc_fwd for callers that cannot allocate device memory. (No counterpart in the reference: it is
 * an accommodation of this library's own scratch requirement, not a replacement for a reference call.)
 * Who needs it: code that gets every byte of device memory from someone else -- a custom node or worker inside the
 * host framework's memory accounting, a C caller without an allocator -- or that may not pay cudaMalloc's
 * device-wide synchronisation. The partial-statistics buffer is a fixed 1 MiB array inside the library, present on
 * every GPU that loads it, for the life of the process.
 *   - Same arguments and results as cgs_groupnorm_nhwc_fwd, minus the scratch pair.
 *   - Limit: N * (chunks + 1) * groups * 8 bytes (= cgs_groupnorm_nhwc_scratch_bytes) must be <= 1 MiB, else
 *     CGS_SCRATCH_TOO_SMALL and nothing is launched.
 *   - The first call on a device resolves the array's address on that device; make it during set-up, before any
 *     stream capture begins.
 *   - All calls on a device share the array; it is never cleared and need not be, but calls whose kernels overlap
 *     on the GPU (other streams, other threads) race on it -- serialise them.
 *   - The tensors must live on the device that is current at the call (checked for x: CGS_BAD_ARGUMENT otherwise);
 *     the address look-up is safe when several host threads make a device's first call at once.
 *   - Allocates and frees nothing.
 */
int cgs_groupnorm_nhwc_noalloc_fwd(const void* x, const void* add_nc, const void* gamma, const void* beta, void* y,
                                   int N, int C, int HW, int groups, float eps, int silu, void* stream);

/*
 * Batched 2-D transpose with an optional residual add: out[n, c, r] = in[n, r, c] (+ residual[n, c, r]).
 * Replaces the NCHW <-> token-major rearranges at the edges of SpatialTransformer.forward and the `x + x_in` that
 * follows the second one (comfy/ldm/modules/attention.py:628, 636, 639).
 *   in [N, R, C], out / residual [N, C, R] bf16 contiguous; R, C multiples of 8; out must not alias in.
 */
int cgs_transpose_add_fwd(const void* in, const void* residual, void* out, int N, int R, int C, void* stream);

/*
 * out = a + b (+ bias[channel]), bf16, n elements (n % 8 == 0, 16-byte aligned), for two tensors that share one dense
 * memory layout. Replaces the ResBlock's `self.skip_connection(x) + h` (comfy/ldm/modules/diffusionmodules/
 * openaimodel.py:264) and the SpatialTransformer's `x + x_in` (comfy/ldm/modules/attention.py:639). bias (may be NULL)
 * is the [channels] bias of the convolution(s) that produced the operands -- torch.nn.Conv2d's separate bias pass
 * (openaimodel.py:181, 217, 228 via ops.Conv2d) folded into this one; channel of element i = (i / inner) % channels,
 * inner = 1 for channels_last (channels % 8 == 0), H*W for NCHW (H*W % 8 == 0). out may alias a or b.
 */
int cgs_add_fwd(const void* a, const void* b, const void* bias, void* out, long long n, int channels, long long inner,
                void* stream);

/*
 * Channel concatenation of two channels-last images: out[p, 0:ca] = a[p, :], out[p, ca:ca+cb] = b[p, :] for each of
 * `pixels` = N*H*W pixels; bf16, ca and cb multiples of 8. Replaces `th.cat([h, hs.pop()], dim=1)`
 * (comfy/ldm/modules/diffusionmodules/openaimodel.py:869) in the channels-last pipeline. bias_a (NULL or [ca] bf16) is
 * added to the `a` part: the bias of the Upsample convolution (openaimodel.py:94-95) that produced it.
 */
int cgs_cat_channels_nhwc_fwd(const void* a, const void* b, const void* bias_a, void* out, long long pixels,
                              int ca, int cb, void* stream);

/*
 * Nearest-neighbour 2x up-sampling of a channels-last image: x [N, H, W, C] -> out [N, 2H, 2W, C], bf16, C % 8 == 0.
 * Replaces F.interpolate(x, size=(2H, 2W), mode="nearest") in Upsample.forward (openaimodel.py:86-93).
 */
int cgs_upsample_nearest2x_nhwc_fwd(const void* x, void* out, int N, int H, int W, int C, void* stream);

/*
 * GEGLU projection (G1): y[m, n] = (x W^T + b)[m, n] * gelu_erf((x W^T + b)[m, N + n]).
 * Replaces GEGLU.forward: Linear -> chunk(2) -> F.gelu -> mul
 * (comfy/ldm/modules/attention.py:56-63).
 *   x [M, K] (lda), W [2N, K] (ldw; rows [0,N) value half, [N,2N) gate half, exactly the
 *   reference's proj.weight), bias [2N] or NULL, y [M, N] (ldy); bf16; K, lda, ldw, ldy
 *   multiples of 8; N multiple of 128.
 */
int cgs_geglu_gemm_fwd(const void* x, const void* w, const void* bias, void* y,
                       long long M, int N, int K,
                       long long lda, long long ldw, long long ldy, void* stream);

/*
 * Linear: y = x W^T (+ bias) (+ residual). Replaces ops.Linear.forward -> F.linear
 * (comfy/ops.py:40-52) for to_q/to_k/to_v/to_out/ff.net.2/proj_in/proj_out, optionally with
 * the block's in-place residual add fused (comfy/ldm/modules/attention.py:527,567,572).
 *   x [M, K], W [N, K], bias [N] or NULL, residual [M, N] (ldr) or NULL, y [M, N]; bf16;
 *   N multiple of 32; y may alias residual.
 */
int cgs_linear_fwd(const void* x, const void* w, const void* bias, const void* residual, void* y,
                   long long M, int N, int K,
                   long long lda, long long ldw, long long ldy, long long ldr, void* stream);

/*
 * UNet input scaling: out[c, b, :] = bf16(x[b, :] / sqrt(sigma[b]^2 + sigma_data^2)), c < copies.
 * Replaces EPS.calculate_input (comfy/model_sampling.py:6-8), the dtype cast
 * (comfy/model_base.py:86-88) and the cond/uncond torch.cat (comfy/samplers.py:187).
 *   x [batch, per_sample] fp32; sigma [batch] fp32; out [copies, batch, per_sample] bf16.
 */
int cgs_unet_input_fwd(const float* x, const float* sigma, void* out_bf16,
                       long long batch, long long per_sample, int copies,
                       float sigma_data, void* stream);

/*
 * eps/v -> denoised, then CFG combine (C1).
 * Replaces EPS/V_PREDICTION.calculate_denoised (comfy/model_sampling.py:10-12, 27-29), the
 * calc_cond_batch accumulate/divide with full-frame conds (comfy/samplers.py:220-226) and
 * cfg_function's default branch (comfy/samplers.py:240).
 *   denoised = D_u + (D_c - D_u) * cond_scale; out_uncond may be NULL (then denoised = D_c).
 *   x, denoised [batch, per_sample] fp32; model outputs bf16 (out_is_bf16 = 1) or fp32;
 *   sigma [batch] fp32; pred_type 0 = EPS, 1 = V_PREDICTION. Bit-identical to the reference's
 *   fp32 eager composition.
 */
int cgs_cfg_denoise_fwd(const float* x, const void* out_uncond, const void* out_cond,
                        int out_is_bf16, const float* sigma, float cond_scale, int pred_type,
                        float sigma_data, float* denoised,
                        long long batch, long long per_sample, void* stream);

/*
 * Euler update (S1): x_next = x + (x - denoised) / sigma_hat * dt.
 * Replaces to_d + the Euler step of sample_euler (comfy/k_diffusion/sampling.py:45-47,137-143).
 * x_next may alias x.
 */
int cgs_sampler_step_euler(const float* x, const float* denoised, float sigma_hat, float dt,
                           float* x_next, long long n, void* stream);

/*
 * DPM-Solver++(2M) update (S1): x_next = c_x*x - c_d*(c_cur*denoised - c_old*old_denoised),
 * or c_x*x - c_d*denoised when old_denoised is NULL (first step / last step).
 * Replaces the update of sample_dpmpp_2m (comfy/k_diffusion/sampling.py:583-592); the
 * coefficients are the reference's sigma_fn(t_next)/sigma_fn(t), (-h).expm1(), 1 + 1/(2r),
 * 1/(2r), computed by the host once per schedule. x_next may alias x.
 */
int cgs_sampler_step_dpmpp2m(const float* x, const float* denoised, const float* old_denoised,
                             float c_x, float c_d, float c_cur, float c_old,
                             float* x_next, long long n, void* stream);

/*
 * The whole tail in one launch: model output -> denoised (CFG) -> sampler update.
 *   sampler 1 = Euler (c0 = sigma_hat, c1 = dt); 2 = DPM++2M single-term form (c0 = c_x,
 *   c1 = c_d); 3 = DPM++2M two-term form (c0..c3 = c_x, c_d, c_cur, c_old; old_denoised
 *   required). Writes denoised and x_next; x_next may alias x; denoised must not alias
 *   old_denoised.
 */
int cgs_cfg_sampler_tail_fwd(const float* x, const void* out_uncond, const void* out_cond,
                             int out_is_bf16, const float* sigma, float cond_scale,
                             int pred_type, float sigma_data, const float* old_denoised,
                             int sampler, float c0, float c1, float c2, float c3,
                             float* denoised, float* x_next,
                             long long batch, long long per_sample, void* stream);

/*
 * cgs_cfg_sampler_tail_fwd with the step's sampler mode and coefficients read from device memory:
 * step[0] = sampler mode (1, 2 or 3, stored as a float), step[1..4] = c0..c3. The launch is then identical for
 * every step of a schedule, which lets the host replay one CUDA graph of the whole denoise step (UNet input ->
 * UNet -> this kernel) instead of re-deriving the per-step Python scalars of comfy/k_diffusion/sampling.py:583-592.
 * old_denoised must be a valid buffer (read only in mode 3).
 */
int cgs_cfg_sampler_tail_dev_fwd(const float* x, const void* out_uncond, const void* out_cond,
                                 int out_is_bf16, const float* sigma, float cond_scale,
                                 int pred_type, float sigma_data, const float* old_denoised,
                                 const float* step, float* denoised, float* x_next,
                                 long long batch, long long per_sample, void* stream);

/*
 * Tuning switch for the micro-benchmarks (no counterpart in the reference): which mainloop cgs_geglu_gemm_fwd /
 * cgs_linear_fwd use -- 0 automatic (default), 1 one CTA per M128 x N256 tile, 2 CTA pairs (cta_group::2) on
 * M256 x N256 tiles. Both compute every output element from the same products in the same order: results are identical.
 * Further codes, all result-neutral as well:
 *   10 / 11 / 12   convolution tiles: automatic / 256 columns / 320 columns where Cout % 320 == 0
 *   20 / 21 / 22   the convolution's last, partly filled round of tiles: whole tiles (default) / 64-column groups dealt
 *                  out evenly / halves or quarters, at most one per worker
 *   30 / 31 / 32   the same for cgs_linear_fwd: 64-column groups / whole tiles / halves or quarters (default)
 * Returns CGS_BAD_ARGUMENT for any other value. Process-wide; meant for A/B measurements (tools/kernel_bench.py,
 * bench.py --gemm-set), not for production code.
 */
int cgs_gemm_set_variant(int variant);

/*
 * 3x3 / stride 1 / padding 1 convolution over channels-last bf16 images as an implicit GEMM on the tensor cores, with
 * bias and a residual image added in the epilogue. Replaces the ResBlock and Upsample convolutions
 *   h = self.in_layers(x); h = self.out_layers(h); return self.skip_connection(x) + h
 * (comfy/ldm/modules/diffusionmodules/openaimodel.py:83-96 Upsample.conv, :177-232 ResBlock layers, :235-264
 * ResBlock._forward; operator class comfy/ops.py:67-80 Conv2d -> F.conv2d), which the reference runs on cuDNN.
 *   x        [n_img, H, W, Cin]  bf16, `ldx` elements between consecutive pixels (>= Cin; rows and images dense)
 *   w        [Cout, 3, 3, Cin]   bf16 contiguous: the channels-last storage of the reference's [Cout, Cin, 3, 3] weight
 *   bias     [Cout] bf16 or NULL; residual [n_img, H, W, Cout] bf16 (pixel stride ldr) or NULL
 *   y        [n_img, H, W, Cout] bf16, pixel stride ldy; y[p, o] = sum_{r,s,c} x[p + (r-1, s-1), c] * w[o, r, s, c]
 *            (zero outside the image) + bias[o] + residual[p, o], fp32 accumulation, one rounding to bf16
 * Cin and the pixel strides must be multiples of 8, Cout of 32. cgs_gemm_set_variant(10 + v) picks the tile shape for
 * Cout % 320 == 0 (v = 0 automatic, 1 tiles of 256 columns, 2 tiles of 320 columns); results do not depend on it.
 */
int cgs_conv3x3_nhwc_fwd(const void* x, const void* w, const void* bias, const void* residual, void* y,
                         int n_img, int H, int W, int Cin, int Cout, long long ldx, long long ldy, long long ldr,
                         void* stream);

/*
 * Upsample.forward with a convolution: x = F.interpolate(x, scale 2, mode="nearest"); x = self.conv(x)
 * (comfy/ldm/modules/diffusionmodules/openaimodel.py:83-96), without the up-sampled image. Output pixel
 * (2i + a, 2j + b) reads only the 2 x 2 input pixels (i + a - 1 + dr, j + b - 1 + ds), dr, ds in {0, 1}; the 3x3 taps
 * that land on the same input pixel are added up beforehand:
 *   w4[a * 2 + b][o][dr][ds][c] = sum over r in R(a, dr), s in R(b, ds) of w[o][c][r][s]     (bf16, summed in fp32)
 *   R(0, 0) = {0}, R(0, 1) = {1, 2}, R(1, 0) = {0, 1}, R(1, 1) = {2}
 * so every phase is a 2 x 2 convolution of x on the same implicit GEMM: 16 instead of 36 multiply-adds per input
 * pixel, and 5/4 of the output's bytes less traffic. Algebraically equal to the reference's result; the pre-summed
 * weights are rounded to bf16 once more (relative 2^-9 per weight, below the rounding of the bf16 output).
 *   x [n_img, H, W, Cin] (pixel stride ldx), w4 [4][Cout][2][2][Cin] contiguous, bias [Cout] or NULL,
 *   y [n_img, 2H, 2W, Cout] (pixel stride ldy). Shape rules as for cgs_conv3x3_nhwc_fwd.
 */
int cgs_upsample2x_conv3x3_nhwc_fwd(const void* x, const void* w4, const void* bias, void* y, int n_img, int H,
                                    int W, int Cin, int Cout, long long ldx, long long ldy, void* stream);

/*
 * Row softmax in place: x[r, 0:cols] <- softmax(scale * x[r, 0:cols]), bf16 storage, fp32 arithmetic. The normalisation
 * step of the VAE decoder's mid-block attention (AttnBlock, one head of d = 512), which runs the reference's
 *   s1 = torch.bmm(q, k) * scale;  s2 = torch.nn.functional.softmax(s1, dim=2);  r1 = torch.bmm(v, s2.permute(0, 2, 1))
 * (comfy/ldm/modules/diffusionmodules/model.py:160-196 slice_attention / normal_attention, in the compute dtype) with the
 * two matrix products on cgs_linear_fwd.
 *   x [rows, cols] bf16, `ld` elements between rows; cols a multiple of 8, at most 32768; scale > 0.
 */
int cgs_softmax_rows_fwd(void* x, long long rows, int cols, long long ld, float scale, void* stream);

/*
 * The same tail for a cond / uncond pair that is split across two GPUs (multi-GPU placement):
 * the exchange of the two halves of the model output and cfg_function + sampler update (comfy/samplers.py:234-247,
 * comfy/k_diffusion/sampling.py:137-143, 583-592) in ONE launch, instead of a collective followed by the tail.
 * This rank evaluated `out_local` (bf16; the cond half when local_is_cond != 0, else the uncond half). Every CTA
 * pushes its chunk of out_local into the partner's receive buffer `peer_data` (a peer mapping of the partner's
 * `recv_data`; plain global stores over NVLink), raises `peer_flags[chunk]`, waits for `recv_flags[chunk]` and then
 * reads the partner's half from its own `recv_data`.
 *   recv_data / peer_data   2 slots x (batch * per_sample) bf16 (slot = sequence number & 1)
 *   recv_flags / peer_flags int32 per chunk of 1024 latent elements (`flag_capacity` of them), zeroed once
 *   step[0..4] as in cgs_cfg_sampler_tail_dev_fwd; step[5] = bit pattern of the int32 sequence number of this
 *   exchange, counted 0, 1, 2, ... identically on both ranks for the life of the buffers.
 * All chunks (batch * per_sample / 1024) must be resident at once: CGS_UNSUPPORTED_SHAPE beyond 8 x SM count.
 * A partner that does not arrive within ~30 s traps (the launch fails) instead of hanging the GPU.
 */
int cgs_cfg_sampler_tail_peer_fwd(const float* x, const void* out_local, int local_is_cond,
                                  const void* recv_data, void* peer_data, const int* recv_flags, int* peer_flags,
                                  long long flag_capacity, const float* sigma, float cond_scale, int pred_type,
                                  float sigma_data, const float* old_denoised, const float* step, float* denoised,
                                  float* x_next, long long batch, long long per_sample, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CGS_B200_H */
