/*
 * b200dce.h — C ABI of the B200-native (sm_100a) Zero-DCE / Zero-DCE++ / SCI
 * inference hot path.
 *
 * This is the drop-in boundary: plain pointers and sizes, no torch types.  Each
 * entry point cites the reference code it replaces (paths relative to the
 * reference repository root, SKKUAutoLab/ETSS-07-ImageEnhance).
 *
 * Conventions
 *   - All image tensors are NCHW, contiguous, fp32, values nominally in [0,1],
 *     device pointers in the current CUDA context.  16-byte alignment is
 *     required for every tensor base pointer (B200_ERR_MISALIGNED otherwise).
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default
 *     stream).  Every call only ENQUEUES work on that stream: the library never
 *     allocates, frees or synchronises, and holds no mutable global state, so it
 *     is re-entrant and thread safe.  The caller owns every buffer, including
 *     the workspace, and must keep them alive until the stream has drained.
 *   - Return value: 0 (B200_OK) or a negative B200_ERR_* code;
 *     b200_strerror() gives the text.  Launch failures are reported from
 *     cudaGetLastError() (B200_ERR_LAUNCH).
 *   - There is NO CPU fallback: on a machine without an sm_100 device every
 *     compute entry point returns B200_ERR_LAUNCH / B200_ERR_NO_DEVICE.
 */
#ifndef B200DCE_H_
#define B200DCE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_OK 0
#define B200_ERR_NULL (-1)        /* a required pointer is NULL */
#define B200_ERR_BAD_SHAPE (-2)   /* non-positive dims, H or W not divisible by scale_factor, ... */
#define B200_ERR_MISALIGNED (-3)  /* a tensor base pointer is not 16-byte aligned */
#define B200_ERR_UNSUPPORTED (-4) /* hyper-parameter outside what the kernels implement */
#define B200_ERR_WORKSPACE (-5)   /* workspace too small (see *_workspace_bytes) */
#define B200_ERR_LAUNCH (-6)      /* CUDA launch error */
#define B200_ERR_NO_DEVICE (-7)   /* no sm_100 CUDA device is current */
#define B200_ERR_TOO_LARGE (-8)   /* the problem needs more scratch than a *_static entry point's built-in array holds */

/* precision selector for the Zero-DCE conv stack */
#define B200_PREC_FP32 0   /* fp32 CUDA-core direct convolution (FFMA2), fp32 accumulate            */
#define B200_PREC_BF16 1   /* bf16 operands on tcgen05 tensor cores, fp32 accumulate in TMEM         */
#define B200_PREC_BF16X3 2 /* 3-term split-bf16 (hi*hi + lo*hi + hi*lo) on tcgen05, fp32 accumulate in TMEM: ~fp32 accuracy */
#define B200_PREC_BF16X3_MMA 3 /* the same split on warp-level mma.sync (round-1 kernels; kept for A/B) */

int b200_version(void);
const char* b200_strerror(int code);
/* 0 if the current device is sm_100 (B200), else B200_ERR_NO_DEVICE. */
int b200_check_device(void);

/* ------------------------------------------------------------------------- *
 * Light-enhancement curve  y <- y + a*(y*y - y), `iters` times.
 * ------------------------------------------------------------------------- */

/* Shared map: the same C-channel map `a` [B,C,H,W] is used for every iteration.
 * Replaces src/mon/vision/enhance/llie/zero_dce/zero_dcepp.py:201-203
 * (upstream src/mon_extra/vision/enhance/llie/zero_dce++/model.py:51-62).
 * Bit-exact against ATen on CPU (mul, sub, mul, add with no contraction). */
int b200_curve_shared(const float* x, const float* a, float* y, int B, int C, int H, int W, int iters, void* stream);

/* GCENet's enhancement loop with the brightness attention map: bam = (1 - max(R,G,B))^gamma, then `iters` times
 *   bright = e*(1 - bam); dark = e*bam; e = bright + dark + adjust*(dark*dark - dark)
 * with one map `adjust` [B,3,H,W] for every iteration; `enhanced` [B,3,H,W] is the loop's result BEFORE the guided
 * filter.  Replaces src/mon/vision/enhance/llie/gcenet/gcenet.py:469-484 and
 * src/mon/nn/modules/prior/image.py:269-272 (BrightnessAttentionMap without the optional median blur).
 * gamma == 0 selects the reference's plain branch (gcenet.py:470-475 = b200_curve_shared); bam / bright / dark are
 * then not written.  bam [B,1,H,W], bright / dark [B,3,H,W] (the last iteration's; the reference's debug outputs)
 * may be NULL.  All arithmetic but the pow is rounded as ATen rounds it; 16-byte aligned bases. */
int b200_gce_curve(const float* image, const float* adjust, float* enhanced, float* bam, float* bright, float* dark,
                   int B, int H, int W, int iters, float gamma, void* stream);

/* GuidedFilter.forward(image = x, guide = y) at full resolution: out [B,C,H,W] = (box(A)/N) * x + box(b)/N with
 * A = cov(x,y) / (var(x) + eps), b = mean_y - A * mean_x over (2 * box_radius + 1)^2 windows clipped at the borders.
 * Replaces src/mon/vision/filtering/guided_filter.py:99-146 (box filter: box_filter.py:26-117, whose window sums are
 * differences of fp32-rounded running sums accumulated in double: repeated here exactly), the step GCENet applies to
 * the result of its curve loop (gcenet.py:391,485: GuidedFilter(radius=3, eps=1e-4)).  NOTE the reference class passes
 * its `radius` to BoxFilter positionally, i.e. as KERNEL SIZE: box_radius = (radius - 1) / 2 (1 for GCENet).
 * x is [B,Cx,H,W] with Cx == C or Cx == 1 (guided_filter.py:124); box_radius >= 1, H, W in [2r+1, 4096]. */
size_t b200_guided_filter_workspace_bytes(int B, int C, int H, int W);
int b200_guided_filter(const float* x, const float* y, float* out, int B, int Cx, int C, int H, int W, int box_radius,
                       float eps, void* workspace, size_t workspace_bytes, void* stream);

/* Per-iteration maps: `a` is [B, C*iters, H, W]; iteration i uses channels
 * [i*C, (i+1)*C).  Replaces zero_dce.py:183-188 (torch.split + loop). */
int b200_curve_per_iter(const float* x, const float* a, float* y, int B, int C, int H, int W, int iters, void* stream);

/* ------------------------------------------------------------------------- *
 * Zero-DCE++  (ZeroDCEpp_RE.forward, zero_dcepp.py:182-208; for scale_factor
 * != 1 the intended semantics of upstream zero_dce++/model.py:39,64-82).
 * Architecture: in_channels=3, num_channels=32 (the reference default).
 * ------------------------------------------------------------------------- */

/* Number of floats in the packed weight buffer. */
size_t b200_dcepp_packed_floats(void);

/* Re-pack the 28 state_dict tensors (device pointers, fp32, the layouts of
 * e_conv{i}.dw_conv.weight [Cin,1,3,3], .dw_conv.bias [Cin], .pw_conv.weight
 * [Cout,Cin,1,1], .pw_conv.bias [Cout]; src/mon/nn/modules/conv/dsconv.py:122-149)
 * into `packed` (device, b200_dcepp_packed_floats() floats).  The four arrays
 * are HOST arrays of 7 device pointers, layer order e_conv1..7. */
int b200_dcepp_pack_weights(const float* const* dw_w, const float* const* dw_b, const float* const* pw_w,
                            const float* const* pw_b, float* packed, void* stream);

size_t b200_dcepp_workspace_bytes(int B, int H, int W, int scale_factor);

/* x [B,3,H,W] -> enhanced [B,3,H,W] and (optional, may be NULL) adjust
 * [B,3,H,W] (the full-resolution curve map the reference returns).
 * scale_factor: integer >= 1; H and W must be divisible by it (the reference
 * raises a broadcast error otherwise).  Stages:
 *   a3  bilinear downsample, align_corners=False (zero_dcepp.py:186-188) — fused into layer 1's load
 *   a4  7x [depthwise 3x3 + pointwise 1x1] + ReLU/tanh (zero_dcepp.py:190-196)
 *   a5  bilinear upsample, align_corners=True (upstream model.py:39,80)   \  one fused full-resolution pass:
 *   a6  8-iteration curve (zero_dcepp.py:201-203)                         /  reads x once, writes each output once
 * Two launches for scale_factor >= 4 and W % 4 == 0 (layer 1; then ONE persistent launch whose work queue holds the
 * tiles of layers 2..7 and the blocks of the full-resolution pass), three otherwise; both orders of launch give
 * bit-identical results. */
int b200_dcepp_forward(const float* x, const float* packed, float* enhanced, float* adjust_or_null, int B, int H, int W,
                       int scale_factor, int iters, void* workspace, size_t workspace_bytes, void* stream);

/* Measurement aid: b200_dcepp_forward, which additionally records the caller's cudaEvent_t `mid_event` (as void*)
 * on `stream` in front of the forward's LAST launch -- the one that carries the HBM-bound full-resolution pass
 * (scale_factor >= 4: the persistent launch with layers 2..7 + the full-resolution blocks; otherwise the separate
 * full-resolution kernel) -- so that launch can be timed with CUDA events for the roofline (bench.py). */
int b200_dcepp_forward_marked(const float* x, const float* packed, float* enhanced, float* adjust_or_null, int B, int H,
                              int W, int scale_factor, int iters, void* workspace, size_t workspace_bytes, void* stream,
                              void* mid_event);

/* b200_dcepp_forward WITHOUT a workspace argument and without any allocation: the scratch (feature maps, low-resolution
 * curve map, tile flags) lives in a file-scope `static __device__` array of B200_DCEPP_STATIC_SCRATCH_BYTES inside the
 * library, which every GPU that loads the library holds until the process exits.  This entry point has NO counterpart
 * in the reference; it was added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  It calls no cudaMalloc / cudaFree and no
 * allocating library routine; inputs and outputs are the caller's.  Same results as b200_dcepp_forward bit for bit.
 *   - Limit: b200_dcepp_workspace_bytes(B, H, W, scale_factor) must be <= B200_DCEPP_STATIC_SCRATCH_BYTES (16 MiB:
 *     one 1080p frame at scale_factor 12 needs 9.0 MiB, one 4K frame at 12 needs 35.8 MiB and does not fit); beyond
 *     that it returns B200_ERR_TOO_LARGE and launches nothing -- use b200_dcepp_forward with a workspace.
 *   - The array is shared by every call in the process (per GPU) and keeps the previous call's contents; each call
 *     clears the counters and flags it uses itself.  Calls that can overlap on the GPU (different streams, or
 *     several host threads on different streams) RACE on it: keep all calls of this entry point on one stream, or
 *     order them with events. */
#define B200_DCEPP_STATIC_SCRATCH_BYTES (16u << 20)
size_t b200_dcepp_static_scratch_bytes(void);
int b200_dcepp_forward_static(const float* x, const float* packed, float* enhanced, float* adjust_or_null, int B, int H,
                              int W, int scale_factor, int iters, void* stream);

/* Only the low-resolution curve-estimation net (stages a3+a4): writes a_lr
 * [B,3,H/sf,W/sf].  Exposed for tests and for callers that cache the map. */
int b200_dcepp_lowres(const float* x, const float* packed, float* a_lr, int B, int H, int W, int scale_factor,
                      void* workspace, size_t workspace_bytes, void* stream);

/* Only stages a5+a6: a_lr [B,3,h,w] -> enhanced / adjust at [B,3,H,W]. */
int b200_dcepp_apply(const float* x, const float* a_lr, float* enhanced, float* adjust_or_null, int B, int H, int W,
                     int h, int w, int iters, void* stream);

/* ------------------------------------------------------------------------- *
 * Zero-DCE  (ZeroDCE_RE.forward, zero_dce.py:173-191): 7x conv3x3 (3->32,
 * 32->32 x3, 64->32 x2, 64->3*iters) with ReLU / tanh and the per-iteration
 * curve.  in_channels=3, num_channels=32, 1 <= iters <= 8.
 * ------------------------------------------------------------------------- */
size_t b200_dce_packed_bytes(int precision, int iters);

/* w[i]: e_conv{i+1}.weight (OIHW fp32, device), b[i]: e_conv{i+1}.bias.
 * HOST arrays of 7 device pointers. */
int b200_dce_pack_weights(const float* const* w, const float* const* b, void* packed, int precision, int iters,
                          void* stream);

size_t b200_dce_workspace_bytes(int B, int H, int W, int precision);

/* x [B,3,H,W] -> adjust [B,3*iters,H,W] (tanh output, zero_dce.py:182) and
 * enhanced [B,3,H,W] (zero_dce.py:186-188).  Both are required outputs of the
 * reference; `adjust` may be NULL to skip materialising it. */
int b200_dce_forward(const float* x, const void* packed, float* adjust_or_null, float* enhanced, int B, int H, int W,
                     int iters, int precision, void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------- *
 * SCI  (Finetunemodel.forward, src/mon_extra/vision/enhance/llie/sci/model.py:155-159
 * over EnhanceNetwork.forward :36-45, layers=1, channels=3, BatchNorm in eval mode).
 * ------------------------------------------------------------------------- */
#define B200_SCI_PACKED_FLOATS 252

/* HOST-side pack (no CUDA): all pointers are HOST memory.  conv weights are
 * [3,3,3,3] OIHW, biases [3]; bn_* are the BatchNorm2d affine/running stats
 * (folded into the second conv).  `packed_host` receives 252 floats that
 * b200_sci_forward passes to the kernel by value (constant bank). */
int b200_sci_pack_weights_host(const float* in_w, const float* in_b, const float* mid_w, const float* mid_b,
                               const float* bn_gamma, const float* bn_beta, const float* bn_mean, const float* bn_var,
                               float bn_eps, const float* out_w, const float* out_b, float* packed_host);

/* x [B,3,H,W] -> illu [B,3,H,W] (may be NULL) and r [B,3,H,W].
 * `packed_host` is HOST memory (B200_SCI_PACKED_FLOATS floats). */
int b200_sci_forward(const float* x, const float* packed_host, float* illu_or_null, float* r, int B, int H, int W,
                     void* stream);

/* ------------------------------------------------------------------------- *
 * RUAS ("next" row SURVEY §8f-4): upstream Network.forward,
 * src/mon_extra/vision/enhance/llie/ruas/model.py:202-206 = EnhanceNetwork
 * (:117-125, three IEM iterations :84-95) then DenoiseNetwork (:142-149); cells =
 * SearchBlock (:46-66) with the IEM / NRM genotypes of genotypes.py:15-16.
 *
 * `packed_host` (HOST memory, B200_RUAS_PACKED_FLOATS floats) is the plain
 * concatenation, conv weights OIHW then bias, of
 *   for i in 0..2  enhance_net.iems.i.cell.{c1_r.op, c2_d.op, c2_r.op, c3_d.op, c4.op, c5}     (387 floats each)
 *   denoise_net.stem                                                                           (168)
 *   for i in 0..2  denoise_net.nrms.i.{c1_d.op, c1_r.op, c2_d.op, c3_d.op, c3_r.op, c5}        (648 each)
 *   denoise_net.activate.0                                                                     (165)
 * (skip_connect slots own no parameters).
 * ------------------------------------------------------------------------- */
#define B200_RUAS_PACKED_FLOATS 3438

/* Two 6-channel feature maps when the denoise net runs, else 0. */
size_t b200_ruas_workspace_bytes(int B, int H, int W, int with_denoise);

/* x [B,3,H,W] -> u [3][B,3,H,W] (u_list[0..2]), t [3][B,3,H,W] (t_list, may be
 * NULL); with `denoised_or_null` != NULL also denoised [B,3,H,W] (u_list[3]) and
 * optionally the noise map n [B,3,H,W]. */
int b200_ruas_forward(const float* x, const float* packed_host, float* u, float* t_or_null, float* denoised_or_null,
                      float* noise_or_null, int B, int H, int W, void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------- *
 * "next" row (partial): the core of ZeroDCE-V, ZeroDCEV.forward
 * src/mon/vision/enhance/llie/zero_dce/zero_dce_v.py:172-197 up to `image_v_fixed_lr`:
 *   image_v    = V channel of the RGB image (= max over channels)            :175-176
 *   image_v_lr = F.interpolate(image_v, size=(S, S))  (nearest)              :177,204-206
 *   adjust     = tanh(DCE-Net(image_v_lr)), in_channels = 1, `iters` maps    :179-187
 *   image_v_fixed_lr = `iters` curve steps on image_v_lr                     :189-193
 * fp32 CUDA-core convolutions (every product in fp32).  NOT built: the guided-filter upsample (:194,208-218), the
 * HSV decomposition / recomposition (kornia) and the division by the global maximum (:195-197).
 * w[i], b[i]: e_conv{i+1}.weight [cout, cin, 3, 3] / .bias with cin(1) = 1, cout(7) = iters (HOST arrays of 7
 * device pointers).  Outputs: image_v [B,1,H,W], image_v_lr [B,1,S,S], adjust [B,iters,S,S], image_v_fixed_lr [B,1,S,S].
 * ------------------------------------------------------------------------- */
size_t b200_dcev_packed_floats(void);
int b200_dcev_pack_weights(const float* const* w, const float* const* b, float* packed, int iters, void* stream);
size_t b200_dcev_workspace_bytes(int B, int S);
int b200_dcev_core(const float* image_rgb, const float* packed, float* image_v, float* image_v_lr, float* adjust,
                   float* image_v_fixed_lr, int B, int H, int W, int S, int iters, void* workspace,
                   size_t workspace_bytes, void* stream);

/* Tail of ZeroDCEV.forward: image_v_fixed [B,1,H,W] = clip(FastGuidedFilter(radius 1, eps 1e-8)(image_v_lr,
 * image_v_fixed_lr, image_v), 0, 1) and enhanced [B,3,H,W] = hsv_to_rgb(H, S of image_rgb, V = image_v_fixed) divided
 * by the frame's maximum.  Replaces src/mon/vision/enhance/llie/zero_dce/zero_dce_v.py:193-197,208-230,
 * src/mon/vision/filtering/guided_filter.py:171-203 and box_filter.py:26-117 (cumulative-sum box filter: the window
 * sums are differences of fp32-rounded prefixes accumulated in double, repeated here exactly -- A = cov / (var + eps) is
 * ill-conditioned in flat windows), bicubic align_corners=True resize of A and b as ATen does it; the HSV conversions
 * are kornia's (src/mon/core/image/color.py:169-190,262-277 delegate to it) restated from their published form.
 * The reference runs with batch 1 only: a batch is B batch-1 calls (per-frame maximum).  3 < S <= 4096
 * (B200_ERR_UNSUPPORTED otherwise). */
size_t b200_dcev_tail_workspace_bytes(int B, int S);
int b200_dcev_tail(const float* image_rgb, const float* image_v, const float* image_v_lr, const float* image_v_fixed_lr,
                   float* image_v_fixed, float* enhanced, int B, int H, int W, int S, void* workspace,
                   size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------- *
 * "next" rows: u8 HWC <-> fp32 NCHW frame conversion.
 * Replaces src/mon/core/image/utils.py:801-848 (to_image_tensor(normalize=True)
 * -> normalize_image_by_range, src/mon/core/image/photometry.py:262-263:
 * (image - 0.0) * ratio + 0.0 with ratio = 1/255, i.e. ONE fp32 multiply by
 * fl32(1/255) -- bit-exact against the reference, pinned by
 * oracle/pin_against_reference.py / tests/golden/u8_wire.npz) and :766-800
 * (to_image_nparray(denormalize=True): *255 in fp32 then astype(uint8), i.e.
 * truncation; values are clamped to [0,1] first, where the reference's cast is
 * implementation-defined -- identical for the enhancers' [0,1] outputs).
 * Any H, W (a scalar kernel takes over when H*W % 4 != 0 or a pointer is not
 * word / 16-byte aligned).
 * ------------------------------------------------------------------------- */
int b200_u8hwc_to_f32nchw(const uint8_t* src, float* dst, int B, int H, int W, void* stream);
int b200_f32nchw_to_u8hwc(const float* src, uint8_t* dst, int B, int H, int W, void* stream);

/* ------------------------------------------------------------------------- *
 * "next" rows: the resize of VisionModel.infer
 * (src/mon/vision/model.py:100-121): core.resize(v, divisible_by=32) before the
 * forward and core.resize(v, (h0, w0)) after it =
 * src/mon/core/image/geometry/transform.py:104-235 -> kornia resize ->
 * F.interpolate(size=(Hout, Wout), mode="bilinear", align_corners=False,
 * antialias=False).  src [planes,Hin,Win] -> dst [planes,Hout,Wout], fp32.
 * ------------------------------------------------------------------------- */
int b200_resize_bilinear(const float* src, float* dst, int planes, int Hin, int Win, int Hout, int Wout, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200DCE_H_ */
