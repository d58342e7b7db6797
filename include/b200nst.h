/*
 * b200nst.h -- C ABI of libb200nst.so: the B200 (sm_100a) per-frame stylisation hot path.
 *
 * The reference (TrentMahaffey/NeuralStyleTransferV1) has no FFI of its own: the seam is the
 * Python nn.Module protocol.  Each entry point below names the reference interface it stands
 * behind.  All pointers are raw; device pointers are marked [dev], host pointers [host].
 * Every function returns 0 on success or a negative nst_status; the message is available from
 * nst_last_error() (thread local).  No function synchronises the device or the stream unless
 * stated; all work is enqueued on the caller's cudaStream_t (passed as void*).
 *
 * A handle is bound to one device and one model; it is not re-entrant.  Distinct handles are
 * independent (one per GPU / worker).
 */
#ifndef B200NST_H_
#define B200NST_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct nst_handle nst_handle;

enum nst_status {
  NST_OK = 0,
  NST_ERR_INVALID = -1,      /* bad argument / unsupported shape */
  NST_ERR_CUDA = -2,         /* CUDA runtime error (message has the cudaError string) */
  NST_ERR_WEIGHTS = -3,      /* missing / mis-shaped tensor in nst_load_weights */
  NST_ERR_STATE = -4,        /* call order violated (e.g. forward before weights) */
  NST_ERR_WORKSPACE = -5     /* workspace too small */
};

/* architectures */
enum nst_arch {
  NST_ARCH_JOHNSON = 0,      /* transformer_net.TransformerNet        (transformer_net.py:4-41)     */
  NST_ARCH_NST = 1,          /* transformer_net_nst.TransformerNet    (transformer_net_nst.py:62-127) */
  NST_ARCH_RECONET = 2       /* model.ReCoNet(frn=False)              (model.py:107-116)            */
};

/* conv operand precision.  Accumulation is fp32 in every mode.  Raw conv outputs are stored as fp16 in NST_PREC_F16 / BF16
 * (fp32 between two error-compensated layers) and as fp32 in F16X3 / FP32_DIRECT. */
enum nst_precision {
  NST_PREC_F16 = 0,          /* fp16 MMA operands, 1 tcgen05.mma term  (default; meets max|d|<=2, PSNR>=45);
                                ReCoNet's three encoder convs run as F16X3 inside this mode (tanh output is x127.5 steep) */
  NST_PREC_F16X3 = 1,        /* fp16 hi+lo split operands, 3 terms (fp32-class accuracy)              */
  NST_PREC_BF16 = 2,         /* bf16 operands, 1 term (north_star wording; misses the tolerance)      */
  NST_PREC_FP32_DIRECT = 3   /* no tensor cores: fp32 FFMA direct-conv kernels (debug / validation)   */
};

/* composite mode of the mask alpha-composite (pipeline.py:2040-2043) */
enum nst_composite_mode { NST_COMPOSITE_KEEP = 0, NST_COMPOSITE_REPLACE = 1 };

/* ---- lifetime ------------------------------------------------------------------------------ */

/* Replaces: `TransformerNet().to(device)` / `ReCoNet().to(device)` (pipeline.py:602,617).
 * Fails with NST_ERR_CUDA when `device` is not an sm_100 GPU: there is no CPU fallback. */
int nst_create(int device, int arch, int precision, nst_handle** out);
void nst_destroy(nst_handle* h);

/* Replaces: `model.load_state_dict(state)` (pipeline.py:568, :91).
 * names[i] are the reference state_dict keys; ptrs[i] are [host] fp32, contiguous,
 * in the reference layout (Conv2d OIHW, ConvTranspose2d IOHW, norm weight/bias [C]);
 * shapes is 4 int64 per tensor (unused trailing dims = 1).  Unknown keys are ignored and counted
 * in *n_unexpected; absent keys are counted in *n_missing (both may be NULL).  Weights are
 * repacked once into library-owned device memory (fp16 hi/lo K-major tiles + fp32 copies).
 * Synchronises the device. */
int nst_load_weights(nst_handle* h, int n_tensors, const char* const* names,
                     const float* const* ptrs, const int64_t* shapes,
                     int* n_missing, int* n_unexpected);

/* Output spatial size for an input of H x W: Johnson/ReCoNet 4*ceil(ceil(H/2)/2), NST H. */
int nst_output_shape(const nst_handle* h, int H, int W, int* Ho, int* Wo);

/* Bytes of [dev] workspace a forward with batch N, H x W needs (activations + statistics). */
int nst_workspace_bytes(nst_handle* h, int N, int H, int W, size_t* bytes);

/* ---- forward ------------------------------------------------------------------------------- */

/* Replaces: `y = model(x_in)` (pipeline.py:1474; transformer_net.py:29-41).
 * x: [dev] fp32 NCHW [N,3,H,W], already io-preset-normalised; y: [dev] fp32 NCHW [N,3,Ho,Wo].
 * workspace: [dev], >= nst_workspace_bytes, 256-byte aligned. */
int nst_forward_f32_nchw(nst_handle* h, const float* x, float* y, int N, int H, int W,
                         void* workspace, size_t workspace_bytes, void* stream);

/* Fused uint8 path; replaces ToTensor -> io_preset pre -> model -> io_preset post -> clamp ->
 * mask composite -> uniform blend -> ToPILImage pack (pipeline.py:1086-1099,1447-1486,
 * 2040-2043,2088-2094; lib.py:27-51).
 *   net_in[c]  = in_scale[c] * u8[src_c] + in_shift[c]                       (src_c = 2-c when swap_rb)
 *   out01[c]   = clamp((net_out[c'] / out_div[c]) * out_scale[c] + out_shift[c], 0, 1)   (c' = 2-c when swap_rb)
 *   if mask:            a = mask[n,h,w]; out01 = clamp(keep ? a*out01+(1-a)*orig : (1-a)*out01+a*orig)
 *   if 0 <= blend < 1:  out01 = clamp(blend*out01 + (1-blend)*orig)          (after the mask composite, pipeline.py:2088-2092;
 *                                                                             orig = u8/255; any other blend disables)
 *   y_u8 = trunc(out01 * 255)
 * Every step after the network is byte work and is evaluated with the reference's own fp32 operation sequence (true division,
 * separately rounded multiplies and adds; `blend` is a double because Python computes 1.0 - blend in double before torch
 * narrows it), so y_u8 is bit-identical to the reference's pack of the same fp32 network output.
 * x: [dev] uint8 NHWC [N,H,W,3]; y: [dev] uint8 NHWC [N,Ho,Wo,3] with (Ho,Wo) = nst_output_shape(H,W) -- equal to (H,W) when
 * H,W are multiples of 4 (always for NST).  Otherwise the frame comes back Ho x Wo like the reference module's output and the
 * caller resizes it (pipeline.py:1512-1516; nst_forward_u8_nhwc_fit does both); mask / blend then require the resize first (NST_ERR_INVALID).
 * mask: [dev] fp32 [N,H,W] or NULL. */
int nst_forward_u8_nhwc(nst_handle* h, const uint8_t* x, uint8_t* y, int N, int H, int W,
                        const float in_scale[3], const float in_shift[3],
                        const float out_div[3], const float out_scale[3], const float out_shift[3], int swap_rb,
                        double blend, const float* mask, int composite_mode,
                        void* workspace, size_t workspace_bytes, void* stream);

/* Same fused uint8 path, but the result is always [N,H,W,3]: when H or W is not a multiple of 4 the generator's larger out01 frame is
 * resized back with the reference's bilinear `F.interpolate(out01, size=(H0, W0), align_corners=False)` BEFORE the mask composite,
 * blend and pack (pipeline.py:1512-1516 "lock stylized to content size").  This is the call a frame loop makes
 * (neuralstyletransferv1_b200/frame_pipeline.py); y: [dev] uint8 NHWC [N,H,W,3]. */
int nst_forward_u8_nhwc_fit(nst_handle* h, const uint8_t* x, uint8_t* y, int N, int H, int W,
                            const float in_scale[3], const float in_shift[3],
                            const float out_div[3], const float out_scale[3], const float out_shift[3], int swap_rb,
                            double blend, const float* mask, int composite_mode,
                            void* workspace, size_t workspace_bytes, void* stream);

/* ---- compositing --------------------------------------------------------------------------- */

/* Replaces: region_blend.composite_regions / composite_regions_advanced
 * (region_blend.py:1049-1108, 1589-1679), single scale:
 *   out = clamp( sum_r m_r * (sum_j w[r][j] * S[idx[r][j]]) / max(sum_r m_r, 1e-6), 0, 1)
 * srcs: K [dev] pointers to uint8 NHWC [H,W,3] frames (value/255); index -1 selects `original`.
 * masks: [dev] fp32 [R,H,W].  region_idx / region_w: [host] R*J entries (J = terms per region).
 * out: [dev] uint8 NHWC (truncating pack).  R <= 16, J <= 4, K <= 8. */
int nst_composite_regions_u8(const uint8_t* const* srcs, int K, const uint8_t* original,
                             const float* masks, int R, int J,
                             const int* region_idx, const float* region_w,
                             uint8_t* out, int H, int W, void* stream);

/* Same, fp32 CHW planes in 0..1 in and out (the reference's tensor form). */
int nst_composite_regions_f32(const float* const* srcs, int K, const float* original,
                              const float* masks, int R, int J,
                              const int* region_idx, const float* region_w,
                              float* out, int H, int W, void* stream);

/* Batched form of the two above (one launch for N frames), replacing the per-frame Python loop around composite_regions /
 * composite_regions_advanced (pipeline.py:1630-1700; region_blend.py:1049-1108, 1589-1679 including the per-scale bilinear upscale of
 * :1662-1668).  srcs[k]: [dev] uint8 [N, src_hw[2k], src_hw[2k+1], 3] (src_hw NULL = every source is H x W); a smaller source is
 * upscaled per pixel like F.interpolate(bilinear, align_corners=False).  masks: [dev] [R,H,W] shared by the batch
 * (masks_per_frame = 0) or [N,R,H,W]; mask_type 0 = fp32, 1 = fp16, 2 = uint8 (value / 255).  region_idx: -1 = original, -2 = unused
 * padding entry of a region with fewer than J terms. */
int nst_composite_regions_batch_u8(const uint8_t* const* srcs, const int* src_hw, int K, const uint8_t* original,
                                   const void* masks, int mask_type, int masks_per_frame, int R, int J,
                                   const int* region_idx, const float* region_w,
                                   uint8_t* out, int N, int H, int W, void* stream);

/* Replaces: region_blend.composite_from_crops + place_crop (region_blend.py:2186-2294, 2080-2109) -- the compositing step of crop-based
 * region styling (pipeline.py:1293-1400).  Region r covers the padded bounding box bbox[4r..4r+3] = (x1, y1, x2, y2) of the H x W canvas and
 * blends nterms[r] <= 4 sources: term_src[4r+j] is a [dev] fp32 [3, term_hw[2(4r+j)], term_hw[2(4r+j)+1]] styled crop (0..1), or NULL for
 * the original frame's crop; a crop whose size differs from the box is resized bilinearly like the reference does.  crop_masks[r]:
 * [dev] fp32 [mask_hw[2r], mask_hw[2r+1]].  canvas += (sum_j w_j * src_j) * mask; weight += mask; where weight < 0.1 the original fills
 * the gap (or, without an original, three max-pool dilations 5 / 11 / 21 -- `scratch` = 8*H*W floats [dev] is then required);
 * out = clamp(canvas / max(weight, 1e-6), 0, 1) as fp32 [3,H,W]; out_u8 (optional) receives its truncating uint8 HWC pack. */
int nst_composite_from_crops_f32(int R, const int* bbox, const float* const* crop_masks, const int* mask_hw, const int* nterms,
                                 const float* const* term_src, const int* term_hw, const float* term_w,
                                 const float* original, float* out, uint8_t* out_u8, float* scratch, int H, int W, void* stream);

/* Replaces: the tail of the per-frame loop on an fp32 out01 frame (pipeline.py:1512-1516 resize to the content size, :2040-2043 mask
 * composite, :2088-2092 blend, :2094 ToPILImage pack).  src01: [dev] fp32 [N,3,Hi,Wi] in 0..1; out: [dev] uint8 [N,H,W,3]; original
 * ([dev] uint8 [N,H,W,3]) is needed when mask ([dev] fp32 [N,H,W]) or a blend in [0,1) is given.  Hi x Wi == H x W is the identity resize. */
int nst_resize_post_pack_u8(const float* src01, int N, int Hi, int Wi, const uint8_t* original, const float* mask, double blend,
                            int composite_mode, uint8_t* out, int H, int W, void* stream);

/* ---- region masks --------------------------------------------------------------------------- */

/* Replaces: region_blend.voronoi_masks + feather_mask / gaussian_blur_mask (region_blend.py:174-236, 69-102) for given seed points:
 * mask_i = GaussianBlur_reflect(argmin_j dist(pixel, point_j) == i), sigma = feather / 3.  points_xy: [host] R (x, y) pairs (the host
 * draws them exactly like `_generate_balanced_points`, neuralstyletransferv1_b200/region_masks.py); region_weights: [host] R relative
 * region sizes or NULL.  masks: [dev] fp32 [R,H,W]; scratch: [dev] H*W bytes + R*H*W floats.  The reference convolves with the 2-D
 * outer-product kernel, this is the separable form: equal to ~1e-6. */
int nst_voronoi_masks_f32(const float* points_xy, const float* region_weights, int R, int feather, float* masks, float* scratch,
                          int H, int W, void* stream);

/* Replaces: the other generators of region_blend.generate_region_masks (region_blend.py:109-172, 307-521, 925-981) for scalars the host
 * has drawn with the reference's `random.Random` sequence (neuralstyletransferv1_b200/region_masks.py: pattern_spec).
 * mode 0 = rectangles (grid_masks, fractal_quad_masks): rects = R x (y1, y2, x1, x2) [host];
 * mode 1 = diagonal (params6[0] = 1: x + y, 0: (W-1-x) + y), 2 = radial (cx, cy, rotation), 3 = waves (frequency, amplitude,
 * direction 0/1/2 = horizontal / vertical / diagonal, phase), 4 = spiral (cx, cy, tightness, rotation, max(H, W)), 5 = concentric (cx, cy):
 * a scalar field per pixel evaluated with the reference's fp32 operation sequence, normalised over the frame (norm 0 = none, 1 = divide by
 * the maximum, 2 = (v - min) / (max - min + 1e-6)), then mask_i = feather(lo[i] <= field < hi[i]); params6 / lo / hi: [host] fp32.
 * masks: [dev] fp32 [R,H,W]; scratch: [dev] nst_mask_scratch_floats(R, H, W) floats.  atan2f / sinf are the device's (<= 2 ulp from the
 * host's): a pixel whose field lies within an ulp of a band edge may land in the neighbouring band. */
int nst_pattern_masks_f32(int mode, int norm, const float* params6, const float* lo, const float* hi, const int* rects, int R, int feather,
                          float* masks, float* scratch, int H, int W, void* stream);

/* Replaces: region_blend.rotate_all_masks (region_blend.py:25-66): every mask through cv2.warpAffine(getRotationMatrix2D((W/2, H/2), angle, 1),
 * INTER_LINEAR, BORDER_REPLICATE) -- OpenCV's 1/1024 fixed-point coordinates rounded to 1/32 pixel, fp32 table weights, bit-exact -- and
 * divided by the clamped sum of the rotated masks, in one pass.  masks / out: [dev] fp32 [R,H,W], not aliased. */
int nst_rotate_masks_f32(const float* masks, float* out, int R, int H, int W, double angle_degrees, void* stream);

/* Replaces: region_blend.warp_all_masks_organic / warp_mask_organic / _generate_flow_field / _simplex_noise_2d (region_blend.py:604-810).
 * mode 0 = blob, 1 = tentacle, 2 = wave, 3 = pulse.  Region r is displaced by its own flow field (noise modes: two 2-octave sin / cos
 * fields evaluated in double like numpy, min / max normalised over the frame; offsets: [host] 8 doubles per region = (offset_x, offset_y)
 * of octave 0, 1 of the x field then of the y field, drawn by the host from numpy's default_rng like the reference) and resampled with
 * cv2.remap(INTER_LINEAR, BORDER_REFLECT) arithmetic; pixels whose coverage sum is < 0.1 are filled by max-pool dilations 5 / 11 / 21 / 41
 * (passes are skipped on the device once no gap is left); finally m / clamp(sum, 1e-6).  time_offset = frame_idx * speed * 0.02.
 * masks / out: [dev] fp32 [R,H,W], not aliased; scratch: [dev] nst_mask_scratch_floats(R, H, W) floats. */
int nst_warp_masks_organic_f32(const float* masks, float* out, int R, int H, int W, int mode, double amplitude, double frequency,
                               double time_offset, const double* offsets, float* scratch, void* stream);

/* Replaces: region_blend.feather_mask / gaussian_blur_mask (region_blend.py:69-102) applied to soft masks -- the re-feathering that
 * blend_by_regions does after a rotation (feather / 2) or an organic warp (max(5, feather / 4)) (region_blend.py:1761-1770): reflect-padded
 * Gaussian, sigma = feather / 3, separable form (equal to the reference's 2-D kernel to ~1e-6), in place.  masks: [dev] fp32 [R,H,W];
 * scratch: [dev] R*H*W floats; feather <= 0 is a no-op. */
int nst_feather_masks_f32(float* masks, int R, int feather, float* scratch, int H, int W, void* stream);

/* Replaces: the reduction of region_blend.compute_mask_bbox (region_blend.py:1969-1994), the first step of prepare_region_crops (:2019-2059),
 * for R masks at once: boxes_dev[4r..4r+3] = (min x, min y, max x, max y) over the pixels with mask > threshold (compared in double, like numpy's
 * fp32 array against the Python float 0.01), or
 * (INT_MAX, INT_MAX, -1, -1) for an empty mask (the caller then takes the full frame, like the reference).  masks: [dev] fp32 [R,H,W];
 * boxes_dev: [dev] 4*R ints.  The host adds +1 to the maxima and the padding (neuralstyletransferv1_b200/region_crops.py). */
int nst_mask_bboxes(const float* masks, int R, int H, int W, double threshold, int* boxes_dev, void* stream);

/* Scratch (in floats) the two mask calls above need. */
size_t nst_mask_scratch_floats(int R, int H, int W);

/* ---- temporal post-processing --------------------------------------------------------------- */

/* Replaces: `_warp_with_flow` + the flow-guided EMA lerp (pipeline.py:425-439, 1918-1931): warp the previous fused frame along the
 * optical flow with cv2.remap's bilinear rule (map quantised to 1/32 pixel, replicate border) and blend
 * out = clip(alpha * curr + (1 - alpha) * warped, 0, 1).  All images [dev] fp32 HWC [H,W,3] in 0..1, flow [dev] fp32 [H,W,2] = (dx, dy);
 * alpha is clamped to [0,1].  out may alias curr, not prev.  (The flow estimate itself stays in OpenCV) */
int nst_flow_ema_f32(const float* curr, const float* prev, const float* flow, float* out, int H, int W, double alpha, void* stream);

/* Replaces: the LAB lightness / chroma EMA of pipeline.py:1938-1975 for N consecutive frames in one launch (each pixel's EMA is
 * independent, so one thread carries its (L, a, b) state through the chunk).  frames / out: [dev] uint8 [N,H,W,3] RGB.  rgb2lab / lab2rgb:
 * [dev] uint32 [2^24] tables of Pillow's `convert("LAB")` / `convert("RGB")` (index c0 << 16 | c1 << 8 | c2 -> c0' | c1' << 8 | c2' << 16),
 * filled by the host from Pillow itself (neuralstyletransferv1_b200/temporal.py).  prev_L / prev_a / prev_b: [dev] fp32 [H,W] running
 * state, read unless `init` (first frame of a clip: prev = current) and always written back.
 *   L_sm = fp32(alpha) * L + fp32(1 - alpha) * prev_L;  prev_L = L_sm;  L8 = trunc(clip(L_sm, 0, 255));  same for a, b with chroma_alpha. */
int nst_lab_ema_u8(const uint8_t* frames, uint8_t* out, int N, int H, int W, const uint32_t* rgb2lab, const uint32_t* lab2rgb,
                   float* prev_L, float* prev_a, float* prev_b, int init, int smooth_lightness, int smooth_chroma,
                   double smooth_alpha, double chroma_alpha, void* stream);

/* Replaces: the blend stage of the per-frame loop on a frame that is already uint8 (after the LAB smoothing): `to_tensor(out_img)`
 * (pipeline.py:1981), the mask composite (:2040-2043), the uniform blend that still applies after it (:2088-2092) and the ToPILImage pack
 * (:2094), N frames per launch.  styled / original / out: [dev] uint8 [N,H,W,3]; mask: [dev] fp32 [N,H,W] or NULL; blend outside [0, 1)
 * switches the uniform blend off; composite_mode 0 = keep (alpha * S + (1 - alpha) * O), 1 = replace.  Bit-exact (true /255, separately
 * rounded multiplies and adds). */
int nst_blend_pack_u8(const uint8_t* styled, const uint8_t* original, const float* mask, double blend, int composite_mode, uint8_t* out,
                      int N, int H, int W, void* stream);

/* Replaces: the motion-adaptive blend of pipeline.py:2072-2086 (MOTION_NORM / MIN_ALPHA / GAUSS_SIGMA of :1075-1077 are arguments):
 *   m = cv2.GaussianBlur(clip(|flow| / motion_norm, 0, 1), (0, 0), sigma);  alpha = blend - (blend - min_alpha) * m;
 *   out = trunc(clip(alpha * styled/255 + (1 - alpha) * original/255, 0, 1) * 255)
 * with OpenCV's fp32 filter arithmetic (row pass: sequential FMA; column pass: symmetric pairs, FMA; scalar tails of the last W % 4 resp.
 * W % 8 columns without FMA -- the AVX2 build's order).  styled / original / out: [dev] uint8 [H,W,3]; flow: [dev] fp32 [H,W,2];
 * scratch: [dev] H*W floats. */
int nst_motion_blend_u8(const uint8_t* styled, const uint8_t* original, const float* flow, uint8_t* out, float* scratch, int H, int W,
                        double blend, double min_alpha, double motion_norm, double sigma, void* stream);

/* No counterpart in the reference: added for callers that cannot allocate device memory (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same computation and same results as nst_motion_blend_u8,
 * but the H*W-float scratch plane is a file-scope `static __device__` array inside the library (csrc/temporal.cu) instead of a caller
 * buffer: this call performs no cudaMalloc / cudaFree and calls no allocating routine.  Limits and rules:
 *   - H * W <= NST_MOTION_STATIC_MAX_PIXELS (2^21 = 2 097 152 pixels: up to 1920x1088); larger frames return NST_ERR_INVALID -- use
 *     nst_motion_blend_u8 with caller scratch.  Every GPU that loads the library holds the 8 MiB array until the process exits.
 *   - the device must be current; the first call on each device resolves the array's device address (cudaGetSymbolAddress, cached, no
 *     allocation) -- make that first call outside a graph capture.
 *   - every call shares the one array per device and it keeps the last call's contents (each call overwrites all the entries it reads, so
 *     nothing needs clearing): calls that overlap on the same GPU (different streams) race; serialise them or use nst_motion_blend_u8. */
#define NST_MOTION_STATIC_MAX_PIXELS (1 << 21)
int nst_motion_blend_u8_noalloc(const uint8_t* styled, const uint8_t* original, const float* flow, uint8_t* out, int H, int W,
                                double blend, double min_alpha, double motion_norm, double sigma, void* stream);

/* ---- introspection ------------------------------------------------------------------------- */

/* Number of conv layers of the model, and per-layer facts of the last plan (after a forward or
 * nst_workspace_bytes): algorithmic FLOPs (2*Ho*Wo*Cout*Cin*k*k*N), algorithmic bytes, and which
 * kernel runs it (0 = direct fp32, 1 = tcgen05 with per-thread operand gather, 2 = tcgen05 with TMA-fed operands).
 * Arrays are caller-allocated, length >= n. */
int nst_num_layers(const nst_handle* h);
int nst_layer_info(const nst_handle* h, int layer, char* name, int name_cap,
                   double* flops, double* bytes, int* kernel_kind);

/* Time each layer of one forward with CUDA events on `stream` (synchronises).  ms has
 * nst_num_layers entries. */
int nst_profile_layers_u8(nst_handle* h, const uint8_t* x, uint8_t* y, int N, int H, int W,
                          void* workspace, size_t workspace_bytes, void* stream, float* ms);

/* Host-only: write a human-readable table of the per-layer tiling the planner picks for this shape
 * (one line per layer) into out[cap].  Needs no GPU. */
int nst_describe_plan(int arch, int precision, int N, int H, int W, char* out, int cap);

/* Number of kernels this library has launched on behalf of `h` since creation. */
long long nst_launch_count(const nst_handle* h);

const char* nst_last_error(void);
const char* nst_version(void);

#ifdef __cplusplus
}
#endif
#endif /* B200NST_H_ */
