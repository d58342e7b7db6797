// Shared device-side definitions of the B200 stylisation hot path.
//
// One "conv layer" of this library is the reference's   [upsample] -> pad -> conv2d   group
// (transformer_net.py:44-54, :79-99; transformer_net_nst.py:12-59; model.py:5-15) together with
// the InstanceNorm-apply / ReLU / residual-add that PRECEDE it in the graph (folded into the
// operand load, "prologue") and the InstanceNorm statistics that FOLLOW it (reduced in the
// epilogue).  Activations live in HBM as "raw" conv outputs (pre-norm), fp16 by default (fp32 for
// the tensors around an error-compensated F16X3 layer and in the validation modes), in the
// channel-planar layout [N][C/8][H][W][8] ("planar-8": one 16-byte slot = 8 channels of a pixel,
// slots of a plane contiguous along W -- the unit both the TMA boxes and the UMMA core matrices
// use), plus per-CTA partial (sum, sum of squares) tables that `stats_finalize_kernel` folds in a
// fixed order into the per-(n,c) InstanceNorm (scale, shift) the consumer layer applies.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace nst {

// IN_F32_NHWC / OUT_F32_NHWC_STATS name the intermediate-activation formats (the names predate the
// planar-8 layout); their element type is fp16 when ConvParams::act16 / out16 is set (the default
// NST_PREC_F16 mode stores raw conv outputs as fp16: half the HBM traffic, parity cost ~2 dB PSNR --
// oracle/precision_study.py STORAGE=1), else fp32.
enum InFmt { IN_F32_NHWC = 0, IN_F32_NCHW3 = 1, IN_U8_NHWC3 = 2 };
enum OutFmt { OUT_F32_NHWC_STATS = 0, OUT_F32_NCHW3 = 1, OUT_U8_NHWC3 = 2,
              OUT_F32_01_NCHW3 = 3 };   // out01 = clamp(post-affine(net output)) as fp32 planes: input of the resize-back pass (postproc.cu)
enum MapMode { MAP_IDENTITY = 0, MAP_UP2 = 1, MAP_ZERO_INSERT2 = 2, MAP_PREPAD_REFLECT = 3 };
enum PadMode { PAD_REFLECT = 0, PAD_ZERO = 1, PAD_REPLICATE = 2 };

constexpr float kInEps = 1e-5f;   // torch.nn.InstanceNorm2d default eps

struct ConvParams {
  // ---- source ----
  const void* src;          // IN_F32_NHWC: act planar-8 [N,Cin/8,Hs,Ws,8]; IN_F32_NCHW3: float [N,3,Hs,Ws]; IN_U8_NHWC3: u8 [N,Hs,Ws,3]
  const void* resid;        // act planar-8 [N,Cin/8,Hs,Ws,8] added after norm(+relu), or null
  void* side_out;           // act planar-8 [N,Cin/8,Hs,Ws,8]: the transformed input is stored here (or null)
  int act16;                // element type of the input-side "act" tensors (src, resid, side_out): 1 = fp16, 0 = fp32
  int out16;                // element type of `out` (OUT_F32_NHWC_STATS): 1 = fp16, 0 = fp32
  const float2* in_coeff;   // [N][Cin] InstanceNorm (scale, shift) of src from stats_finalize_kernel, or null (no norm)
  int in_fmt, pre_relu, post_relu;
  float in_scale[3], in_shift[3];
  int swap_rb;
  int N, Hs, Ws, Cin;
  // ---- conv-space -> source mapping ----
  int Hc, Wc;               // extent of the virtual conv input (after up2 / zero-insert / pre-pad)
  int map_mode, prepad, pad_mode;
  int pad_t, pad_l;
  int ks, stride;
  // ---- output ----
  int Ho, Wo, Cout;         // produced extent
  int oy_off, ox_off;       // produced (0,0) sits at this conv-output coordinate (NST crop)
  int co_off, co_pass;      // channel split of a phase layer whose 4*Cout columns exceed TMEM: this launch produces output
                            // channels [co_off, co_off + co_pass) (co_pass = Cout, co_off = 0 for every other layer)
  int phase_mode;           // sub-pixel form of upsample x2 + conv3: (Ho, Wo) is the LOW-RES tile grid, GEMM column = phase*Cout + co,
                            // phase (a,b) of grid position (i,j) is output pixel (2i+a, 2j+b) of the [N, 2Ho, 2Wo, Cout] tensor
  int out_colsplit;         // store every output row as [even columns | odd columns]: the single reader is a stride-2 TMA-fed layer whose
                            // column-parity views are then contiguous in memory (pixel x sits at slot (x & 1) * ((Wo + 1) / 2) + (x >> 1))
  int in_colsplit;          // the source tensor is stored that way
  int out_fmt;
  void* out;                // OUT_F32_NHWC_STATS: act planar-8 [N,Cout/8,Ho,Wo,8]; OUT_F32_NCHW3: float [N,3,Ho,Wo]; OUT_U8: u8 [N,Ho,Wo,3]
  double* out_part;         // [parts][N][Cout][2] per-CTA partial (sum, sumsq); part = blockIdx.x (tcgen05 kernel) or 0 (direct kernel, atomics)
  const float* bias;        // [Cout] (only applied for the image-producing formats)
  int tanh_out;
  float out_div[3], out_scale[3], out_shift[3];     // out01 = clamp((y / div) * scale + shift, 0, 1), each op rounded like the reference's
  const uint8_t* orig;      // u8 NHWC original frame for blend / mask composite (or null)
  const float* mask;        // [N,Ho,Wo] alpha or null
  float blend, blend_comp;  // uniform blend in [0,1) and float(1.0 - blend) computed in double like the reference's Python; a blend
                            // outside [0,1) disables (pipeline.py:2089 `if 0.0 <= blend < 1.0`)
  int composite_mode;
};

// element index of channel c of pixel (y, x) of sample n in a planar-8 tensor [N][C/8][H][W][8]
__host__ __device__ __forceinline__ size_t act_index(int n, int c, int y, int x, int C, int H, int W) {
  return ((((size_t)n * (C >> 3) + (c >> 3)) * H + y) * W + x) * 8 + (c & 7);
}
__device__ __forceinline__ float ld_act(const void* p, size_t i, int act16) {
  return act16 ? __half2float(((const __half*)p)[i]) : ((const float*)p)[i];
}
__device__ __forceinline__ void st_act(void* p, size_t i, float v, int act16) {
  if (act16) ((__half*)p)[i] = __float2half_rn(v); else ((float*)p)[i] = v;
}

// Map one conv-space coordinate u (already minus the conv padding) to a source coordinate.
// Returns -1 when the tap reads a structural zero (zero padding / zero insertion).
__device__ __forceinline__ int map_coord(int u, int Hc, int Hs, int map_mode, int prepad, int pad_mode) {
  if (pad_mode == PAD_REFLECT) {
    if (u < 0) u = -u;
    if (u >= Hc) u = 2 * (Hc - 1) - u;
    u = max(0, min(u, Hc - 1));            // rows of partial tiles beyond the image: any finite value
  } else if (pad_mode == PAD_REPLICATE) {
    u = max(0, min(u, Hc - 1));
  } else {
    if (u < 0 || u >= Hc) return -1;
  }
  if (map_mode == MAP_IDENTITY) return u;
  if (map_mode == MAP_UP2) return u >> 1;                       // nearest: src = floor(dst/2)
  if (map_mode == MAP_ZERO_INSERT2) return (u & 1) ? -1 : (u >> 1);
  // MAP_PREPAD_REFLECT: conv space is ReflectionPad2d(prepad) of the source
  int v = u - prepad;
  if (v < 0) v = -v;
  if (v >= Hs) v = 2 * (Hs - 1) - v;
  return max(0, min(v, Hs - 1));
}

// InstanceNorm scale/shift of channel c of sample n (precomputed by stats_finalize_kernel).
__device__ __forceinline__ void norm_coeff(const ConvParams& p, int n, int c, float& scale, float& shift) {
  if (p.in_coeff == nullptr) { scale = 1.f; shift = 0.f; return; }
  const float2 cf = p.in_coeff[(size_t)n * p.Cin + c];
  scale = cf.x; shift = cf.y;
}

// float(u8) / 255.0f, correctly rounded (ToTensor, pipeline.py:1099) in three instructions instead of a division: q = k * fl(1/255),
// one Newton correction with the exact remainder.  Bit-identical to the IEEE division for every k in 0..255 (checked exhaustively,
// tests/test_oracle.py::test_u8_to_unit_shortcut_is_exact).
__device__ __forceinline__ float u8_to_unit(float k) {
  const float r = __uint_as_float(0x3b808081u);          // fl32(1 / 255)
  const float q = __fmul_rn(k, r);
  return fmaf(fmaf(-q, 255.0f, k), r, q);
}

// Final image epilogue shared by both conv kernels: v[0..2] are the three net outputs (+bias).
// Byte work: every step is the reference's own fp32 operation, individually rounded (no FMA contraction, true
// division), so the uint8 result is bit-identical to pipeline.py:1451-1486 -> :2040-2043 -> :2089-2092 -> ToPILImage
// (`pic.mul(255).byte()`, truncation) for the same fp32 network output.
__device__ __forceinline__ void store_image(const ConvParams& p, int n, int oy, int ox, float v0, float v1, float v2) {
  if (p.tanh_out) { v0 = tanhf(v0); v1 = tanhf(v1); v2 = tanhf(v2); }
  const size_t pix = ((size_t)n * p.Ho + oy) * p.Wo + ox;
  if (p.out_fmt == OUT_F32_NCHW3) {
    float* o = (float*)p.out;
    const size_t plane = (size_t)p.Ho * p.Wo;
    const size_t base = (size_t)n * 3 * plane + (size_t)oy * p.Wo + ox;
    o[base] = v0; o[base + plane] = v1; o[base + 2 * plane] = v2;
    return;
  }
  float s[3] = {v0, v1, v2};
  if (p.out_fmt == OUT_F32_01_NCHW3) {
    float* o = (float*)p.out;
    const size_t plane = (size_t)p.Ho * p.Wo;
    const size_t base = (size_t)n * 3 * plane + (size_t)oy * p.Wo + ox;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float y = p.swap_rb ? s[2 - c] : s[c];
      const float o01 = __fadd_rn(__fmul_rn(__fdiv_rn(y, p.out_div[c]), p.out_scale[c]), p.out_shift[c]);
      o[base + (size_t)c * plane] = fminf(fmaxf(o01, 0.f), 1.f);
    }
    return;
  }
  uint8_t* o = (uint8_t*)p.out + pix * 3;
  float a = -1.f;
  if (p.mask != nullptr) a = p.mask[pix];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float y = p.swap_rb ? s[2 - c] : s[c];
    // out01 = clamp((y / div) * scale + shift): x/1, x*1 and x+0 are exact, so each preset runs exactly its own op sequence
    float o01 = __fadd_rn(__fmul_rn(__fdiv_rn(y, p.out_div[c]), p.out_scale[c]), p.out_shift[c]);
    o01 = fminf(fmaxf(o01, 0.f), 1.f);
    if (p.orig != nullptr) {
      const float og = u8_to_unit((float)p.orig[pix * 3 + c]);                // ToTensor (pipeline.py:1099)
      if (p.mask != nullptr) {                                               // mask alpha-composite (pipeline.py:2040-2043)
        const float na = __fsub_rn(1.0f, a);
        o01 = (p.composite_mode == 0) ? __fadd_rn(__fmul_rn(a, o01), __fmul_rn(na, og)) : __fadd_rn(__fmul_rn(na, o01), __fmul_rn(a, og));
        o01 = fminf(fmaxf(o01, 0.f), 1.f);
      }
      if (p.blend >= 0.f && p.blend < 1.f) {      // "uniform global blend still applies after region composite" (pipeline.py:2088-2092)
        o01 = __fadd_rn(__fmul_rn(p.blend, o01), __fmul_rn(p.blend_comp, og));
        o01 = fminf(fmaxf(o01, 0.f), 1.f);
      }
    }
    o[c] = (uint8_t)__fmul_rn(o01, 255.0f);                                  // truncation (ToPILImage)
  }
}

}  // namespace nst
