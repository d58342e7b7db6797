// Shared device-side definitions of the B200 stylisation hot path.
//
// One "conv layer" of this library is the reference's   [upsample] -> pad -> conv2d   group
// (transformer_net.py:44-54, :79-99; transformer_net_nst.py:12-59; model.py:5-15) together with
// the InstanceNorm-apply / ReLU / residual-add that PRECEDE it in the graph (folded into the
// operand load, "prologue") and the InstanceNorm statistics that FOLLOW it (reduced in the
// epilogue).  Activations live in HBM as fp32 NHWC "raw" conv outputs (pre-norm) plus a
// [N][C][2] double table of (sum, sum of squares).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace nst {

// IN_F32_NHWC / OUT_F32_NHWC_STATS name the intermediate-activation formats; their element type is
// fp32, or fp16 when ConvParams::act16 is set (the default NST_PREC_F16 mode stores raw conv outputs
// as fp16: half the HBM traffic, parity cost ~2 dB PSNR -- oracle/precision_study.py STORAGE=1).
enum InFmt { IN_F32_NHWC = 0, IN_F32_NCHW3 = 1, IN_U8_NHWC3 = 2 };
enum OutFmt { OUT_F32_NHWC_STATS = 0, OUT_F32_NCHW3 = 1, OUT_U8_NHWC3 = 2 };
enum MapMode { MAP_IDENTITY = 0, MAP_UP2 = 1, MAP_ZERO_INSERT2 = 2, MAP_PREPAD_REFLECT = 3 };
enum PadMode { PAD_REFLECT = 0, PAD_ZERO = 1, PAD_REPLICATE = 2 };

constexpr float kInEps = 1e-5f;   // torch.nn.InstanceNorm2d default eps

struct ConvParams {
  // ---- source ----
  const void* src;          // IN_F32_NHWC: act [N,Hs,Ws,Cin]; IN_F32_NCHW3: float [N,3,Hs,Ws]; IN_U8_NHWC3: u8 [N,Hs,Ws,3]
  const void* resid;        // act NHWC [N,Hs,Ws,Cin] added after norm(+relu), or null
  void* side_out;           // act NHWC [N,Hs,Ws,Cin]: the transformed input is stored here (or null)
  int act16;                // element type of the "act" tensors (src, resid, side_out, out): 1 = fp16, 0 = fp32
  const double* in_stats;   // [N][Cin][2] (sum, sumsq) of src, or null (no norm)
  const float* in_gamma;    // [Cin]
  const float* in_beta;     // [Cin]
  float in_inv_count;       // 1 / (Hs*Ws)
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
  int out_fmt;
  void* out;                // OUT_F32_NHWC_STATS: float [N,Ho,Wo,Cout]; OUT_F32_NCHW3: float [N,3,Ho,Wo]; OUT_U8: u8 [N,Ho,Wo,3]
  double* out_stats;        // [N][Cout][2]
  const float* bias;        // [Cout] (only applied for the image-producing formats)
  int tanh_out;
  float out_scale[3], out_shift[3];
  const uint8_t* orig;      // u8 NHWC original frame for blend / mask composite (or null)
  const float* mask;        // [N,Ho,Wo] alpha or null
  float blend;              // uniform blend in [0,1); >= 1 disables
  int composite_mode;
};

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

// InstanceNorm scale/shift of channel c of sample n from the (sum, sumsq) table.
__device__ __forceinline__ void norm_coeff(const ConvParams& p, int n, int c, float& scale, float& shift) {
  if (p.in_stats == nullptr) { scale = 1.f; shift = 0.f; return; }
  const double* st = p.in_stats + ((size_t)n * p.Cin + c) * 2;
  double mean = st[0] * (double)p.in_inv_count;
  double var = st[1] * (double)p.in_inv_count - mean * mean;
  if (var < 0.0) var = 0.0;
  float inv = rsqrtf((float)var + kInEps);
  scale = p.in_gamma[c] * inv;
  shift = p.in_beta[c] - (float)mean * scale;
}

// Final image epilogue shared by both conv kernels: v[0..2] are the three net outputs (+bias).
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
  uint8_t* o = (uint8_t*)p.out + pix * 3;
  float a = -1.f;
  if (p.mask != nullptr) a = p.mask[pix];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float y = p.swap_rb ? s[2 - c] : s[c];
    float o01 = fminf(fmaxf(y * p.out_scale[c] + p.out_shift[c], 0.f), 1.f);
    if (p.orig != nullptr && (p.mask != nullptr || p.blend < 1.f)) {
      float og = (float)p.orig[pix * 3 + c] / 255.0f;            // ToTensor (pipeline.py:1099)
      if (p.mask != nullptr) {
        o01 = (p.composite_mode == 0) ? (a * o01 + (1.f - a) * og) : ((1.f - a) * o01 + a * og);
      } else {
        o01 = p.blend * o01 + (1.f - p.blend) * og;
      }
      o01 = fminf(fmaxf(o01, 0.f), 1.f);
    }
    o[c] = (uint8_t)(o01 * 255.0f);                              // truncation (ToPILImage)
  }
}

}  // namespace nst
