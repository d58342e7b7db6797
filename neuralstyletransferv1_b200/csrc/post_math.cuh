// Per-pixel arithmetic of the region-mask and temporal post-processing kernels (csrc/masks.cu, csrc/temporal.cu), written once as
// host+device inline functions.  The device kernels are the product; tests/hostcheck compiles the SAME functions with g++ to check
// the operation order against the reference's fixtures on a machine without a GPU (test infrastructure, never shipped in the .so).
//
// Every step is the reference's own operation, individually rounded: torch / numpy evaluate tensor expressions op by op, OpenCV's
// warpAffine / remap work in 1/32-pixel fixed point with fp32 table weights.  Nothing here may be contracted into an FMA unless the
// reference does so itself (cv2.GaussianBlur's AVX2 row / column filters do; see motion blend).
#pragma once
#include <cmath>
#include <cstdint>

#ifdef __CUDACC__
#define PM_HD __host__ __device__ __forceinline__
#else
#define PM_HD inline
#endif

namespace nst {
namespace pm {

#ifdef __CUDA_ARCH__
PM_HD float mul(float a, float b) { return __fmul_rn(a, b); }
PM_HD float add(float a, float b) { return __fadd_rn(a, b); }
PM_HD float sub(float a, float b) { return __fsub_rn(a, b); }
PM_HD float div(float a, float b) { return __fdiv_rn(a, b); }
PM_HD float fsqrt(float a) { return __fsqrt_rn(a); }
PM_HD double dmul(double a, double b) { return __dmul_rn(a, b); }
PM_HD double dadd(double a, double b) { return __dadd_rn(a, b); }
PM_HD double dsub(double a, double b) { return __dsub_rn(a, b); }
PM_HD int rint_f(float a) { return __float2int_rn(a); }
PM_HD int rint_d(double a) { return __double2int_rn(a); }
#else
PM_HD float mul(float a, float b) { return a * b; }
PM_HD float add(float a, float b) { return a + b; }
PM_HD float sub(float a, float b) { return a - b; }
PM_HD float div(float a, float b) { return a / b; }
PM_HD float fsqrt(float a) { return sqrtf(a); }
PM_HD double dmul(double a, double b) { return a * b; }
PM_HD double dadd(double a, double b) { return a + b; }
PM_HD double dsub(double a, double b) { return a - b; }
PM_HD int rint_f(float a) { return (int)lrintf(a); }
PM_HD int rint_d(double a) { return (int)lrint(a); }
#endif
PM_HD int imin(int a, int b) { return a < b ? a : b; }
PM_HD int imax(int a, int b) { return a > b ? a : b; }

// ---------------------------------------------------------------------------------------------
// region patterns (region_blend.py:109-521): one scalar field per pixel, thresholded into bands
// ---------------------------------------------------------------------------------------------
enum { PAT_RECTS = 0, PAT_DIAGONAL = 1, PAT_RADIAL = 2, PAT_WAVES = 3, PAT_SPIRAL = 4, PAT_CONCENTRIC = 5 };
enum { NORM_NONE = 0, NORM_DIV_MAX = 1, NORM_MINMAX = 2 };
constexpr int PAT_MAX_R = 16;
struct PatternParams {
  int mode, norm, R, H, W;
  float p[6];                       // mode-specific scalars, already narrowed to fp32 the way torch narrows Python scalars
  float lo[PAT_MAX_R], hi[PAT_MAX_R];   // band i = lo[i] <= field < hi[i]
  int rect[PAT_MAX_R * 4];          // PAT_RECTS: (y1, y2, x1, x2) per region
};
constexpr float PI32 = 3.14159274101257324f;       // fp32(math.pi)
constexpr float TWOPI32 = 6.28318548202514648f;    // fp32(2 * math.pi)

// torch.remainder for floats (ATen BinaryOpsKernel: fmod, then shifted into the divisor's sign)
PM_HD float remainder_f(float a, float b) {
  float m = fmodf(a, b);
  if (m != 0.f && ((b < 0.f) != (m < 0.f))) m = add(m, b);
  return m;
}

// the un-normalised field of the analytic patterns
PM_HD float pattern_field(const PatternParams& q, int y, int x) {
  switch (q.mode) {
    case PAT_DIAGONAL:                                                              // region_blend.py:150-159
      return q.p[0] != 0.f ? add((float)x, (float)y) : add(sub((float)(q.W - 1), (float)x), (float)y);
    case PAT_RADIAL: {                                                              // :377-389  p = (cx, cy, rotation)
      const float yc = sub((float)y, q.p[1]), xc = sub((float)x, q.p[0]);
      const float ang = add(add(atan2f(yc, xc), PI32), q.p[2]);
      return remainder_f(ang, TWOPI32);
    }
    case PAT_WAVES: {                                                               // :414-436  p = (frequency, amplitude, direction, phase)
      const float yn = div((float)y, (float)q.H), xn = div((float)x, (float)q.W);
      const int dir = (int)q.p[2];
      const float t = dir == 0 ? xn : (dir == 1 ? yn : div(add(xn, yn), 2.f));
      const float wave = mul(sinf(add(mul(mul(mul(t, q.p[0]), 2.f), PI32), q.p[3])), q.p[1]);
      return add(dir == 0 ? yn : (dir == 1 ? xn : t), wave);
    }
    case PAT_SPIRAL: {                                                              // :459-476  p = (cx, cy, tightness, rotation, max(H, W))
      const float yc = sub((float)y, q.p[1]), xc = sub((float)x, q.p[0]);
      const float r = fsqrt(add(mul(xc, xc), mul(yc, yc)));
      const float theta = add(add(atan2f(yc, xc), PI32), q.p[3]);
      const float s = remainder_f(add(theta, mul(mul(mul(div(r, q.p[4]), q.p[2]), 2.f), PI32)), TWOPI32);
      return div(s, TWOPI32);
    }
    case PAT_CONCENTRIC: {                                                          // :497-510  p = (cx, cy)
      const float yc = sub((float)y, q.p[1]), xc = sub((float)x, q.p[0]);
      return fsqrt(add(mul(xc, xc), mul(yc, yc)));
    }
    default: return 0.f;
  }
}
// normalisation over the whole frame + band lookup; 255 = the pixel belongs to no region (e.g. field == 1.0 exactly)
PM_HD int pattern_label(const PatternParams& q, float v, float vmin, float vmax) {
  if (q.norm == NORM_DIV_MAX) v = div(v, vmax);
  else if (q.norm == NORM_MINMAX) v = div(sub(v, vmin), add(sub(vmax, vmin), 1e-6f));
  for (int i = 0; i < q.R; ++i)
    if (v >= q.lo[i] && v < q.hi[i]) return i;
  return 255;
}
PM_HD int rect_label(const PatternParams& q, int y, int x) {
  for (int i = 0; i < q.R; ++i)
    if (y >= q.rect[4 * i] && y < q.rect[4 * i + 1] && x >= q.rect[4 * i + 2] && x < q.rect[4 * i + 3]) return i;
  return 255;
}

// ---------------------------------------------------------------------------------------------
// OpenCV bilinear sampling in 1/32-pixel fixed point (imgwarp.cpp: INTER_BITS = 5, fp32 BilinearTab)
// ---------------------------------------------------------------------------------------------
enum { BORDER_REPLICATE_ = 0, BORDER_REFLECT_ = 1 };
PM_HD int border_idx(int i, int n, int mode) {
  if (mode == BORDER_REPLICATE_) return imin(imax(i, 0), n - 1);
  if (n == 1) return 0;
  while (i < 0 || i >= n) i = i < 0 ? -i - 1 : 2 * n - 1 - i;           // BORDER_REFLECT: fedcba|abcdefgh|hgfedcb
  return i;
}
// value of a single-channel fp32 image at fixed-point (X, Y) = 32 * coordinate
PM_HD float sample_fix5(const float* img, int H, int W, int X, int Y, int border) {
  const int ix = X >> 5, iy = Y >> 5;
  const float fx = mul((float)(X & 31), 1.f / 32.f), fy = mul((float)(Y & 31), 1.f / 32.f);
  const float w00 = mul(sub(1.f, fy), sub(1.f, fx)), w01 = mul(sub(1.f, fy), fx), w10 = mul(fy, sub(1.f, fx)), w11 = mul(fy, fx);
  const int x0 = border_idx(ix, W, border), x1 = border_idx(ix + 1, W, border);
  const int y0 = border_idx(iy, H, border), y1 = border_idx(iy + 1, H, border);
  const float p00 = img[(size_t)y0 * W + x0], p01 = img[(size_t)y0 * W + x1], p10 = img[(size_t)y1 * W + x0], p11 = img[(size_t)y1 * W + x1];
  return add(add(add(mul(p00, w00), mul(p01, w01)), mul(p10, w10)), mul(p11, w11));
}
// cv2.warpAffine's source coordinate (WarpAffineInvoker: AB_BITS = 10; adelta / bdelta per column, X0 / Y0 per row, then >> 5);
// m = the INVERTED 2x3 matrix in double
PM_HD void affine_fix5(const double* m, int x, int y, int& X, int& Y) {
  const int adelta = rint_d(dmul(dmul(m[0], (double)x), 1024.0)), bdelta = rint_d(dmul(dmul(m[3], (double)x), 1024.0));
  const int X0 = rint_d(dmul(dadd(dmul(m[1], (double)y), m[2]), 1024.0)) + 16;
  const int Y0 = rint_d(dmul(dadd(dmul(m[4], (double)y), m[5]), 1024.0)) + 16;
  X = (X0 + adelta) >> 5;
  Y = (Y0 + bdelta) >> 5;
}
// cv2.getRotationMatrix2D((W / 2, H / 2), angle, 1.0) followed by warpAffine's own inversion (no WARP_INVERSE_MAP), all in double (host)
inline void rotation_inverse_matrix(int W, int H, double angle_degrees, double* M) {
  const float cx = (float)(W / 2.0), cy = (float)(H / 2.0);               // the centre is a Point2f
  const double ang = angle_degrees * (3.1415926535897932384626433832795 / 180.0);
  const double alpha = std::cos(ang), beta = std::sin(ang);
  M[0] = alpha; M[1] = beta; M[2] = (1 - alpha) * cx - beta * cy;
  M[3] = -beta; M[4] = alpha; M[5] = beta * cx + (1 - alpha) * cy;
  double D = M[0] * M[4] - M[1] * M[3];
  D = D != 0 ? 1. / D : 0;
  const double A11 = M[4] * D, A22 = M[0] * D;
  M[0] = A11; M[1] *= -D; M[3] *= -D; M[4] = A22;
  const double b1 = -M[0] * M[2] - M[1] * M[5], b2 = -M[3] * M[2] - M[4] * M[5];
  M[2] = b1; M[5] = b2;
}
// cv2.remap with fp32 maps: the map is quantised with cvRound(map * 32)
PM_HD void map_fix5(float mx, float my, int& X, int& Y) {
  X = rint_f(mul(mx, 32.f));
  Y = rint_f(mul(my, 32.f));
}

// ---------------------------------------------------------------------------------------------
// organic warp (region_blend.py:604-734): multi-octave sin/cos noise evaluated in double like numpy, accumulated in fp32
// ---------------------------------------------------------------------------------------------
enum { ORG_BLOB = 0, ORG_TENTACLE = 1, ORG_WAVE = 2, ORG_PULSE = 3 };
// np.linspace(0, stop, num)[j]
PM_HD double linspace_at(double stop, int num, int j) {
  if (num > 1 && j == num - 1) return stop;
  const double step = num > 1 ? stop / (double)(num - 1) : stop;        // numpy: step = delta / div (num == 1: div = 1, single value 0)
  return dadd(dmul((double)j, step), 0.0);
}
// _simplex_noise_2d before the min/max normalisation: `result / total_amplitude` (fp32); off = (offset_x, offset_y) per octave
PM_HD float simplex_raw(double xx, double yy, const double* off, int octaves) {
  float result = 0.f;
  double amplitude = 1.0, total = 0.0, fm = 1.0;
  for (int o = 0; o < octaves; ++o) {
    const double ox = off[2 * o], oy = off[2 * o + 1];
    double noise = dmul(sin(dadd(dmul(xx, fm), ox)), cos(dadd(dmul(yy, fm), oy)));
    noise = dadd(noise, dmul(sin(dadd(dmul(dmul(dadd(xx, yy), fm), 0.7), dmul(ox, 0.8))), 0.5));
    noise = dadd(noise, dmul(cos(dadd(dmul(dmul(dsub(xx, yy), fm), 0.5), dmul(oy, 0.6))), 0.3));
    result = (float)dadd((double)result, dmul(noise, amplitude));        // fp32 array += fp64 array
    total += amplitude;
    amplitude *= 0.5;
    fm *= 2.0;
  }
  return div(result, (float)total);
}
// (noise - min) / (max - min + 1e-6) * 2 - 1
PM_HD float simplex_flow(float raw, float mn, float mx) {
  const float n = div(sub(raw, mn), add(sub(mx, mn), 1e-6f));
  return sub(mul(n, 2.f), 1.f);
}

// ---------------------------------------------------------------------------------------------
// motion-adaptive blend (pipeline.py:2072-2086): cv2.GaussianBlur(fp32) = AVX2 row filter (sequential FMA) then symmetric column
// filter (pairs summed first, FMA); the scalar tails of both loops (last W % 4 resp. W % 8 columns) round each product separately
// ---------------------------------------------------------------------------------------------
// cv2.GaussianBlur(m, (0, 0), sigma) on CV_32F: ksize = cvRound(sigma * 4 * 2 + 1) | 1; getGaussianKernel computes and normalises the
// coefficients in double and narrows them to fp32 (host).  Returns k, or 0 when it exceeds cap.
inline int gaussian_kernel_f32(double sigma, float* kern, int cap) {
  const int k = (int)lrint(sigma * 4 * 2 + 1) | 1;
  if (k > cap) return 0;
  double t[256], sum = 0.0;
  const double scale2x = -0.5 / (sigma * sigma);
  for (int i = 0; i < k; ++i) { const double x = i - (k - 1) * 0.5; t[i] = std::exp(scale2x * x * x); sum += t[i]; }
  for (int i = 0; i < k; ++i) kern[i] = (float)(t[i] * (1.0 / sum));
  return k;
}
PM_HD int reflect101(int i, int n) {
  if (n == 1) return 0;
  while (i < 0 || i >= n) i = i < 0 ? -i : 2 * (n - 1) - i;
  return i;
}
PM_HD float motion_m(const float* flow, size_t pix, float norm_div) {     // clip(|flow| / MOTION_NORM, 0, 1)
  const float fx = flow[2 * pix], fy = flow[2 * pix + 1];
  const float mag = fsqrt(add(mul(fx, fx), mul(fy, fy)));
  return fminf(fmaxf(div(mag, norm_div), 0.f), 1.f);
}
// row filter output at (y, x): s = k0*S0; s = fma(S_t, k_t, s) ... ; columns at or beyond W & ~3 are the scalar tail (no FMA)
PM_HD float motion_row_at(const float* flow, int W, int y, int x, const float* kern, int k, float norm_div) {
  const int pad = k / 2;
  const bool fma_path = x < (W & ~3);
  float acc = 0.f;
  for (int t = 0; t < k; ++t) {
    const float m = motion_m(flow, (size_t)y * W + reflect101(x + t - pad, W), norm_div);
    acc = t == 0 ? mul(m, kern[0]) : (fma_path ? fmaf(m, kern[t], acc) : add(acc, mul(m, kern[t])));
  }
  return acc;
}
// symmetric column filter: s = kc*S[y]; s = fma(S[y+j] + S[y-j], k_{c+j}, s); columns at or beyond W & ~7 are the scalar tail
PM_HD float motion_col_at(const float* rowf, int H, int W, int y, int x, const float* kern, int k) {
  const int pad = k / 2;
  const bool fma_path = x < (W & ~7);
  float acc = mul(rowf[(size_t)y * W + x], kern[pad]);
  for (int j = 1; j <= pad; ++j) {
    const float pair = add(rowf[(size_t)reflect101(y + j, H) * W + x], rowf[(size_t)reflect101(y - j, H) * W + x]);
    acc = fma_path ? fmaf(pair, kern[pad + j], acc) : add(acc, mul(pair, kern[pad + j]));
  }
  return acc;
}
// float(u8) / 255, correctly rounded (ToTensor); the device form is nst_kernels.cuh's u8_to_unit (three instructions, no division)
PM_HD float unit255(float k) {
#ifdef __CUDA_ARCH__
  const float r = __uint_as_float(0x3b808081u);
  const float q = __fmul_rn(k, r);
  return fmaf(fmaf(-q, 255.0f, k), r, q);
#else
  return k / 255.0f;
#endif
}
// alpha_map = max_alpha - (max_alpha - min_alpha) * m;  out = pack(clip(alpha * styled + (1 - alpha) * original, 0, 1))
PM_HD void motion_fuse(float m, float max_alpha, float alpha_span, const uint8_t* styled, const uint8_t* original, uint8_t* out) {
  const float alpha = sub(max_alpha, mul(alpha_span, m));
  const float ac = sub(1.f, alpha);
  for (int c = 0; c < 3; ++c) {
    const float fused = fminf(fmaxf(add(mul(alpha, unit255((float)styled[c])), mul(ac, unit255((float)original[c]))), 0.f), 1.f);
    out[c] = (uint8_t)mul(fused, 255.f);                                      // ToPILImage: mul(255).byte()
  }
}

// mask composite (pipeline.py:2040-2043), then the uniform blend that still applies after it (:2088-2092), then the ToPILImage pack
// (:2094), for one channel of a frame that is already uint8 (`to_tensor(out_img)`, :1981): blend outside [0, 1) = off
PM_HD uint8_t blend_pack_u8(uint8_t styled, uint8_t original, bool has_mask, float a, int composite_mode, float blend, float blend_comp) {
  float o01 = unit255((float)styled);
  const float og = unit255((float)original);
  if (has_mask) {
    const float na = sub(1.0f, a);
    o01 = composite_mode == 0 ? add(mul(a, o01), mul(na, og)) : add(mul(na, o01), mul(a, og));
    o01 = fminf(fmaxf(o01, 0.f), 1.f);
  }
  if (blend >= 0.f && blend < 1.f) {
    o01 = add(mul(blend, o01), mul(blend_comp, og));
    o01 = fminf(fmaxf(o01, 0.f), 1.f);
  }
  return (uint8_t)mul(o01, 255.f);
}

// ---------------------------------------------------------------------------------------------
// LAB EMA (pipeline.py:1938-1975): one frame of one pixel; tables map (c0 << 16 | c1 << 8 | c2) -> c0' | c1' << 8 | c2' << 16
// ---------------------------------------------------------------------------------------------
struct LabState { float L, a, b; };
PM_HD uint32_t lab_ema_step(uint32_t lab, LabState& st, bool first, bool smooth_L, bool smooth_ab, float aL, float aL_comp, float aC, float aC_comp) {
  float L = (float)(lab & 255u), A = (float)((lab >> 8) & 255u), B = (float)((lab >> 16) & 255u);
  if (smooth_L) {                                                             // prev_L = L.copy() on the first frame (:1948-1949)
    st.L = add(mul(aL, L), mul(aL_comp, first ? L : st.L));                   // the state keeps the un-clipped value (:1956)
    L = fminf(fmaxf(st.L, 0.f), 255.f);
  }
  if (smooth_ab) {
    st.a = add(mul(aC, A), mul(aC_comp, first ? A : st.a));
    st.b = add(mul(aC, B), mul(aC_comp, first ? B : st.b));
    A = fminf(fmaxf(st.a, 0.f), 255.f); B = fminf(fmaxf(st.b, 0.f), 255.f);
  }
  return ((uint32_t)L << 16) | ((uint32_t)A << 8) | (uint32_t)B;              // astype(uint8) truncates; index into lab2rgb
}

}  // namespace pm
}  // namespace nst
