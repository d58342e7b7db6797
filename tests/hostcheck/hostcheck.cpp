// TEST INFRASTRUCTURE ONLY.  Compiles the per-pixel functions of neuralstyletransferv1_b200/csrc/post_math.cuh for the HOST and loops
// them over small frames, so that `pytest -m "not gpu"` can check the operation order of the region-mask / temporal kernels against the
// reference's fixtures on a machine without a GPU.  Built by tests/test_hostcheck.py into tests/hostcheck/_hostcheck.so (git-ignored);
// nothing in the product links or loads it, and the product library contains none of these loops.
#include <algorithm>
#include <cstring>
#include <vector>

#include "../../neuralstyletransferv1_b200/csrc/post_math.cuh"

using namespace nst::pm;

extern "C" {

void hc_pattern_labels(int mode, int norm, const float* p6, const float* lo, const float* hi, const int* rects, int R, int H, int W, uint8_t* label) {
  PatternParams q;
  memset(&q, 0, sizeof(q));
  q.mode = mode; q.norm = norm; q.R = R; q.H = H; q.W = W;
  for (int i = 0; i < 6 && p6; ++i) q.p[i] = p6[i];
  for (int i = 0; i < R; ++i) {
    if (lo) { q.lo[i] = lo[i]; q.hi[i] = hi[i]; }
    if (rects) for (int j = 0; j < 4; ++j) q.rect[4 * i + j] = rects[4 * i + j];
  }
  if (mode == PAT_RECTS) {
    for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) label[y * W + x] = (uint8_t)rect_label(q, y, x);
    return;
  }
  std::vector<float> f((size_t)H * W);
  float mn = INFINITY, mx = -INFINITY;
  for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) { const float v = pattern_field(q, y, x); f[y * W + x] = v; mn = std::min(mn, v); mx = std::max(mx, v); }
  for (int i = 0; i < H * W; ++i) label[i] = (uint8_t)pattern_label(q, f[i], mn, mx);
}

void hc_rotate_masks(const float* in, float* out, int R, int H, int W, double angle) {
  double m[6];
  rotation_inverse_matrix(W, H, angle, m);
  const size_t hw = (size_t)H * W;
  for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) {
    int X, Y;
    affine_fix5(m, x, y, X, Y);
    float v[PAT_MAX_R], sum = 0.f;
    for (int r = 0; r < R; ++r) { v[r] = sample_fix5(in + r * hw, H, W, X, Y, BORDER_REPLICATE_); sum = add(sum, v[r]); }
    sum = std::max(sum, 1e-6f);
    for (int r = 0; r < R; ++r) out[r * hw + (size_t)y * W + x] = div(v[r], sum);
  }
}

// returns the number of gap pixels before the fill
int hc_warp_masks_organic(const float* in, float* out, int R, int H, int W, int mode, double amplitude, double frequency, double time_offset,
                          const double* off) {
  const size_t hw = (size_t)H * W;
  const double freq = mode == ORG_TENTACLE ? frequency * 2 : frequency;
  const double max_disp_d = std::max(H, W) * amplitude;
  const float max_disp = (float)max_disp_d;
  const bool noisy = mode == ORG_BLOB || mode == ORG_TENTACLE;
  std::vector<float> raw(noisy ? 2 * R * hw : 0), mm(4 * R);
  if (noisy)
    for (int rf = 0; rf < 2 * R; ++rf) {
      float mn = INFINITY, mx = -INFINITY;
      for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) {
        const float v = simplex_raw(linspace_at(freq, W, x), linspace_at(freq, H, y), off + rf * 4, 2);
        raw[rf * hw + (size_t)y * W + x] = v; mn = std::min(mn, v); mx = std::max(mx, v);
      }
      mm[2 * rf] = mn; mm[2 * rf + 1] = mx;
    }
  std::vector<uint8_t> gap(hw);
  int gaps = 0;
  for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) {
    const size_t pix = (size_t)y * W + x;
    float mx = 0.f, my = 0.f, sum = 0.f;
    if (mode == ORG_WAVE) {
      const double fx = sin(dadd(linspace_at(M_PI * freq, H, y), dmul(time_offset, 2.0)));
      const double fy = cos(dadd(linspace_at(M_PI * freq, W, x), dmul(time_offset, 1.5)));
      mx = (float)dadd((double)x, dmul(fx, max_disp_d)); my = (float)dadd((double)y, dmul(fy, max_disp_d));
    } else if (mode == ORG_PULSE) {
      const long long yi = y - H / 2, xi = x - W / 2;
      const double r = dadd(sqrt((double)(xi * xi + yi * yi)), 1e-6);
      const double theta = atan2((double)yi, (double)xi);
      const double pulse = dadd(dmul(sin(dsub(dmul(r, 0.05), dmul(time_offset, 3.0))), 0.5), 0.5);
      mx = (float)dadd((double)x, dmul(dmul(cos(theta), pulse), max_disp_d)); my = (float)dadd((double)y, dmul(dmul(sin(theta), pulse), max_disp_d));
    }
    for (int r = 0; r < R; ++r) {
      if (noisy) {
        float fx = simplex_flow(raw[(2 * r) * hw + pix], mm[4 * r], mm[4 * r + 1]);
        float fy = simplex_flow(raw[(2 * r + 1) * hw + pix], mm[4 * r + 2], mm[4 * r + 3]);
        if (mode == ORG_TENTACLE) fy = (float)dadd((double)fy, dmul(sin(dadd(dmul(dmul(linspace_at(1.0, H, y), M_PI), 3.0), time_offset)), 0.5));
        mx = add((float)x, mul(fx, max_disp)); my = add((float)y, mul(fy, max_disp));
      }
      int X, Y;
      map_fix5(mx, my, X, Y);
      const float v = sample_fix5(in + r * hw, H, W, X, Y, BORDER_REFLECT_);
      out[r * hw + pix] = v; sum = add(sum, v);
    }
    gap[pix] = sum < 0.1f; gaps += gap[pix];
  }
  const int before = gaps;
  const int ks[4] = {5, 11, 21, 41};
  std::vector<float> tmp(R * hw);
  for (int p = 0; p < 4 && gaps > 0; ++p) {
    const int pad = ks[p] / 2;
    for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) {
      const size_t pix = (size_t)y * W + x;
      if (!gap[pix]) continue;
      for (int r = 0; r < R; ++r) {
        float best = -INFINITY;
        for (int yy = std::max(y - pad, 0); yy <= std::min(y + pad, H - 1); ++yy)
          for (int xx = std::max(x - pad, 0); xx <= std::min(x + pad, W - 1); ++xx) best = std::max(best, out[r * hw + (size_t)yy * W + xx]);
        tmp[r * hw + pix] = best;
      }
    }
    gaps = 0;
    for (size_t pix = 0; pix < hw; ++pix) {
      if (!gap[pix]) continue;
      float sum = 0.f;
      for (int r = 0; r < R; ++r) { out[r * hw + pix] = tmp[r * hw + pix]; sum = add(sum, tmp[r * hw + pix]); }
      gap[pix] = sum < 0.1f; gaps += gap[pix];
    }
  }
  for (size_t pix = 0; pix < hw; ++pix) {
    float sum = 0.f;
    for (int r = 0; r < R; ++r) sum = add(sum, out[r * hw + pix]);
    sum = std::max(sum, 1e-6f);
    for (int r = 0; r < R; ++r) out[r * hw + pix] = div(out[r * hw + pix], sum);
  }
  return before;
}

void hc_lab_ema(const uint8_t* frames, uint8_t* out, int N, int H, int W, const uint32_t* rgb2lab, const uint32_t* lab2rgb, int smooth_L, int smooth_ab,
                double smooth_alpha, double chroma_alpha) {
  const size_t hw = (size_t)H * W;
  for (size_t pix = 0; pix < hw; ++pix) {
    LabState st = {0.f, 0.f, 0.f};
    for (int n = 0; n < N; ++n) {
      const uint8_t* px = frames + (n * hw + pix) * 3;
      const uint32_t lab = rgb2lab[((uint32_t)px[0] << 16) | ((uint32_t)px[1] << 8) | px[2]];
      const uint32_t rgb = lab2rgb[lab_ema_step(lab, st, n == 0, smooth_L, smooth_ab, (float)smooth_alpha, (float)(1.0 - smooth_alpha), (float)chroma_alpha,
                                                (float)(1.0 - chroma_alpha))];
      uint8_t* o = out + (n * hw + pix) * 3;
      o[0] = rgb & 255u; o[1] = (rgb >> 8) & 255u; o[2] = (rgb >> 16) & 255u;
    }
  }
}

void hc_motion_blend(const uint8_t* styled, const uint8_t* original, const float* flow, uint8_t* out, float* alpha_map, int H, int W, double blend,
                     double min_alpha, double motion_norm, double sigma) {
  float kern[64];
  const int k = gaussian_kernel_f32(sigma, kern, 64);
  std::vector<float> rowf((size_t)H * W);
  for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) rowf[(size_t)y * W + x] = motion_row_at(flow, W, y, x, kern, k, (float)motion_norm);
  for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) {
    const size_t pix = (size_t)y * W + x;
    const float m = motion_col_at(rowf.data(), H, W, y, x, kern, k);
    if (alpha_map) alpha_map[pix] = sub((float)blend, mul((float)(blend - min_alpha), m));
    motion_fuse(m, (float)blend, (float)(blend - min_alpha), styled + pix * 3, original + pix * 3, out + pix * 3);
  }
}

void hc_blend_pack(const uint8_t* styled, const uint8_t* original, const float* mask, double blend, int composite_mode, uint8_t* out, size_t npix) {
  const bool on = blend >= 0.0 && blend < 1.0;
  for (size_t i = 0; i < npix; ++i)
    for (int c = 0; c < 3; ++c)
      out[i * 3 + c] = blend_pack_u8(styled[i * 3 + c], original[i * 3 + c], mask != nullptr, mask ? mask[i] : 0.f, composite_mode, on ? (float)blend : 2.f,
                                     (float)(1.0 - blend));
}

}  // extern "C"
