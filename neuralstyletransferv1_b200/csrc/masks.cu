// Region masks on the device, second part: every pattern of region_blend.generate_region_masks, the per-frame
// mask animation (rotation, organic warp, gap fill, renormalisation).  All HBM-bound one-pass kernels over the R mask planes; the
// per-pixel arithmetic lives in post_math.cuh (shared with the host-side operation-order check in tests/hostcheck).
//
//   pattern_field / pattern_label   grid, diagonal, fractal, radial, waves, spiral, concentric (region_blend.py:109-521) -> labels,
//                                   then the Gaussian feather of postproc.cu (feather_labels)
//   rotate_masks                    rotate_all_masks (region_blend.py:25-66): cv2.warpAffine(INTER_LINEAR, BORDER_REPLICATE) of every
//                                   mask + division by the clamped sum, one pass
//   organic_noise / organic_warp    warp_all_masks_organic (region_blend.py:604-810): noise flow field per region (double sin / cos
//                                   like numpy), cv2.remap(INTER_LINEAR, BORDER_REFLECT), coverage sum, gap detection
//   gap_fill / gap_commit           the max-pool 5 / 11 / 21 / 41 dilation fill of region_blend.py:772-802, skipped on the device when the
//                                   gap counter is zero (no host round trip)
//   normalize_masks                 m / clamp(sum, 1e-6)
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstring>

#include "api_util.cuh"
#include "post_math.cuh"

namespace nst {
using namespace pm;

// ordered float min / max through integer atomics (values may be negative)
__device__ __forceinline__ void atomic_min_f(float* addr, float v) {
  v += 0.f;                                              // -0 -> +0: its bit pattern would order below every negative
  if (v >= 0.f) atomicMin(reinterpret_cast<int*>(addr), __float_as_int(v));
  else atomicMax(reinterpret_cast<unsigned*>(addr), __float_as_uint(v));
}
__device__ __forceinline__ void atomic_max_f(float* addr, float v) {
  v += 0.f;
  if (v >= 0.f) atomicMax(reinterpret_cast<int*>(addr), __float_as_int(v));
  else atomicMin(reinterpret_cast<unsigned*>(addr), __float_as_uint(v));
}
// block-wide min / max, then one atomic pair per block; mm = (min, max)
__device__ __forceinline__ void block_minmax(float lo, float hi, float* mm) {
  __shared__ float s_lo[8], s_hi[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  const int w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  if ((threadIdx.x & 31) == 0) { s_lo[w] = lo; s_hi[w] = hi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int i = 1; i < nw; ++i) { lo = fminf(lo, s_lo[i]); hi = fmaxf(hi, s_hi[i]); }
    atomic_min_f(mm, lo);
    atomic_max_f(mm + 1, hi);
  }
  __syncthreads();
}
__global__ void init_minmax_kernel(float* mm, int pairs, int* counters, int ncounters) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < pairs) { mm[2 * i] = INFINITY; mm[2 * i + 1] = -INFINITY; }
  if (i < ncounters) counters[i] = 0;
}

// ---------------------------------------------------------------------------------------------
// patterns
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) pattern_field_kernel(const __grid_constant__ PatternParams q, float* field, float* mm) {
  const int hw = q.H * q.W;
  float lo = INFINITY, hi = -INFINITY;
  for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < hw; pix += gridDim.x * blockDim.x) {
    const int y = pix / q.W, x = pix - y * q.W;
    const float v = pattern_field(q, y, x);
    field[pix] = v;
    lo = fminf(lo, v); hi = fmaxf(hi, v);
  }
  block_minmax(lo, hi, mm);
}
__global__ void __launch_bounds__(256) pattern_label_kernel(const __grid_constant__ PatternParams q, const float* field, const float* mm, uint8_t* label) {
  const int hw = q.H * q.W;
  const float vmin = q.mode == PAT_RECTS ? 0.f : mm[0], vmax = q.mode == PAT_RECTS ? 0.f : mm[1];
  for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < hw; pix += gridDim.x * blockDim.x) {
    const int y = pix / q.W, x = pix - y * q.W;
    label[pix] = (uint8_t)(q.mode == PAT_RECTS ? rect_label(q, y, x) : pattern_label(q, field[pix], vmin, vmax));
  }
}

// ---------------------------------------------------------------------------------------------
// rotation
// ---------------------------------------------------------------------------------------------
struct RotateArgs { const float* in; float* out; int R, H, W; double m[6]; };
__global__ void __launch_bounds__(256) rotate_masks_kernel(const __grid_constant__ RotateArgs a) {
  const int hw = a.H * a.W;
  for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < hw; pix += gridDim.x * blockDim.x) {
    const int y = pix / a.W, x = pix - y * a.W;
    int X, Y;
    affine_fix5(a.m, x, y, X, Y);
    float v[PAT_MAX_R];
    float sum = 0.f;
#pragma unroll 4
    for (int r = 0; r < a.R; ++r) {
      v[r] = sample_fix5(a.in + (size_t)r * hw, a.H, a.W, X, Y, BORDER_REPLICATE_);
      sum = add(sum, v[r]);
    }
    sum = fmaxf(sum, 1e-6f);                                                  // mask_sum.clamp(min=1e-6)
    for (int r = 0; r < a.R; ++r) a.out[(size_t)r * hw + pix] = div(v[r], sum);
  }
}

// ---------------------------------------------------------------------------------------------
// organic warp
// ---------------------------------------------------------------------------------------------
struct OrganicArgs {
  const float* in; float* warped;       // [R,H,W]
  float* raw;                           // [R,2,H,W] un-normalised noise (blob / tentacle)
  float* mm;                            // [R,2,2] (min, max) of each noise field
  uint8_t* gap; int* counters;          // [H,W], gap-pixel count per fill pass
  int R, H, W, mode;
  double freq, time_offset, max_disp_d; float max_disp;      // max_disp: fp32 against the fp32 noise flow, fp64 against the analytic flows
  double off[PAT_MAX_R * 8];            // per region: field x (octave 0: ox, oy; octave 1: ox, oy), field y (same)
  double* tab;                          // [2R planes][2 octaves][5 W + 5 H] column / row trig tables (organic_tables_kernel)
};
// The noise of _simplex_noise_2d is, per octave,  sin(X1) cos(Y1) + 0.5 sin(X2 + Y2) + 0.3 cos(X3 - Y3)  with X* functions of the column and
// Y* of the row only (np.linspace grids).  Evaluating three double-precision sin / cos per octave and pixel made this the slowest kernel of
// the library (B200's fp64 rate: 0.94 ms per 1080p mask set); instead the 10 column- and row-functions are tabulated once per launch
// (organic_tables_kernel: (5 W + 5 H) trig evaluations per plane and octave) and the pixel combines them with the angle-addition identities
// -- 5 fp64 multiplies and 4 adds per octave.  sin(X2 + Y2) computed as sin X2 cos Y2 + cos X2 sin Y2 differs from numpy's
// sin(fl(X2 + Y2)) by ~1e-13 (arguments reach 1e3), which moves the fp32-narrowed noise by one ulp about twice per million values; after the
// 1/32-pixel quantisation of the map that is an isolated pixel per several frames (tests allow 8 of 5040 / 64 of 2 M; measured: none).
constexpr int ORG_TAB = 10;     // per (plane, octave): x tables [5][W] then y tables [5][H]
__global__ void __launch_bounds__(256) organic_tables_kernel(const __grid_constant__ OrganicArgs a) {
  const int rf = blockIdx.y >> 1, o = blockIdx.y & 1;                           // plane = region * 2 + field; two octaves
  const double fm = o ? 2.0 : 1.0, ox = a.off[rf * 4 + 2 * o], oy = a.off[rf * 4 + 2 * o + 1];
  double* tx = a.tab + (size_t)blockIdx.y * 5 * (a.W + a.H);
  double* ty = tx + 5 * (size_t)a.W;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < a.W + a.H; i += gridDim.x * blockDim.x) {
    if (i < a.W) {
      const double xx = linspace_at(a.freq, a.W, i);
      const double x1 = dadd(dmul(xx, fm), ox), x2 = dadd(dmul(dmul(xx, fm), 0.7), dmul(ox, 0.8)), x3 = dadd(dmul(dmul(xx, fm), 0.5), dmul(oy, 0.6));
      tx[i] = sin(x1); tx[a.W + i] = sin(x2); tx[2 * a.W + i] = cos(x2); tx[3 * a.W + i] = cos(x3); tx[4 * a.W + i] = sin(x3);
    } else {
      const int j = i - a.W;
      const double yy = linspace_at(a.freq, a.H, j);
      const double y1 = dadd(dmul(yy, fm), oy), y2 = dmul(dmul(yy, fm), 0.7), y3 = dmul(dmul(yy, fm), 0.5);
      ty[j] = cos(y1); ty[a.H + j] = cos(y2); ty[2 * a.H + j] = sin(y2); ty[3 * a.H + j] = cos(y3); ty[4 * a.H + j] = sin(y3);
    }
  }
}
__global__ void __launch_bounds__(256) organic_noise_kernel(const __grid_constant__ OrganicArgs a) {
  const int hw = a.H * a.W;
  const int rf = blockIdx.y;                                                  // region * 2 + field
  float lo = INFINITY, hi = -INFINITY;
  for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < hw; pix += gridDim.x * blockDim.x) {
    const int y = pix / a.W, x = pix - y * a.W;
    float result = 0.f;
    double amplitude = 1.0, total = 0.0;
#pragma unroll
    for (int o = 0; o < 2; ++o) {
      const double* tx = a.tab + (size_t)(rf * 2 + o) * 5 * (a.W + a.H);
      const double* ty = tx + 5 * (size_t)a.W;
      const double t1 = dmul(tx[x], ty[y]);
      const double t2 = dadd(dmul(tx[a.W + x], ty[a.H + y]), dmul(tx[2 * a.W + x], ty[2 * a.H + y]));          // sin(X2 + Y2)
      const double t3 = dadd(dmul(tx[3 * a.W + x], ty[3 * a.H + y]), dmul(tx[4 * a.W + x], ty[4 * a.H + y]));  // cos(X3 - Y3)
      const double noise = dadd(dadd(t1, dmul(t2, 0.5)), dmul(t3, 0.3));
      result = (float)dadd((double)result, dmul(noise, amplitude));             // fp32 array += fp64 array
      total += amplitude;
      amplitude *= 0.5;
    }
    const float v = div(result, (float)total);
    a.raw[(size_t)rf * hw + pix] = v;
    lo = fminf(lo, v); hi = fmaxf(hi, v);
  }
  block_minmax(lo, hi, a.mm + 2 * rf);
}
__global__ void __launch_bounds__(256) organic_warp_kernel(const __grid_constant__ OrganicArgs a) {
  const int hw = a.H * a.W;
  int gaps = 0;
  for (int base = blockIdx.x * blockDim.x; base < hw; base += gridDim.x * blockDim.x) {
    const int pix = base + threadIdx.x;
    int is_gap = 0;
    if (pix < hw) {
      const int y = pix / a.W, x = pix - y * a.W;
      float sum = 0.f;
      float mx = 0.f, my = 0.f;
      if (a.mode == ORG_WAVE) {                                               // region_blend.py:699-704: the flow is fp64 until the map is narrowed
        const double fx = sin(dadd(linspace_at(M_PI * a.freq, a.H, y), dmul(a.time_offset, 2.0)));
        const double fy = cos(dadd(linspace_at(M_PI * a.freq, a.W, x), dmul(a.time_offset, 1.5)));
        mx = (float)dadd((double)x, dmul(fx, a.max_disp_d));
        my = (float)dadd((double)y, dmul(fy, a.max_disp_d));
      } else if (a.mode == ORG_PULSE) {                                       // :705-714
        const long long yi = y - a.H / 2, xi = x - a.W / 2;
        const double r = dadd(sqrt((double)(xi * xi + yi * yi)), 1e-6);
        const double theta = atan2((double)yi, (double)xi);
        const double pulse = dadd(dmul(sin(dsub(dmul(r, 0.05), dmul(a.time_offset, 3.0))), 0.5), 0.5);
        mx = (float)dadd((double)x, dmul(dmul(cos(theta), pulse), a.max_disp_d));
        my = (float)dadd((double)y, dmul(dmul(sin(theta), pulse), a.max_disp_d));
      }
      for (int r = 0; r < a.R; ++r) {
        if (a.mode == ORG_BLOB || a.mode == ORG_TENTACLE) {
          const float* mmr = a.mm + 4 * r;
          float fx = simplex_flow(a.raw[(size_t)(2 * r) * hw + pix], mmr[0], mmr[1]);
          float fy = simplex_flow(a.raw[(size_t)(2 * r + 1) * hw + pix], mmr[2], mmr[3]);
          if (a.mode == ORG_TENTACLE)                                         // :696-698  flow_y += sin(y * pi * 3 + t) * 0.5
            fy = (float)dadd((double)fy, dmul(sin(dadd(dmul(dmul(linspace_at(1.0, a.H, y), M_PI), 3.0), a.time_offset)), 0.5));
          mx = add((float)x, mul(fx, a.max_disp));
          my = add((float)y, mul(fy, a.max_disp));
        }
        int X, Y;
        map_fix5(mx, my, X, Y);
        const float v = sample_fix5(a.in + (size_t)r * hw, a.H, a.W, X, Y, BORDER_REFLECT_);
        a.warped[(size_t)r * hw + pix] = v;
        sum = add(sum, v);
      }
      is_gap = sum < 0.1f;                                                    // MIN_COVERAGE (region_blend.py:769)
      a.gap[pix] = (uint8_t)is_gap;
    }
    gaps += __syncthreads_count(is_gap);
  }
  if (threadIdx.x == 0 && gaps) atomicAdd(a.counters, gaps);
}

// One dilation pass (region_blend.py:781-797), split in two so that every window reads the masks of the previous pass: `fill` writes
// max_pool2d(m, k, stride 1, pad k/2) of the gap pixels into tmp, `commit` copies them back, recomputes the coverage and the gap map.
struct GapArgs { float* masks; float* tmp; uint8_t* gap; int* counters; int R, H, W, k, pass; };
__global__ void __launch_bounds__(256) gap_fill_kernel(const __grid_constant__ GapArgs a) {
  if (a.counters[a.pass] == 0) return;
  const int hw = a.H * a.W, pad = a.k / 2;
  for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < hw; pix += gridDim.x * blockDim.x) {
    if (!a.gap[pix]) continue;
    const int y = pix / a.W, x = pix - y * a.W;
    const int y0 = max(y - pad, 0), y1 = min(y + pad, a.H - 1), x0 = max(x - pad, 0), x1 = min(x + pad, a.W - 1);
    for (int r = 0; r < a.R; ++r) {
      const float* pl = a.masks + (size_t)r * hw;
      float best = -INFINITY;
      for (int yy = y0; yy <= y1; ++yy)
        for (int xx = x0; xx <= x1; ++xx) best = fmaxf(best, pl[(size_t)yy * a.W + xx]);
      a.tmp[(size_t)r * hw + pix] = best;                                     // m * (1 - gap) + dilated * gap at a gap pixel
    }
  }
}
__global__ void __launch_bounds__(256) gap_commit_kernel(const __grid_constant__ GapArgs a) {
  if (a.counters[a.pass] == 0) return;
  const int hw = a.H * a.W;
  int gaps = 0;
  for (int base = blockIdx.x * blockDim.x; base < hw; base += gridDim.x * blockDim.x) {
    const int pix = base + threadIdx.x;
    int is_gap = 0;
    if (pix < hw && a.gap[pix]) {
      float sum = 0.f;
      for (int r = 0; r < a.R; ++r) {
        const float v = a.tmp[(size_t)r * hw + pix];
        a.masks[(size_t)r * hw + pix] = v;
        sum = add(sum, v);
      }
      is_gap = sum < 0.1f;
      a.gap[pix] = (uint8_t)is_gap;
    }
    gaps += __syncthreads_count(is_gap);
  }
  if (threadIdx.x == 0 && gaps) atomicAdd(a.counters + a.pass + 1, gaps);
}
__global__ void __launch_bounds__(256) normalize_masks_kernel(const float* in, float* out, int R, int H, int W) {
  const int hw = H * W;
  for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < hw; pix += gridDim.x * blockDim.x) {
    float v[PAT_MAX_R];
    float sum = 0.f;
    for (int r = 0; r < R; ++r) { v[r] = in[(size_t)r * hw + pix]; sum = add(sum, v[r]); }
    sum = fmaxf(sum, 1e-6f);
    for (int r = 0; r < R; ++r) out[(size_t)r * hw + pix] = div(v[r], sum);
  }
}

// ---------------------------------------------------------------------------------------------
// tight bounding boxes (region_blend.compute_mask_bbox, region_blend.py:1969-1994)
// ---------------------------------------------------------------------------------------------
// box[4r .. 4r+3] = (min x, min y, max x, max y) over the pixels of mask r above the threshold; one atomic quartet per block and region
__global__ void mask_bbox_init_kernel(int* box, int R) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < R) { box[4 * r] = INT_MAX; box[4 * r + 1] = INT_MAX; box[4 * r + 2] = -1; box[4 * r + 3] = -1; }
}
__global__ void __launch_bounds__(256) mask_bbox_kernel(const float* masks, int* box, int H, int W, double threshold) {
  const int hw = H * W, r = blockIdx.y;
  const float* pl = masks + (size_t)r * hw;
  int x0 = INT_MAX, y0 = INT_MAX, x1 = -1, y1 = -1;
  for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < hw; pix += gridDim.x * blockDim.x) {
    if ((double)pl[pix] > threshold) {                  // numpy: fp32 array > Python double
      const int y = pix / W, x = pix - y * W;
      x0 = min(x0, x); y0 = min(y0, y); x1 = max(x1, x); y1 = max(y1, y);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    x0 = min(x0, __shfl_xor_sync(0xffffffffu, x0, o)); y0 = min(y0, __shfl_xor_sync(0xffffffffu, y0, o));
    x1 = max(x1, __shfl_xor_sync(0xffffffffu, x1, o)); y1 = max(y1, __shfl_xor_sync(0xffffffffu, y1, o));
  }
  if ((threadIdx.x & 31) == 0 && x1 >= 0) {
    atomicMin(box + 4 * r, x0); atomicMin(box + 4 * r + 1, y0); atomicMax(box + 4 * r + 2, x1); atomicMax(box + 4 * r + 3, y1);
  }
}

static inline int grid_for(int hw) { return std::min((hw + 255) / 256, 148 * 16); }

}  // namespace nst

using namespace nst;

extern "C" int nst_pattern_masks_f32(int mode, int norm, const float* params6, const float* lo, const float* hi, const int* rects, int R,
                                     int feather, float* masks, float* scratch, int H, int W, void* stream) {
  if (!masks || !scratch || R < 1 || R > PAT_MAX_R || H < 2 || W < 2 || mode < PAT_RECTS || mode > PAT_CONCENTRIC || norm < NORM_NONE || norm > NORM_MINMAX)
    return api_fail(NST_ERR_INVALID, "pattern_masks: bad argument (1 <= R <= %d, mode 0..5)", PAT_MAX_R);
  if (mode == PAT_RECTS ? rects == nullptr : (params6 == nullptr || lo == nullptr || hi == nullptr))
    return api_fail(NST_ERR_INVALID, "pattern_masks: mode %d needs %s", mode, mode == PAT_RECTS ? "rects" : "params, lo and hi");
  cudaStream_t st = (cudaStream_t)stream;
  PatternParams q;
  memset(&q, 0, sizeof(q));
  q.mode = mode; q.norm = norm; q.R = R; q.H = H; q.W = W;
  for (int i = 0; i < 6 && params6; ++i) q.p[i] = params6[i];
  for (int i = 0; i < R; ++i) {
    if (lo) { q.lo[i] = lo[i]; q.hi[i] = hi[i]; }
    if (rects) for (int j = 0; j < 4; ++j) q.rect[4 * i + j] = rects[4 * i + j];
  }
  // scratch: [2 floats min/max + pad to 4][H*W floats field][H*W bytes label][R*H*W floats blur temporary]
  const size_t hw = (size_t)H * W;
  float* mm = scratch;
  float* field = scratch + 4;
  uint8_t* label = reinterpret_cast<uint8_t*>(field + hw);
  float* tmp = field + hw + (hw + 3) / 4;
  if (mode != PAT_RECTS) {
    init_minmax_kernel<<<1, 32, 0, st>>>(mm, 1, nullptr, 0);
    pattern_field_kernel<<<grid_for((int)hw), 256, 0, st>>>(q, field, mm);
  }
  pattern_label_kernel<<<grid_for((int)hw), 256, 0, st>>>(q, field, mm, label);
  NST_CUDA_TRY(cudaGetLastError());
  return feather_labels(label, R, feather, masks, tmp, H, W, st, "pattern_masks");
}

extern "C" int nst_rotate_masks_f32(const float* masks, float* out, int R, int H, int W, double angle_degrees, void* stream) {
  if (!masks || !out || masks == out || R < 1 || R > PAT_MAX_R || H < 1 || W < 1)
    return api_fail(NST_ERR_INVALID, "rotate_masks: bad argument (1 <= R <= %d, out must not alias masks)", PAT_MAX_R);
  RotateArgs a;
  a.in = masks; a.out = out; a.R = R; a.H = H; a.W = W;
  rotation_inverse_matrix(W, H, angle_degrees, a.m);
  rotate_masks_kernel<<<grid_for(H * W), 256, 0, (cudaStream_t)stream>>>(a);
  NST_CUDA_TRY(cudaGetLastError());
  return NST_OK;
}

// floats of scratch in front of the trig tables, rounded up so that the tables are 8-byte aligned (the scratch base is a cudaMalloc /
// torch allocation: at least 256-byte aligned)
static size_t organic_table_offset_floats(int R, size_t hw) { return (4 * PAT_MAX_R + 8 + 3 * (size_t)R * hw + (hw + 3) / 4 + 1) & ~(size_t)1; }

extern "C" int nst_warp_masks_organic_f32(const float* masks, float* out, int R, int H, int W, int mode, double amplitude, double frequency,
                                          double time_offset, const double* offsets, float* scratch, void* stream) {
  if (!masks || !out || !scratch || masks == out || R < 1 || R > PAT_MAX_R || H < 2 || W < 2 || mode < ORG_BLOB || mode > ORG_PULSE)
    return api_fail(NST_ERR_INVALID, "warp_masks_organic: bad argument (1 <= R <= %d, mode 0..3, out must not alias masks)", PAT_MAX_R);
  const bool noisy = mode == ORG_BLOB || mode == ORG_TENTACLE;
  if (noisy && offsets == nullptr) return api_fail(NST_ERR_INVALID, "warp_masks_organic: blob / tentacle need the 8 noise offsets of every region");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t hw = (size_t)H * W;
  // scratch: [4R floats min/max][8 ints counters][2R*H*W floats noise][R*H*W floats fill temporary][H*W bytes gap map][trig tables, doubles]
  OrganicArgs a;
  memset(&a, 0, sizeof(a));
  a.in = masks; a.warped = out; a.R = R; a.H = H; a.W = W; a.mode = mode;
  a.mm = scratch;
  a.counters = reinterpret_cast<int*>(scratch + 4 * PAT_MAX_R);
  a.raw = scratch + 4 * PAT_MAX_R + 8;
  float* tmp = a.raw + 2 * (size_t)R * hw;
  a.gap = reinterpret_cast<uint8_t*>(tmp + (size_t)R * hw);
  a.tab = reinterpret_cast<double*>(scratch + organic_table_offset_floats(R, hw));
  a.freq = mode == ORG_TENTACLE ? frequency * 2 : frequency;              // region_blend.py:693
  a.time_offset = time_offset;
  a.max_disp_d = std::max(H, W) * amplitude;
  a.max_disp = (float)a.max_disp_d;                       // :720; numpy narrows the Python float against fp32 arrays
  if (noisy) memcpy(a.off, offsets, sizeof(double) * 8 * R);
  init_minmax_kernel<<<1, 64, 0, st>>>(a.mm, 2 * R, a.counters, 8);
  if (noisy) {
    organic_tables_kernel<<<dim3((W + H + 255) / 256, 4 * R), 256, 0, st>>>(a);
    organic_noise_kernel<<<dim3(grid_for((int)hw), 2 * R), 256, 0, st>>>(a);
  }
  organic_warp_kernel<<<grid_for((int)hw), 256, 0, st>>>(a);
  const int ks[4] = {5, 11, 21, 41};
  for (int p = 0; p < 4; ++p) {
    GapArgs g;
    g.masks = out; g.tmp = tmp; g.gap = a.gap; g.counters = a.counters; g.R = R; g.H = H; g.W = W; g.k = ks[p]; g.pass = p;
    gap_fill_kernel<<<grid_for((int)hw), 256, 0, st>>>(g);
    gap_commit_kernel<<<grid_for((int)hw), 256, 0, st>>>(g);
  }
  normalize_masks_kernel<<<grid_for((int)hw), 256, 0, st>>>(out, out, R, H, W);
  NST_CUDA_TRY(cudaGetLastError());
  return NST_OK;
}

/* floats of scratch the two calls above need */
extern "C" size_t nst_mask_scratch_floats(int R, int H, int W) {
  const size_t hw = (size_t)H * W;
  const size_t pattern = 4 + hw + (hw + 3) / 4 + (size_t)R * hw;
  const size_t organic = organic_table_offset_floats(R, hw) + 2 * (size_t)(4 * R) * 5 * (W + H);      // tables are doubles
  return std::max(pattern, organic);
}

extern "C" int nst_mask_bboxes(const float* masks, int R, int H, int W, double threshold, int* boxes_dev, void* stream) {
  if (!masks || !boxes_dev || R < 1 || H < 1 || W < 1) return api_fail(NST_ERR_INVALID, "mask_bboxes: null / empty argument");
  cudaStream_t st = (cudaStream_t)stream;
  mask_bbox_init_kernel<<<(R + 63) / 64, 64, 0, st>>>(boxes_dev, R);
  mask_bbox_kernel<<<dim3(std::min((H * W + 255) / 256, 148 * 4), R), 256, 0, st>>>(masks, boxes_dev, H, W, threshold);
  NST_CUDA_TRY(cudaGetLastError());
  return NST_OK;
}
