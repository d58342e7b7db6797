// Temporal post-processing on the device, second part: the LAB lightness / chroma EMA and the motion-adaptive
// blend that follow the flow-guided EMA of postproc.cu in the reference's per-frame loop.
//
//   lab_ema        pipeline.py:1938-1975.  Pillow's RGB <-> LAB conversion is littleCMS with interpolated 16-bit tables -- a pure
//                  function of the uint8 triple, so the device reads it from two 2^24-entry tables the host fills ONCE by running
//                  Pillow's own transform over all colours (neuralstyletransferv1_b200/temporal.py).  The EMA is independent per
//                  pixel: one thread walks its pixel through a whole chunk of N frames with the running (L, a, b) state in
//                  registers, so the state planes are read and written once per chunk instead of once per frame.
//   motion_blend   pipeline.py:2072-2086.  alpha = blend - (blend - MIN_ALPHA) * GaussianBlur(clip(|flow| / MOTION_NORM, 0, 1), sigma 3),
//                  out = pack(clip(alpha * styled + (1 - alpha) * original, 0, 1)).  Two launches: flow magnitude + row filter, then
//                  column filter + blend + truncating pack.  cv2.GaussianBlur's fp32 operation order is documented in post_math.cuh.
#include <algorithm>
#include <cmath>
#include <cstring>

#include "api_util.cuh"
#include "post_math.cuh"

namespace nst {
using namespace pm;

struct LabEmaArgs {
  const uint8_t* frames; uint8_t* out;          // [N,H,W,3]
  const uint32_t* rgb2lab; const uint32_t* lab2rgb;   // [2^24]: index (c0 << 16 | c1 << 8 | c2) -> c0' | c1' << 8 | c2' << 16
  float* prev_L; float* prev_a; float* prev_b;  // [H,W] running state (in / out)
  int N, H, W, init, smooth_L, smooth_ab;
  float aL, aL_comp, aC, aC_comp;               // fp32(alpha), fp32(1 - alpha): numpy narrows the Python doubles against fp32 arrays
};
// The two table look-ups of a frame are random 4-byte gathers (the tables are 64 MB each): walking the frames one by one chains
// 2 * N dependent memory latencies per thread.  Only the EMA arithmetic is sequential over the frames, so the kernel works on blocks of
// LAB_UNROLL frames: all RGB->LAB gathers of the block in flight together, then the (register-only) EMA recurrence, then all LAB->RGB
// gathers together, then the stores.
constexpr int LAB_UNROLL = 8;
__global__ void __launch_bounds__(256) lab_ema_kernel(const __grid_constant__ LabEmaArgs a) {
  const int hw = a.H * a.W;
  for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < hw; pix += gridDim.x * blockDim.x) {
    LabState st = {0.f, 0.f, 0.f};
    if (!a.init) {
      if (a.smooth_L) st.L = a.prev_L[pix];
      if (a.smooth_ab) { st.a = a.prev_a[pix]; st.b = a.prev_b[pix]; }
    }
    for (int n0 = 0; n0 < a.N; n0 += LAB_UNROLL) {
      uint32_t v[LAB_UNROLL];
#pragma unroll
      for (int u = 0; u < LAB_UNROLL; ++u) {
        if (n0 + u < a.N) {
          const uint8_t* px = a.frames + ((size_t)(n0 + u) * hw + pix) * 3;
          v[u] = ((uint32_t)px[0] << 16) | ((uint32_t)px[1] << 8) | px[2];
        }
      }
#pragma unroll
      for (int u = 0; u < LAB_UNROLL; ++u)
        if (n0 + u < a.N) v[u] = __ldg(a.rgb2lab + v[u]);
#pragma unroll
      for (int u = 0; u < LAB_UNROLL; ++u)
        if (n0 + u < a.N) v[u] = lab_ema_step(v[u], st, a.init && n0 + u == 0, a.smooth_L, a.smooth_ab, a.aL, a.aL_comp, a.aC, a.aC_comp);
#pragma unroll
      for (int u = 0; u < LAB_UNROLL; ++u)
        if (n0 + u < a.N) v[u] = __ldg(a.lab2rgb + v[u]);
#pragma unroll
      for (int u = 0; u < LAB_UNROLL; ++u) {
        if (n0 + u < a.N) {
          uint8_t* o = a.out + ((size_t)(n0 + u) * hw + pix) * 3;
          o[0] = (uint8_t)(v[u] & 255u); o[1] = (uint8_t)((v[u] >> 8) & 255u); o[2] = (uint8_t)((v[u] >> 16) & 255u);
        }
      }
    }
    if (a.smooth_L) a.prev_L[pix] = st.L;
    if (a.smooth_ab) { a.prev_a[pix] = st.a; a.prev_b[pix] = st.b; }
  }
}

constexpr int MB_MAX_K = 63;
struct MotionBlendArgs {
  const uint8_t* styled; const uint8_t* original; const float* flow; uint8_t* out; float* rowf;   // rowf: [H,W] row-filtered motion
  int H, W, k;
  float norm_div;                     // MOTION_NORM
  float max_alpha, alpha_span;        // fp32(blend), fp32(blend - MIN_ALPHA)
  float kern[MB_MAX_K];
};
// Row pass: one block = 256 consecutive pixels of one row.  The clipped flow magnitude of the segment and its 2 * (k / 2) neighbours is computed
// ONCE into shared memory (the per-pixel form evaluated sqrt + division for each of the k taps: 25x redundant at sigma 3), then every thread
// runs the same tap sequence as before -- bit-identical.
__global__ void __launch_bounds__(256) motion_row_kernel(const __grid_constant__ MotionBlendArgs a) {
  __shared__ float sm[256 + MB_MAX_K];
  const int pad = a.k / 2, y = blockIdx.y, x0 = blockIdx.x * 256;
  for (int i = threadIdx.x; i < 256 + 2 * pad; i += 256) {
    const int xx = reflect101(x0 + i - pad, a.W);                    // (columns beyond the row's end are never read by a live thread)
    sm[i] = motion_m(a.flow, (size_t)y * a.W + xx, a.norm_div);
  }
  __syncthreads();
  const int x = x0 + threadIdx.x;
  if (x >= a.W) return;
  const bool fma_path = x < (a.W & ~3);
  float acc = mul(sm[threadIdx.x], a.kern[0]);
  for (int t = 1; t < a.k; ++t) acc = fma_path ? fmaf(sm[threadIdx.x + t], a.kern[t], acc) : add(acc, mul(sm[threadIdx.x + t], a.kern[t]));
  a.rowf[(size_t)y * a.W + x] = acc;
}
__global__ void __launch_bounds__(256) motion_col_blend_kernel(const __grid_constant__ MotionBlendArgs a) {
  const int hw = a.H * a.W;
  for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < hw; pix += gridDim.x * blockDim.x) {
    const int y = pix / a.W, x = pix - y * a.W;
    const float m = motion_col_at(a.rowf, a.H, a.W, y, x, a.kern, a.k);
    motion_fuse(m, a.max_alpha, a.alpha_span, a.styled + (size_t)pix * 3, a.original + (size_t)pix * 3, a.out + (size_t)pix * 3);
  }
}

struct BlendPackArgs { const uint8_t* styled; const uint8_t* original; const float* mask; uint8_t* out; size_t npix; int composite_mode; float blend, blend_comp; };
__global__ void __launch_bounds__(256) blend_pack_kernel(const __grid_constant__ BlendPackArgs a) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.npix; i += (size_t)gridDim.x * blockDim.x) {
    const float al = a.mask ? a.mask[i] : 0.f;
#pragma unroll
    for (int c = 0; c < 3; ++c)
      a.out[i * 3 + c] = blend_pack_u8(a.styled[i * 3 + c], a.original[i * 3 + c], a.mask != nullptr, al, a.composite_mode, a.blend, a.blend_comp);
  }
}

// four pixels per thread: 12 bytes of each frame as three 32-bit words, one 16-byte mask load, three 32-bit stores (npix % 4 == 0, aligned)
__global__ void __launch_bounds__(256) blend_pack_x4_kernel(const __grid_constant__ BlendPackArgs a) {
  const size_t quads = a.npix / 4;
  for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < quads; q += (size_t)gridDim.x * blockDim.x) {
    const uint32_t* s32 = reinterpret_cast<const uint32_t*>(a.styled + q * 12);
    const uint32_t* o32 = reinterpret_cast<const uint32_t*>(a.original + q * 12);
    const uint32_t sw[3] = {s32[0], s32[1], s32[2]}, ow[3] = {o32[0], o32[1], o32[2]};
    float al[4] = {0.f, 0.f, 0.f, 0.f};
    if (a.mask) { const float4 t = *reinterpret_cast<const float4*>(a.mask + q * 4); al[0] = t.x; al[1] = t.y; al[2] = t.z; al[3] = t.w; }
    uint32_t outw[3] = {0u, 0u, 0u};
#pragma unroll
    for (int j = 0; j < 12; ++j) {
      const uint8_t b = blend_pack_u8((uint8_t)((sw[j >> 2] >> ((j & 3) * 8)) & 0xFFu), (uint8_t)((ow[j >> 2] >> ((j & 3) * 8)) & 0xFFu), a.mask != nullptr,
                                      al[j / 3], a.composite_mode, a.blend, a.blend_comp);
      outw[j >> 2] |= (uint32_t)b << ((j & 3) * 8);
    }
    uint32_t* d32 = reinterpret_cast<uint32_t*>(a.out + q * 12);
    d32[0] = outw[0]; d32[1] = outw[1]; d32[2] = outw[2];
  }
}

static inline int grid_for(int hw) { return std::min((hw + 255) / 256, 148 * 16); }

}  // namespace nst

using namespace nst;

extern "C" int nst_lab_ema_u8(const uint8_t* frames, uint8_t* out, int N, int H, int W, const uint32_t* rgb2lab, const uint32_t* lab2rgb,
                              float* prev_L, float* prev_a, float* prev_b, int init, int smooth_lightness, int smooth_chroma,
                              double smooth_alpha, double chroma_alpha, void* stream) {
  if (!frames || !out || !rgb2lab || !lab2rgb || N < 1 || H < 1 || W < 1) return api_fail(NST_ERR_INVALID, "lab_ema: null / empty argument");
  if ((smooth_lightness && !prev_L) || (smooth_chroma && (!prev_a || !prev_b))) return api_fail(NST_ERR_INVALID, "lab_ema: the enabled channels need their state planes");
  LabEmaArgs a;
  a.frames = frames; a.out = out; a.rgb2lab = rgb2lab; a.lab2rgb = lab2rgb; a.prev_L = prev_L; a.prev_a = prev_a; a.prev_b = prev_b;
  a.N = N; a.H = H; a.W = W; a.init = init != 0; a.smooth_L = smooth_lightness != 0; a.smooth_ab = smooth_chroma != 0;
  a.aL = (float)smooth_alpha; a.aL_comp = (float)(1.0 - smooth_alpha);
  a.aC = (float)chroma_alpha; a.aC_comp = (float)(1.0 - chroma_alpha);
  lab_ema_kernel<<<grid_for(H * W), 256, 0, (cudaStream_t)stream>>>(a);
  NST_CUDA_TRY(cudaGetLastError());
  return NST_OK;
}

extern "C" int nst_blend_pack_u8(const uint8_t* styled, const uint8_t* original, const float* mask, double blend, int composite_mode, uint8_t* out,
                                 int N, int H, int W, void* stream) {
  if (!styled || !original || !out || N < 1 || H < 1 || W < 1) return api_fail(NST_ERR_INVALID, "blend_pack: null / empty argument");
  BlendPackArgs a;
  a.styled = styled; a.original = original; a.mask = mask; a.out = out; a.npix = (size_t)N * H * W; a.composite_mode = composite_mode;
  const bool on = blend >= 0.0 && blend < 1.0;
  a.blend = on ? (float)blend : 2.f; a.blend_comp = (float)(1.0 - blend);       // Python computes 1.0 - blend in double before torch narrows it
  const auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };
  if (a.npix % 4 == 0 && al16(styled) && al16(original) && al16(out) && (!mask || al16(mask)))
    blend_pack_x4_kernel<<<(int)std::min<size_t>((a.npix / 4 + 255) / 256, (size_t)148 * 16), 256, 0, (cudaStream_t)stream>>>(a);
  else
    blend_pack_kernel<<<(int)std::min<size_t>((a.npix + 255) / 256, (size_t)148 * 16), 256, 0, (cudaStream_t)stream>>>(a);
  NST_CUDA_TRY(cudaGetLastError());
  return NST_OK;
}

extern "C" int nst_motion_blend_u8(const uint8_t* styled, const uint8_t* original, const float* flow, uint8_t* out, float* scratch, int H, int W,
                                   double blend, double min_alpha, double motion_norm, double sigma, void* stream) {
  if (!styled || !original || !flow || !out || !scratch || H < 1 || W < 1) return api_fail(NST_ERR_INVALID, "motion_blend: null / empty argument");
  if (!(sigma > 0) || !(motion_norm > 0)) return api_fail(NST_ERR_INVALID, "motion_blend: sigma and motion_norm must be positive");
  MotionBlendArgs a;
  memset(&a, 0, sizeof(a));
  const int k = gaussian_kernel_f32(sigma, a.kern, MB_MAX_K);
  if (k == 0) return api_fail(NST_ERR_INVALID, "motion_blend: sigma %.2f needs more than %d taps", sigma, MB_MAX_K);
  a.styled = styled; a.original = original; a.flow = flow; a.out = out; a.rowf = scratch; a.H = H; a.W = W; a.k = k;
  a.norm_div = (float)motion_norm;
  a.max_alpha = (float)blend; a.alpha_span = (float)(blend - min_alpha);
  cudaStream_t st = (cudaStream_t)stream;
  motion_row_kernel<<<dim3((W + 255) / 256, H), 256, 0, st>>>(a);
  motion_col_blend_kernel<<<grid_for(H * W), 256, 0, st>>>(a);
  NST_CUDA_TRY(cudaGetLastError());
  return NST_OK;
}

// ---------------------------------------------------------------------------------------------
// allocation-free variant of nst_motion_blend_u8 (no counterpart in the reference)
// ---------------------------------------------------------------------------------------------
// The row-filtered motion plane lives in a file-scope device array instead of caller scratch, for callers that cannot allocate
// (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  Every GPU that loads the library holds the array
// (NST_MOTION_STATIC_MAX_PIXELS floats = 8 MiB) until the process exits.  `g_motion_rowf` is only a host-side stand-in in host code:
// the device address is resolved once per device with cudaGetSymbolAddress and cached.
namespace nst {
constexpr int MOTION_STATIC_MAX_PIXELS = NST_MOTION_STATIC_MAX_PIXELS;
static __device__ __align__(256) float g_motion_rowf[MOTION_STATIC_MAX_PIXELS];
constexpr int MOTION_MAX_DEVICES = 64;
static float* g_motion_rowf_addr[MOTION_MAX_DEVICES];       // device address of g_motion_rowf on each device, resolved on first use
}  // namespace nst

extern "C" int nst_motion_blend_u8_noalloc(const uint8_t* styled, const uint8_t* original, const float* flow, uint8_t* out, int H, int W,
                                           double blend, double min_alpha, double motion_norm, double sigma, void* stream) {
  if (H < 1 || W < 1) return api_fail(NST_ERR_INVALID, "motion_blend_noalloc: empty frame");
  if ((long long)H * W > (long long)MOTION_STATIC_MAX_PIXELS)
    return api_fail(NST_ERR_INVALID, "motion_blend_noalloc: %d x %d = %lld pixels exceed the static scratch of %d pixels; use nst_motion_blend_u8 with caller scratch",
                    H, W, (long long)H * W, MOTION_STATIC_MAX_PIXELS);
  int dev = -1;
  NST_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= MOTION_MAX_DEVICES) return api_fail(NST_ERR_INVALID, "motion_blend_noalloc: device ordinal %d out of range", dev);
  float* rowf = __atomic_load_n(&g_motion_rowf_addr[dev], __ATOMIC_ACQUIRE);
  if (rowf == nullptr) {                     // once per device, with that device current (two racing threads resolve the same address)
    NST_CUDA_TRY(cudaGetSymbolAddress((void**)&rowf, g_motion_rowf));
    if (rowf == nullptr) return api_fail(NST_ERR_CUDA, "motion_blend_noalloc: cudaGetSymbolAddress returned a null address");
    __atomic_store_n(&g_motion_rowf_addr[dev], rowf, __ATOMIC_RELEASE);
  }
  // the row pass writes every one of the H*W entries the column pass reads: nothing in the array has to start at zero
  return nst_motion_blend_u8(styled, original, flow, out, rowf, H, W, blend, min_alpha, motion_norm, sigma, stream);
}
