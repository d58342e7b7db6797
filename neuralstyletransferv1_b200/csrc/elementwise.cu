// Memory-bound companions of the conv kernels: multi-style region compositing.
//
// Replaces region_blend.composite_regions / composite_regions_advanced
// (region_blend.py:1049-1108, 1589-1679):
//   out = clamp( sum_r m_r * (sum_j w[r][j] * S[idx[r][j]]) / max(sum_r m_r, 1e-6), 0, 1 )
// One thread per pixel, every source read exactly once, coalesced; HBM-bound by construction.
#include "nst_kernels.cuh"

namespace nst {

constexpr int COMP_MAX_R = 16;
constexpr int COMP_MAX_J = 4;
constexpr int COMP_MAX_K = 8;

struct CompositeArgs {
  const void* srcs[COMP_MAX_K];
  const void* original;
  const float* masks;       // [R,H,W]
  void* out;
  int K, R, J, H, W;
  int idx[COMP_MAX_R * COMP_MAX_J];
  float w[COMP_MAX_R * COMP_MAX_J];
};

// U8 = true : sources / output uint8 NHWC (value/255, truncating pack)
// U8 = false: sources / output fp32 CHW planes in 0..1
template <bool U8>
__global__ void __launch_bounds__(256) composite_kernel(const __grid_constant__ CompositeArgs a) {
  const size_t hw = (size_t)a.H * a.W;
  for (size_t pix = (size_t)blockIdx.x * blockDim.x + threadIdx.x; pix < hw; pix += (size_t)gridDim.x * blockDim.x) {
    float src[COMP_MAX_K + 1][3];
#pragma unroll
    for (int k = 0; k <= COMP_MAX_K; ++k) {
      const void* sp = (k < COMP_MAX_K) ? ((k < a.K) ? a.srcs[k] : nullptr) : a.original;
      if (sp != nullptr) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          if (U8) src[k][c] = (float)((const uint8_t*)sp)[pix * 3 + c] / 255.0f;
          else src[k][c] = ((const float*)sp)[(size_t)c * hw + pix];
        }
      } else {
        src[k][0] = src[k][1] = src[k][2] = 0.f;
      }
    }
    float acc[3] = {0.f, 0.f, 0.f};
    float wsum = 0.f;
    for (int r = 0; r < a.R; ++r) {
      const float m = a.masks[(size_t)r * hw + pix];
      float rb[3] = {0.f, 0.f, 0.f};
      for (int j = 0; j < a.J; ++j) {
        const int id = a.idx[r * a.J + j];
        const float wt = a.w[r * a.J + j];
        if (wt == 0.f && a.J > 1) continue;
        const int k = id < 0 ? COMP_MAX_K : id;
#pragma unroll
        for (int kk = 0; kk <= COMP_MAX_K; ++kk) {
          if (kk == k) { rb[0] += wt * src[kk][0]; rb[1] += wt * src[kk][1]; rb[2] += wt * src[kk][2]; }
        }
      }
      acc[0] += rb[0] * m; acc[1] += rb[1] * m; acc[2] += rb[2] * m;
      wsum += m;
    }
    wsum = fmaxf(wsum, 1e-6f);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float o = fminf(fmaxf(acc[c] / wsum, 0.f), 1.f);
      if (U8) ((uint8_t*)a.out)[pix * 3 + c] = (uint8_t)(o * 255.0f);
      else ((float*)a.out)[(size_t)c * hw + pix] = o;
    }
  }
}

// uint8 fast path: 4 pixels (12 bytes = three 32-bit words per source, one float4 per mask) per thread; same arithmetic.
__global__ void __launch_bounds__(256) composite_u8x4_kernel(const __grid_constant__ CompositeArgs a) {
  const size_t hw = (size_t)a.H * a.W;
  const size_t nq = hw >> 2;
  for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < nq; q += (size_t)gridDim.x * blockDim.x) {
    float acc[12], wsum[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < 12; ++i) acc[i] = 0.f;
    for (int r = 0; r < a.R; ++r) {
      const float4 m4 = __ldg(reinterpret_cast<const float4*>(a.masks + (size_t)r * hw) + q);
      const float m[4] = {m4.x, m4.y, m4.z, m4.w};
      float rb[12];
#pragma unroll
      for (int i = 0; i < 12; ++i) rb[i] = 0.f;
      for (int j = 0; j < a.J; ++j) {
        const int id = a.idx[r * a.J + j];
        const float wt = a.w[r * a.J + j];
        if (wt == 0.f && a.J > 1) continue;
        const uint32_t* sp = reinterpret_cast<const uint32_t*>(id < 0 ? a.original : a.srcs[id]) + q * 3;   // L1/L2 serve the re-reads
        const uint32_t w3[3] = {__ldg(sp), __ldg(sp + 1), __ldg(sp + 2)};
#pragma unroll
        for (int i = 0; i < 12; ++i) rb[i] += wt * ((float)((w3[i >> 2] >> (8 * (i & 3))) & 0xFFu) / 255.0f);
      }
#pragma unroll
      for (int i = 0; i < 12; ++i) acc[i] += rb[i] * m[i / 3];
#pragma unroll
      for (int i = 0; i < 4; ++i) wsum[i] += m[i];
    }
    uint32_t o3[3] = {0u, 0u, 0u};
#pragma unroll
    for (int i = 0; i < 12; ++i) {
      const float o = fminf(fmaxf(acc[i] / fmaxf(wsum[i / 3], 1e-6f), 0.f), 1.f);
      o3[i >> 2] |= (uint32_t)(uint8_t)(o * 255.0f) << (8 * (i & 3));
    }
    uint32_t* op = reinterpret_cast<uint32_t*>(a.out) + q * 3;
    op[0] = o3[0]; op[1] = o3[1]; op[2] = o3[2];
  }
}

// ---------------------------------------------------------------------------------------------
// InstanceNorm statistics: per-CTA partial (sum, sumsq) -> per-(n, c) (scale, shift) of the consumer layer
// ---------------------------------------------------------------------------------------------
// torch.nn.InstanceNorm2d(affine=True) in eval with instance statistics (transformer_net.py:9,66-68; model.py:30-35):
//   y = (x - mean) / sqrt(var_biased + 1e-5) * gamma + beta  =  x * scale + shift.
// The conv epilogue leaves one partial per CTA and sample (conv_tc.cu) -- the CTAs whose contiguous tile run intersects sample n
// are summed here in index order, in double, so the result does not depend on scheduling.  grid_parts == 1: a single part that
// holds the whole sum (direct kernel).
struct StatsFinalizeArgs {
  const double* part;       // [parts][N][C][2]
  const float* gamma;       // [C]
  const float* beta;        // [C]
  float2* coeff;            // [N][C]
  int rot;                  // round-robin assignment: positions of full round i are rotated by i * rot (TcPlan::rr_rot)
  int rr;                   // tile -> CTA assignment of the producing launch: 1 = CTA b takes tiles b, b + parts, ...; 0 = contiguous runs
  int parts, groups, N, C;  // part table [parts * groups][N][C][2]: slot (CTA b, epilogue group g) = b * groups + g
  int tiles_per_img;        // tiles of one sample; CTA b owns tiles [b*T/parts, (b+1)*T/parts), T = N * tiles_per_img; its group g drains
                            // the tiles at odd/even offsets from the start of that run (offset % groups == g)
  double inv_count;         // 1 / (H*W)
};

// One warp per (n, c): lane l adds the slots of CTAs l, l + 32, ... in index order, then a fixed butterfly joins the 32 lanes -- the
// summation tree depends only on (parts, groups), never on scheduling.  Which slots were written is decided first, for a batch of
// CTAs at a time, so that the loads are independent and in flight together (this kernel sits between two dependent layers: the
// first version chained ~10 L2 round trips per lane and took 9 us).
__device__ __forceinline__ bool slot_written(const StatsFinalizeArgs& a, int b, int g, int lo, int hi, int T) {
  if (a.rr && a.rot) {                   // round r: tile r*parts + (b + r*rot) % parts (identity in the last, partial round)
    // sample n's tiles [lo, hi) span rounds r0..r1; every round strictly between them lies inside the sample, so CTA b has a
    // tile there -- only the two boundary rounds need the membership test
    const int r0 = lo / a.parts, r1 = (hi - 1) / a.parts;
    const int ri = r0 + 1 + ((g - (r0 + 1)) % a.groups + a.groups) % a.groups;      // first interior round drained by group g
    if (ri < r1) return true;
    for (int k = 0; k < 2; ++k) {
      const int r = k ? r1 : r0;
      if (r % a.groups != g || (k && r1 == r0)) continue;
      const int tl = (r + 1) * a.parts <= T ? r * a.parts + (b + r * a.rot) % a.parts : r * a.parts + b;
      if (tl >= lo && tl < hi && tl < T) return true;
    }
    return false;
  }
  int from, to;                          // [from, to): indices i of CTA b's tile sequence that fall into sample n
  if (a.rr) {                            // CTA b drains tiles b + i * parts
    from = lo > b ? (lo - b + a.parts - 1) / a.parts : 0;
    to = hi > b ? (hi - b + a.parts - 1) / a.parts : 0;
  } else {                               // CTA b drains tiles [b*T/parts, (b+1)*T/parts)
    const int t0 = (int)((long long)b * T / a.parts), t1 = (int)((long long)(b + 1) * T / a.parts);
    from = (lo > t0 ? lo : t0) - t0;
    to = (hi < t1 ? hi : t1) - t0;
  }
  const int first = from + ((g - from) % a.groups + a.groups) % a.groups;      // epilogue group g drains the indices with i % groups == g
  return first < to;
}

__global__ void __launch_bounds__(128) stats_finalize_kernel(const __grid_constant__ StatsFinalizeArgs a) {
  const int i = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (i >= a.N * a.C) return;
  const int n = i / a.C, c = i - n * a.C;
  double s1 = 0.0, s2 = 0.0;
  if (a.parts == 1 && a.groups == 1) {
    if (lane == 0) { const double* st = a.part + ((size_t)n * a.C + c) * 2; s1 = st[0]; s2 = st[1]; }
  } else {
    const int T = a.N * a.tiles_per_img, lo = n * a.tiles_per_img, hi = lo + a.tiles_per_img;
    constexpr int KB = 5, KG = 2;          // CTAs per lane and epilogue groups per batch: 5 * 32 = 160 >= 148 SMs in one batch
    for (int b0 = lane; b0 < a.parts; b0 += 32 * KB) {
      double2 v[KB][KG];
#pragma unroll
      for (int k = 0; k < KB; ++k)
#pragma unroll
        for (int g = 0; g < KG; ++g) {
          const int b = b0 + 32 * k;
          v[k][g] = make_double2(0.0, 0.0);
          if (b < a.parts && g < a.groups && slot_written(a, b, g, lo, hi, T))
            v[k][g] = *reinterpret_cast<const double2*>(a.part + ((((size_t)b * a.groups + g) * a.N + n) * a.C + c) * 2);
        }
#pragma unroll
      for (int k = 0; k < KB; ++k)
#pragma unroll
        for (int g = 0; g < KG; ++g) { s1 += v[k][g].x; s2 += v[k][g].y; }
    }
  }
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    s1 += __shfl_xor_sync(0xffffffffu, s1, off);
    s2 += __shfl_xor_sync(0xffffffffu, s2, off);
  }
  if (lane != 0) return;
  const double mean = s1 * a.inv_count;
  double var = s2 * a.inv_count - mean * mean;
  if (var < 0.0) var = 0.0;
  const float inv = (float)(1.0 / sqrt(var + (double)kInEps));
  const float scale = a.gamma[c] * inv;
  a.coeff[i] = make_float2(scale, a.beta[c] - (float)mean * scale);
}

cudaError_t launch_stats_finalize(const double* part, int parts, int groups, int rr, int rot, int N, int C, int tiles_per_img, double inv_count,
                                  const float* gamma, const float* beta, float2* coeff, cudaStream_t stream) {
  StatsFinalizeArgs a;
  a.part = part; a.gamma = gamma; a.beta = beta; a.coeff = coeff;
  a.parts = parts; a.groups = groups; a.rr = rr; a.rot = rot; a.N = N; a.C = C; a.tiles_per_img = tiles_per_img; a.inv_count = inv_count;
  stats_finalize_kernel<<<(N * C + 3) / 4, 128, 0, stream>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_composite(const CompositeArgs& a, bool u8, cudaStream_t stream) {
  if (u8 && ((size_t)a.H * a.W) % 4 == 0) {
    bool aligned = ((uintptr_t)a.out % 4 == 0) && ((uintptr_t)a.masks % 16 == 0) && (a.original == nullptr || (uintptr_t)a.original % 4 == 0);
    for (int k = 0; k < a.K; ++k) aligned = aligned && ((uintptr_t)a.srcs[k] % 4 == 0);
    if (aligned) {
      const size_t nq = (size_t)a.H * a.W / 4;
      int blocks = (int)((nq + 255) / 256);
      if (blocks > 148 * 8) blocks = 148 * 8;
      composite_u8x4_kernel<<<blocks, 256, 0, stream>>>(a);
      return cudaGetLastError();
    }
  }
  const size_t hw = (size_t)a.H * a.W;
  int blocks = (int)((hw + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (u8) composite_kernel<true><<<blocks, 256, 0, stream>>>(a);
  else composite_kernel<false><<<blocks, 256, 0, stream>>>(a);
  return cudaGetLastError();
}

}  // namespace nst
