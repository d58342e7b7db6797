// Direct fp32 convolution: the validation / bring-up kernel of the hot path.
//
// Same layer contract as the tcgen05 kernel (conv_tc.cu): prologue-fused InstanceNorm apply +
// ReLU + residual, reflection / zero padding, nearest-upsample and zero-insert folded into the
// operand fetch, statistics reduced in the epilogue -- but with plain FFMA accumulation, no
// shared-memory tiling and no tensor cores.  It exists so that (a) every layer shape of the three
// generator families has a CUDA path from day one, (b) the tcgen05 kernel can be checked against
// an independent GPU implementation, (c) precision NST_PREC_FP32_DIRECT gives an fp32-exact mode.
// It is NOT the benchmarked path.
#include "nst_kernels.cuh"

namespace nst {

constexpr int DIRECT_THREADS = 128;
constexpr int DIRECT_CO = 16;       // output channels accumulated per pass

// Fetch one transformed input element (n, sy, sx, c).  smem_coeff holds (scale, shift) per channel.
__device__ __forceinline__ float fetch_in(const ConvParams& p, int n, int sy, int sx, int c,
                                          const float2* coeff) {
  if (p.in_fmt == IN_F32_NHWC) {
    const size_t idx = act_index(n, c, sy, sx, p.Cin, p.Hs, p.Ws);
    float v = ld_act(p.src, idx, p.act16);
    if (p.in_coeff != nullptr) v = v * coeff[c].x + coeff[c].y;
    if (p.pre_relu) v = fmaxf(v, 0.f);
    if (p.resid != nullptr) v += ld_act(p.resid, idx, p.act16);
    if (p.post_relu) v = fmaxf(v, 0.f);
    return v;
  }
  if (p.in_fmt == IN_F32_NCHW3) {
    return ((const float*)p.src)[(((size_t)n * 3 + c) * p.Hs + sy) * p.Ws + sx];
  }
  const int cs = p.swap_rb ? 2 - c : c;
  const float u = (float)((const uint8_t*)p.src)[(((size_t)n * p.Hs + sy) * p.Ws + sx) * 3 + cs];
  return u * p.in_scale[c] + p.in_shift[c];
}

// wdir: [ks*ks][Cin][Cout] fp32
__global__ void __launch_bounds__(DIRECT_THREADS)
conv_direct_kernel(const __grid_constant__ ConvParams p, const float* __restrict__ wdir) {
  extern __shared__ float2 coeff[];     // [Cin]
  const int n = blockIdx.y;
  for (int c = threadIdx.x; c < p.Cin; c += DIRECT_THREADS) {
    float s, h;
    norm_coeff(p, n, c, s, h);
    coeff[c] = make_float2(s, h);
  }
  __syncthreads();

  const int pos = blockIdx.x * DIRECT_THREADS + threadIdx.x;
  const bool valid = pos < p.Ho * p.Wo;
  const int oy = valid ? pos / p.Wo : 0;
  const int ox = valid ? pos - oy * p.Wo : 0;
  const int lane = threadIdx.x & 31;

  // materialise the transformed input (residual trunk) -- identity-mapped stride-1 layers only
  if (p.side_out != nullptr && valid) {
    for (int c = 0; c < p.Cin; ++c) {
      st_act(p.side_out, act_index(n, c, oy, ox, p.Cin, p.Hs, p.Ws), fetch_in(p, n, oy, ox, c, coeff), p.act16);
    }
  }

  float img[3] = {0.f, 0.f, 0.f};
  for (int co0 = 0; co0 < p.Cout; co0 += DIRECT_CO) {
    float acc[DIRECT_CO];
#pragma unroll
    for (int j = 0; j < DIRECT_CO; ++j) acc[j] = 0.f;
    if (valid) {
      for (int r = 0; r < p.ks; ++r) {
        const int sy = map_coord((oy + p.oy_off) * p.stride + r - p.pad_t, p.Hc, p.Hs, p.map_mode, p.prepad, p.pad_mode);
        if (sy < 0) continue;
        for (int s = 0; s < p.ks; ++s) {
          const int sx = map_coord((ox + p.ox_off) * p.stride + s - p.pad_l, p.Wc, p.Ws, p.map_mode, p.prepad, p.pad_mode);
          if (sx < 0) continue;
          const float* wt = wdir + (size_t)(r * p.ks + s) * p.Cin * p.Cout + co0;
          for (int c = 0; c < p.Cin; ++c) {
            const float x = fetch_in(p, n, sy, sx, c, coeff);
#pragma unroll
            for (int j = 0; j < DIRECT_CO; ++j) {
              if (co0 + j < p.Cout) acc[j] = fmaf(x, wt[(size_t)c * p.Cout + j], acc[j]);
            }
          }
        }
      }
    }
    if (p.out_fmt == OUT_F32_NHWC_STATS) {
      if (valid) {
#pragma unroll
        for (int j = 0; j < DIRECT_CO; ++j) if (co0 + j < p.Cout) st_act(p.out, act_index(n, co0 + j, oy, ox, p.Cout, p.Ho, p.Wo), acc[j], p.out16);
      }
#pragma unroll
      for (int j = 0; j < DIRECT_CO; ++j) {
        float s1 = valid ? acc[j] : 0.f;
        float s2 = s1 * s1;
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
          s1 += __shfl_xor_sync(0xffffffffu, s1, off);
          s2 += __shfl_xor_sync(0xffffffffu, s2, off);
        }
        if (lane == 0 && co0 + j < p.Cout) {
          double* st = p.out_part + ((size_t)n * p.Cout + co0 + j) * 2;      // single part, zeroed by the host before the launch
          atomicAdd(st, (double)s1);
          atomicAdd(st + 1, (double)s2);
        }
      }
    } else {
#pragma unroll
      for (int j = 0; j < 3; ++j) if (co0 + j < p.Cout && co0 == 0) img[j] = acc[j] + p.bias[j];
    }
  }
  if (p.out_fmt != OUT_F32_NHWC_STATS && valid) store_image(p, n, oy, ox, img[0], img[1], img[2]);
}

cudaError_t launch_conv_direct(const ConvParams& p, const float* wdir, cudaStream_t stream) {
  dim3 grid((p.Ho * p.Wo + DIRECT_THREADS - 1) / DIRECT_THREADS, p.N);
  conv_direct_kernel<<<grid, DIRECT_THREADS, sizeof(float2) * p.Cin, stream>>>(p, wdir);
  return cudaGetLastError();
}

}  // namespace nst
