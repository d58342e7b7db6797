// Memory-bound companions of the generators, second set: everything the reference does to a frame between the network
// output and the encoder, as coalesced one-pass kernels (HBM roofline).
//
//   resize_post_pack   pipeline.py:1512-1516 (bilinear "lock stylised to content size") -> :2040-2043 mask composite ->
//                      :2088-2092 uniform blend -> :2094 ToPILImage pack, for frame sizes that are not multiples of 4
//   composite_batch    region_blend.composite_regions / composite_regions_advanced (region_blend.py:1049-1108, 1589-1679)
//                      over a BATCH of frames in one launch, masks fp32 / fp16 / uint8, per-scale bilinear upscale of
//                      half-resolution sources (:1662-1668)
//   composite_crops    region_blend.composite_from_crops + place_crop (region_blend.py:2080-2109, 2186-2294): per-region padded
//                      bounding boxes accumulated on a canvas, gaps (weight < 0.1) filled with the original, normalised
//   crops_dilate       the no-original branch of the same function (:2271-2289): max-pool 5 / 11 / 21 gap fill
//   flow_ema           pipeline.py:425-439 `_warp_with_flow` (cv2.remap, bilinear, replicate border) + :1927-1931 EMA lerp
//
// Every arithmetic step keeps the reference's fp32 operation order (separately rounded multiplies and adds where torch / numpy
// evaluate tensor expressions op by op; fused multiply-adds where ATen's vectorised bilinear kernel contracts them -- checked
// bit-for-bit against torch 2.11 CPU in tests/test_oracle.py), so the uint8 results are the reference's.
#include "nst_kernels.cuh"

namespace nst {

// ---------------------------------------------------------------------------------------------
// torch.nn.functional.interpolate(mode="bilinear", align_corners=False) source index / weight of one output coordinate
// (ATen UpSampleKernel: area_pixel_compute_source_index + guard_index_and_lambda, fp32, the multiply-add contracted)
// ---------------------------------------------------------------------------------------------
struct Lerp { int i0, i1; float w0, w1; };
__device__ __forceinline__ Lerp bilinear_coord(int dst, int in_size, int out_size) {
  Lerp l;
  const float scale = (float)in_size / (float)out_size;
  float real = fmaf(scale, (float)dst + 0.5f, -0.5f);
  real = fmaxf(real, 0.f);
  l.i0 = min((int)floorf(real), in_size - 1);
  l.w1 = fminf(fmaxf(real - (float)l.i0, 0.f), 1.f);
  l.i1 = l.i0 + (l.i0 < in_size - 1 ? 1 : 0);
  l.w0 = 1.f - l.w1;
  return l;
}
// wy0 * (wx0*a + wx1*b) + wy1 * (wx0*c + wx1*d) in ATen's contraction: fma(w0, v0, w1 * v1) at both levels
__device__ __forceinline__ float bilinear_mix(const Lerp& ly, const Lerp& lx, float a, float b, float c, float d) {
  const float top = fmaf(lx.w0, a, __fmul_rn(lx.w1, b));
  const float bot = fmaf(lx.w0, c, __fmul_rn(lx.w1, d));
  return fmaf(ly.w0, top, __fmul_rn(ly.w1, bot));
}

// the tail of store_image (nst_kernels.cuh): out01 -> mask composite -> blend -> truncating pack, one channel
__device__ __forceinline__ uint8_t post_pack(float o01, float og, bool has_mask, float a, int composite_mode, float blend, float blend_comp) {
  if (has_mask) {
    const float na = __fsub_rn(1.0f, a);
    o01 = (composite_mode == 0) ? __fadd_rn(__fmul_rn(a, o01), __fmul_rn(na, og)) : __fadd_rn(__fmul_rn(na, o01), __fmul_rn(a, og));
    o01 = fminf(fmaxf(o01, 0.f), 1.f);
  }
  if (blend >= 0.f && blend < 1.f) {
    o01 = __fadd_rn(__fmul_rn(blend, o01), __fmul_rn(blend_comp, og));
    o01 = fminf(fmaxf(o01, 0.f), 1.f);
  }
  return (uint8_t)__fmul_rn(o01, 255.0f);
}

struct ResizePackArgs {
  const float* src01;       // [N,3,Hi,Wi] clamped out01 planes
  const uint8_t* orig;      // [N,H,W,3] original frames (mask / blend) or null
  const float* mask;        // [N,H,W] or null
  uint8_t* out;             // [N,H,W,3]
  int N, Hi, Wi, H, W;
  float blend, blend_comp;
  int composite_mode;
};

__global__ void __launch_bounds__(256) resize_post_pack_kernel(const __grid_constant__ ResizePackArgs a) {
  const size_t hw = (size_t)a.H * a.W, total = hw * a.N;
  const size_t in_plane = (size_t)a.Hi * a.Wi;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int n = (int)(i / hw);
    const int pix = (int)(i - (size_t)n * hw);
    const int y = pix / a.W, x = pix - y * a.W;
    const Lerp ly = bilinear_coord(y, a.Hi, a.H), lx = bilinear_coord(x, a.Wi, a.W);
    const float* s = a.src01 + (size_t)n * 3 * in_plane;
    const float al = a.mask ? a.mask[i] : 0.f;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float* pl = s + (size_t)c * in_plane;
      const float v = bilinear_mix(ly, lx, pl[(size_t)ly.i0 * a.Wi + lx.i0], pl[(size_t)ly.i0 * a.Wi + lx.i1], pl[(size_t)ly.i1 * a.Wi + lx.i0],
                                   pl[(size_t)ly.i1 * a.Wi + lx.i1]);
      const float og = a.orig ? u8_to_unit((float)a.orig[i * 3 + c]) : 0.f;
      a.out[i * 3 + c] = post_pack(v, og, a.mask != nullptr && a.orig != nullptr, al, a.composite_mode, a.orig ? a.blend : 2.f, a.blend_comp);
    }
  }
}

// Identity size (the frame already has the content size: every multiple-of-4 frame, and the pack stage of post_chain.PostChain), H*W a multiple
// of 4: four pixels per thread -- three 16-byte plane loads, 12 bytes of original, one 16-byte mask load, three 32-bit stores.  The bilinear
// weights of an identity resize are exactly (1, 0), so the resized value IS the source value: same bytes as the general kernel.
__global__ void __launch_bounds__(256) post_pack_x4_kernel(const __grid_constant__ ResizePackArgs a) {
  const size_t hw = (size_t)a.H * a.W, quads = hw / 4, total = quads * a.N;
  const bool has_mask = a.mask != nullptr && a.orig != nullptr;
  const float blend = a.orig ? a.blend : 2.f;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const size_t n = i / quads, q = i - n * quads;
    const float* s = a.src01 + n * 3 * hw + q * 4;
    float v[3][4];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float4 t = *reinterpret_cast<const float4*>(s + (size_t)c * hw);
      v[c][0] = t.x; v[c][1] = t.y; v[c][2] = t.z; v[c][3] = t.w;
    }
    uint32_t ow[3] = {0u, 0u, 0u};
    if (a.orig) {
      const uint32_t* o32 = reinterpret_cast<const uint32_t*>(a.orig + (n * hw + q * 4) * 3);
      ow[0] = o32[0]; ow[1] = o32[1]; ow[2] = o32[2];
    }
    float al[4] = {0.f, 0.f, 0.f, 0.f};
    if (a.mask) { const float4 t = *reinterpret_cast<const float4*>(a.mask + n * hw + q * 4); al[0] = t.x; al[1] = t.y; al[2] = t.z; al[3] = t.w; }
    uint32_t outw[3] = {0u, 0u, 0u};
#pragma unroll
    for (int j = 0; j < 12; ++j) {                       // byte j of the 12-byte group = pixel j / 3, channel j % 3
      const float og = a.orig ? u8_to_unit((float)((ow[j >> 2] >> ((j & 3) * 8)) & 0xFFu)) : 0.f;
      const uint8_t b = post_pack(v[j % 3][j / 3], og, has_mask, al[j / 3], a.composite_mode, blend, a.blend_comp);
      outw[j >> 2] |= (uint32_t)b << ((j & 3) * 8);
    }
    uint32_t* d32 = reinterpret_cast<uint32_t*>(a.out + (n * hw + q * 4) * 3);
    d32[0] = outw[0]; d32[1] = outw[1]; d32[2] = outw[2];
  }
}

cudaError_t launch_resize_post_pack(const ResizePackArgs& a, cudaStream_t stream) {
  const size_t total = (size_t)a.N * a.H * a.W;
  const auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };
  if (a.Hi == a.H && a.Wi == a.W && ((size_t)a.H * a.W) % 4 == 0 && al16(a.src01) && al16(a.out) && (!a.orig || al16(a.orig)) && (!a.mask || al16(a.mask))) {
    const size_t quads = total / 4;
    post_pack_x4_kernel<<<(int)std::min<size_t>((quads + 255) / 256, (size_t)148 * 16), 256, 0, stream>>>(a);
    return cudaGetLastError();
  }
  int blocks = (int)std::min<size_t>((total + 255) / 256, (size_t)148 * 16);
  resize_post_pack_kernel<<<blocks, 256, 0, stream>>>(a);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// batched region composite
// ---------------------------------------------------------------------------------------------
constexpr int CB_MAX_R = 16, CB_MAX_J = 4, CB_MAX_K = 8;
enum MaskType { MASK_F32 = 0, MASK_F16 = 1, MASK_U8 = 2 };

struct CompositeBatchArgs {
  const void* srcs[CB_MAX_K];      // K sources, each [N, Hs_k, Ws_k, 3] uint8 (value / 255) -- a source smaller than H x W is
  int src_h[CB_MAX_K], src_w[CB_MAX_K];   // upscaled bilinearly per pixel (region_blend.py:1662-1668)
  const uint8_t* original;         // [N,H,W,3] or null
  const void* masks;               // [R,H,W] (mask_frame_stride 0) or [N,R,H,W], element type `mask_type`
  uint8_t* out;                    // [N,H,W,3]
  int N, K, R, J, H, W;
  int mask_type;
  size_t mask_frame_stride;        // elements between the mask sets of consecutive frames (0 = one set for the whole batch)
  int idx[CB_MAX_R * CB_MAX_J];
  float w[CB_MAX_R * CB_MAX_J];
};

template <int MT>
__device__ __forceinline__ float load_mask(const void* base, size_t i) {
  if (MT == MASK_F32) return __ldg((const float*)base + i);
  if (MT == MASK_F16) return __half2float(__ldg((const __half*)base + i));
  return u8_to_unit((float)__ldg((const uint8_t*)base + i));
}

// one thread = one pixel of one frame; every full-resolution source byte is read once (coalesced 3-byte runs), masks once per frame
template <int MT>
__global__ void __launch_bounds__(256) composite_batch_kernel(const __grid_constant__ CompositeBatchArgs a) {
  const size_t hw = (size_t)a.H * a.W, total = hw * a.N;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int n = (int)(i / hw);
    const size_t pix = i - (size_t)n * hw;
    const int y = (int)(pix / a.W), x = (int)(pix - (size_t)y * a.W);
    float acc[3] = {0.f, 0.f, 0.f}, wsum = 0.f;
    const size_t mbase = (size_t)n * a.mask_frame_stride + pix;
    for (int r = 0; r < a.R; ++r) {
      const float m = load_mask<MT>(a.masks, mbase + (size_t)r * hw);
      float rb[3] = {0.f, 0.f, 0.f};
      for (int j = 0; j < a.J; ++j) {
        const int id = a.idx[r * a.J + j];
        const float wt = a.w[r * a.J + j];
        if (id < -1) continue;                                    // padding entry of a region with fewer than J terms
        float s[3];
        if (id < 0) {
#pragma unroll
          for (int c = 0; c < 3; ++c) s[c] = u8_to_unit((float)__ldg(a.original + i * 3 + c));
        } else {
          const uint8_t* sp = (const uint8_t*)a.srcs[id];
          const int sh = a.src_h[id], sw = a.src_w[id];
          if (sh == a.H && sw == a.W) {
#pragma unroll
            for (int c = 0; c < 3; ++c) s[c] = u8_to_unit((float)__ldg(sp + i * 3 + c));
          } else {
            const Lerp ly = bilinear_coord(y, sh, a.H), lx = bilinear_coord(x, sw, a.W);
            const uint8_t* f = sp + (size_t)n * sh * sw * 3;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
              const float v00 = u8_to_unit((float)__ldg(f + ((size_t)ly.i0 * sw + lx.i0) * 3 + c));
              const float v01 = u8_to_unit((float)__ldg(f + ((size_t)ly.i0 * sw + lx.i1) * 3 + c));
              const float v10 = u8_to_unit((float)__ldg(f + ((size_t)ly.i1 * sw + lx.i0) * 3 + c));
              const float v11 = u8_to_unit((float)__ldg(f + ((size_t)ly.i1 * sw + lx.i1) * 3 + c));
              s[c] = bilinear_mix(ly, lx, v00, v01, v10, v11);
            }
          }
        }
#pragma unroll
        for (int c = 0; c < 3; ++c) rb[c] = __fadd_rn(rb[c], __fmul_rn(wt, s[c]));          // region_blend += weight * source
      }
#pragma unroll
      for (int c = 0; c < 3; ++c) acc[c] = __fadd_rn(acc[c], __fmul_rn(rb[c], m));           // result += region_blend * mask
      wsum = __fadd_rn(wsum, m);
    }
    wsum = fmaxf(wsum, 1e-6f);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float o = fminf(fmaxf(__fdiv_rn(acc[c], wsum), 0.f), 1.f);
      a.out[i * 3 + c] = (uint8_t)__fmul_rn(o, 255.0f);
    }
  }
}

// Fast path (every source at full resolution, H*W a multiple of 4, 16-byte aligned rows): 4 pixels per thread -- three 32-bit words per
// source, one 16 / 8 / 4-byte mask load per region, three 32-bit stores.  Same arithmetic, same order, as the scalar kernel.
template <int MT>
__device__ __forceinline__ void load_mask4(const void* base, size_t i, float (&m)[4]) {
  if (MT == MASK_F32) { const float4 v = __ldg(reinterpret_cast<const float4*>((const float*)base + i)); m[0] = v.x; m[1] = v.y; m[2] = v.z; m[3] = v.w; }
  else if (MT == MASK_F16) {
    const uint2 v = __ldg(reinterpret_cast<const uint2*>((const __half*)base + i));
    const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&v.x)), b = __half22float2(*reinterpret_cast<const __half2*>(&v.y));
    m[0] = a.x; m[1] = a.y; m[2] = b.x; m[3] = b.y;
  } else {
    const uint32_t v = __ldg(reinterpret_cast<const uint32_t*>((const uint8_t*)base + i));
#pragma unroll
    for (int k = 0; k < 4; ++k) m[k] = u8_to_unit((float)((v >> (8 * k)) & 0xFFu));
  }
}
template <int MT>
__global__ void __launch_bounds__(256) composite_batch_x4_kernel(const __grid_constant__ CompositeBatchArgs a) {
  const size_t hw = (size_t)a.H * a.W, hwq = hw >> 2, total = hwq * a.N;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int n = (int)(i / hwq);
    const size_t q = i - (size_t)n * hwq;                       // group of 4 pixels inside the frame
    const size_t word0 = ((size_t)n * hw + q * 4) * 3 / 4;       // first 32-bit word of the group's 12 bytes
    float acc[12], wsum[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int k = 0; k < 12; ++k) acc[k] = 0.f;
    for (int r = 0; r < a.R; ++r) {
      float m[4];
      load_mask4<MT>(a.masks, (size_t)n * a.mask_frame_stride + (size_t)r * hw + q * 4, m);
      float rb[12];
#pragma unroll
      for (int k = 0; k < 12; ++k) rb[k] = 0.f;
      for (int j = 0; j < a.J; ++j) {
        const int id = a.idx[r * a.J + j];
        const float wt = a.w[r * a.J + j];
        if (id < -1) continue;
        const uint32_t* sp = reinterpret_cast<const uint32_t*>(id < 0 ? (const void*)a.original : a.srcs[id]) + word0;
        const uint32_t w3[3] = {__ldg(sp), __ldg(sp + 1), __ldg(sp + 2)};
#pragma unroll
        for (int k = 0; k < 12; ++k)
          rb[k] = __fadd_rn(rb[k], __fmul_rn(wt, u8_to_unit((float)((w3[k >> 2] >> (8 * (k & 3))) & 0xFFu))));
      }
#pragma unroll
      for (int k = 0; k < 12; ++k) acc[k] = __fadd_rn(acc[k], __fmul_rn(rb[k], m[k / 3]));
#pragma unroll
      for (int k = 0; k < 4; ++k) wsum[k] = __fadd_rn(wsum[k], m[k]);
    }
    uint32_t o3[3] = {0u, 0u, 0u};
#pragma unroll
    for (int k = 0; k < 12; ++k) {
      const float o = fminf(fmaxf(__fdiv_rn(acc[k], fmaxf(wsum[k / 3], 1e-6f)), 0.f), 1.f);
      o3[k >> 2] |= (uint32_t)(uint8_t)__fmul_rn(o, 255.0f) << (8 * (k & 3));
    }
    uint32_t* op = reinterpret_cast<uint32_t*>(a.out) + word0;
    op[0] = o3[0]; op[1] = o3[1]; op[2] = o3[2];
  }
}

cudaError_t launch_composite_batch(const CompositeBatchArgs& a, cudaStream_t stream) {
  const size_t hw = (size_t)a.H * a.W;
  bool fast = hw % 4 == 0 && ((uintptr_t)a.out % 4 == 0) && ((uintptr_t)a.masks % 16 == 0) && (a.original == nullptr || (uintptr_t)a.original % 4 == 0) &&
              (a.mask_frame_stride % 4 == 0);
  for (int k = 0; k < a.K; ++k) fast = fast && a.src_h[k] == a.H && a.src_w[k] == a.W && ((uintptr_t)a.srcs[k] % 4 == 0);
  if (fast) {
    const size_t total4 = (size_t)a.N * (hw / 4);
    const int blocks4 = (int)std::min<size_t>((total4 + 255) / 256, (size_t)148 * 16);
    if (a.mask_type == MASK_F32) composite_batch_x4_kernel<MASK_F32><<<blocks4, 256, 0, stream>>>(a);
    else if (a.mask_type == MASK_F16) composite_batch_x4_kernel<MASK_F16><<<blocks4, 256, 0, stream>>>(a);
    else composite_batch_x4_kernel<MASK_U8><<<blocks4, 256, 0, stream>>>(a);
    return cudaGetLastError();
  }
  const size_t total = (size_t)a.N * a.H * a.W;
  const int blocks = (int)std::min<size_t>((total + 255) / 256, (size_t)148 * 16);
  if (a.mask_type == MASK_F32) composite_batch_kernel<MASK_F32><<<blocks, 256, 0, stream>>>(a);
  else if (a.mask_type == MASK_F16) composite_batch_kernel<MASK_F16><<<blocks, 256, 0, stream>>>(a);
  else composite_batch_kernel<MASK_U8><<<blocks, 256, 0, stream>>>(a);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// crop-based composite (region_blend.py:2186-2294)
// ---------------------------------------------------------------------------------------------
struct CropTerm { const float* src; int sh, sw; float weight; };   // src: fp32 [3, sh, sw] styled crop, or null = the original's crop
struct CropRegion { int x1, y1, x2, y2; const float* mask; int mh, mw; int nterms; CropTerm term[CB_MAX_J]; };
struct CompositeCropsArgs {
  CropRegion reg[CB_MAX_R];
  int R, H, W;
  const float* original;    // fp32 [3,H,W] or null
  float* canvas;            // out: fp32 [3,H,W]   (final frame when `finish`, else the un-normalised accumulator)
  float* weight;            // out when !finish: fp32 [H,W] accumulated mask weight
  uint8_t* out_u8;          // optional: uint8 [H,W,3] truncating pack of the final frame (finish only)
  int finish;               // 1: gap fill with the original (if any), normalise, clamp;  0: leave canvas / weight for the dilation passes
};

__global__ void __launch_bounds__(256) composite_crops_kernel(const __grid_constant__ CompositeCropsArgs a) {
  const int hw = a.H * a.W;
  for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < hw; pix += gridDim.x * blockDim.x) {
    const int y = pix / a.W, x = pix - y * a.W;
    float cv[3] = {0.f, 0.f, 0.f}, ws = 0.f;
    for (int r = 0; r < a.R; ++r) {
      const CropRegion& g = a.reg[r];
      if (x < g.x1 || x >= g.x2 || y < g.y1 || y >= g.y2) continue;
      const int ch = g.y2 - g.y1, cw = g.x2 - g.x1, cy = y - g.y1, cx = x - g.x1;
      float rb[3] = {0.f, 0.f, 0.f};
      for (int j = 0; j < g.nterms; ++j) {
        const CropTerm& tm = g.term[j];
        float s[3];
        if (tm.src == nullptr) {                                   // extract_crop(original, padded_bbox)
#pragma unroll
          for (int c = 0; c < 3; ++c) s[c] = a.original[(size_t)c * hw + pix];
        } else if (tm.sh == ch && tm.sw == cw) {
#pragma unroll
          for (int c = 0; c < 3; ++c) s[c] = tm.src[((size_t)c * ch + cy) * cw + cx];
        } else {                                                   // F.interpolate(source, size=(crop_h, crop_w), bilinear)
          const Lerp ly = bilinear_coord(cy, tm.sh, ch), lx = bilinear_coord(cx, tm.sw, cw);
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            const float* pl = tm.src + (size_t)c * tm.sh * tm.sw;
            s[c] = bilinear_mix(ly, lx, pl[ly.i0 * tm.sw + lx.i0], pl[ly.i0 * tm.sw + lx.i1], pl[ly.i1 * tm.sw + lx.i0], pl[ly.i1 * tm.sw + lx.i1]);
          }
        }
#pragma unroll
        for (int c = 0; c < 3; ++c) rb[c] = __fadd_rn(rb[c], __fmul_rn(tm.weight, s[c]));
      }
      float m;
      if (g.mh == ch && g.mw == cw) m = g.mask[(size_t)cy * cw + cx];
      else {
        const Lerp ly = bilinear_coord(cy, g.mh, ch), lx = bilinear_coord(cx, g.mw, cw);
        m = bilinear_mix(ly, lx, g.mask[ly.i0 * g.mw + lx.i0], g.mask[ly.i0 * g.mw + lx.i1], g.mask[ly.i1 * g.mw + lx.i0], g.mask[ly.i1 * g.mw + lx.i1]);
      }
#pragma unroll
      for (int c = 0; c < 3; ++c) cv[c] = __fadd_rn(cv[c], __fmul_rn(rb[c], m));           // place_crop: canvas += crop * mask
      ws = __fadd_rn(ws, m);
    }
    if (!a.finish) {
#pragma unroll
      for (int c = 0; c < 3; ++c) a.canvas[(size_t)c * hw + pix] = cv[c];
      a.weight[pix] = ws;
      continue;
    }
    if (a.original != nullptr) {                                   // canvas + original * gap; weight + gap (gap = weight < 0.1)
      const float gap = ws < 0.1f ? 1.f : 0.f;
#pragma unroll
      for (int c = 0; c < 3; ++c) cv[c] = __fadd_rn(cv[c], __fmul_rn(a.original[(size_t)c * hw + pix], gap));
      ws = __fadd_rn(ws, gap);
    }
    ws = fmaxf(ws, 1e-6f);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float o = fminf(fmaxf(__fdiv_rn(cv[c], ws), 0.f), 1.f);
      a.canvas[(size_t)c * hw + pix] = o;
      if (a.out_u8) a.out_u8[(size_t)pix * 3 + c] = (uint8_t)__fmul_rn(o, 255.0f);
    }
  }
}

// One pass of the no-original gap fill: F.max_pool2d(k, stride 1, pad k/2) of canvas (per channel) and weight, applied where
// weight < 0.1:  canvas = canvas*(1-gap) + dilated*gap.  in -> out (ping-pong buffers); `last` also normalises and clamps.
struct CropsDilateArgs {
  const float* canvas_in; const float* weight_in;
  float* canvas_out; float* weight_out;
  uint8_t* out_u8;
  int H, W, k, last;
};
__global__ void __launch_bounds__(256) crops_dilate_kernel(const __grid_constant__ CropsDilateArgs a) {
  const int hw = a.H * a.W, rad = a.k / 2;
  for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < hw; pix += gridDim.x * blockDim.x) {
    const int y = pix / a.W, x = pix - y * a.W;
    float ws = a.weight_in[pix];
    float cv[3] = {a.canvas_in[pix], a.canvas_in[(size_t)hw + pix], a.canvas_in[2 * (size_t)hw + pix]};
    if (ws < 0.1f) {
      float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
      for (int yy = max(0, y - rad); yy <= min(a.H - 1, y + rad); ++yy)
        for (int xx = max(0, x - rad); xx <= min(a.W - 1, x + rad); ++xx) {
          const int q = yy * a.W + xx;
          mx[0] = fmaxf(mx[0], a.canvas_in[q]); mx[1] = fmaxf(mx[1], a.canvas_in[(size_t)hw + q]); mx[2] = fmaxf(mx[2], a.canvas_in[2 * (size_t)hw + q]);
          mx[3] = fmaxf(mx[3], a.weight_in[q]);
        }
      // canvas * (1 - 1) + dilated * 1, weight likewise (the reference's arithmetic with gap = 1)
#pragma unroll
      for (int c = 0; c < 3; ++c) cv[c] = __fadd_rn(__fmul_rn(cv[c], 0.f), mx[c]);
      ws = __fadd_rn(__fmul_rn(ws, 0.f), mx[3]);
    }
    if (a.last) {
      const float wn = fmaxf(ws, 1e-6f);
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float o = fminf(fmaxf(__fdiv_rn(cv[c], wn), 0.f), 1.f);
        a.canvas_out[(size_t)c * hw + pix] = o;
        if (a.out_u8) a.out_u8[(size_t)pix * 3 + c] = (uint8_t)__fmul_rn(o, 255.0f);
      }
    } else {
#pragma unroll
      for (int c = 0; c < 3; ++c) a.canvas_out[(size_t)c * hw + pix] = cv[c];
      a.weight_out[pix] = ws;
    }
  }
}

cudaError_t launch_composite_crops(const CompositeCropsArgs& a, cudaStream_t stream) {
  const int hw = a.H * a.W;
  composite_crops_kernel<<<std::min((hw + 255) / 256, 148 * 16), 256, 0, stream>>>(a);
  return cudaGetLastError();
}
cudaError_t launch_crops_dilate(const CropsDilateArgs& a, cudaStream_t stream) {
  const int hw = a.H * a.W;
  crops_dilate_kernel<<<std::min((hw + 255) / 256, 148 * 16), 256, 0, stream>>>(a);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// flow-guided EMA (pipeline.py:425-439, 1918-1931)
// ---------------------------------------------------------------------------------------------
// cv2.remap(prev, x + dx, y + dy, INTER_LINEAR, BORDER_REPLICATE) on float32 HWC images: OpenCV quantises the map to 1/32 pixel
// (INTER_BITS = 5: sx = cvRound(map * 32), integer part sx >> 5, fraction (sx & 31) / 32) and blends the four taps with the
// fp32 table weights (1-fy)(1-fx), (1-fy)fx, fy(1-fx), fy*fx summed left to right; then
// fused = clip(a * curr + (1 - a) * warped, 0, 1) in fp32 (numpy), `a` and `1 - a` narrowed from Python doubles.
struct FlowEmaArgs {
  const float* curr;        // [H,W,3] current styled frame, 0..1
  const float* prev;        // [H,W,3] previous fused frame
  const float* flow;        // [H,W,2] (dx, dy)
  float* out;               // [H,W,3] fused frame (may alias curr)
  int H, W;
  float alpha, alpha_comp;
};
__global__ void __launch_bounds__(256) flow_ema_kernel(const __grid_constant__ FlowEmaArgs a) {
  const int hw = a.H * a.W;
  for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < hw; pix += gridDim.x * blockDim.x) {
    const int y = pix / a.W, x = pix - y * a.W;
    const float2 f = reinterpret_cast<const float2*>(a.flow)[pix];
    const float mx = __fadd_rn((float)x, f.x), my = __fadd_rn((float)y, f.y);
    const int sx = __float2int_rn(__fmul_rn(mx, 32.f)), sy = __float2int_rn(__fmul_rn(my, 32.f));
    const int ix = sx >> 5, iy = sy >> 5;
    const float fx = (float)(sx & 31) * (1.f / 32.f), fy = (float)(sy & 31) * (1.f / 32.f);
    const float w00 = __fmul_rn(1.f - fy, 1.f - fx), w01 = __fmul_rn(1.f - fy, fx), w10 = __fmul_rn(fy, 1.f - fx), w11 = __fmul_rn(fy, fx);
    const int x0 = min(max(ix, 0), a.W - 1), x1 = min(max(ix + 1, 0), a.W - 1);
    const int y0 = min(max(iy, 0), a.H - 1), y1 = min(max(iy + 1, 0), a.H - 1);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float p00 = a.prev[((size_t)y0 * a.W + x0) * 3 + c], p01 = a.prev[((size_t)y0 * a.W + x1) * 3 + c];
      const float p10 = a.prev[((size_t)y1 * a.W + x0) * 3 + c], p11 = a.prev[((size_t)y1 * a.W + x1) * 3 + c];
      const float wv = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(p00, w00), __fmul_rn(p01, w01)), __fmul_rn(p10, w10)), __fmul_rn(p11, w11));
      const float fused = __fadd_rn(__fmul_rn(a.alpha, a.curr[(size_t)pix * 3 + c]), __fmul_rn(a.alpha_comp, wv));
      a.out[(size_t)pix * 3 + c] = fminf(fmaxf(fused, 0.f), 1.f);
    }
  }
}
cudaError_t launch_flow_ema(const FlowEmaArgs& a, cudaStream_t stream) {
  const int hw = a.H * a.W;
  flow_ema_kernel<<<std::min((hw + 255) / 256, 148 * 16), 256, 0, stream>>>(a);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// region masks: Voronoi cells + Gaussian feather (region_blend.py:69-102, 174-236)
// ---------------------------------------------------------------------------------------------
// label(y, x) = argmin_i sqrt((x - px_i)^2 + (y - py_i)^2) / (sqrt(weight_i) + 1e-6)   (first minimum wins, like torch.argmin);
// mask_i = feather(label == i): reflect-padded Gaussian blur, sigma = feather / 3, kernel int(6 sigma + 1) made odd.  The reference
// convolves with the 2-D outer-product kernel; the separable form here sums the same products in another order (|diff| ~ 1e-6).
constexpr int VOR_MAX_R = 16, BLUR_MAX_K = 255;
struct VoronoiArgs {
  float px[VOR_MAX_R], py[VOR_MAX_R], inv_w[VOR_MAX_R];   // inv_w = 1 / (sqrt(weight) + 1e-6), or 0 = unweighted
  int R, H, W, weighted;
  uint8_t* label;           // [H,W]
};
__global__ void __launch_bounds__(256) voronoi_label_kernel(const __grid_constant__ VoronoiArgs a) {
  const int hw = a.H * a.W;
  for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < hw; pix += gridDim.x * blockDim.x) {
    const int y = pix / a.W, x = pix - y * a.W;
    float best = INFINITY; int bi = 0;
    for (int i = 0; i < a.R; ++i) {
      const float dx = __fsub_rn((float)x, a.px[i]), dy = __fsub_rn((float)y, a.py[i]);
      float d = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
      if (a.weighted) d = __fdiv_rn(d, a.inv_w[i]);          // inv_w holds sqrt(weight) + 1e-6 in the weighted case
      if (d < best) { best = d; bi = i; }
    }
    a.label[pix] = (uint8_t)bi;
  }
}
struct BlurArgs {
  const uint8_t* label;     // horizontal pass input (label == r is the hard mask), or null: the horizontal pass reads src (soft masks)
  const float* src;         // vertical pass input [R,H,W]
  float* dst;               // [R,H,W]
  int R, H, W, k, vertical;
  float w[BLUR_MAX_K];
};
__device__ __forceinline__ int reflect_idx(int i, int n) {        // F.pad(mode='reflect'): -1 -> 1, n -> n - 2
  if (i < 0) i = -i;
  if (i >= n) i = 2 * (n - 1) - i;
  return i;
}
// Horizontal pass: one block = 256 consecutive pixels of one row of one region; the segment and its 2 * pad reflected neighbours are staged in
// shared memory once, then every thread runs the k taps from there (LDS + FMA per tap instead of index arithmetic + a global load).
// Vertical pass: one thread per pixel; rows at least `pad` away from the top / bottom edge step a pointer instead of reflecting an index.
// Both keep the tap order t = 0 .. k-1 of the original loops: bit-identical results.
__global__ void __launch_bounds__(256) blur_pass_kernel(const __grid_constant__ BlurArgs a) {
  const size_t hw = (size_t)a.H * a.W;
  const int pad = a.k / 2;
  if (!a.vertical) {
    __shared__ float sm[256 + BLUR_MAX_K];
    const int r = blockIdx.z, y = blockIdx.y, x0 = blockIdx.x * 256;
    for (int i = threadIdx.x; i < 256 + 2 * pad; i += 256) {
      int xx = x0 + i - pad;
      if (xx >= a.W + pad) { sm[i] = 0.f; continue; }                    // beyond what any live thread of the last segment reads
      xx = reflect_idx(xx, a.W);
      sm[i] = a.label ? (a.label[(size_t)y * a.W + xx] == r ? 1.f : 0.f) : a.src[(size_t)r * hw + (size_t)y * a.W + xx];
    }
    __syncthreads();
    const int x = x0 + threadIdx.x;
    if (x >= a.W) return;
    float acc = 0.f;
    for (int t = 0; t < a.k; ++t) acc = fmaf(a.w[t], sm[threadIdx.x + t], acc);
    a.dst[(size_t)r * hw + (size_t)y * a.W + x] = acc;
    return;
  }
  const size_t total = hw * a.R;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / hw);
    const int pix = (int)(i - (size_t)r * hw);
    const int y = pix / a.W, x = pix - y * a.W;
    const float* pl = a.src + (size_t)r * hw;
    float acc = 0.f;
    if (y >= pad && y + pad < a.H) {
      const float* p = pl + (size_t)(y - pad) * a.W + x;
      for (int t = 0; t < a.k; ++t, p += a.W) acc = fmaf(a.w[t], *p, acc);
    } else {
      for (int t = 0; t < a.k; ++t) acc = fmaf(a.w[t], pl[(size_t)reflect_idx(y + t - pad, a.H) * a.W + x], acc);
    }
    a.dst[i] = acc;
  }
}
cudaError_t launch_voronoi_label(const VoronoiArgs& a, cudaStream_t stream) {
  voronoi_label_kernel<<<std::min((a.H * a.W + 255) / 256, 148 * 16), 256, 0, stream>>>(a);
  return cudaGetLastError();
}
cudaError_t launch_blur_pass(const BlurArgs& a, cudaStream_t stream) {
  if (!a.vertical) {
    blur_pass_kernel<<<dim3((a.W + 255) / 256, a.H, a.R), 256, 0, stream>>>(a);
    return cudaGetLastError();
  }
  const size_t total = (size_t)a.R * a.H * a.W;
  blur_pass_kernel<<<(int)std::min<size_t>((total + 255) / 256, (size_t)148 * 16), 256, 0, stream>>>(a);
  return cudaGetLastError();
}

}  // namespace nst
