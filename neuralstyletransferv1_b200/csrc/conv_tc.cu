// tcgen05 / TMEM implicit-GEMM convolution for sm_100a -- the hot kernel of the stylisation path.
//
// Replaces, per layer, the reference's  [F.interpolate] -> ReflectionPad2d / zero pad -> Conv2d
// (transformer_net.py:44-54,79-99; transformer_net_nst.py:12-59; model.py:5-15) plus the
// InstanceNorm/ReLU/residual ops around it (transformer_net.py:30-39,71-76; model.py:30-35,57-60).
//
// One persistent CTA per SM, 16 warps (x 128 registers = the whole register file), warp-specialised:
//   warp 0      weight loader : cp.async.bulk (TMA bulk copy, UBLKCP) of pre-packed fp16 weight
//                               tiles global -> shared, mbarrier complete_tx ring
//   warp 1      MMA issuer    : one thread issues tcgen05.mma (M=128, N=Cout, K=16) for every
//                               (channel chunk, filter tap, K-step, precision term, M block);
//                               accumulators in TMEM (double buffered across tiles)
//   warps 2-5   epilogue      : tcgen05.ld TMEM -> registers, store raw fp32 NHWC output,
//                               per-(n,c) sum / sum-of-squares by warp-shuffle butterfly,
//                               or bias/tanh/affine/clamp/blend/uint8 pack for the image layer
//   warps 6-15  operand producers (8 or 10 take part, TcPlan::prod_warps) : gather the input halo tile from HBM with reflection /
//                               zero padding, nearest-upsample, zero-insertion and stride-2 parity
//                               split folded into the index math; apply the previous layer's
//                               InstanceNorm scale/shift + ReLU + residual add in registers;
//                               convert to fp16 (hi [+ lo]) and write UMMA core-matrix layout
//
// Shared-memory operand layout (no swizzle, K-major): plane[g][pos] = 8 channels (16 B) of halo
// position `pos`; 8 consecutive positions form one 128-byte core matrix (SBO = 128 B), the two
// 8-channel halves of a K=16 step are LBO = plane stride apart.  Because the layout is linear in
// `pos`, the A operand of filter tap (r,s) is the same buffer with the descriptor start address
// advanced by (r*P + s) * 16 bytes -- the halo tile is loaded and transformed once per tile and
// reused by all k*k taps.
#include "conv_tc.cuh"

namespace nst {

// ---------------------------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* ptr) {
  return (uint32_t)__cvta_generic_to_shared(ptr);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ uint32_t mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok;
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// Bounded wait: a pipeline bug must never hang the GPU.  After ~4 s the CTA records a code in the
// (host-mapped) error word and traps; the host turns that into NST_ERR_CUDA.
__device__ __noinline__ void mbar_wait_slow(uint32_t bar, uint32_t parity, unsigned int* err, unsigned int code) {
  const unsigned long long t0 = global_ns();
  unsigned int spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (((++spins) & 0xFFu) == 0u && global_ns() - t0 > 4000000000ull) {
      if (err != nullptr) {
        atomicExch(err, 0x80000000u | (code << 16) | (blockIdx.x & 0xFFFFu));
        __threadfence_system();
      }
      __trap();
    }
  }
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, unsigned int* err, unsigned int code) {
  if (mbar_try_wait(bar, parity)) return;
  mbar_wait_slow(bar, parity, err, code);
}
// Debug instrumentation (NST_TC_PROF=1): cycles each role spends blocked on each barrier kind.
struct WaitProf {
  unsigned long long* buf;
  long long acc[8];
  long long t_start;
  __device__ __forceinline__ void init(unsigned long long* b) {
    buf = b;
    if (buf) { for (int i = 0; i < 8; ++i) acc[i] = 0; t_start = clock64(); }
  }
  __device__ __forceinline__ void flush(int role, bool leader) {
    if (buf && leader) {
      for (int i = 1; i < 8; ++i) if (acc[i]) atomicAdd(buf + i, (unsigned long long)acc[i]);
      if (acc[0]) atomicAdd(buf + 13, (unsigned long long)acc[0]);
      atomicAdd(buf + 8 + role, (unsigned long long)(clock64() - t_start));
    }
  }
};
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, unsigned int* err, unsigned int code, WaitProf& wp) {
  if (wp.buf) {
    const long long t0 = clock64();
    mbar_wait(bar, parity, err, code);
    wp.acc[code] += clock64() - t0;
  } else {
    mbar_wait(bar, parity, err, code);
  }
}
// Barrier granularity (A/B-measured on one box, tools/abn.sh): one lane polling for its warp (+ __syncwarp) is 5-10 % SLOWER on every
// layer than all 32 lanes executing try_wait (the hardware parks the whole warp on the barrier), and one arrival per warp instead
// of one per thread is neutral -- so both default to the per-thread form.
#ifndef NST_WARP_POLL
#define NST_WARP_POLL 0
#endif
#ifndef NST_WARP_ARRIVE
#define NST_WARP_ARRIVE 0
#endif
constexpr uint32_t ARRIVALS_PER_WARP = NST_WARP_ARRIVE ? 1u : 32u;
__device__ __forceinline__ void mbar_wait_warp(uint32_t bar, uint32_t parity, unsigned int* err, unsigned int code, WaitProf& wp) {
#if NST_WARP_POLL
  if ((threadIdx.x & 31) == 0) mbar_wait(bar, parity, err, code, wp);
  __syncwarp();
#else
  mbar_wait(bar, parity, err, code, wp);
#endif
}
__device__ __forceinline__ void mbar_wait_warp(uint32_t bar, uint32_t parity, unsigned int* err, unsigned int code) {
#if NST_WARP_POLL
  if ((threadIdx.x & 31) == 0) mbar_wait(bar, parity, err, code);
  __syncwarp();
#else
  mbar_wait(bar, parity, err, code);
#endif
}
__device__ __forceinline__ void mbar_arrive_warp(uint32_t bar) {
#if NST_WARP_ARRIVE
  __syncwarp();
  if ((threadIdx.x & 31) == 0) mbar_arrive(bar);
#else
  mbar_arrive(bar);
#endif
}
__device__ __forceinline__ uint32_t elect_one_sync() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred;
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst_smem), "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
// TMA tiled load of one 4-D box {8 channels, box cols, box rows, 1 plane} of a planar-8 activation tensor
// (tensor map dims: 8 x W x H x (N * C/8)); out-of-bounds elements arrive as zeros and count towards complete_tx.
__device__ __forceinline__ void tma_load_4d(uint32_t dst_smem, const CUtensorMap* tmap, int c1, int c2, int c3, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
               ::"r"(dst_smem), "l"(tmap), "r"(0), "r"(c1), "r"(c2), "r"(c3), "r"(bar)
               : "memory");
}
// same box through a "wide" map: 3-D {2*W 8-byte elements, H, planes}; c1 is the first slot column
__device__ __forceinline__ void tma_load_3d_wide(uint32_t dst_smem, const CUtensorMap* tmap, int c1, int c2, int c3, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
               ::"r"(dst_smem), "l"(tmap), "r"(2 * c1), "r"(c2), "r"(c3), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void tma_load_act(bool wide, uint32_t dst_smem, const CUtensorMap* tmap, int c1, int c2, int c3, uint32_t bar) {
  if (wide) tma_load_3d_wide(dst_smem, tmap, c1, c2, c3, bar); else tma_load_4d(dst_smem, tmap, c1, c2, c3, bar);
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]; kind::f16 covers fp16 and bf16 operands with fp32 accumulation.
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// 16 consecutive accumulator columns, no wait: the caller overlaps the load with work on the previous fragment (tmem_ld_wait before use)
__device__ __forceinline__ void tmem_ld16_nowait(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
// 32 consecutive accumulator columns of this thread's TMEM lane; the caller waits (tmem_ld_wait) before using them
__device__ __forceinline__ void tmem_ld32_nowait(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// UMMA shared-memory matrix descriptor, SWIZZLE_NONE, K-major (cute::UMMA::SmemDescriptor):
// [0,14) start>>4, [16,30) LBO>>4, [32,46) SBO>>4, [46,48) version=1, [61,64) layout=0.
__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes, int swap = 0) {
  if (swap) { const uint32_t x = lbo_bytes; lbo_bytes = sbo_bytes; sbo_bytes = x; }
  return (uint64_t)((addr >> 4) & 0x3FFFu) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): c_format F32, a/b format, K-major both, N>>3, M>>4.
__device__ __forceinline__ uint32_t make_idesc(int npad, int is_bf16) {
  const uint32_t fmt = is_bf16 ? 1u : 0u;
  return (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(npad >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}

__device__ __forceinline__ uint32_t pack2(float a, float b, int is_bf16) {
  uint32_t r;
  if (is_bf16) {
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  } else {
    // saturate to +-65504 so an out-of-range activation can never become inf (inf * 0 = NaN in the MMA)
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  }
  return r;
}
// same with ReLU folded into the conversion (max commutes with round-to-nearest, 0 is exact)
__device__ __forceinline__ uint32_t pack2_relu(float a, float b, int is_bf16) {
  uint32_t r;
  if (is_bf16) asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  else asm("cvt.rn.relu.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  return r;
}
__device__ __forceinline__ float2 unpack2(uint32_t u, int is_bf16) {
  if (is_bf16) return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u));
  return __half22float2(*reinterpret_cast<__half2*>(&u));
}

struct TileCoord { int n, ty0, tx0; };
// `tile` is the CTA's position in the launch-wide sequence: CTA b visits b, b + G, b + 2G, ... (G = grid size).  Inside every FULL round
// of G tiles the positions are rotated by round * rr_rot, so a CTA does not meet the same tile row / column every round when G and
// the tile grid resonate (NST trunk: 37 tile rows, G = 148 = 4 * 37 -- the CTAs that owned the border rows finished 37 % late).
__device__ __forceinline__ TileCoord decode_tile(const TcPlan& t, int tile) {
  if (t.rr_rot) {
    const int G = (int)gridDim.x, round = tile / G;
    if ((round + 1) * G <= t.num_tiles) tile = round * G + (tile - round * G + round * t.rr_rot) % G;
  }
  const int per_img = t.tiles_x * t.tiles_y;
  TileCoord c;
  c.n = tile / per_img;
  const int r = tile - c.n * per_img;
  const int txi = r / t.tiles_y;                     // column-major: the next tile is the one below
  c.ty0 = (r - txi * t.tiles_y) * t.TH;
  c.tx0 = txi * t.TW;
  return c;
}

// ---------------------------------------------------------------------------------------------
// operand producer: one (halo position, 8-channel group) item
// ---------------------------------------------------------------------------------------------
// Source-table entry of one halo slot: -1 = structural zero / filler, else the in-frame source pixel index
// sy*Ws+sx (< 2^30), with bit 30 set when this tile owns the materialised ("side") value.
constexpr int SIDE_BIT = 1 << 30;

__device__ __forceinline__ int locate(const ConvParams& p, const TcPlan& t, const TileCoord& tc, int q) {
  int hy, hx;
  bool filler = false;
  // divisions by the (runtime) pitch as float multiplies: exact for q < 2^20 (error 2^-24 * q / P << 0.5 / P)
  const float inv_p = 1.0f / (float)t.P;
  if (p.stride == 1) {
    hy = (int)(((float)q + 0.5f) * inv_p);
    hx = q - hy * t.P;
    filler = hy >= t.halo_h;
  } else {
    const int sub = (int)(((float)q + 0.5f) * (1.0f / (float)t.SUBPOS));
    const int rem = q - sub * t.SUBPOS;
    const int i = (int)(((float)rem + 0.5f) * inv_p);
    const int j = rem - i * t.P;
    hy = 2 * i + (sub >> 1);
    hx = 2 * j + (sub & 1);
    filler = sub >= 4 || hy >= t.halo_h || hx >= t.halo_w;
  }
  if (filler) return -1;
  const int uy = (tc.ty0 + p.oy_off) * p.stride + hy - p.pad_t;
  const int ux = (tc.tx0 + p.ox_off) * p.stride + hx - p.pad_l;
  const int sy = map_coord(uy, p.Hc, p.Hs, p.map_mode, p.prepad, p.pad_mode);
  const int sx = map_coord(ux, p.Wc, p.Ws, p.map_mode, p.prepad, p.pad_mode);
  if (sy < 0 || sx < 0) return -1;
  int e = sy * p.Ws + sx;
  if (p.side_out != nullptr) {
    // interior positions of this tile own the materialised value (identity-mapped stride-1 layers)
    const int iy = hy - p.pad_t, ix = hx - p.pad_l;
    if (iy >= 0 && iy < t.TH && ix >= 0 && ix < t.TW && tc.ty0 + iy < p.Ho && tc.tx0 + ix < p.Wo) e |= SIDE_BIT;
  }
  return e;
}

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---------------------------------------------------------------------------------------------
// MMA issue: one weight stage (`n_ent` taps) against the current operand stage.  Descriptor low words
// are (addr >> 4) | (LBO >> 4) << 16; every step is a 32-bit add in 16-byte units.  K16 / MB are
// compile-time so the 2*K16*MB descriptors of a tap are straight-line uniform-register code.
// ---------------------------------------------------------------------------------------------
struct IssueArgs {
  uint32_t tmem_acc, a_lo, b_lo, b_ent16, a_kstep16, b_kstep16, a_term16, b_term16, npad, idesc;
  const uint32_t* ent_tab;
  int n_ent;
  uint32_t acc_on;
};
constexpr uint32_t DESC_HI = (128u >> 4) | (1u << 14);     // SBO = 128 B, descriptor version 1, SWIZZLE_NONE

// NTERMS = 3: error-compensated operands (hi*hi, lo*hi, hi*lo).  SWAP: weights are the A operand (M = 128 output channels), the halo
// tile is the B operand (N = MB*128 positions) -- one MMA per (tap, K step) instead of MB.
template <int K16, int MB, int NTERMS, bool SWAP>
__device__ __forceinline__ void issue_group(const IssueArgs& a) {
  uint32_t acc = a.acc_on;
  uint32_t b_lo = a.b_lo;
#pragma unroll 2
  for (int e = 0; e < a.n_ent; ++e, b_lo += a.b_ent16) {
    const uint32_t a_lo = a.a_lo + a.ent_tab[e];
#pragma unroll
    for (int j = 0; j < K16; ++j) {
#pragma unroll
      for (int term = 0; term < NTERMS; ++term) {
        const uint64_t bdesc = ((uint64_t)DESC_HI << 32) | (b_lo + (uint32_t)j * a.b_kstep16 + (term == 2 ? a.b_term16 : 0u));
        const uint32_t a_j = a_lo + (uint32_t)j * a.a_kstep16 + (term == 1 ? a.a_term16 : 0u);
        if (SWAP) {
          umma_f16(a.tmem_acc, bdesc, ((uint64_t)DESC_HI << 32) | a_j, a.idesc, acc);
        } else {
#pragma unroll
          for (int mb = 0; mb < MB; ++mb)
            umma_f16(a.tmem_acc + (uint32_t)mb * a.npad, ((uint64_t)DESC_HI << 32) | (a_j + (uint32_t)mb * 128u), bdesc, a.idesc, acc);
        }
        acc = 1u;
      }
    }
  }
}

__device__ __noinline__ void issue_group_generic(const IssueArgs& a, int k16, int MB, int nterms) {
  uint32_t acc = a.acc_on;
  uint32_t b_lo = a.b_lo;
  for (int e = 0; e < a.n_ent; ++e, b_lo += a.b_ent16) {
    const uint32_t a_lo = a.a_lo + a.ent_tab[e];
    for (int j = 0; j < k16; ++j) {
      for (int term = 0; term < nterms; ++term) {                      // hi*hi, lo*hi, hi*lo
        const uint64_t bdesc = ((uint64_t)DESC_HI << 32) | (b_lo + (uint32_t)j * a.b_kstep16 + (term == 2 ? a.b_term16 : 0u));
        const uint32_t a_j = a_lo + (uint32_t)j * a.a_kstep16 + (term == 1 ? a.a_term16 : 0u);
        for (int mb = 0; mb < MB; ++mb)
          umma_f16(a.tmem_acc + (uint32_t)mb * a.npad, ((uint64_t)DESC_HI << 32) | (a_j + (uint32_t)mb * 128u), bdesc, a.idesc, acc);
        acc = 1u;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// pipelined operand producer for fp16 NHWC activations (the default path of every layer but the first)
// ---------------------------------------------------------------------------------------------
// Each producer thread owns a fixed 8-channel plane g and the halo slots ptid/planes + k*(threads/planes);
// its work is a flat sequence of batches (tile, channel chunk, batch-in-chunk) of PU slots.  The
// global loads of batch i+1 are issued BEFORE batch i is transformed and written to shared memory,
// across chunk and tile boundaries, so ~PU*16..32 B per thread are always in flight (the first version
// waited for every batch: ncu showed the producers stalled on long_scoreboard ~45 % of the time).
// PU = halo slots per batch per thread: 8 without a residual operand, 4 with (register budget)

template <bool HAS_RESID, int PU>
struct ProdBatch {
  int ent[PU];
  uint4 a[PU];
  uint4 r[HAS_RESID ? PU : 1];
};

struct ProdCursor {
  int tile, ch, b;
  uint32_t tb;         // which of the two source tables belongs to `tile`
};

template <bool HAS_RESID, int PU>
__device__ __forceinline__ void prod_load(ProdBatch<HAS_RESID, PU>& B, const ProdCursor& c, bool valid, const ConvParams& p, const TcPlan& t,
                                          const int* srcpix, int tbl_stride, int ptid, int pshift, int g, int nitems, int PT) {
  const __half* __restrict__ src = (const __half*)p.src;
  const __half* __restrict__ res = (const __half*)p.resid;
  const int* tbl = srcpix + c.tb * tbl_stride;
  // planar-8: slot (n, 8-channel plane, pixel) sits at ((n * Cin/8 + plane) * Hs*Ws + pixel) * 8 elements
  const size_t plane_base = ((size_t)(c.tile / (t.tiles_x * t.tiles_y)) * (p.Cin >> 3) + (size_t)(c.ch * t.planes + g)) * ((size_t)p.Hs * p.Ws);
#pragma unroll
  for (int u = 0; u < PU; ++u) {
    const int idx = ptid + (c.b * PU + u) * PT;
    B.ent[u] = (valid && idx < nitems && !(t.dbg & 2)) ? tbl[idx >> pshift] : -2;
  }
#pragma unroll
  for (int u = 0; u < PU; ++u) {
    B.a[u] = make_uint4(0u, 0u, 0u, 0u);
    if (HAS_RESID) B.r[u] = make_uint4(0u, 0u, 0u, 0u);
    if (B.ent[u] >= 0) {
      const size_t off = (plane_base + (size_t)(B.ent[u] & (SIDE_BIT - 1))) * 8;
      if (!(t.dbg & 1024)) {
      B.a[u] = __ldg(reinterpret_cast<const uint4*>(src + off));
      if (HAS_RESID) B.r[u] = __ldg(reinterpret_cast<const uint4*>(res + off));
      }
    }
  }
}

// MODE: 0 = ReLU positions taken from the layer at run time, 1 = no ReLU anywhere, 2 = only a final ReLU (folded into the pack)
// BF16 / SIDE: operand type and presence of a side output as compile-time facts (a run-time is_bf16 costs a branch per pack)
template <bool HAS_RESID, int PU, int MODE, bool BF16, bool SIDE>
__device__ void produce_act16(const ConvParams& p, const TcPlan& t, uint8_t* smem_a, int* srcpix, float2* coeff, uint32_t bar_a_full0,
                              uint32_t bar_a_empty0, unsigned int* err, unsigned long long* prof, int first_tile, int tile_step, int end_tile, int ptid) {
  const int PT = t.prod_warps * 32;          // producer threads of this layer (8 or 10 warps, chosen per layer on the host)
  const int planes = t.planes;
  const int pshift = planes == 4 ? 2 : (planes == 2 ? 1 : 0);
  const int g = ptid & (planes - 1);
  const int nitems = t.NPOS * planes;
  const int tbl_stride = (t.NPOS + 3) & ~3;
  const int nB = (nitems + PU * PT - 1) / (PU * PT);     // batches per chunk
  constexpr int is_bf16 = BF16 ? 1 : 0;
  const float lo1 = (p.pre_relu || (!HAS_RESID && p.post_relu)) ? 0.f : -INFINITY;   // without a residual add the two ReLUs coincide
  const float lo2 = p.post_relu ? 0.f : -INFINITY;
  __half* side = (__half*)p.side_out;

  auto build_table = [&](int tile, uint32_t tb) {
    if (tile >= end_tile) return;
    const TileCoord tc = decode_tile(t, tile);
    int* tbl = srcpix + tb * tbl_stride;
    for (int q = ptid; q < t.NPOS; q += PT) tbl[q] = locate(p, t, tc, q);
  };
  auto advance = [&](ProdCursor& c) {
    if (++c.b == nB) {
      c.b = 0;
      if (++c.ch == t.nchunks) { c.ch = 0; c.tile += tile_step; c.tb ^= 1u; }
    }
  };

  uint32_t sa = 0, pa = 0;
  int cur_n = -1;
  float sc[8], sh[8];
  uint8_t* plane_hi = smem_a;
  WaitProf wprof; wprof.init(prof);

  // tile start: everyone is done with the table of the previous tile -> build the next tile's (it is
  // first read by the prefetch that follows), refresh the per-sample norm coefficients
  auto tile_begin = [&](const ProdCursor& c) {
    if (c.b != 0 || c.ch != 0) return;
    named_bar_sync(1, PT);
    const int n = c.tile / (t.tiles_x * t.tiles_y);
    if (n != cur_n) {
      for (int ch = ptid; ch < p.Cin; ch += PT) {
        float s_, h_;
        norm_coeff(p, n, ch, s_, h_);
        coeff[ch] = make_float2(s_, h_);
      }
      cur_n = n;
    }
    build_table(c.tile + tile_step, c.tb ^ 1u);
    named_bar_sync(2, PT);
  };
  auto process = [&](const ProdBatch<HAS_RESID, PU>& B, const ProdCursor& c) {
    if (c.b == 0) {
      const int c0 = c.ch * t.CK + g * 8;
#pragma unroll
      for (int i = 0; i < 8; ++i) { const float2 cf = coeff[c0 + i]; sc[i] = cf.x; sh[i] = cf.y; }
      mbar_wait_warp(bar_a_empty0 + 8u * sa, pa ^ 1u, err, 6, wprof);
      plane_hi = smem_a + (size_t)sa * t.a_stage_bytes + (size_t)g * t.a_plane_bytes;
    }
    const size_t plane_base = ((size_t)(c.tile / (t.tiles_x * t.tiles_y)) * (p.Cin >> 3) + (size_t)(c.ch * t.planes + g)) * ((size_t)p.Hs * p.Ws);
    // straight-line over the PU slots (no per-slot branches, so the PU dependency chains interleave): slots with nothing to do
    // (e == -2) or structural zeros (e == -1) run on zero data and are masked at the store
    uint32_t hi[PU][4];
#pragma unroll
    for (int u = 0; u < PU; ++u) {
      const uint32_t aw[4] = {B.a[u].x, B.a[u].y, B.a[u].z, B.a[u].w};
      uint32_t rw[4] = {0u, 0u, 0u, 0u};
      if (HAS_RESID) { rw[0] = B.r[u].x; rw[1] = B.r[u].y; rw[2] = B.r[u].z; rw[3] = B.r[u].w; }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 av = unpack2(aw[i], 0);
        float v0 = fmaf(av.x, sc[2 * i], sh[2 * i]);
        float v1 = fmaf(av.y, sc[2 * i + 1], sh[2 * i + 1]);
        if (MODE == 0) { v0 = fmaxf(v0, lo1); v1 = fmaxf(v1, lo1); }
        if (HAS_RESID) {
          const float2 rv = unpack2(rw[i], 0);
          v0 += rv.x; v1 += rv.y;
          if (MODE == 0) { v0 = fmaxf(v0, lo2); v1 = fmaxf(v1, lo2); }
        }
        hi[u][i] = B.ent[u] >= 0 ? (MODE == 2 ? pack2_relu(v0, v1, is_bf16) : pack2(v0, v1, is_bf16)) : 0u;
      }
    }
#pragma unroll
    for (int u = 0; u < PU; ++u) {
      const int e = B.ent[u];
      const int q = (ptid + (c.b * PU + u) * PT) >> pshift;
      if (SIDE) {
        if (e >= 0 && (e & SIDE_BIT)) {
          // the materialised block input is stored as fp16 -- the same rounding the MMA operand gets
          uint32_t sv[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) { if (is_bf16) { const float2 f = unpack2(hi[u][i], 1); sv[i] = pack2(f.x, f.y, 0); } else sv[i] = hi[u][i]; }
          *reinterpret_cast<uint4*>(side + (plane_base + (size_t)(e & (SIDE_BIT - 1))) * 8) = make_uint4(sv[0], sv[1], sv[2], sv[3]);
        }
      }
      if (e != -2) *reinterpret_cast<uint4*>(plane_hi + (size_t)q * 16) = make_uint4(hi[u][0], hi[u][1], hi[u][2], hi[u][3]);
    }
    if (c.b == nB - 1) {
      fence_proxy_async_smem();
      mbar_arrive_warp(bar_a_full0 + 8u * sa);
      if (++sa == (uint32_t)t.a_stages) { sa = 0; pa ^= 1u; }
    }
  };

  ProdCursor cur;
  cur.tile = first_tile; cur.ch = 0; cur.b = 0; cur.tb = 0u;
  build_table(first_tile, 0u);
  named_bar_sync(2, PT);
  ProdBatch<HAS_RESID, PU> B0, B1;
  prod_load<HAS_RESID, PU>(B0, cur, cur.tile < end_tile, p, t, srcpix, tbl_stride, ptid, pshift, g, nitems, PT);
  while (cur.tile < end_tile) {
    ProdCursor nxt = cur;
    advance(nxt);
    tile_begin(cur);
    prod_load<HAS_RESID, PU>(B1, nxt, nxt.tile < end_tile, p, t, srcpix, tbl_stride, ptid, pshift, g, nitems, PT);
    process(B0, cur);
    cur = nxt;
    if (cur.tile >= end_tile) break;
    advance(nxt);
    tile_begin(cur);
    prod_load<HAS_RESID, PU>(B0, nxt, nxt.tile < end_tile, p, t, srcpix, tbl_stride, ptid, pshift, g, nitems, PT);
    process(B1, cur);
    cur = nxt;
  }
  wprof.flush(3, ptid == 0);
}

// ---------------------------------------------------------------------------------------------
// TMA-fed operand path (TcPlan::tma_mode): fp16 planar-8 activations, one MMA term
// ---------------------------------------------------------------------------------------------
// The halo tile of a channel chunk reaches shared memory as raw fp16 through cp.async.bulk.tensor -- one box
// {8 channels, P columns, halo rows} per 8-channel plane, which IS the UMMA K-major core-matrix layout
// [plane][position][8 x f16] because the planar-8 tensor stores a plane's slots contiguously along W.  Stride-2 layers load
// the four (row, column) parity sub-planes through four parity views of the source (tensor maps with doubled strides), so
// filter taps stay unit-stride in the position index.  Out-of-image elements arrive as zeros.  The transform warps then apply
// the previous layer's InstanceNorm scale/shift + ReLU (+ residual from a second staged tile) in place, shared -> shared, with
// no index tables and no global gathers; tiles that touch the image border get a second short pass that mirrors (reflection
// / replicate padding) or zeroes (zero padding) the out-of-image slots from the already transformed in-image ones.

// conv-space coordinate of halo slot (0, 0) of a tile
__device__ __forceinline__ void tile_origin(const ConvParams& p, const TileCoord& tc, int& u0, int& v0) {
  u0 = (tc.ty0 + p.oy_off) * p.stride - p.pad_t;
  v0 = (tc.tx0 + p.ox_off) * p.stride - p.pad_l;
}

// Loader: one thread walks (tile, channel chunk) in MMA order and keeps the operand ring full.
__device__ void tma_issue_role(const TcArgs& args, uint8_t* smem_a, uint32_t bar_tma_full0, uint32_t bar_a_empty0, int first_tile, int tile_step, int end_tile) {
  const ConvParams& p = args.p;
  const TcPlan& t = args.t;
  WaitProf wprof; wprof.init(args.prof);
  const int nmaps = p.stride == 2 ? 4 : 1;
  for (int i = 0; i < nmaps; ++i) tma_prefetch_desc(&args.tmap[i]);
  if (p.resid != nullptr) tma_prefetch_desc(&args.tmap_res);
  const uint32_t box_bytes = (uint32_t)t.tma_rows * (uint32_t)t.P * 16u;
  const uint32_t stage_tx = box_bytes * (uint32_t)t.planes * (p.stride == 2 ? 4u : 1u) * (p.resid != nullptr ? 2u : 1u);
  const int c8 = p.Cin >> 3;
  const bool wide = t.tma_wide != 0;
  uint32_t sa = 0, pa = 0;
  for (int tile = first_tile; tile < end_tile; tile += tile_step) {
    const TileCoord tc = decode_tile(t, tile);
    int u0, v0;
    tile_origin(p, tc, u0, v0);
    for (int ch = 0; ch < t.nchunks; ++ch) {
      mbar_wait(bar_a_empty0 + 8u * sa, pa ^ 1u, args.err_flag, 6, wprof);
      const uint32_t full = bar_tma_full0 + 8u * sa;
      mbar_arrive_expect_tx(full, stage_tx);
      const uint32_t stage = smem_u32(smem_a) + sa * t.a_stage_bytes;
      const int plane0 = tc.n * c8 + ch * t.planes;
      for (int g = 0; g < t.planes; ++g) {
        const uint32_t dst = stage + (uint32_t)g * t.a_plane_bytes;
        if (p.stride == 1) {
          tma_load_act(wide, dst, &args.tmap[0], v0, u0, plane0 + g, full);
          if (p.resid != nullptr) tma_load_act(wide, dst + t.a_res_off, &args.tmap_res, v0, u0, plane0 + g, full);
        } else {
#pragma unroll
          for (int sub = 0; sub < 4; ++sub) {
            // halo slot (2i + py, 2j + px) is conv-space (u0 + py + 2i, v0 + px + 2j): row i / column j of the parity view
            // ((u0 + py) & 1, (v0 + px) & 1) starting at floor((u0 + py) / 2), floor((v0 + px) / 2)  (>> is arithmetic)
            const int uu = u0 + (sub >> 1), vv = v0 + (sub & 1);
            tma_load_act(wide, dst + (uint32_t)(sub * t.SUBPOS) * 16u, &args.tmap[((uu & 1) << 1) | (vv & 1)], vv >> 1, uu >> 1, plane0 + g, full);
          }
        }
      }
      if (++sa == (uint32_t)t.a_stages) { sa = 0; pa ^= 1u; }
    }
  }
  wprof.flush(4, true);
}

// halo slot index -> (row, column) inside the halo; false for the filler slots behind a stride-2 parity sub-plane
__device__ __forceinline__ bool slot_rc(const TcPlan& t, int stride, int q, float inv_p, float inv_sub, int& hy, int& hx) {
  if (stride == 1) {
    hy = (int)(((float)q + 0.5f) * inv_p);
    hx = q - hy * t.P;
    return true;
  }
  const int sub = (int)(((float)q + 0.5f) * inv_sub);
  const int rem = q - sub * t.SUBPOS;
  const int i = (int)(((float)rem + 0.5f) * inv_p);
  const int j = rem - i * t.P;
  hy = 2 * i + (sub >> 1);
  hx = 2 * j + (sub & 1);
  return i < t.tma_rows;
}
__device__ __forceinline__ int rc_slot(const TcPlan& t, int stride, int hy, int hx) {
  if (stride == 1) return hy * t.P + hx;
  return (((hy & 1) << 1) | (hx & 1)) * t.SUBPOS + (hy >> 1) * t.P + (hx >> 1);
}
// padding rule on one conv-space coordinate: the in-image coordinate whose value the out-of-image coordinate u takes, or -1
// (zero padding; and coordinates of a partial tile so far outside that no produced output reads them)
__device__ __forceinline__ int pad_donor(int u, int extent, int pad_mode) {
  if (u >= 0 && u < extent) return u;
  if (pad_mode == PAD_REFLECT) { const int r = u < 0 ? -u : 2 * (extent - 1) - u; return (r >= 0 && r < extent) ? r : -1; }
  if (pad_mode == PAD_REPLICATE) return u < 0 ? 0 : extent - 1;
  return -1;
}

// MODE / HAS_RESID / SIDE as in produce_act16
template <bool HAS_RESID, int MODE, bool SIDE>
__device__ void transform_act16(const TcArgs& args, uint8_t* smem_a, float2* coeff, uint32_t bar_tma_full0, uint32_t bar_a_full0, int first_tile, int tile_step, int end_tile, int ptid) {
  const ConvParams& p = args.p;
  const TcPlan& t = args.t;
  constexpr int PU = 4;
  const int PT = t.prod_warps * 32;
  const int planes = t.planes;
  const int g = (ptid >> 3) & (planes - 1);                 // 8 consecutive lanes sweep one 128-byte core matrix of one plane
  const int blk0 = ptid / (8 * planes), blk_step = PT / (8 * planes);
  const int l8 = ptid & 7;
  const int nblk = (t.npos_used + 7) >> 3;
  const float lo1 = (p.pre_relu || (!HAS_RESID && p.post_relu)) ? 0.f : -INFINITY;
  const float lo2 = p.post_relu ? 0.f : -INFINITY;
  const float inv_p = 1.0f / (float)t.P, inv_sub = p.stride == 2 ? 1.0f / (float)t.SUBPOS : 0.f;
  const int halo_w = p.stride == 1 ? t.P : t.halo_w;
  const int per_img = t.tiles_x * t.tiles_y;
  __half* side = (__half*)p.side_out;
  // loop invariants in registers (the plan sits behind a generic pointer; asm volatile "memory" clobbers would force reloads)
  const int kP = t.P, kTH = t.TH, kTW = t.TW, kHo = p.Ho, kWo = p.Wo, kWs = p.Ws, kPadT = p.pad_t, kPadL = p.pad_l;
  WaitProf wprof; wprof.init(args.prof);
  uint32_t sa = 0, pa = 0;
  int cur_n = -1;
  float sc[8], sh[8];
  // (The residual operand is staged by TMA next to the source tile.  Reading it from HBM in the transform instead -- registers, one loop
  // iteration ahead -- was 30 % slower: the ~2 us load latency under load needs a prefetch distance of whole stages, which only the
  // shared-memory ring provides.)
  const int nit = ((nblk + blk_step - 1) / blk_step + PU - 1) / PU;      // loop iterations per (tile, chunk), the same for every thread
  // The slots a thread visits are the same in every tile: their in-plane offset hy * Ws + hx and whether they are tile-interior
  // (side output) are computed once (the planner keeps nit <= RIT for layers with a residual operand or a side output).
  constexpr int RIT = 2;
  int soff[RIT][PU];
  uint32_t live_m = 0u, inner_m = 0u;
  if (SIDE) {
#pragma unroll
    for (int it = 0; it < RIT; ++it)
#pragma unroll
      for (int u = 0; u < PU; ++u) {
        const int blk = blk0 + (it * PU + u) * blk_step;
        const int q = blk * 8 + l8;
        const int hy = (int)(((float)q + 0.5f) * inv_p), hx = q - hy * kP;
        soff[it][u] = hy * kWs + hx;
        if (blk < nblk) live_m |= 1u << (it * PU + u);
        const int iy = hy - kPadT, ix = hx - kPadL;
        if ((iy >= 0) & (iy < kTH) & (ix >= 0) & (ix < kTW)) inner_m |= 1u << (it * PU + u);
      }
  }
  for (int tile = first_tile; tile < end_tile; tile += tile_step) {
    const TileCoord tc = decode_tile(t, tile);
    int u0, v0;
    tile_origin(p, tc, u0, v0);
    if (tc.n != cur_n) {                                      // per-sample InstanceNorm coefficients -> shared table
      named_bar_sync(1, PT);
      for (int ch = ptid; ch < p.Cin; ch += PT) {
        float s_, h_;
        norm_coeff(p, tc.n, ch, s_, h_);
        coeff[ch] = make_float2(s_, h_);
      }
      cur_n = tc.n;
      named_bar_sync(2, PT);
    }
    const bool border = u0 < 0 || v0 < 0 || u0 + t.halo_h > p.Hc || v0 + halo_w > p.Wc;
    for (int ch = 0; ch < t.nchunks; ++ch) {
      {
        const float4* cf4 = reinterpret_cast<const float4*>(coeff + ch * t.CK + g * 8);      // two (scale, shift) pairs per 16-byte load
#pragma unroll
        for (int i = 0; i < 4; ++i) { const float4 cf = cf4[i]; sc[2 * i] = cf.x; sh[2 * i] = cf.y; sc[2 * i + 1] = cf.z; sh[2 * i + 1] = cf.w; }
      }
      mbar_wait_warp(bar_tma_full0 + 8u * sa, pa, args.err_flag, 7, wprof);
      uint8_t* plane = smem_a + (size_t)sa * t.a_stage_bytes + (size_t)g * t.a_plane_bytes;
      const uint8_t* rplane = plane + t.a_res_off;
      // side output: slot (hy, hx) of plane (n, ch*planes + g) is source pixel (u0 + hy, v0 + hx) (identity-mapped stride-1 layers)
      const long long side_base = SIDE ? (((long long)tc.n * (p.Cin >> 3) + (ch * planes + g)) * p.Hs + u0) * p.Ws + v0 : 0;   // in slots; may be < 0
      const bool edge_tile = tc.ty0 + kTH > kHo || tc.tx0 + kTW > kWo;       // tile overhangs the image: interior slots need the bound check
      auto process_iter = [&](const int itx) {
        const int blk = blk0 + itx * PU * blk_step;
        uint4 a[PU], r[PU];
#pragma unroll
        for (int u = 0; u < PU; ++u) {
          const int q = (blk + u * blk_step) * 8 + l8;
          const bool live = blk + u * blk_step < nblk;
          a[u] = live ? *reinterpret_cast<const uint4*>(plane + (size_t)q * 16) : make_uint4(0u, 0u, 0u, 0u);
          if (HAS_RESID) r[u] = live ? *reinterpret_cast<const uint4*>(rplane + (size_t)q * 16) : make_uint4(0u, 0u, 0u, 0u);
        }
#pragma unroll
        for (int u = 0; u < PU; ++u) {
          const int q = (blk + u * blk_step) * 8 + l8;
          const uint32_t aw[4] = {a[u].x, a[u].y, a[u].z, a[u].w};
          uint32_t rw[4] = {0u, 0u, 0u, 0u};
          if (HAS_RESID) { rw[0] = r[u].x; rw[1] = r[u].y; rw[2] = r[u].z; rw[3] = r[u].w; }
          uint32_t hi[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float2 av = unpack2(aw[i], 0);
            float x0 = fmaf(av.x, sc[2 * i], sh[2 * i]);
            float x1 = fmaf(av.y, sc[2 * i + 1], sh[2 * i + 1]);
            if (MODE == 0) { x0 = fmaxf(x0, lo1); x1 = fmaxf(x1, lo1); }
            if (HAS_RESID) {
              const float2 rv = unpack2(rw[i], 0);
              x0 += rv.x; x1 += rv.y;
              if (MODE == 0) { x0 = fmaxf(x0, lo2); x1 = fmaxf(x1, lo2); }
            }
            hi[i] = MODE == 2 ? pack2_relu(x0, x1, 0) : pack2(x0, x1, 0);
          }
          if (blk + u * blk_step < nblk) {
            *reinterpret_cast<uint4*>(plane + (size_t)q * 16) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            if (SIDE) {
              bool own = ((inner_m >> ((itx % RIT) * PU + u)) & 1u) != 0u;
              if (edge_tile && own) {
                const int hy = (int)(((float)q + 0.5f) * inv_p);
                own = (tc.ty0 + hy - kPadT < kHo) & (tc.tx0 + (q - hy * kP) - kPadL < kWo);
              }
              if (own) *reinterpret_cast<uint4*>(side + (size_t)(side_base + soff[itx % RIT][u]) * 8) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            }
          }
        }
      };
      if constexpr (SIDE) {
#pragma unroll
        for (int itx = 0; itx < RIT; ++itx) { if (itx < nit) process_iter(itx); }
      } else {
        for (int itx = 0; itx < nit; ++itx) process_iter(itx);
      }
      if (border) {
        // Out-of-image slots form a frame of `top` / `bot` rows and `left` / `right` columns around the in-image rectangle: only those are
        // visited (one row or column of slots for a 3x3 layer), each taking its value from the mirrored / replicated in-image slot.
        named_bar_sync(1, PT);                 // every in-image slot of the stage is transformed
        const int top = min(max(-u0, 0), t.halo_h), bot = min(max(u0 + t.halo_h - p.Hc, 0), t.halo_h - top);
        const int left = min(max(-v0, 0), halo_w), right = min(max(v0 + halo_w - p.Wc, 0), halo_w - left);
        const int n_rows = (top + bot) * halo_w, lr = left + right, n_oob = n_rows + (t.halo_h - top - bot) * lr;
        for (int item = ptid; item < n_oob * planes; item += PT) {
          const int gi = item % planes, idx = item / planes;
          int hy, hx;
          if (idx < n_rows) { const int r = idx / halo_w; hx = idx - r * halo_w; hy = r < top ? r : t.halo_h - bot + (r - top); }
          else { const int j = idx - n_rows, r = j / lr, k = j - r * lr; hy = top + r; hx = k < left ? k : halo_w - right + (k - left); }
          uint8_t* pl = smem_a + (size_t)sa * t.a_stage_bytes + (size_t)gi * t.a_plane_bytes;
          uint4 val = make_uint4(0u, 0u, 0u, 0u);
          const int du = pad_donor(u0 + hy, p.Hc, p.pad_mode), dv = pad_donor(v0 + hx, p.Wc, p.pad_mode);
          if (du >= 0 && dv >= 0) {
            const int dy = du - u0, dx = dv - v0;
            if (dy < 0 || dy >= t.halo_h || dx < 0 || dx >= halo_w) continue;      // partial-tile overhang: nothing reads this slot
            val = *reinterpret_cast<const uint4*>(pl + (size_t)rc_slot(t, p.stride, dy, dx) * 16);
          } else if (p.pad_mode != PAD_ZERO) continue;
          *reinterpret_cast<uint4*>(pl + (size_t)rc_slot(t, p.stride, hy, hx) * 16) = val;
        }
      }
      fence_proxy_async_smem();       // generic-proxy smem writes -> visible to tcgen05.mma (async proxy)
      mbar_arrive_warp(bar_a_full0 + 8u * sa);
      if (++sa == (uint32_t)t.a_stages) { sa = 0; pa ^= 1u; }
    }
  }
  wprof.flush(3, ptid == 0);
}

// ---------------------------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------------------------

// ---------------------------------------------------------------------------------------------
// MMA issuer role.  The whole warp walks the pipeline converged and one elected lane issues (the
// CUTLASS idiom): ptxas then keeps every descriptor in uniform registers.  The role is a template
// on (K16, MB) dispatched ONCE per kernel: measured with the other roles switched off, the same
// loop with a per-weight-stage runtime dispatch (jump table + descriptor set-up from constant
// memory) cost ~106 cycles per N=128 MMA against 70 for tools/umma_pattern.cu -- the tensor pipe's
// issue queue is short, so every bubble at a stage boundary is exposed.
// ---------------------------------------------------------------------------------------------
struct MmaCtx {
  uint32_t smem_a, smem_b, bar0, tmem_base;
  const uint32_t* ent_tab;
  unsigned int* err;
  unsigned long long* prof;
  int first_tile, tile_step, end_tile;
};

template <int K16, int MB_, int NTERMS = 1, bool SWAP = false>
__device__ __noinline__ void mma_role(const TcPlan& t, const MmaCtx c) {
  const uint32_t idesc = make_idesc(SWAP ? t.MB * 128 : t.NPAD, t.is_bf16);
  const uint32_t a_lbo16 = (t.pair_mode ? 16u * (uint32_t)t.pair_mode : t.a_plane_bytes) >> 4;   // pair: next slot, quad: two slots on
  const uint32_t a_lo_base = ((c.smem_a >> 4) & 0x3FFFu) | (a_lbo16 << 16);
  const uint32_t b_lo_base = ((c.smem_b >> 4) & 0x3FFFu) | ((uint32_t)t.NPAD << 16);   // LBO = NPAD*16 B
  const uint32_t a_kstep16 = (2u * t.a_plane_bytes) >> 4;
  const uint32_t b_kstep16 = 2u * (uint32_t)t.NPAD;
  const uint32_t a_stage16 = t.a_stage_bytes >> 4, b_stage16 = t.b_stage_bytes >> 4;
  const uint32_t a_term16 = t.a_term_bytes >> 4, b_term16 = t.b_term_bytes >> 4, b_ent16 = t.b_ent_bytes >> 4;
  const int MB = MB_ ? MB_ : t.MB, NPAD = t.NPAD, k16 = t.k16_per_ent, nterms = t.nterms, tg = t.tg, nent = t.nent;
  const int nchunks = t.nchunks, ngroups = t.ngroups, num_tiles = c.end_tile, dbg = t.dbg;
  const bool resident = t.b_resident != 0;
  const uint32_t a_stages = (uint32_t)t.a_stages, b_stages = (uint32_t)t.b_stages;
  const uint32_t A_FULL0 = c.bar0, A_EMPTY0 = c.bar0 + 8u * TC_MAX_A_STAGES, B_FULL0 = c.bar0 + 8u * (2 * TC_MAX_A_STAGES),
                 B_EMPTY0 = B_FULL0 + 8u * TC_MAX_B_STAGES, ACC_FULL0 = B_FULL0 + 8u * (2 * TC_MAX_B_STAGES), ACC_EMPTY0 = ACC_FULL0 + 16u;
  WaitProf wprof; wprof.init(c.prof);
  uint32_t sa = 0, pa = 0, sb = 0, pb = 0, it = 0;
  for (int tile = c.first_tile; tile < num_tiles; tile += c.tile_step, ++it) {
    const uint32_t buf = it & 1u;
    if (!(dbg & 32)) mbar_wait_warp(ACC_EMPTY0 + 8u * buf, ((it >> 1) & 1u) ^ 1u, c.err, 2, wprof);
    tc_fence_after();
    const uint32_t tmem_acc = c.tmem_base + buf * (uint32_t)(MB * NPAD);
    uint32_t acc_on = 0u;
    for (int ch = 0; ch < nchunks; ++ch) {
      if (!(dbg & 16)) mbar_wait_warp(A_FULL0 + 8u * sa, pa, c.err, 3, wprof);
      tc_fence_after();
      const uint32_t a_stage_lo = a_lo_base + sa * a_stage16;
      int ent = 0;
      for (int g = 0; g < ngroups; ++g) {
        if (!(dbg & 64)) mbar_wait_warp(B_FULL0 + 8u * sb, resident ? 0u : pb, c.err, 4, wprof);
        tc_fence_after();
        if (elect_one_sync()) {
          const IssueArgs ia = {tmem_acc, a_stage_lo, b_lo_base + sb * b_stage16, b_ent16, a_kstep16, b_kstep16, a_term16, b_term16, (uint32_t)NPAD, idesc,
                                c.ent_tab + ent, min(tg, nent - ent), acc_on};
          if (K16 != 0) issue_group<(K16 ? K16 : 1), (MB_ ? MB_ : 1), NTERMS, SWAP>(ia);
          else issue_group_generic(ia, k16, MB, nterms);
          if (!resident) umma_commit(B_EMPTY0 + 8u * sb);
        }
        __syncwarp();
        acc_on = 1u;
        ent += tg;
        if (++sb == b_stages) { sb = 0; pb ^= 1u; }
      }
      if (elect_one_sync()) umma_commit(A_EMPTY0 + 8u * sa);
      __syncwarp();
      if (++sa == a_stages) { sa = 0; pa ^= 1u; }
    }
    if (elect_one_sync()) umma_commit(ACC_FULL0 + 8u * buf);
    __syncwarp();
  }
  wprof.flush(1, (threadIdx.x & 31) == 0);
}

// Sub-pixel (phase) layers: every entry multiplies its own accumulator column range [col, col+n) with its own N.
// The entry table is compile-time (it mirrors PH_* in nst_api.cu), so a weight stage unrolls into straight-line
// uniform-register code; a runtime-table loop cost ~290 cycles per MMA (descriptors built in vector registers).
template <int KIND, int K16, int MB, int E0, int E1>
__device__ __forceinline__ void issue_phase(uint32_t tmem_acc, uint32_t a_stage_lo, uint32_t b_stage_addr, uint32_t a_kstep16, uint32_t P,
                                            uint32_t cout, uint32_t planes, uint32_t npad, uint32_t idesc0, uint32_t acc_first) {
  // KIND 1: upsample + reflect pad + conv3 (11 entries); KIND 2: ConvTranspose2d k3 s2 p1 op1 (5 entries)
  constexpr int R[11]  = {1, KIND == 1 ? 0 : 2, KIND == 1 ? 2 : 1, 1, KIND == 1 ? 1 : 2, 1, 1, 0, 0, 2, 2};
  constexpr int S[11]  = {1, 1, KIND == 1 ? 1 : 2, KIND == 1 ? 0 : 2, KIND == 1 ? 0 : 2, 2, 2, 0, 2, 0, 2};
  constexpr int P0[11] = {0, KIND == 1 ? 0 : 2, KIND == 1 ? 2 : 1, KIND == 1 ? 0 : 3, KIND == 1 ? 2 : 3, 1, 3, 0, 1, 2, 3};
  constexpr int NP[11] = {4, 2, KIND == 1 ? 2 : 1, 1, 1, 1, 1, 1, 1, 1, 1};
  uint32_t units = 0;
#pragma unroll
  for (int e = E0; e < E1; ++e) {
    const uint32_t n = (uint32_t)NP[e] * cout, col = (uint32_t)P0[e] * cout;
    const uint32_t a_lo = a_stage_lo + (uint32_t)R[e] * P + (uint32_t)S[e];
    const uint32_t b_lo = (b_stage_addr + units * planes * cout) | (n << 16);           // LBO = n*16 bytes
    const uint32_t idesc = idesc0 | ((n >> 3) << 17);
#pragma unroll
    for (int j = 0; j < K16; ++j) {
      const uint64_t bdesc = ((uint64_t)DESC_HI << 32) | (b_lo + (uint32_t)j * 2u * n);
#pragma unroll
      for (int mb = 0; mb < MB; ++mb)
        umma_f16(tmem_acc + (uint32_t)mb * npad + col, ((uint64_t)DESC_HI << 32) | (a_lo + (uint32_t)j * a_kstep16 + (uint32_t)mb * 128u), bdesc, idesc,
                 (e == 0 && j == 0) ? acc_first : 1u);
    }
    units += (uint32_t)NP[e];
  }
}

template <int KIND, int K16, int MB_>
__device__ __noinline__ void mma_role_phase(const TcPlan& t, const MmaCtx c) {
  const uint32_t fmt = t.is_bf16 ? 1u : 0u;
  const uint32_t idesc0 = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(128 >> 4) << 24);
  const uint32_t a_lo_base = ((c.smem_a >> 4) & 0x3FFFu) | ((t.a_plane_bytes >> 4) << 16);
  const uint32_t b_addr_base = (c.smem_b >> 4) & 0x3FFFu;
  const uint32_t a_kstep16 = (2u * t.a_plane_bytes) >> 4;
  const uint32_t a_stage16 = t.a_stage_bytes >> 4, b_stage16 = t.b_stage_bytes >> 4;
  const int MB = MB_ ? MB_ : t.MB, NPAD = t.NPAD, k16 = t.k16_per_ent;
  const uint32_t P = (uint32_t)t.P, cout = (uint32_t)(t.NPAD >> 2), planes = (uint32_t)t.planes;
  const int nchunks = t.nchunks, ngroups = t.ngroups, num_tiles = c.end_tile;
  const uint32_t a_stages = (uint32_t)t.a_stages, b_stages = (uint32_t)t.b_stages;
  const uint32_t A_FULL0 = c.bar0, A_EMPTY0 = c.bar0 + 8u * TC_MAX_A_STAGES, B_FULL0 = c.bar0 + 8u * (2 * TC_MAX_A_STAGES),
                 B_EMPTY0 = B_FULL0 + 8u * TC_MAX_B_STAGES, ACC_FULL0 = B_FULL0 + 8u * (2 * TC_MAX_B_STAGES), ACC_EMPTY0 = ACC_FULL0 + 16u;
  const uint32_t* ent_info = c.ent_tab + 32;
  int gfirst[5];
#pragma unroll
  for (int i = 0; i < 5; ++i) gfirst[i] = t.grp_first[i];
  WaitProf wprof; wprof.init(c.prof);
  uint32_t sa = 0, pa = 0, sb = 0, pb = 0, it = 0;
  for (int tile = c.first_tile; tile < num_tiles; tile += c.tile_step, ++it) {
    const uint32_t buf = it & 1u;
    mbar_wait_warp(ACC_EMPTY0 + 8u * buf, ((it >> 1) & 1u) ^ 1u, c.err, 2, wprof);
    tc_fence_after();
    const uint32_t tmem_acc = c.tmem_base + buf * (uint32_t)(MB * NPAD);
    for (int ch = 0; ch < nchunks; ++ch) {
      mbar_wait_warp(A_FULL0 + 8u * sa, pa, c.err, 3, wprof);
      tc_fence_after();
      const uint32_t a_stage_lo = a_lo_base + sa * a_stage16;
      for (int g = 0; g < ngroups; ++g) {
        mbar_wait_warp(B_FULL0 + 8u * sb, pb, c.err, 4, wprof);
        tc_fence_after();
        if (elect_one_sync()) {
          const uint32_t b_stage_addr = b_addr_base + sb * b_stage16;
          if (K16 != 0) {
            constexpr int G1 = KIND == 1 ? 3 : 2, NE = KIND == 1 ? 11 : 5;
            if (g == 0) issue_phase<KIND, (K16 ? K16 : 1), (MB_ ? MB_ : 1), 0, G1>(tmem_acc, a_stage_lo, b_stage_addr, a_kstep16, P, cout, planes, (uint32_t)NPAD, idesc0, ch > 0 ? 1u : 0u);
            else issue_phase<KIND, (K16 ? K16 : 1), (MB_ ? MB_ : 1), G1, NE>(tmem_acc, a_stage_lo, b_stage_addr, a_kstep16, P, cout, planes, (uint32_t)NPAD, idesc0, 1u);
          } else {
          const int e1 = g == 0 ? gfirst[1] : gfirst[2];
          for (int e = (g == 0 ? gfirst[0] : gfirst[1]); e < e1; ++e) {
            const uint32_t info = ent_info[e];
            const uint32_t col = info & 511u, n = ((info >> 9) & 31u) * 16u;
            const uint32_t a_lo = a_stage_lo + c.ent_tab[e];
            const uint32_t b_lo = (b_stage_addr + (info >> 14)) | (n << 16);          // LBO = n*16 bytes
            const uint32_t idesc = idesc0 | ((n >> 3) << 17);
            const uint32_t acc = (ch > 0 || e > 0) ? 1u : 0u;
            for (int j = 0; j < k16; ++j) {
              const uint64_t bdesc = ((uint64_t)DESC_HI << 32) | (b_lo + (uint32_t)j * 2u * n);
              for (int mb = 0; mb < MB; ++mb)
                umma_f16(tmem_acc + (uint32_t)(mb * NPAD) + col, ((uint64_t)DESC_HI << 32) | (a_lo + (uint32_t)j * a_kstep16 + (uint32_t)mb * 128u), bdesc,
                         idesc, (j > 0) ? 1u : acc);
            }
          }
          }
          umma_commit(B_EMPTY0 + 8u * sb);
        }
        __syncwarp();
        if (++sb == b_stages) { sb = 0; pb ^= 1u; }
      }
      if (elect_one_sync()) umma_commit(A_EMPTY0 + 8u * sa);
      __syncwarp();
      if (++sa == a_stages) { sa = 0; pa ^= 1u; }
    }
    if (elect_one_sync()) umma_commit(ACC_FULL0 + 8u * buf);
    __syncwarp();
  }
  wprof.flush(1, (threadIdx.x & 31) == 0);
}

// ---------------------------------------------------------------------------------------------
// epilogue of the swapped (weights = A operand) form: TMEM lane = output channel, column = tile position
// ---------------------------------------------------------------------------------------------
// Warp `quad` owns channels [32*quad, 32*quad + 32): each thread reads its channel's accumulator row 32 positions at a time, adds
// the valid positions into its own (sum, sum of squares) -- no cross-lane reduction -- and the 8 lanes that share an 8-channel
// plane transpose their (channel x position) fragments with 12 shuffles per 16 positions so that every lane ends up with all 8
// channels of two positions = two 16-byte planar-8 slots, stored straight to HBM (8 lanes x 32 B = 256 contiguous bytes per plane).
// Running sums over the CTA's tile run are double registers; a sample change (and the end) writes them to this CTA's slot of the
// partial table (stats_finalize_kernel folds the CTAs in index order: bit-reproducible).
__device__ __forceinline__ void transpose8x8(uint32_t (&r)[8], int l8) {
#pragma unroll
  for (int s = 4; s >= 1; s >>= 1) {
    const bool up = (l8 & s) != 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (i & s) continue;
      const uint32_t send = up ? r[i] : r[i + s];
      const uint32_t recv = __shfl_xor_sync(0xffffffffu, send, s);
      if (up) r[i] = recv; else r[i + s] = recv;
    }
  }
}

__device__ __noinline__ void epilogue_swap(const TcArgs& args, uint32_t tmem_base, uint32_t acc_full0, uint32_t acc_empty0, int first_tile, int tile_step,
                                           int end_tile, int quad, int lane, int eg, int EG) {
  const ConvParams& p = args.p;
  const TcPlan& t = args.t;
  const int co = quad * 32 + lane;                 // this thread's output channel = its TMEM lane
  const int l8 = lane & 7, g8 = lane >> 3;
  const int ncols = t.MB * 128;                    // positions per tile (accumulator columns)
  // loop invariants in registers: `args` lives in the kernel-parameter bank behind a generic pointer here, and every asm volatile
  // ("memory") in the loop would otherwise force a reload (ncu: long-scoreboard stalls on exactly these loads)
  const int P = t.P, TH = t.TH, TW = t.TW, Ho = p.Ho, Wo = p.Wo;
  const float inv_p = 1.0f / (float)P;
  const size_t plane_slots = (size_t)Ho * Wo;      // 16-byte slots per 8-channel plane
  uint4* const out_base = reinterpret_cast<uint4*>(p.out);
  const int c8 = p.Cout >> 3;
  const bool csplit = p.out_colsplit != 0;
  double sum_d = 0.0, sq_d = 0.0;
  int cur_n = -1;
  auto flush = [&]() {
    if (cur_n < 0) return;
    double* st = p.out_part + ((((size_t)blockIdx.x * EG + eg) * p.N + cur_n) * p.Cout + co) * 2;
    st[0] = sum_d; st[1] = sq_d;
    sum_d = 0.0; sq_d = 0.0;
  };
  WaitProf wprof; wprof.init(args.prof);
  uint32_t it = (uint32_t)eg;
  for (int tile = first_tile + eg * tile_step; tile < ((t.dbg & 32) ? 0 : end_tile); tile += EG * tile_step, it += EG) {
    const TileCoord tc = decode_tile(t, tile);
    if (tc.n != cur_n) { flush(); cur_n = tc.n; }
    const uint32_t buf = it & 1u;
    mbar_wait_warp(acc_full0 + 8u * buf, (it >> 1) & 1u, args.err_flag, 5, wprof);
    tc_fence_after();
    const uint32_t tbase = tmem_base + ((uint32_t)(quad * 32) << 16) + buf * (uint32_t)ncols;
    uint4* const oplane = out_base + ((size_t)tc.n * c8 + (size_t)(quad * 4 + g8)) * plane_slots;
    float ts = 0.f, tq = 0.f;
    for (int c0 = 0; c0 < ncols; c0 += 32) {
      uint32_t v[32];
      tmem_ld32_nowait(tbase + (uint32_t)c0, v);
      // while the load is in flight: validity / output slot of position c0 + lane (the same for every lane's channel)
      const int m = c0 + lane;
      const int ty = (int)(((float)m + 0.5f) * inv_p);
      const int tx = m - ty * P;
      const int oy = tc.ty0 + ty, ox = tc.tx0 + tx;
      const bool valid = (ty < TH) & (tx < TW) & (oy < Ho) & (ox < Wo);
      const int pix = valid ? oy * Wo + (csplit ? (ox & 1) * ((Wo + 1) >> 1) + (ox >> 1) : ox) : -1;
      const uint32_t vmask = __ballot_sync(0xffffffffu, valid);
      const long long tA = wprof.buf ? clock64() : 0;
      tmem_ld_wait();
      const long long tB = wprof.buf ? clock64() : 0;
      if (c0 + 32 >= ncols) { tc_fence_before(); mbar_arrive_warp(acc_empty0 + 8u * buf); }     // last read of this accumulator buffer
      uint32_t h[16];
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        const float a = __uint_as_float(v[2 * k]), b = __uint_as_float(v[2 * k + 1]);
        const float am = ((vmask >> (2 * k)) & 1u) ? a : 0.f, bm = ((vmask >> (2 * k + 1)) & 1u) ? b : 0.f;
        ts += am; tq = fmaf(am, am, tq);
        ts += bm; tq = fmaf(bm, bm, tq);
        h[k] = pack2(a, b, 0);
      }
#pragma unroll
      for (int blk = 0; blk < 2; ++blk) {
        uint32_t r[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) r[k] = h[blk * 8 + k];
        transpose8x8(r, l8);                              // r[k] = (channel k of the plane) x (positions 2*l8, 2*l8 + 1) of this block
        const int pix0 = __shfl_sync(0xffffffffu, pix, blk * 16 + 2 * l8);
        const int pix1 = __shfl_sync(0xffffffffu, pix, blk * 16 + 2 * l8 + 1);
        if (pix0 >= 0)
          oplane[pix0] = make_uint4(__byte_perm(r[0], r[1], 0x5410), __byte_perm(r[2], r[3], 0x5410), __byte_perm(r[4], r[5], 0x5410), __byte_perm(r[6], r[7], 0x5410));
        if (pix1 >= 0)
          oplane[pix1] = make_uint4(__byte_perm(r[0], r[1], 0x7632), __byte_perm(r[2], r[3], 0x7632), __byte_perm(r[4], r[5], 0x7632), __byte_perm(r[6], r[7], 0x7632));
      }
      if (wprof.buf) wprof.acc[0] += tB - tA;                                              // debug: cycles waiting for the TMEM load
    }
    sum_d += (double)ts; sq_d += (double)tq;
  }
  flush();
  wprof.flush(2, eg == 0 && quad == (TC_FIRST_EPI_WARP & 3) && lane == 0);
}

constexpr int NBARS = 3 * TC_MAX_A_STAGES + 2 * TC_MAX_B_STAGES + 4;     // A_FULL, A_EMPTY, B_FULL, B_EMPTY, ACC_FULL/EMPTY x2, TMA_FULL

__global__ void __launch_bounds__(TC_THREADS, 1)
conv_tc_kernel(const __grid_constant__ TcArgs args) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const ConvParams& p = args.p;
  const TcPlan& t = args.t;
  unsigned int* err = args.err_flag;

  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem_a + (size_t)t.a_stages * t.a_stage_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_b + (size_t)t.b_stages * t.b_stage_bytes);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + NBARS);
  uint32_t* ent_tab = tmem_slot + 4;                                         // [TC_MAX_ENT rounded to 84] tap offsets (MMA warp)
  int* srcpix = reinterpret_cast<int*>(ent_tab + 84);                       // [2][NPOS] source tables of the current / next tile (gather producers only)
  float2* coeff = reinterpret_cast<float2*>(srcpix + t.tbl_ints);             // [Cin] InstanceNorm (scale, shift) of sample n
  float* fold_ring0 = reinterpret_cast<float*>(coeff + 256);                   // [256][TC_FOLD_PITCH] (fold_s layers) | double [4 warps][stat_ch][2] statistics

  const uint32_t bar0 = smem_u32(bars);
  auto A_FULL = [&](int s) { return bar0 + 8u * (uint32_t)s; };
  auto A_EMPTY = [&](int s) { return bar0 + 8u * (uint32_t)(TC_MAX_A_STAGES + s); };
  auto B_FULL = [&](int s) { return bar0 + 8u * (uint32_t)(2 * TC_MAX_A_STAGES + s); };
  auto B_EMPTY = [&](int s) { return bar0 + 8u * (uint32_t)(2 * TC_MAX_A_STAGES + TC_MAX_B_STAGES + s); };
  auto ACC_FULL = [&](int s) { return bar0 + 8u * (uint32_t)(2 * TC_MAX_A_STAGES + 2 * TC_MAX_B_STAGES + s); };
  auto ACC_EMPTY = [&](int s) { return bar0 + 8u * (uint32_t)(2 * TC_MAX_A_STAGES + 2 * TC_MAX_B_STAGES + 2 + s); };
  auto TMA_FULL = [&](int s) { return bar0 + 8u * (uint32_t)(2 * TC_MAX_A_STAGES + 2 * TC_MAX_B_STAGES + 4 + s); };

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const long long prof_c0 = args.prof ? clock64() : 0;
  const unsigned long long prof_t0 = args.prof ? global_ns() : 0ull;
  const int first_prod_warp = TC_FIRST_PROD_WARP + 4 * (t.epi_groups - 1);   // a second epilogue group takes warps 6-9 (TMA-fed layers only)

  if (threadIdx.x == 0) {
    for (int s = 0; s < TC_MAX_A_STAGES; ++s) { mbar_init(A_FULL(s), (uint32_t)t.prod_warps * ARRIVALS_PER_WARP); mbar_init(A_EMPTY(s), 1); }
    for (int s = 0; s < TC_MAX_B_STAGES; ++s) { mbar_init(B_FULL(s), 1); mbar_init(B_EMPTY(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(ACC_FULL(s), 1); mbar_init(ACC_EMPTY(s), 4 * ARRIVALS_PER_WARP); }
    for (int s = 0; s < TC_MAX_A_STAGES; ++s) mbar_init(TMA_FULL(s), 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(smem_u32(tmem_slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // Each CTA walks a contiguous run of tiles; tiles are numbered column-major inside a frame (decode_tile), so consecutive tiles
  // of a CTA are vertical neighbours (halo rows are shared, statistics are flushed once per frame run, L2 sees nearby rows).
  // Tile -> CTA assignment (static, so the statistics' partial sums are reproducible).  rr_tiles: CTA b takes tiles b, b + grid, ...
  // -- the tiles in flight at any moment are neighbours (their halos hit L2 together) and the expensive ones (image borders: a whole
  // tile column at the left / right edge) spread evenly; with contiguous runs the CTAs that own a border column finish 20-35 % late
  // (tcspan, tools/prof2.sh).  Otherwise: contiguous runs, column-major inside a frame.
  const int first_tile = t.rr_tiles ? (int)blockIdx.x : (int)((long long)blockIdx.x * t.num_tiles / gridDim.x);
  const int end_tile = t.rr_tiles ? t.num_tiles : (int)((long long)(blockIdx.x + 1) * t.num_tiles / gridDim.x);
  const int tile_step = t.rr_tiles ? (int)gridDim.x : 1;

  if (warp == 0) {
    // =========================== weight loader ===========================
    if (lane == 0 && !(t.dbg & 64)) {
      WaitProf wprof; wprof.init(args.prof);
      uint32_t sb = 0, pb = 0;
      const uint8_t* wp = reinterpret_cast<const uint8_t*>(args.wpack);
      const int nstage_per_tile = t.nchunks * t.ngroups;
      if (t.b_resident) {
        for (int g = 0; g < nstage_per_tile; ++g) {
          mbar_arrive_expect_tx(B_FULL(g), t.b_stage_bytes);
          bulk_g2s(smem_u32(smem_b + (size_t)g * t.b_stage_bytes), wp + (size_t)g * t.b_stage_bytes, t.b_stage_bytes, B_FULL(g));
        }
      } else
      for (int tile = first_tile; tile < end_tile; tile += tile_step) {
        for (int g = 0; g < nstage_per_tile; ++g) {
          mbar_wait(B_EMPTY(sb), pb ^ 1u, err, 1, wprof);
          if (t.dbg & 8) { mbar_arrive(B_FULL(sb)); if (++sb == (uint32_t)t.b_stages) { sb = 0; pb ^= 1u; } continue; }
          mbar_arrive_expect_tx(B_FULL(sb), t.b_stage_bytes);
          bulk_g2s(smem_u32(smem_b + (size_t)sb * t.b_stage_bytes), wp + (size_t)g * t.b_stage_bytes,
                   t.b_stage_bytes, B_FULL(sb));
          if (++sb == (uint32_t)t.b_stages) { sb = 0; pb ^= 1u; }
        }
      }
      wprof.flush(0, true);
    }
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    for (int i = lane; i < t.nent; i += 32) ent_tab[i] = (uint32_t)t.ent_off[i];
    if (t.phase_mode && lane < 16) ent_tab[32 + lane] = t.ent_info[lane];
    __syncwarp();
    const MmaCtx mc = {smem_u32(smem_a), smem_u32(smem_b), bar0, tmem_base, ent_tab, err, args.prof, first_tile, tile_step, end_tile};
    if (t.phase_mode) {
      if (t.phase_mode == 1 && t.k16_per_ent == 2 && t.MB == 1) mma_role_phase<1, 2, 1>(t, mc);
      else if (t.phase_mode == 1 && t.k16_per_ent == 2 && t.MB == 2) mma_role_phase<1, 2, 2>(t, mc);
      else if (t.phase_mode == 2 && t.k16_per_ent == 2 && t.MB == 1) mma_role_phase<2, 2, 1>(t, mc);
      else if (t.phase_mode == 2 && t.k16_per_ent == 2 && t.MB == 2) mma_role_phase<2, 2, 2>(t, mc);
      else mma_role_phase<1, 0, 0>(t, mc);
    }
    else if (t.swap_ab && t.k16_per_ent == 2 && t.MB == 2) mma_role<2, 2, 1, true>(t, mc);
    else if (t.swap_ab && t.k16_per_ent == 2) mma_role<2, 1, 1, true>(t, mc);
    else if (t.swap_ab && t.MB == 2) mma_role<1, 2, 1, true>(t, mc);
    else if (t.swap_ab) mma_role<1, 1, 1, true>(t, mc);
    else if (t.nterms == 3 && t.k16_per_ent == 1 && t.MB == 4) mma_role<1, 4, 3>(t, mc);
    else if (t.nterms == 3 && t.k16_per_ent == 1 && t.MB == 2) mma_role<1, 2, 3>(t, mc);
    else if (t.nterms == 3 && t.k16_per_ent == 1 && t.MB == 1) mma_role<1, 1, 3>(t, mc);
    else if (t.nterms == 3 && t.k16_per_ent == 2 && t.MB == 1) mma_role<2, 1, 3>(t, mc);
    else if (t.nterms == 3 && t.k16_per_ent == 2 && t.MB == 2) mma_role<2, 2, 3>(t, mc);
    else if (t.nterms == 1 && t.k16_per_ent == 2 && t.MB == 2) mma_role<2, 2>(t, mc);
    else if (t.nterms == 1 && t.k16_per_ent == 2 && t.MB == 4) mma_role<2, 4>(t, mc);
    else if (t.nterms == 1 && t.k16_per_ent == 2 && t.MB == 8) mma_role<2, 8>(t, mc);
    else if (t.nterms == 1 && t.k16_per_ent == 2 && t.MB == 1) mma_role<2, 1>(t, mc);
    else if (t.nterms == 1 && t.k16_per_ent == 1 && t.MB == 8) mma_role<1, 8>(t, mc);
    else if (t.nterms == 1 && t.k16_per_ent == 1 && t.MB == 4) mma_role<1, 4>(t, mc);
    else if (t.nterms == 1 && t.k16_per_ent == 1 && t.MB == 2) mma_role<1, 2>(t, mc);
    else if (t.nterms == 1 && t.k16_per_ent == 1 && t.MB == 1) mma_role<1, 1>(t, mc);
    else mma_role<0, 0>(t, mc);
  } else if (warp < first_prod_warp) {
    // =========================== epilogue ===========================
    // EG = 1: one group of four warps drains every tile.  EG = 2 (epilogue-bound TMA-fed layers): warps 2-5 drain the even tiles of
    // the CTA's run (accumulator buffer 0), warps 6-9 the odd ones (buffer 1), each with its own scratch and named barriers.
    const int EG = t.epi_groups, eg = (warp - TC_FIRST_EPI_WARP) >> 2;
    const int ewarp = (warp - TC_FIRST_EPI_WARP) & 3;   // warp inside its group
    const int quad = warp & 3;                        // TMEM lane quadrant this warp may read
    float* const fold_ring = fold_ring0 + (size_t)eg * (t.epi_scratch_bytes >> 2);
    const int bar_a = 3 + 3 * eg, bar_b = 4 + 3 * eg, bar_c = 5 + 3 * eg;
    const int njc = t.NPAD >> 4;
    const int my_chan = ((lane & 16) ? 8 : 0) + ((lane & 8) ? 4 : 0) + ((lane & 4) ? 2 : 0) + ((lane & 2) ? 1 : 0);
    if (t.swap_ab) {
      epilogue_swap(args, tmem_base, ACC_FULL(0), ACC_EMPTY(0), first_tile, tile_step, end_tile, quad, lane, eg, EG);
    } else if (t.fold_s == 2) {
      // Image layer with the horizontal taps folded into N, pitch 32:  TMEM row m holds T[m][s*3+co] = sum_{r,ci} in[m + r*32][ci] w[co][ci][r][s]
      // and out[m][co] = sum_s T[m + s][s*3 + co].  A warp's 32 rows are exactly one tile row (tx = lane) and only tx < TW <= 24 is an
      // output, so row m + s is lane + s of the same warp: the sum is 24 shuffles, no shared memory and no barrier.
      const float b0 = p.bias[0], b1 = p.bias[1], b2 = p.bias[2];
      uint32_t it = (uint32_t)eg;
      WaitProf wprof; wprof.init(args.prof);
      for (int tile = first_tile + eg * tile_step; tile < ((t.dbg & 32) ? 0 : end_tile); tile += EG * tile_step, it += EG) {
        const TileCoord tc = decode_tile(t, tile);
        const uint32_t buf = it & 1u;
        mbar_wait_warp(ACC_FULL(buf), (it >> 1) & 1u, err, 5, wprof);
        tc_fence_after();
        const int ox = tc.tx0 + lane;
        const bool col_ok = lane < t.TW && ox < p.Wo;
        for (int mb = 0; mb < t.MB; ++mb) {
          const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + buf * (uint32_t)(t.MB * t.NPAD) + (uint32_t)(mb * t.NPAD);
          uint32_t v[32];
          tmem_ld32_nowait(taddr, v);
          tmem_ld_wait();
          if (mb == t.MB - 1) { tc_fence_before(); mbar_arrive_warp(ACC_EMPTY(buf)); }
          float a0 = b0 + __uint_as_float(v[0]), a1 = b1 + __uint_as_float(v[1]), a2 = b2 + __uint_as_float(v[2]);
#pragma unroll
          for (int sx = 1; sx < 9; ++sx) {
            a0 += __shfl_down_sync(0xffffffffu, __uint_as_float(v[3 * sx]), sx);
            a1 += __shfl_down_sync(0xffffffffu, __uint_as_float(v[3 * sx + 1]), sx);
            a2 += __shfl_down_sync(0xffffffffu, __uint_as_float(v[3 * sx + 2]), sx);
          }
          const int ty = mb * 4 + quad;                  // 128 rows of an M block = 4 tile rows of 32 positions
          const int oy = tc.ty0 + ty;
          if (col_ok && ty < t.TH && oy < p.Ho) store_image(p, tc.n, oy, ox, a0, a1, a2);
        }
      }
      wprof.flush(2, warp == TC_FIRST_EPI_WARP && lane == 0);
    } else if (t.fold_s) {
      // Image layer with the horizontal taps folded into N:  TMEM holds T[m][s*3+co] = sum_{r,ci} in[m + r*P][ci] w[co][ci][r][s];
      // out[m][co] = sum_s T[m + s][s*3 + co].  Rows are exchanged through a ring of 128 + 8 rows in shared memory (row R at
      // slot R mod 136: block mb never overwrites the last 8 rows of block mb - 1), then every thread finishes the output 8 rows
      // behind its own TMEM row.
      const int tid = quad * 32 + lane;             // TMEM lane (row of the M block) this thread reads (= its slot of the group's ring)
      const float b0 = p.bias[0], b1 = p.bias[1], b2 = p.bias[2];
      uint32_t it = (uint32_t)eg;
      WaitProf wprof; wprof.init(args.prof);
      for (int tile = first_tile + eg * tile_step; tile < ((t.dbg & 32) ? 0 : end_tile); tile += EG * tile_step, it += EG) {
        const TileCoord tc = decode_tile(t, tile);
        const uint32_t buf = it & 1u;
        mbar_wait_warp(ACC_FULL(buf), (it >> 1) & 1u, err, 5, wprof);
        tc_fence_after();
        for (int mb = 0; mb < t.MB; ++mb) {
          const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + buf * (uint32_t)(t.MB * t.NPAD) + (uint32_t)(mb * t.NPAD);
          float v0[16], v1[16];
          tmem_ld16(taddr, v0);
          tmem_ld16(taddr + 16u, v1);
          if (mb == t.MB - 1) { tc_fence_before(); mbar_arrive_warp(ACC_EMPTY(buf)); }
          named_bar_sync(bar_a, 128);                       // readers of the previous block are done with this half's neighbour rows
          float* row = fold_ring + (size_t)((mb * 128 + tid) % TC_FOLD_ROWS) * TC_FOLD_PITCH;
#pragma unroll
          for (int i = 0; i < 16; ++i) row[i] = v0[i];
#pragma unroll
          for (int i = 0; i < 11; ++i) row[16 + i] = v1[i];
          named_bar_sync(bar_b, 128);
          const int m = mb * 128 + tid - 8;
          if (m >= 0) {
            const int ty = m / t.P, tx = m - ty * t.P;
            const int oy = tc.ty0 + ty, ox = tc.tx0 + tx;
            if (ty < t.TH && tx < t.TW && oy < p.Ho && ox < p.Wo) {
              float a0 = b0, a1 = b1, a2 = b2;
#pragma unroll
              for (int sx = 0; sx < 9; ++sx) {
                const int mm = m + sx;
                const float* rr = fold_ring + (size_t)(mm % TC_FOLD_ROWS) * TC_FOLD_PITCH + sx * 3;
                a0 += rr[0]; a1 += rr[1]; a2 += rr[2];
              }
              store_image(p, tc.n, oy, ox, a0, a1, a2);
            }
          }
        }
      }
      wprof.flush(2, warp == TC_FIRST_EPI_WARP && lane == 0);
    } else {
      // Per-warp running (sum, sum of squares) of every output channel live in shared memory as DOUBLES ([stat_ch][2] per warp):
      // a tile contributes fp32 sums over <= MB*128 rows, the running sum over the whole CTA run is double, so E[x^2] - mean^2
      // stays well conditioned for |mean| >> sigma.  When the sample changes (and at the end) the four warps' sums are added in
      // a fixed order and written -- no atomics -- to this CTA's slot of the partial table; stats_finalize_kernel folds the CTAs
      // of a sample in index order, so the statistics (hence the whole forward) are bit-reproducible run to run.
      double* wstat = reinterpret_cast<double*>(fold_ring) + (size_t)ewarp * t.stat_ch * 2;
      for (int c = lane; c < t.stat_ch * 2; c += 32) wstat[c] = 0.0;
      __syncwarp();
      int cur_n = -1;
      const int etid = ewarp * 32 + lane;
      auto flush = [&]() {
        if (cur_n < 0 || p.out_fmt != OUT_F32_NHWC_STATS) return;
        named_bar_sync(bar_c, 128);             // every warp of the group has added its last tile of sample cur_n
        double* all = reinterpret_cast<double*>(fold_ring);
        for (int c = etid; c < p.co_pass; c += 128) {
          double s1 = 0.0, s2 = 0.0;
#pragma unroll
          for (int w = 0; w < 4; ++w) { s1 += all[((size_t)w * t.stat_ch + c) * 2]; s2 += all[((size_t)w * t.stat_ch + c) * 2 + 1]; }
          double* st = p.out_part + ((((size_t)blockIdx.x * EG + eg) * p.N + cur_n) * p.Cout + p.co_off + c) * 2;
          st[0] = s1; st[1] = s2;
#pragma unroll
          for (int w = 0; w < 4; ++w) { all[((size_t)w * t.stat_ch + c) * 2] = 0.0; all[((size_t)w * t.stat_ch + c) * 2 + 1] = 0.0; }
        }
        named_bar_sync(bar_c, 128);
      };
      const float inv_p = 1.0f / (float)t.P;
      uint32_t it = (uint32_t)eg;
      WaitProf wprof; wprof.init(args.prof);
      for (int tile = first_tile + eg * tile_step; tile < ((t.dbg & 32) ? 0 : end_tile); tile += EG * tile_step, it += EG) {
        const TileCoord tc = decode_tile(t, tile);
        if (tc.n != cur_n) { flush(); cur_n = tc.n; }
        const uint32_t buf = it & 1u;
        mbar_wait_warp(ACC_FULL(buf), (it >> 1) & 1u, err, 5, wprof);
        tc_fence_after();
        const uint32_t tbase = tmem_base + ((uint32_t)(quad * 32) << 16) + buf * (uint32_t)(t.MB * t.NPAD);
        if (p.out_fmt == OUT_F32_NHWC_STATS && !p.phase_mode && !(t.dbg & (4 | 65536))) {
          // Plain layers, software-pipelined: the accumulator is drained as a flat sequence of (16-channel group, M block) fragments and
          // the tcgen05.ld of fragment i + 1 is in flight while fragment i is packed, stored and added to the statistics (with four
          // epilogue warps -- one per scheduler -- nothing else hides the TMEM latency; tcprof had this role 96 % busy on the first
          // layer and the MMA warp 45 % of its time waiting for a free accumulator on the stride-2 32->64 layer).  The accumulator is
          // released as soon as its last fragment is in registers.  Same per-lane summation order as the unpipelined form below.
          const int MBn = t.MB;                                       // (small frames plan MB = 3, 5, ...: no shift / mask tricks)
          const int nit = njc * MBn;
          const int Wout = p.Wo;
          const size_t plane_elems = (size_t)p.Ho * Wout * 8;
          float s[16], q[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) { s[i] = 0.f; q[i] = 0.f; }
          int icg = 0, imb = 0, pcg = 0, pmb = 0;                     // (channel group, M block) of the next fragment to load / to process
          auto issue = [&](uint32_t (&r)[16]) {
            tmem_ld16_nowait(tbase + (uint32_t)(imb * t.NPAD + (icg << 4)), r);
            if (++imb == MBn) { imb = 0; ++icg; }
          };
          auto release = [&]() { tc_fence_before(); mbar_arrive_warp(ACC_EMPTY(buf)); };
          auto process = [&](const uint32_t (&r)[16]) {
            const int mb = pmb, co0 = pcg << 4;
            if (++pmb == MBn) { pmb = 0; ++pcg; }
            const int m = mb * 128 + quad * 32 + lane;
            const int ty = (int)(((float)m + 0.5f) * inv_p);
            const int tx = m - ty * t.P;
            const int oy = tc.ty0 + ty, ox = tc.tx0 + tx;
            if (ty < t.TH && tx < t.TW && oy < p.Ho && ox < p.Wo) {
              float v[16];
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
              const size_t plane0 = ((size_t)tc.n * (p.Cout >> 3) + (size_t)((p.co_off + co0) >> 3)) * plane_elems;
              const size_t pix = (size_t)oy * Wout + (p.out_colsplit ? (ox & 1) * ((Wout + 1) >> 1) + (ox >> 1) : ox);
              if (p.out16) {
                __half* o = (__half*)p.out + plane0 + pix * 8;
#pragma unroll
                for (int hp = 0; hp < 2; ++hp)
                  *reinterpret_cast<uint4*>(o + (size_t)hp * plane_elems) =
                      make_uint4(pack2(v[8 * hp], v[8 * hp + 1], 0), pack2(v[8 * hp + 2], v[8 * hp + 3], 0),
                                 pack2(v[8 * hp + 4], v[8 * hp + 5], 0), pack2(v[8 * hp + 6], v[8 * hp + 7], 0));
              } else {
                float* o = (float*)p.out + plane0 + pix * 8;
#pragma unroll
                for (int hp = 0; hp < 2; ++hp) {
                  float4* o4 = reinterpret_cast<float4*>(o + (size_t)hp * plane_elems);
                  o4[0] = make_float4(v[8 * hp], v[8 * hp + 1], v[8 * hp + 2], v[8 * hp + 3]);
                  o4[1] = make_float4(v[8 * hp + 4], v[8 * hp + 5], v[8 * hp + 6], v[8 * hp + 7]);
                }
              }
#pragma unroll
              for (int j = 0; j < 16; ++j) { s[j] += v[j]; q[j] = fmaf(v[j], v[j], q[j]); }
            }
            if (mb == MBn - 1) {                    // last M block of this channel group: one butterfly over the 32 lanes
#pragma unroll
              for (int w = 8, off = 16; w >= 1; w >>= 1, off >>= 1) {
                const bool up = (lane & off) != 0;
#pragma unroll
                for (int j = 0; j < w; ++j) {
                  const float send_s = up ? s[j] : s[j + w];
                  const float send_q = up ? q[j] : q[j + w];
                  const float keep_s = up ? s[j + w] : s[j];
                  const float keep_q = up ? q[j + w] : q[j];
                  s[j] = keep_s + __shfl_xor_sync(0xffffffffu, send_s, off);
                  q[j] = keep_q + __shfl_xor_sync(0xffffffffu, send_q, off);
                }
              }
              const float ts = s[0] + __shfl_xor_sync(0xffffffffu, s[0], 1);
              const float tq = q[0] + __shfl_xor_sync(0xffffffffu, q[0], 1);
              if ((lane & 1) == 0) {
                double2* w2 = reinterpret_cast<double2*>(wstat) + co0 + my_chan;
                double2 cur = *w2;
                cur.x += (double)ts; cur.y += (double)tq;
                *w2 = cur;
              }
#pragma unroll
              for (int j = 0; j < 16; ++j) { s[j] = 0.f; q[j] = 0.f; }
            }
          };
          uint32_t ra[16], rb[16];
          issue(ra);
          for (int i = 0; i < nit; i += 2) {
            tmem_ld_wait();
            if (i + 1 < nit) issue(rb); else release();
            process(ra);
            if (i + 1 < nit) {
              tmem_ld_wait();
              if (i + 2 < nit) issue(ra); else release();
              process(rb);
            }
          }
        } else
        if (p.out_fmt == OUT_F32_NHWC_STATS) {
          const int nmb = (t.dbg & 4) ? 0 : t.MB;
          const int npair = p.phase_mode ? 2 : 1;                           // sub-pixel layers: column = phase*Cout + co, phase = a*2 + b;
                                                                            // the two b phases of a row phase a are adjacent pixels
          const int ncg = p.phase_mode ? (p.co_pass >> 4) : njc;               // 16-channel groups of the OUTPUT tensor
          const int Hout = p.phase_mode ? 2 * p.Ho : p.Ho, Wout = p.phase_mode ? 2 * p.Wo : p.Wo;
          const size_t plane_elems = (size_t)Hout * Wout * 8;                // one 8-channel plane of the planar-8 output
          for (int cg = 0; cg < ncg; ++cg) {
            // 16 channels at a time: the rows of every M block (and of the four sub-pixel phases, which feed the same
            // channel statistics) are accumulated per lane first, then ONE butterfly (16 -> 8 -> 4 -> 2 -> 1 live
            // channels per step) reduces over the 32 lanes.
            float s[16], q[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) { s[i] = 0.f; q[i] = 0.f; }
            const int co0 = cg * 16;
            const size_t plane0 = ((size_t)tc.n * (p.Cout >> 3) + (size_t)((p.co_off + co0) >> 3)) * plane_elems;
            for (int pa = 0; pa < npair; ++pa) {
            for (int mb = 0; mb < nmb; ++mb) {
              const int m = mb * 128 + quad * 32 + lane;
              const int ty = (int)(((float)m + 0.5f) * inv_p);
              const int tx = m - ty * t.P;
              const int oy = tc.ty0 + ty, ox = tc.tx0 + tx;
              const bool valid = ty < t.TH && tx < t.TW && oy < p.Ho && ox < p.Wo;
              float v[16], w[16];
              if (p.phase_mode) {
                tmem_ld16(tbase + (uint32_t)(mb * t.NPAD + (2 * pa) * p.co_pass + co0), v);          // pixel (2oy+pa, 2ox)
                tmem_ld16(tbase + (uint32_t)(mb * t.NPAD + (2 * pa + 1) * p.co_pass + co0), w);      // pixel (2oy+pa, 2ox+1)
              } else {
                tmem_ld16(tbase + (uint32_t)(mb * t.NPAD + co0), v);
              }
              if (valid) {
                // planar-8 store: 8 channels of a pixel = one slot; a warp's 32 rows are 32 consecutive pixels of an output row
                // (phase layers: 64), so each plane receives full 128-byte lines
                const size_t pix = p.phase_mode ? ((size_t)(2 * oy + pa) * Wout + 2 * ox)
                                                : ((size_t)oy * Wout + (p.out_colsplit ? (ox & 1) * ((Wout + 1) >> 1) + (ox >> 1) : ox));
                if (p.out16) {
                  __half* o = (__half*)p.out + plane0 + pix * 8;
#pragma unroll
                  for (int hp = 0; hp < 2; ++hp) {                           // the two 8-channel planes of this 16-channel group
                    uint4* o4 = reinterpret_cast<uint4*>(o + (size_t)hp * plane_elems);
                    o4[0] = make_uint4(pack2(v[8 * hp], v[8 * hp + 1], 0), pack2(v[8 * hp + 2], v[8 * hp + 3], 0),
                                       pack2(v[8 * hp + 4], v[8 * hp + 5], 0), pack2(v[8 * hp + 6], v[8 * hp + 7], 0));
                    if (p.phase_mode)
                      o4[1] = make_uint4(pack2(w[8 * hp], w[8 * hp + 1], 0), pack2(w[8 * hp + 2], w[8 * hp + 3], 0),
                                         pack2(w[8 * hp + 4], w[8 * hp + 5], 0), pack2(w[8 * hp + 6], w[8 * hp + 7], 0));
                  }
                } else {
                  float* o = (float*)p.out + plane0 + pix * 8;
#pragma unroll
                  for (int hp = 0; hp < 2; ++hp) {
                    float4* o4 = reinterpret_cast<float4*>(o + (size_t)hp * plane_elems);
                    o4[0] = make_float4(v[8 * hp], v[8 * hp + 1], v[8 * hp + 2], v[8 * hp + 3]);
                    o4[1] = make_float4(v[8 * hp + 4], v[8 * hp + 5], v[8 * hp + 6], v[8 * hp + 7]);
                    if (p.phase_mode) {
                      o4[2] = make_float4(w[8 * hp], w[8 * hp + 1], w[8 * hp + 2], w[8 * hp + 3]);
                      o4[3] = make_float4(w[8 * hp + 4], w[8 * hp + 5], w[8 * hp + 6], w[8 * hp + 7]);
                    }
                  }
                }
#pragma unroll
                for (int i = 0; i < 16; ++i) { s[i] += v[i]; q[i] = fmaf(v[i], v[i], q[i]); }
                if (p.phase_mode) {
#pragma unroll
                  for (int i = 0; i < 16; ++i) { s[i] += w[i]; q[i] = fmaf(w[i], w[i], q[i]); }
                }
              }
            }
            }
            if (cg == ncg - 1 && nmb > 0) { tc_fence_before(); mbar_arrive_warp(ACC_EMPTY(buf)); }
#pragma unroll
            for (int w = 8, off = 16; w >= 1; w >>= 1, off >>= 1) {
              const bool up = (lane & off) != 0;
#pragma unroll
              for (int i = 0; i < w; ++i) {
                const float send_s = up ? s[i] : s[i + w];
                const float send_q = up ? q[i] : q[i + w];
                const float keep_s = up ? s[i + w] : s[i];
                const float keep_q = up ? q[i + w] : q[i];
                s[i] = keep_s + __shfl_xor_sync(0xffffffffu, send_s, off);
                q[i] = keep_q + __shfl_xor_sync(0xffffffffu, send_q, off);
              }
            }
            const float ts = s[0] + __shfl_xor_sync(0xffffffffu, s[0], 1);
            const float tq = q[0] + __shfl_xor_sync(0xffffffffu, q[0], 1);
            if ((lane & 1) == 0) {
              double2* w2 = reinterpret_cast<double2*>(wstat) + co0 + my_chan;
              double2 cur = *w2;
              cur.x += (double)ts; cur.y += (double)tq;
              *w2 = cur;
            }
          }
          if (ncg == 0 || nmb == 0) { tc_fence_before(); mbar_arrive_warp(ACC_EMPTY(buf)); }
        } else {
          for (int mb = 0; mb < t.MB; ++mb) {
            const int m = mb * 128 + quad * 32 + lane;
            const int ty = m / t.P, tx = m - ty * t.P;
            const int oy = tc.ty0 + ty, ox = tc.tx0 + tx;
            float v[16];
            tmem_ld16(tbase + (uint32_t)(mb * t.NPAD), v);
            if (ty < t.TH && tx < t.TW && oy < p.Ho && ox < p.Wo) store_image(p, tc.n, oy, ox, v[0] + p.bias[0], v[1] + p.bias[1], v[2] + p.bias[2]);
          }
          tc_fence_before();
          mbar_arrive_warp(ACC_EMPTY(buf));
        }
      }
      flush();
      wprof.flush(2, warp == TC_FIRST_EPI_WARP && lane == 0);
    }
  } else if (t.tma_mode) {
    // =========================== TMA-fed operand path ===========================
    if (warp == TC_FIRST_PROD_WARP + TC_PROD_WARPS - 1) {
      if (lane == 0) tma_issue_role(args, smem_a, TMA_FULL(0), A_EMPTY(0), first_tile, tile_step, end_tile);
    } else if (warp < first_prod_warp + t.prod_warps) {
      const int ptid = threadIdx.x - first_prod_warp * 32;
      const bool has_resid = p.resid != nullptr;
#define NST_XFORM(R, M) do { if (p.side_out != nullptr) transform_act16<R, M, true>(args, smem_a, coeff, TMA_FULL(0), A_FULL(0), first_tile, tile_step, end_tile, ptid); \
                             else transform_act16<R, M, false>(args, smem_a, coeff, TMA_FULL(0), A_FULL(0), first_tile, tile_step, end_tile, ptid); } while (0)
      if (has_resid) {
        if (!p.pre_relu && !p.post_relu) NST_XFORM(true, 1);
        else if (!p.pre_relu) NST_XFORM(true, 2);
        else NST_XFORM(true, 0);
      } else {
        if (p.pre_relu || p.post_relu) NST_XFORM(false, 2); else NST_XFORM(false, 1);
      }
#undef NST_XFORM
    }
  } else if (warp < TC_FIRST_PROD_WARP + t.prod_warps) {
    // =========================== operand producers (per-thread gather) ===========================
    const int PT = t.prod_warps * 32;
    const int ptid = threadIdx.x - TC_FIRST_PROD_WARP * 32;
    const int planes = t.planes;
    const int pshift = planes == 4 ? 2 : (planes == 2 ? 1 : 0);
    const int g = ptid & (planes - 1);
    const int nitems = t.NPOS * planes;
    const int is_bf16 = t.is_bf16;
    const bool two = t.a_tplanes == 2;
    const bool nhwc = p.in_fmt == IN_F32_NHWC;
    const float lo1 = p.pre_relu ? 0.f : -INFINITY;
    const float lo2 = p.post_relu ? 0.f : -INFINITY;
    const float* __restrict__ srcf = (const float*)p.src;
    const bool has_resid = p.resid != nullptr;
    constexpr int U = 4;                    // items in flight per thread
    uint32_t sa = 0, pa = 0;
    int cur_n = -1;
    if (t.dbg & 16) {
    } else if (nhwc && p.act16) {
#define NST_PRODUCE4(R, M, B, S) produce_act16<R, 4, M, B, S>(p, t, smem_a, srcpix, coeff, A_FULL(0), A_EMPTY(0), err, args.prof, first_tile, tile_step, end_tile, ptid)
#define NST_PRODUCE(R, M) do { if (t.is_bf16) NST_PRODUCE4(R, 0, true, true); else if (p.side_out != nullptr) NST_PRODUCE4(R, M, false, true); else NST_PRODUCE4(R, M, false, false); } while (0)
      if (has_resid) {
        if (!p.pre_relu && !p.post_relu) NST_PRODUCE(true, 1);
        else if (!p.pre_relu) NST_PRODUCE(true, 2);
        else NST_PRODUCE(true, 0);
      } else {
        if (p.pre_relu || p.post_relu) NST_PRODUCE(false, 2); else NST_PRODUCE(false, 1);
      }
#undef NST_PRODUCE4
#undef NST_PRODUCE
    } else if (p.in_fmt == IN_U8_NHWC3 && t.pair_mode == 2 && p.stride == 1 && t.a_tplanes == 1 && !(t.dbg & 32768)) {
      // ---- first layer, uint8 frames, quad mode ----
      // One slot = (pixel q, pixel q+1) x (3 channels + pad).  With a single 8-channel plane nothing amortises a per-position
      // source table (it was ~70 % of this role's instructions), so the halo is addressed through a row table [halo_h] and a
      // column table [P] built per tile (reflection / pre-pad folded in), and 8 slots of raw bytes are in flight per thread.
      int* rowv = srcpix;
      int* colv = srcpix + 128;                       // halo_h <= 72 rows
      const float inv_p = 1.0f / (float)t.P;
      constexpr int UB = 8;
      for (int tile = first_tile; tile < end_tile; tile += tile_step) {
        const TileCoord tc = decode_tile(t, tile);
        named_bar_sync(1, PT);          // every producer is done with the previous tile's tables
        if (ptid < t.halo_h) {
          const int sy = map_coord((tc.ty0 + p.oy_off) + ptid - p.pad_t, p.Hc, p.Hs, p.map_mode, p.prepad, p.pad_mode);
          rowv[ptid] = sy < 0 ? -1 : (tc.n * p.Hs + sy) * p.Ws;
        }
        for (int j = ptid - 64; j >= 0 && j < t.P; j += PT - 64)
          colv[j] = map_coord((tc.tx0 + p.ox_off) + j - p.pad_l, p.Wc, p.Ws, p.map_mode, p.prepad, p.pad_mode);
        named_bar_sync(2, PT);
        mbar_wait_warp(A_EMPTY(sa), pa ^ 1u, err, 6);
        uint8_t* plane = smem_a + (size_t)sa * t.a_stage_bytes;
        const uint8_t* __restrict__ src8 = (const uint8_t*)p.src;
        for (int q0 = ptid; q0 < t.NPOS; q0 += UB * PT) {
          uint32_t raw[UB][2];
#pragma unroll
          for (int u = 0; u < UB; ++u) {
            const int q = q0 + u * PT;
            raw[u][0] = raw[u][1] = 0xFF000000u;                       // top byte set = structural zero
            if (q < t.NPOS) {
              const int hy = (int)(((float)q + 0.5f) * inv_p);
              const int hx = q - hy * t.P;
              const int rb = hy < t.halo_h ? rowv[hy] : -1;
#pragma unroll
              for (int j = 0; j < 2; ++j) {
                const int cx = (hx + j < t.P) ? colv[hx + j] : -1;
                if (rb >= 0 && cx >= 0) {
                  const uint8_t* s8 = src8 + (size_t)(rb + cx) * 3;
                  raw[u][j] = (uint32_t)__ldg(s8) | ((uint32_t)__ldg(s8 + 1) << 8) | ((uint32_t)__ldg(s8 + 2) << 16);
                }
              }
            }
          }
#pragma unroll
          for (int u = 0; u < UB; ++u) {
            const int q = q0 + u * PT;
            if (q >= t.NPOS) continue;
            uint32_t hw[4];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              const uint32_t r = raw[u][j];
              const float b0 = (float)(r & 0xFFu), b1 = (float)((r >> 8) & 0xFFu), b2 = (float)((r >> 16) & 0xFFu);
              const bool live = (r >> 24) == 0u;
              const float x0 = (p.swap_rb ? b2 : b0) * p.in_scale[0] + p.in_shift[0];
              const float x1 = b1 * p.in_scale[1] + p.in_shift[1];
              const float x2 = (p.swap_rb ? b0 : b2) * p.in_scale[2] + p.in_shift[2];
              hw[2 * j] = live ? pack2(x0, x1, is_bf16) : 0u;
              hw[2 * j + 1] = live ? pack2(x2, 0.f, is_bf16) : 0u;
            }
            *reinterpret_cast<uint4*>(plane + (size_t)q * 16) = make_uint4(hw[0], hw[1], hw[2], hw[3]);
          }
        }
        fence_proxy_async_smem();
        mbar_arrive_warp(A_FULL(sa));
        if (++sa == (uint32_t)t.a_stages) { sa = 0; pa ^= 1u; }
      }
    } else
    for (int tile = first_tile; tile < end_tile; tile += tile_step) {
      const TileCoord tc = decode_tile(t, tile);
      named_bar_sync(1, PT);          // every producer is done with the previous tile's tables
      if (nhwc && tc.n != cur_n) {
        for (int c = ptid; c < p.Cin; c += PT) {
          float s_, h_;
          norm_coeff(p, tc.n, c, s_, h_);
          coeff[c] = make_float2(s_, h_);
        }
        cur_n = tc.n;
      }
      for (int q = ptid; q < t.NPOS; q += PT) srcpix[q] = locate(p, t, tc, q);
      named_bar_sync(2, PT);
      for (int ch = 0; ch < t.nchunks; ++ch) {
        const int c0 = ch * t.CK + g * 8;
        float sc[8], sh[8];
        if (nhwc) {
#pragma unroll
          for (int i = 0; i < 8; ++i) { const float2 cf = coeff[c0 + i]; sc[i] = cf.x; sh[i] = cf.y; }
        }
        mbar_wait_warp(A_EMPTY(sa), pa ^ 1u, err, 6);
        uint8_t* stage = smem_a + (size_t)sa * t.a_stage_bytes;
        uint8_t* plane_hi = stage + (size_t)g * t.a_plane_bytes;
        uint8_t* plane_lo = plane_hi + t.a_term_bytes;
        const size_t plane_base = ((size_t)tc.n * (p.Cin >> 3) + (size_t)(ch * planes + g)) * ((size_t)p.Hs * p.Ws);   // planar-8 (nhwc sources)

        for (int idx0 = ptid; idx0 < nitems; idx0 += U * PT) {
          int ent[U];
          float x[U][8];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int idx = idx0 + u * PT;
            ent[u] = idx < nitems ? srcpix[idx >> pshift] : -2;       // -2: nothing to do
          }
          if (nhwc) {
            float4 a[U][2], r[U][2];
#pragma unroll
            for (int u = 0; u < U; ++u) {
              a[u][0] = a[u][1] = r[u][0] = r[u][1] = make_float4(0.f, 0.f, 0.f, 0.f);
              if (ent[u] >= 0) {
                const size_t off = (plane_base + (size_t)(ent[u] & (SIDE_BIT - 1))) * 8;
                const float4* s4 = reinterpret_cast<const float4*>(srcf + off);
                a[u][0] = __ldg(s4); a[u][1] = __ldg(s4 + 1);
                if (has_resid) {
                  const float4* r4 = reinterpret_cast<const float4*>((const float*)p.resid + off);
                  r[u][0] = __ldg(r4); r[u][1] = __ldg(r4 + 1);
                }
              }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
              const float av[8] = {a[u][0].x, a[u][0].y, a[u][0].z, a[u][0].w, a[u][1].x, a[u][1].y, a[u][1].z, a[u][1].w};
              const float rv[8] = {r[u][0].x, r[u][0].y, r[u][0].z, r[u][0].w, r[u][1].x, r[u][1].y, r[u][1].z, r[u][1].w};
              const bool live = ent[u] >= 0;
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                float v = fmaxf(fmaf(av[i], sc[i], sh[i]), lo1) + rv[i];
                v = fmaxf(v, lo2);
                x[u][i] = live ? v : 0.f;
              }
              if (live && (ent[u] & SIDE_BIT)) {
                float4* d4 = reinterpret_cast<float4*>((float*)p.side_out + (plane_base + (size_t)(ent[u] & (SIDE_BIT - 1))) * 8);
                d4[0] = make_float4(x[u][0], x[u][1], x[u][2], x[u][3]);
                d4[1] = make_float4(x[u][4], x[u][5], x[u][6], x[u][7]);
              }
            }
          } else {
#pragma unroll
            for (int u = 0; u < U; ++u) {
#pragma unroll
              for (int i = 0; i < 8; ++i) x[u][i] = 0.f;
              // pair mode: the slot holds one pixel (3 of 8 K slots); quad mode: this pixel and its right neighbour, 4 K slots each
              const int npx = t.pair_mode == 2 ? 2 : 1;
              for (int j = 0; j < npx; ++j) {
                int e = ent[u];
                if (j == 1) { const int q1 = idx0 + u * PT + 1; e = (ent[u] != -2 && q1 < t.NPOS) ? srcpix[q1] : -1; }
                if (e < 0) continue;
                float* xo = &x[u][4 * j];
                if (p.in_fmt == IN_F32_NCHW3) {
                  // entry = sy*Ws+sx  ->  plane-separated address
                  const size_t hw = (size_t)p.Hs * p.Ws;
                  const float* s = srcf + ((size_t)tc.n * 3 * hw + e);
                  xo[0] = __ldg(s); xo[1] = __ldg(s + hw); xo[2] = __ldg(s + 2 * hw);
                } else {
                  const uint8_t* s = (const uint8_t*)p.src + ((size_t)tc.n * p.Hs * p.Ws + e) * 3;
                  const float b0 = (float)__ldg(s), b1 = (float)__ldg(s + 1), b2 = (float)__ldg(s + 2);
                  xo[0] = (p.swap_rb ? b2 : b0) * p.in_scale[0] + p.in_shift[0];
                  xo[1] = b1 * p.in_scale[1] + p.in_shift[1];
                  xo[2] = (p.swap_rb ? b0 : b2) * p.in_scale[2] + p.in_shift[2];
                }
              }
            }
          }
#pragma unroll
          for (int u = 0; u < U; ++u) {
            if (ent[u] == -2) continue;
            const int q = (idx0 + u * PT) >> pshift;
            uint32_t hi[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) hi[i] = pack2(x[u][2 * i], x[u][2 * i + 1], is_bf16);
            *reinterpret_cast<uint4*>(plane_hi + (size_t)q * 16) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            if (two) {
              uint32_t lo[4];
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const float2 back = unpack2(hi[i], is_bf16);
                lo[i] = pack2(x[u][2 * i] - back.x, x[u][2 * i + 1] - back.y, is_bf16);
              }
              *reinterpret_cast<uint4*>(plane_lo + (size_t)q * 16) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
            }
          }
        }
        fence_proxy_async_smem();       // generic-proxy smem writes -> visible to tcgen05.mma (async proxy)
        mbar_arrive_warp(A_FULL(sa));
        if (++sa == (uint32_t)t.a_stages) { sa = 0; pa ^= 1u; }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
  if (args.prof != nullptr && threadIdx.x == 0) {
    if (blockIdx.x == 0) {                                                  // debug: effective SM clock of this launch = cycles / ns
      atomicAdd(args.prof + 14, (unsigned long long)(clock64() - prof_c0));
      atomicAdd(args.prof + 15, (unsigned long long)(global_ns() - prof_t0));
    }
    if (args.prof_cta != nullptr) {                                         // per-CTA wall-clock span (launch skew / tail analysis)
      args.prof_cta[2 * blockIdx.x] = prof_t0;
      args.prof_cta[2 * blockIdx.x + 1] = global_ns();
    }
  }
}

// Per-device opt-in to > 48 KB dynamic shared memory (call once per device, from nst_create).
cudaError_t conv_tc_init() {
  return cudaFuncSetAttribute(conv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_LIMIT);
}

int conv_tc_grid(const TcPlan& t, int num_tiles, int num_sms) {
  (void)t;
  return num_tiles < num_sms ? num_tiles : num_sms;
}

cudaError_t launch_conv_tc(const TcArgs& a, int num_sms, cudaStream_t stream) {
  const int grid = conv_tc_grid(a.t, a.t.num_tiles, num_sms);
  conv_tc_kernel<<<grid, TC_THREADS, a.t.smem_bytes, stream>>>(a);
  return cudaGetLastError();
}

}  // namespace nst
