// tcgen05 / TMEM implicit-GEMM convolution for sm_100a -- the hot kernel of the stylisation path.
//
// Replaces, per layer, the reference's  [F.interpolate] -> ReflectionPad2d / zero pad -> Conv2d
// (transformer_net.py:44-54,79-99; transformer_net_nst.py:12-59; model.py:5-15) plus the
// InstanceNorm/ReLU/residual ops around it (transformer_net.py:30-39,71-76; model.py:30-35,57-60).
//
// One persistent CTA per SM, 14 warps, warp-specialised:
//   warp 0      weight loader : cp.async.bulk (TMA bulk copy, UBLKCP) of pre-packed fp16 weight
//                               tiles global -> shared, mbarrier complete_tx ring
//   warp 1      MMA issuer    : one thread issues tcgen05.mma (M=128, N=Cout, K=16) for every
//                               (channel chunk, filter tap, K-step, precision term, M block);
//                               accumulators in TMEM (double buffered across tiles)
//   warps 2-5   epilogue      : tcgen05.ld TMEM -> registers, store raw fp32 NHWC output,
//                               per-(n,c) sum / sum-of-squares by warp-shuffle butterfly,
//                               or bias/tanh/affine/clamp/blend/uint8 pack for the image layer
//   warps 6-13  operand producers : gather the input halo tile from HBM with reflection /
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
__device__ __forceinline__ float2 unpack2(uint32_t u, int is_bf16) {
  if (is_bf16) return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u));
  return __half22float2(*reinterpret_cast<__half2*>(&u));
}

struct TileCoord { int n, ty0, tx0; };
__device__ __forceinline__ TileCoord decode_tile(const TcPlan& t, int tile) {
  const int per_img = t.tiles_x * t.tiles_y;
  TileCoord c;
  c.n = tile / per_img;
  const int r = tile - c.n * per_img;
  const int tyi = r / t.tiles_x;
  c.ty0 = tyi * t.TH;
  c.tx0 = (r - tyi * t.tiles_x) * t.TW;
  return c;
}

// ---------------------------------------------------------------------------------------------
// operand producer: one (halo position, 8-channel group) item
// ---------------------------------------------------------------------------------------------
// Source-table entry of one halo slot: -1 = structural zero / filler, else the source pixel index
// (n*Hs+sy)*Ws+sx (< 2^30), with bit 30 set when this tile owns the materialised ("side") value.
constexpr int SIDE_BIT = 1 << 30;

__device__ __forceinline__ int locate(const ConvParams& p, const TcPlan& t, const TileCoord& tc, int q) {
  int hy, hx;
  bool filler = false;
  if (p.stride == 1) {
    hy = q / t.P;
    hx = q - hy * t.P;
    filler = hy >= t.halo_h;
  } else {
    const int sub = q / t.SUBPOS;
    const int rem = q - sub * t.SUBPOS;
    const int i = rem / t.P;
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
  int e = (tc.n * p.Hs + sy) * p.Ws + sx;
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
// the kernel
// ---------------------------------------------------------------------------------------------
constexpr int NBARS = 2 * TC_MAX_A_STAGES + 2 * TC_MAX_B_STAGES + 4;

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
  int* srcpix = reinterpret_cast<int*>(tmem_slot + 4);                       // [NPOS] source table of the current tile
  float2* coeff = reinterpret_cast<float2*>(srcpix + ((t.NPOS + 3) & ~3));    // [Cin] InstanceNorm (scale, shift) of sample n

  const uint32_t bar0 = smem_u32(bars);
  auto A_FULL = [&](int s) { return bar0 + 8u * (uint32_t)s; };
  auto A_EMPTY = [&](int s) { return bar0 + 8u * (uint32_t)(TC_MAX_A_STAGES + s); };
  auto B_FULL = [&](int s) { return bar0 + 8u * (uint32_t)(2 * TC_MAX_A_STAGES + s); };
  auto B_EMPTY = [&](int s) { return bar0 + 8u * (uint32_t)(2 * TC_MAX_A_STAGES + TC_MAX_B_STAGES + s); };
  auto ACC_FULL = [&](int s) { return bar0 + 8u * (uint32_t)(2 * TC_MAX_A_STAGES + 2 * TC_MAX_B_STAGES + s); };
  auto ACC_EMPTY = [&](int s) { return bar0 + 8u * (uint32_t)(2 * TC_MAX_A_STAGES + 2 * TC_MAX_B_STAGES + 2 + s); };

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < TC_MAX_A_STAGES; ++s) { mbar_init(A_FULL(s), TC_PROD_THREADS); mbar_init(A_EMPTY(s), 1); }
    for (int s = 0; s < TC_MAX_B_STAGES; ++s) { mbar_init(B_FULL(s), 1); mbar_init(B_EMPTY(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(ACC_FULL(s), 1); mbar_init(ACC_EMPTY(s), 128); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(smem_u32(tmem_slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int first_tile = blockIdx.x;
  const int tile_step = gridDim.x;

  if (warp == 0) {
    // =========================== weight loader ===========================
    if (lane == 0) {
      uint32_t sb = 0, pb = 0;
      const uint8_t* wp = reinterpret_cast<const uint8_t*>(args.wpack);
      const int nstage_per_tile = t.nchunks * t.ngroups;
      for (int tile = first_tile; tile < t.num_tiles; tile += tile_step) {
        for (int g = 0; g < nstage_per_tile; ++g) {
          mbar_wait(B_EMPTY(sb), pb ^ 1u, err, 1);
          mbar_arrive_expect_tx(B_FULL(sb), t.b_stage_bytes);
          bulk_g2s(smem_u32(smem_b + (size_t)sb * t.b_stage_bytes), wp + (size_t)g * t.b_stage_bytes,
                   t.b_stage_bytes, B_FULL(sb));
          if (++sb == (uint32_t)t.b_stages) { sb = 0; pb ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    // The whole warp walks the pipeline converged and one elected lane issues (the CUTLASS idiom):
    // ptxas then keeps every descriptor in uniform registers.  Measured with tools/umma_bench.cu: a
    // tcgen05.mma runs at its floor (64 cycles for M=128,N=128,K=16) only when the issuing loop costs
    // less than that; building 64-bit descriptors from scratch per MMA in divergent code cost 170-370
    // cycles per MMA.  So: descriptor high words are constants, low words are (addr>>4)|(LBO>>4)<<16
    // and every step below is one 32-bit add in 16-byte units.
    {
      const uint32_t idesc = make_idesc(t.NPAD, t.is_bf16);
      const uint32_t desc_hi = (128u >> 4) | (1u << 14);                      // SBO = 128 B, descriptor version 1
      const uint32_t a_lbo16 = (t.pair_mode ? 16u : t.a_plane_bytes) >> 4;
      const uint32_t a_lo_base = ((smem_u32(smem_a) >> 4) & 0x3FFFu) | (a_lbo16 << 16);
      const uint32_t b_lo_base = ((smem_u32(smem_b) >> 4) & 0x3FFFu) | ((uint32_t)t.NPAD << 16);   // LBO = NPAD*16 B
      const uint32_t a_kstep16 = (2u * t.a_plane_bytes) >> 4;
      const uint32_t b_kstep16 = 2u * (uint32_t)t.NPAD;
      const uint32_t a_stage16 = t.a_stage_bytes >> 4, b_stage16 = t.b_stage_bytes >> 4;
      const uint32_t a_term16 = t.a_term_bytes >> 4, b_term16 = t.b_term_bytes >> 4, b_ent16 = t.b_ent_bytes >> 4;
      const int MB = t.MB, NPAD = t.NPAD, k16 = t.k16_per_ent, nterms = t.nterms, tg = t.tg, nent = t.nent;
      uint32_t sa = 0, pa = 0, sb = 0, pb = 0, it = 0;
      for (int tile = first_tile; tile < t.num_tiles; tile += tile_step, ++it) {
        const uint32_t buf = it & 1u;
        mbar_wait(ACC_EMPTY(buf), ((it >> 1) & 1u) ^ 1u, err, 2);
        tc_fence_after();
        const uint32_t tmem_acc = tmem_base + buf * (uint32_t)(MB * NPAD);
        uint32_t acc_on = 0u;
        for (int ch = 0; ch < t.nchunks; ++ch) {
          mbar_wait(A_FULL(sa), pa, err, 3);
          tc_fence_after();
          const uint32_t a_stage_lo = a_lo_base + sa * a_stage16;
          int ent = 0;
          for (int g = 0; g < t.ngroups; ++g) {
            mbar_wait(B_FULL(sb), pb, err, 4);
            tc_fence_after();
            if (elect_one_sync()) {
              uint32_t b_lo = b_lo_base + sb * b_stage16;
              const int e_end = min(tg, nent - ent);
              for (int e = 0; e < e_end; ++e, b_lo += b_ent16) {
                const uint32_t a_lo = a_stage_lo + (uint32_t)t.ent_off[ent + e];
                if (nterms == 1) {
                  for (int j = 0; j < k16; ++j) {
                    const uint64_t bdesc = ((uint64_t)desc_hi << 32) | (b_lo + (uint32_t)j * b_kstep16);
                    const uint32_t a_j = a_lo + (uint32_t)j * a_kstep16;
                    for (int mb = 0; mb < MB; ++mb)
                      umma_f16(tmem_acc + (uint32_t)(mb * NPAD), ((uint64_t)desc_hi << 32) | (a_j + (uint32_t)mb * 128u), bdesc, idesc, acc_on);
                    acc_on = 1u;
                  }
                } else {
                  for (int j = 0; j < k16; ++j) {
                    for (int term = 0; term < 3; ++term) {                      // hi*hi, lo*hi, hi*lo
                      const uint64_t bdesc = ((uint64_t)desc_hi << 32) | (b_lo + (uint32_t)j * b_kstep16 + (term == 2 ? b_term16 : 0u));
                      const uint32_t a_j = a_lo + (uint32_t)j * a_kstep16 + (term == 1 ? a_term16 : 0u);
                      for (int mb = 0; mb < MB; ++mb)
                        umma_f16(tmem_acc + (uint32_t)(mb * NPAD), ((uint64_t)desc_hi << 32) | (a_j + (uint32_t)mb * 128u), bdesc, idesc, acc_on);
                      acc_on = 1u;
                    }
                  }
                }
              }
              umma_commit(B_EMPTY(sb));
            }
            __syncwarp();
            acc_on = 1u;
            ent += tg;
            if (++sb == (uint32_t)t.b_stages) { sb = 0; pb ^= 1u; }
          }
          if (elect_one_sync()) umma_commit(A_EMPTY(sa));
          __syncwarp();
          if (++sa == (uint32_t)t.a_stages) { sa = 0; pa ^= 1u; }
        }
        if (elect_one_sync()) umma_commit(ACC_FULL(buf));
        __syncwarp();
      }
    }
  } else if (warp < TC_FIRST_PROD_WARP) {
    // =========================== epilogue ===========================
    const int quad = warp & 3;                        // TMEM lane quadrant this warp may read
    const int njc = t.NPAD >> 4;
    float ssum[12], ssq[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) { ssum[i] = 0.f; ssq[i] = 0.f; }
    const int my_chan = ((lane & 16) ? 8 : 0) + ((lane & 8) ? 4 : 0) + ((lane & 4) ? 2 : 0) + ((lane & 2) ? 1 : 0);
    int cur_n = -1;
    auto flush = [&]() {
      if (cur_n < 0 || p.out_fmt != OUT_F32_NHWC_STATS) return;
#pragma unroll
      for (int jc = 0; jc < 12; ++jc) {
        if (jc < njc) {
          const int c = jc * 16 + my_chan;
          if ((lane & 1) == 0 && c < p.Cout) {
            double* st = p.out_stats + ((size_t)cur_n * p.Cout + c) * 2;
            atomicAdd(st, (double)ssum[jc]);
            atomicAdd(st + 1, (double)ssq[jc]);
          }
          ssum[jc] = 0.f; ssq[jc] = 0.f;
        }
      }
    };
    uint32_t it = 0;
    for (int tile = first_tile; tile < t.num_tiles; tile += tile_step, ++it) {
      const TileCoord tc = decode_tile(t, tile);
      if (tc.n != cur_n) { flush(); cur_n = tc.n; }
      const uint32_t buf = it & 1u;
      mbar_wait(ACC_FULL(buf), (it >> 1) & 1u, err, 5);
      tc_fence_after();
      for (int mb = 0; mb < t.MB; ++mb) {
        const int m = mb * 128 + quad * 32 + lane;
        const int ty = m / t.P;
        const int tx = m - ty * t.P;
        const int oy = tc.ty0 + ty, ox = tc.tx0 + tx;
        const bool valid = ty < t.TH && tx < t.TW && oy < p.Ho && ox < p.Wo;
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + buf * (uint32_t)(t.MB * t.NPAD) + (uint32_t)(mb * t.NPAD);
        if (p.out_fmt == OUT_F32_NHWC_STATS) {
          float* orow = (float*)p.out + (((size_t)tc.n * p.Ho + oy) * p.Wo + ox) * p.Cout;
#pragma unroll
          for (int jc = 0; jc < 12; ++jc) {
            if (jc < njc) {
              float v[16];
              tmem_ld16(taddr + (uint32_t)(jc * 16), v);
              if (valid) {
                float4* o4 = reinterpret_cast<float4*>(orow + jc * 16);
                o4[0] = make_float4(v[0], v[1], v[2], v[3]);
                o4[1] = make_float4(v[4], v[5], v[6], v[7]);
                o4[2] = make_float4(v[8], v[9], v[10], v[11]);
                o4[3] = make_float4(v[12], v[13], v[14], v[15]);
              }
              // per-channel sum / sum of squares over the 32 rows of this warp: butterfly that halves
              // the number of live channels per step (16 -> 8 -> 4 -> 2 -> 1), then one pair add.
              float s[16], q[16];
#pragma unroll
              for (int i = 0; i < 16; ++i) { s[i] = valid ? v[i] : 0.f; q[i] = s[i] * s[i]; }
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
              ssum[jc] += s[0] + __shfl_xor_sync(0xffffffffu, s[0], 1);
              ssq[jc] += q[0] + __shfl_xor_sync(0xffffffffu, q[0], 1);
            }
          }
        } else {
          float v[16];
          tmem_ld16(taddr, v);
          if (valid) store_image(p, tc.n, oy, ox, v[0] + p.bias[0], v[1] + p.bias[1], v[2] + p.bias[2]);
        }
      }
      tc_fence_before();
      mbar_arrive(ACC_EMPTY(buf));
    }
    flush();
  } else {
    // =========================== operand producers ===========================
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
    for (int tile = first_tile; tile < t.num_tiles; tile += tile_step) {
      const TileCoord tc = decode_tile(t, tile);
      named_bar_sync(1, TC_PROD_THREADS);          // every producer is done with the previous tile's tables
      if (nhwc && tc.n != cur_n) {
        for (int c = ptid; c < p.Cin; c += TC_PROD_THREADS) {
          float s_, h_;
          norm_coeff(p, tc.n, c, s_, h_);
          coeff[c] = make_float2(s_, h_);
        }
        cur_n = tc.n;
      }
      for (int q = ptid; q < t.NPOS; q += TC_PROD_THREADS) srcpix[q] = locate(p, t, tc, q);
      named_bar_sync(2, TC_PROD_THREADS);
      for (int ch = 0; ch < t.nchunks; ++ch) {
        const int c0 = ch * t.CK + g * 8;
        float sc[8], sh[8];
        if (nhwc) {
#pragma unroll
          for (int i = 0; i < 8; ++i) { const float2 cf = coeff[c0 + i]; sc[i] = cf.x; sh[i] = cf.y; }
        }
        mbar_wait(A_EMPTY(sa), pa ^ 1u, err, 6);
        uint8_t* stage = smem_a + (size_t)sa * t.a_stage_bytes;
        uint8_t* plane_hi = stage + (size_t)g * t.a_plane_bytes;
        uint8_t* plane_lo = plane_hi + t.a_term_bytes;

        for (int idx0 = ptid; idx0 < nitems; idx0 += U * TC_PROD_THREADS) {
          int ent[U];
          float x[U][8];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int idx = idx0 + u * TC_PROD_THREADS;
            ent[u] = idx < nitems ? srcpix[idx >> pshift] : -2;       // -2: nothing to do
          }
          if (nhwc) {
            float4 a[U][2], r[U][2];
#pragma unroll
            for (int u = 0; u < U; ++u) {
              a[u][0] = a[u][1] = r[u][0] = r[u][1] = make_float4(0.f, 0.f, 0.f, 0.f);
              if (ent[u] >= 0) {
                const size_t off = (size_t)(ent[u] & (SIDE_BIT - 1)) * p.Cin + c0;
                const float4* s4 = reinterpret_cast<const float4*>(srcf + off);
                a[u][0] = __ldg(s4); a[u][1] = __ldg(s4 + 1);
                if (has_resid) {
                  const float4* r4 = reinterpret_cast<const float4*>(p.resid + off);
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
                float4* d4 = reinterpret_cast<float4*>(p.side_out + (size_t)(ent[u] & (SIDE_BIT - 1)) * p.Cin + c0);
                d4[0] = make_float4(x[u][0], x[u][1], x[u][2], x[u][3]);
                d4[1] = make_float4(x[u][4], x[u][5], x[u][6], x[u][7]);
              }
            }
          } else {
#pragma unroll
            for (int u = 0; u < U; ++u) {
#pragma unroll
              for (int i = 0; i < 8; ++i) x[u][i] = 0.f;
              if (ent[u] >= 0) {
                if (p.in_fmt == IN_F32_NCHW3) {
                  // entry = (n*Hs+sy)*Ws+sx  ->  plane-separated address
                  const int hw = p.Hs * p.Ws;
                  const int n = ent[u] / hw;
                  const int rem = ent[u] - n * hw;
                  const float* s = srcf + ((size_t)n * 3 * hw + rem);
                  x[u][0] = __ldg(s); x[u][1] = __ldg(s + hw); x[u][2] = __ldg(s + 2 * (size_t)hw);
                } else {
                  const uint8_t* s = (const uint8_t*)p.src + (size_t)ent[u] * 3;
                  const float b0 = (float)__ldg(s), b1 = (float)__ldg(s + 1), b2 = (float)__ldg(s + 2);
                  x[u][0] = (p.swap_rb ? b2 : b0) * p.in_scale[0] + p.in_shift[0];
                  x[u][1] = b1 * p.in_scale[1] + p.in_shift[1];
                  x[u][2] = (p.swap_rb ? b0 : b2) * p.in_scale[2] + p.in_shift[2];
                }
              }
            }
          }
#pragma unroll
          for (int u = 0; u < U; ++u) {
            if (ent[u] == -2) continue;
            const int q = (idx0 + u * TC_PROD_THREADS) >> pshift;
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
        mbar_arrive(A_FULL(sa));
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
}

// Per-device opt-in to > 48 KB dynamic shared memory (call once per device, from nst_create).
cudaError_t conv_tc_init() {
  return cudaFuncSetAttribute(conv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_LIMIT);
}

cudaError_t launch_conv_tc(const TcArgs& a, int num_sms, cudaStream_t stream) {
  const int grid = a.t.num_tiles < num_sms ? a.t.num_tiles : num_sms;
  conv_tc_kernel<<<grid, TC_THREADS, a.t.smem_bytes, stream>>>(a);
  return cudaGetLastError();
}

}  // namespace nst
