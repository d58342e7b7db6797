// Microbenchmark + correctness probe of tcgen05.mma operand layouts on sm_100a.
//
// Questions it answers (results recorded in DESIGN.md):
//   1. cycles per M=128 x N x K=16 MMA for SWIZZLE_NONE vs SWIZZLE_128B K-major smem operands
//   2. the cost of a descriptor start address that is not 128-byte aligned (SWIZZLE_NONE tap shift)
//   3. whether a SWIZZLE_128B A operand may start at an arbitrary ROW (start += r*128 B, with the
//      descriptor's base_offset field) and still read rows r..r+127 correctly -- the property the
//      implicit-GEMM conv needs to alias one halo tile for all k*k filter taps.
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/umma_bench tools/umma_bench.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(c) : "memory"); }
__device__ __forceinline__ uint32_t mbar_try(uint32_t bar, uint32_t ph) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(ph) : "memory");
  return ok;
}
__device__ __forceinline__ void umma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void umma_ts(uint32_t d, uint32_t a_tmem, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a_tmem), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void commit(uint32_t bar) { asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory"); }

struct Cfg {
  int layout;        // 0 none, 2 sw128
  int N;
  uint32_t a_off;    // byte offset of the A start address from a 1024-aligned base
  uint32_t a_lbo, a_sbo, b_lbo, b_sbo;
  uint32_t a_kstep, b_kstep;   // bytes added to the start address per K=16 step
  int ksteps;        // K steps cycled through
  int nacc;          // accumulators alternated (M blocks)
  uint32_t a_mstep;  // bytes between M blocks
  int base_off_mode; // 0: base_offset = 0, 1: (addr>>7)&7
  int iters;
  int M;             // 64 or 128
  int a_tmem;        // 1: A operand read from TMEM (TS mode)
};

__device__ __forceinline__ uint64_t mk_desc(uint32_t addr, uint32_t lbo, uint32_t sbo, int layout, int base_off_mode) {
  uint64_t d = (uint64_t)((addr >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46);
  if (base_off_mode) d |= (uint64_t)((addr >> 7) & 7u) << 49;
  d |= (uint64_t)layout << 61;
  return d;
}

__global__ void __launch_bounds__(128, 1) bench_kernel(Cfg c, const __half* a_init, const __half* b_init, int a_bytes, int b_bytes,
                                                       unsigned long long* cyc, float* d_out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sa = smem;                 // 96 KB
  uint8_t* sb = smem + 96 * 1024;     // 64 KB
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 160 * 1024);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 2);
  for (int i = threadIdx.x; i < a_bytes / 16; i += blockDim.x) reinterpret_cast<uint4*>(sa)[i] = reinterpret_cast<const uint4*>(a_init)[i];
  for (int i = threadIdx.x; i < b_bytes / 16; i += blockDim.x) reinterpret_cast<uint4*>(sb)[i] = reinterpret_cast<const uint4*>(b_init)[i];
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (threadIdx.x == 0) { mbar_init(smem_u32(bar), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *slot;
  const uint32_t idesc = (1u << 4) | ((uint32_t)(c.N >> 3) << 17) | ((uint32_t)(c.M >> 4) << 24);
  if (threadIdx.x == 0) {
    const uint32_t a0 = smem_u32(sa) + c.a_off, b0 = smem_u32(sb);
    long long t0 = clock64();
    for (int it = 0; it < c.iters; ++it) {
      for (int k = 0; k < c.ksteps; ++k) {
        const uint64_t bd = mk_desc(b0 + k * c.b_kstep, c.b_lbo, c.b_sbo, c.layout, 0);
        for (int m = 0; m < c.nacc; ++m) {
          const uint64_t ad = mk_desc(a0 + k * c.a_kstep + m * c.a_mstep, c.a_lbo, c.a_sbo, c.layout, c.base_off_mode);
          if (c.a_tmem) umma_ts(tmem + m * c.N, tmem + 256 + k * 8, bd, idesc, (it | k) != 0);
          else umma(tmem + m * c.N, ad, bd, idesc, (it | k) != 0);
        }
      }
    }
    long long ti = clock64();
    cyc[gridDim.x + blockIdx.x] = (unsigned long long)(ti - t0);
    commit(smem_u32(bar));
    while (!mbar_try(smem_u32(bar), 0)) {}
    long long t1 = clock64();
    cyc[blockIdx.x] = (unsigned long long)(t1 - t0);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (d_out != nullptr && blockIdx.x == 0) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int j = 0; j < c.N; j += 16) {
      uint32_t r[16];
      const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + j;
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                   : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
                     "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                   : "r"(taddr) : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      for (int i = 0; i < 16; ++i) d_out[(warp * 32 + lane) * c.N + j + i] = __uint_as_float(r[i]);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
  }
}

// ---- issue-loop variants: how the MMA-issuing code is written matters more than the MMA itself ----
__device__ __forceinline__ uint32_t elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred;
}
template <int VARIANT, int N, int NACC>
__global__ void __launch_bounds__(128, 1) issue_kernel(int iters, unsigned long long* cyc) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sa = smem;
  uint8_t* sb = smem + 96 * 1024;
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 160 * 1024);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 2);
  for (int i = threadIdx.x; i < 160 * 1024 / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0x3c003c00u, 0x38003800u, 0x3c003c00u, 0x34003400u);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (threadIdx.x == 0) { mbar_init(smem_u32(bar), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *slot;
  constexpr uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  // SWIZZLE_128B K-major descriptors: hi word is constant, lo word = (addr >> 4) | (LBO>>4 << 16)
  constexpr uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
  if (threadIdx.x < 32) {
    const uint32_t a_lo = ((smem_u32(sa) >> 4) & 0x3FFFu) | (1u << 16);
    const uint32_t b_lo = ((smem_u32(sb) >> 4) & 0x3FFFu) | (1u << 16);
    long long t0 = clock64();
    if (VARIANT == 0) {
      if (threadIdx.x == 0) {
        for (int it = 0; it < iters; ++it) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
#pragma unroll
            for (int m = 0; m < NACC; ++m)
              umma(tmem + m * N, ((uint64_t)desc_hi << 32) | (a_lo + k * 2 + m * 1024), ((uint64_t)desc_hi << 32) | (b_lo + k * 2), idesc, (it | k) != 0);
          }
        }
      }
    } else if (VARIANT == 1) {
      if (elect_one()) {
        for (int it = 0; it < iters; ++it) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
#pragma unroll
            for (int m = 0; m < NACC; ++m)
              umma(tmem + m * N, ((uint64_t)desc_hi << 32) | (a_lo + k * 2 + m * 1024), ((uint64_t)desc_hi << 32) | (b_lo + k * 2), idesc, (it | k) != 0);
          }
        }
      }
    } else if (VARIANT == 3 || VARIANT == 4) {
      // SWIZZLE_NONE K-major: LBO = plane stride (1282*16 B), SBO = 128 B, K step = 2 planes; VARIANT 4 starts 560 B off alignment
      constexpr uint32_t PL = 1282u * 16u;
      constexpr uint32_t hi_a = (128u >> 4) | (1u << 14);
      constexpr uint32_t hi_b = (128u >> 4) | (1u << 14);
      const uint32_t a_lo2 = (((smem_u32(sa) + (VARIANT == 4 ? 560u : 0u)) >> 4) & 0x3FFFu) | ((PL >> 4) << 16);
      const uint32_t b_lo2 = ((smem_u32(sb) >> 4) & 0x3FFFu) | (((uint32_t)N * 16u >> 4) << 16);
      if (elect_one()) {
        for (int it = 0; it < iters; ++it) {
#pragma unroll
          for (int k = 0; k < 2; ++k) {
#pragma unroll
            for (int m = 0; m < 2 * NACC; ++m)
              umma(tmem + (m % NACC) * N, ((uint64_t)hi_a << 32) | (a_lo2 + k * (2 * PL >> 4) + m * 128), ((uint64_t)hi_b << 32) | (b_lo2 + k * (2 * N * 16 >> 4)), idesc, (it | k | m) != 0);
          }
        }
      }
    } else {
      for (int it = 0; it < iters; ++it) {
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
#pragma unroll
            for (int m = 0; m < NACC; ++m)
              umma(tmem + m * N, ((uint64_t)desc_hi << 32) | (a_lo + k * 2 + m * 1024), ((uint64_t)desc_hi << 32) | (b_lo + k * 2), idesc, (it | k) != 0);
          }
        }
        __syncwarp();
      }
    }
    long long ti = clock64();
    if (elect_one()) commit(smem_u32(bar));
    __syncwarp();
    while (!mbar_try(smem_u32(bar), 0)) {}
    long long t1 = clock64();
    if (threadIdx.x == 0) { cyc[blockIdx.x] = (unsigned long long)(t1 - t0); cyc[gridDim.x + blockIdx.x] = (unsigned long long)(ti - t0); }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
  }
}

template <int VARIANT, int N, int NACC>
static void run_issue(unsigned long long* dcyc, int SMEM) {
  CK(cudaFuncSetAttribute(issue_kernel<VARIANT, N, NACC>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
  const int grid = 148, iters = 300;
  for (int rep = 0; rep < 2; ++rep) { issue_kernel<VARIANT, N, NACC><<<grid, 128, SMEM>>>(iters, dcyc); CK(cudaDeviceSynchronize()); }
  std::vector<unsigned long long> h(2 * grid);
  CK(cudaMemcpy(h.data(), dcyc, 16 * grid, cudaMemcpyDeviceToHost));
  double mx = 0, mi = 0; for (int i = 0; i < grid; ++i) { mx = h[i] > mx ? h[i] : mx; mi = h[grid + i] > mi ? h[grid + i] : mi; }
  const double n = (double)iters * 4 * NACC;
  const char* names[] = {"if (tid==0) {loop}", "if (elect) {loop}", "loop { if (elect) {4 MMA} syncwarp }", "SWIZZLE_NONE aligned, if (elect) {loop}", "SWIZZLE_NONE +560B, if (elect) {loop}"};
  printf("issue variant %-38s N=%3d nacc=%d  %.1f cyc/MMA (issue %.1f) floor %d\n", names[VARIANT], N, NACC, mx / n, mi / n, 128 * N / 256);
}

static float frand() { return (float)((rand() % 17) - 8) / 8.0f; }

int main() {
  CK(cudaSetDevice(0));
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  const int SMEM = 161 * 1024;
  CK(cudaFuncSetAttribute(bench_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
  const int A_BYTES = 96 * 1024, B_BYTES = 64 * 1024;
  std::vector<__half> ha(A_BYTES / 2), hb(B_BYTES / 2);
  __half *da, *db; unsigned long long* dcyc; float* dd;
  CK(cudaMalloc(&da, A_BYTES)); CK(cudaMalloc(&db, B_BYTES)); CK(cudaMalloc(&dcyc, 16 * 256)); CK(cudaMalloc(&dd, 128 * 256 * 4));
  int sm_clock = 0; CK(cudaDeviceGetAttribute(&sm_clock, cudaDevAttrClockRate, 0));
  printf("device %s, %d SMs, clock %d kHz\n", prop.name, prop.multiProcessorCount, sm_clock);

  run_issue<0, 64, 1>(dcyc, SMEM); run_issue<1, 64, 1>(dcyc, SMEM); run_issue<2, 64, 1>(dcyc, SMEM);
  run_issue<0, 64, 2>(dcyc, SMEM); run_issue<1, 64, 2>(dcyc, SMEM); run_issue<2, 64, 2>(dcyc, SMEM);
  run_issue<1, 16, 1>(dcyc, SMEM); run_issue<1, 32, 1>(dcyc, SMEM); run_issue<1, 128, 1>(dcyc, SMEM); run_issue<1, 256, 1>(dcyc, SMEM);
  run_issue<2, 16, 1>(dcyc, SMEM); run_issue<2, 32, 1>(dcyc, SMEM); run_issue<2, 128, 1>(dcyc, SMEM); run_issue<2, 256, 1>(dcyc, SMEM);
  run_issue<1, 128, 2>(dcyc, SMEM); run_issue<1, 256, 2>(dcyc, SMEM); run_issue<2, 128, 2>(dcyc, SMEM); run_issue<2, 256, 2>(dcyc, SMEM);
  run_issue<3, 16, 1>(dcyc, SMEM); run_issue<3, 32, 1>(dcyc, SMEM); run_issue<3, 64, 1>(dcyc, SMEM); run_issue<3, 128, 1>(dcyc, SMEM); run_issue<3, 128, 2>(dcyc, SMEM); run_issue<3, 256, 1>(dcyc, SMEM);
  run_issue<4, 16, 1>(dcyc, SMEM); run_issue<4, 32, 1>(dcyc, SMEM); run_issue<4, 64, 1>(dcyc, SMEM); run_issue<4, 128, 1>(dcyc, SMEM); run_issue<4, 128, 2>(dcyc, SMEM); run_issue<4, 256, 1>(dcyc, SMEM);
  return 0;
  // ---------------- throughput ----------------
  srand(1);
  for (auto& v : ha) v = __float2half(frand());
  for (auto& v : hb) v = __float2half(frand());
  CK(cudaMemcpy(da, ha.data(), A_BYTES, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(db, hb.data(), B_BYTES, cudaMemcpyHostToDevice));
  auto run = [&](const char* name, Cfg c, int grid) {
    bench_kernel<<<grid, 128, SMEM>>>(c, da, db, A_BYTES, B_BYTES, dcyc, nullptr);
    CK(cudaDeviceSynchronize());
    bench_kernel<<<grid, 128, SMEM>>>(c, da, db, A_BYTES, B_BYTES, dcyc, nullptr);
    CK(cudaDeviceSynchronize());
    std::vector<unsigned long long> h(grid);
    CK(cudaMemcpy(h.data(), dcyc, 8 * grid, cudaMemcpyDeviceToHost));
    double mx = 0; for (auto v : h) mx = v > mx ? v : mx;
    const double n = (double)c.iters * c.ksteps * c.nacc;
    printf("%-58s N=%3d grid=%3d  %.1f cyc/MMA  (floor %d)\n", name, c.N, grid, mx / n, 128 * c.N / 256);
  };
  auto run2 = [&](const char* name, Cfg c, int grid) {
    for (int rep = 0; rep < 2; ++rep) { bench_kernel<<<grid, 128, SMEM>>>(c, da, db, A_BYTES, B_BYTES, dcyc, nullptr); CK(cudaDeviceSynchronize()); }
    std::vector<unsigned long long> h(2 * grid);
    CK(cudaMemcpy(h.data(), dcyc, 16 * grid, cudaMemcpyDeviceToHost));
    double mx = 0, mi = 0; for (int i = 0; i < grid; ++i) { mx = h[i] > mx ? h[i] : mx; mi = h[grid + i] > mi ? h[grid + i] : mi; }
    const double n = (double)c.iters * c.ksteps * c.nacc;
    printf("%-28s M=%3d N=%3d nacc=%d ksteps=%d iters=%4d grid=%3d  %.1f cyc/MMA (issue %.1f) floor %d\n", name, c.M, c.N, c.nacc, c.ksteps, c.iters, grid, mx / n, mi / n,
           128 * c.N / 256);
  };
  for (int grid : {1, 148}) {
    for (int M : {128, 64}) for (int N : {16, 64, 128, 256}) for (int nacc : {1, 2, 4}) for (int ksteps : {1, 4}) for (int iters : {100, 400}) {
      if (nacc * N > 256 && !(nacc == 2 && N == 256)) continue;
      if (iters == 400 && !(N == 256 || N == 64)) continue;
      Cfg c; memset(&c, 0, sizeof(c));
      c.M = M; c.N = N; c.iters = iters; c.nacc = nacc;
      c.layout = 2; c.a_lbo = 16; c.a_sbo = 1024; c.b_lbo = 16; c.b_sbo = 1024;
      c.a_kstep = 32; c.b_kstep = 32; c.ksteps = ksteps; c.a_mstep = 128 * 128; c.a_off = 0; c.base_off_mode = 0;
      run2("sw128 SS", c, grid);
      if (nacc * N <= 256 && ksteps == 4 && iters == 100) { c.a_tmem = 1; run2("sw128 TS (A in TMEM)", c, grid); }
    }
  }
  // ---------------- correctness of row-shifted SWIZZLE_128B A ----------------
  // logical A_full[pos][k], pos in [0, 384), k in [0,64): phys = pos*128 + ((k/8) ^ (pos%8))*16 + (k%8)*2
  // logical B[n][k], n in [0,N): same swizzle.
  for (int mode = 0; mode < 2; ++mode) {
    for (int shift : {0, 1, 3, 8, 35}) {
      const int N = 64;
      std::vector<float> fa(384 * 64), fb(N * 64);
      for (auto& v : fa) v = frand();
      for (auto& v : fb) v = frand();
      std::fill(ha.begin(), ha.end(), __float2half(0.f));
      std::fill(hb.begin(), hb.end(), __float2half(0.f));
      for (int p = 0; p < 384; ++p) for (int k = 0; k < 64; ++k) ha[(p * 128 + (((k / 8) ^ (p % 8)) * 16) + (k % 8) * 2) / 2] = __float2half(fa[p * 64 + k]);
      for (int n = 0; n < N; ++n) for (int k = 0; k < 64; ++k) hb[(n * 128 + (((k / 8) ^ (n % 8)) * 16) + (k % 8) * 2) / 2] = __float2half(fb[n * 64 + k]);
      CK(cudaMemcpy(da, ha.data(), A_BYTES, cudaMemcpyHostToDevice));
      CK(cudaMemcpy(db, hb.data(), B_BYTES, cudaMemcpyHostToDevice));
      Cfg c; memset(&c, 0, sizeof(c));
      c.M = 128; c.N = N; c.iters = 1; c.nacc = 1; c.layout = 2; c.a_lbo = 16; c.a_sbo = 1024; c.b_lbo = 16; c.b_sbo = 1024;
      c.a_kstep = 32; c.b_kstep = 32; c.ksteps = 4; c.a_off = 128 * shift; c.base_off_mode = mode;
      CK(cudaMemset(dd, 0, 128 * 256 * 4));
      bench_kernel<<<1, 128, SMEM>>>(c, da, db, A_BYTES, B_BYTES, dcyc, dd);
      CK(cudaDeviceSynchronize());
      std::vector<float> hd(128 * N);
      CK(cudaMemcpy(hd.data(), dd, 128 * N * 4, cudaMemcpyDeviceToHost));
      double maxerr = 0;
      for (int m = 0; m < 128; ++m) for (int n = 0; n < N; ++n) {
        double ref = 0;
        for (int k = 0; k < 64; ++k) ref += (double)fa[(m + shift) * 64 + k] * fb[n * 64 + k];
        const double e = fabs(ref - hd[m * N + n]);
        if (e > maxerr) maxerr = e;
      }
      printf("sw128 row-shift correctness: shift=%2d rows, base_offset %s -> max|err| = %.4f %s\n", shift, mode ? "set" : "0  ", maxerr,
             maxerr < 1e-2 ? "OK" : "MISMATCH");
    }
  }
  return 0;
}
