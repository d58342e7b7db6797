// Host/device shared description of the tcgen05 implicit-GEMM conv tiling (see conv_tc.cu).
#pragma once
#include <cuda.h>            // CUtensorMap (type only; the encoder is resolved at run time through cudaGetDriverEntryPoint)
#include "nst_kernels.cuh"

namespace nst {

constexpr int TC_MAX_ENT = 81;          // tap entries per channel chunk (9x9 conv)
constexpr int TC_PROD_WARPS = 10;        // launched; a layer uses TcPlan::prod_warps (8 or 10) of them.  TMA-fed layers: at most 9 transform
                                         // warps, the last launched warp issues the activation TMA boxes
constexpr int TC_PROD_THREADS = TC_PROD_WARPS * 32;
constexpr int TC_FIRST_EPI_WARP = 2;
constexpr int TC_FIRST_PROD_WARP = 6;
constexpr int TC_THREADS = (TC_FIRST_PROD_WARP + TC_PROD_WARPS) * 32;   // 512: 16 warps x 128 registers = the whole register file
constexpr int TC_MAX_A_STAGES = 4;
constexpr int TC_MAX_B_STAGES = 8;
constexpr int TC_SMEM_LIMIT = 227 * 1024;
constexpr int TC_FOLD_PITCH = 29;      // floats per ring row (27 used; odd pitch = conflict-free)
constexpr int TC_FOLD_ROWS = 136;      // one M block + the 8 rows of the previous block its first outputs still read
constexpr int TC_FOLD_RING_BYTES = TC_FOLD_ROWS * TC_FOLD_PITCH * 4;

// GEMM view of one layer:  D[m, co] = sum_{entry e, k} A[m + ent_off[e], k] * W[e][k][co]
//   m      : flat position inside the CTA tile's halo buffer (pitch P), M = MB*128 rows
//   entry  : one filter tap (or a horizontal PAIR of taps for the 3-channel first layer)
// A lives in shared memory as [term][plane = 8-channel group][position][8 x f16] (UMMA no-swizzle
// K-major core matrices: 8 rows x 16 bytes, SBO = 128 B, LBO = plane stride), so a tap shift is a
// plain byte offset on the descriptor start address.
struct TcPlan {
  int TH, TW, P, MB;
  int NPOS;                 // positions per plane
  int SUBPOS;               // stride-2 layers: positions per parity sub-plane
  int halo_h, halo_w;       // input halo extent in conv-padded space
  int tiles_x, tiles_y, num_tiles;     // num_tiles includes the batch
  int CK, nchunks, planes;  // channels per chunk, chunks, 8-channel planes per chunk
  int nent;                 // tap entries per chunk
  int k16_per_ent;          // MMA K-steps per entry
  int pair_mode;            // first layer: K = 2 adjacent pixels x 8 (3 real) channels
  int NPAD;                 // UMMA N (Cout padded to a multiple of 16)
  int tg, ngroups;          // entries per weight stage, stages per chunk
  int a_stages, b_stages;
  int prod_warps;           // operand-producer warps that take part (the rest of TC_PROD_WARPS exit at once)
  int rr_tiles;             // CTA b takes tiles b, b + grid, b + 2*grid, ... instead of one contiguous run
  int rr_rot;               // ... with the positions of full round i rotated by i * rr_rot (chosen on the host to balance border tiles)
  int epi_groups;           // 1, or 2 on epilogue-bound TMA-fed layers: warps 6-9 form a second epilogue group that drains the odd tiles of
                            // the CTA's run (accumulator buffer 1) while warps 2-5 drain the even ones; the transform then has <= 5 warps
  uint32_t epi_scratch_bytes;   // shared-memory scratch of ONE epilogue group (fold ring, or per-warp double channel statistics)
  int b_resident;           // the layer's whole weight set fits the B ring: load every stage once, never recycle
  int nterms;               // 1, or 3 (hi*hi + lo*hi + hi*lo)
  int a_tplanes, b_tplanes; // operand planes kept (1, or 2 = hi and lo)
  int is_bf16;
  int swap_ab;              // Cout = 128 layers: the weights are the MMA's A operand (M = 128 output channels) and the halo tile its B operand
                            // (N = MB*128 positions): one M=128 x N=256 MMA reads 4 KB of weights + 8 KB of activations from shared
                            // memory per 128 cycles (96 B/clk) where two N=128 MMAs read 16 KB (128 B/clk = all of the SM's shared-memory
                            // bandwidth).  The accumulator is then [channel lane][position column]: per-channel statistics are per-thread sums.
  int fold_s;               // image layer (Cout = 3, k9): GEMM N = (horizontal tap s, co) = 27 of 32 columns, entries = the 9 filter
                            // rows; the epilogue sums T[m + s][s*3 + co] over s through a shared-memory ring (9x fewer MMAs)
  int dbg;                  // debug bits (tools/dbg_sweep.sh): switch pipeline roles off one at a time
  // TMA-fed operand path (fp16 planar-8 activations, one MMA term): the loader warp copies the raw halo tile of a channel chunk
  // into the operand stage with cp.async.bulk.tensor (one box per 8-channel plane; four parity boxes per plane on stride-2
  // layers), the transform warps apply InstanceNorm / ReLU / residual IN PLACE (shared -> shared), then mirror the borders.
  int tma_mode;
  int tma_wide;             // tensor maps are 3-D over 8-byte elements (a box row = one contiguous run of P slots) instead of 4-D fp16
  int tma_rows;             // box rows: halo_h (stride 1) or TH + 1 (one stride-2 parity sub-plane)
  int npos_used;            // positions the transform sweeps: halo_h * P (stride 1) or 4 * SUBPOS (stride 2)
  uint32_t a_res_off;       // byte offset of the residual staging planes inside an operand stage (0 = no residual operand)
  int tbl_ints;             // ints reserved for the gather producers' two source tables (0 on the TMA-fed path)
  int stat_ch;              // channels of per-warp statistics scratch (output channels of this launch, rounded up to 16)
  uint32_t a_plane_bytes, a_term_bytes, a_stage_bytes;
  uint32_t b_ent_bytes, b_term_bytes, b_stage_bytes;
  uint32_t smem_bytes;
  int ent_off[TC_MAX_ENT];  // A position offset of each entry
  // sub-pixel (phase) layers: entry e multiplies only the accumulator columns [col, col + n) -- the phases its tap feeds.
  // ent_info[e] = col | (n / 16) << 9 | (B offset inside the weight stage, 16-byte units) << 14; 0 entries = plain layer
  int phase_mode;
  int grp_first[5];         // first entry of each weight stage (phase layers: stages hold different entry counts)
  uint32_t ent_info[16];
};

struct TcArgs {
  // tensor maps first: cp.async.bulk.tensor needs them 64-byte aligned in the (constant-bank) kernel parameter space.
  // [0] source (stride 1) or [py*2+px] the four parity views of the source (stride 2); tmap_res: residual operand.
  alignas(64) CUtensorMap tmap[4];
  alignas(64) CUtensorMap tmap_res;
  ConvParams p;
  TcPlan t;
  const void* wpack;        // [chunk][group][term][entry in group][plane][NPAD][8] 16-bit
  unsigned int* err_flag;   // set non-zero when a pipeline wait times out
  unsigned long long* prof_cta;   // debug: [grid][2] globaltimer at kernel entry / exit of every CTA, or null
  unsigned long long* prof; // debug (NST_TC_PROF=1): [code 1..6] cycles waited per barrier kind, [8+role] role cycles; or null
};

cudaError_t launch_conv_tc(const TcArgs& a, int num_sms, cudaStream_t stream);
int conv_tc_grid(const TcPlan& t, int num_tiles, int num_sms);     // CTAs launched (= parts of the statistics table)
cudaError_t conv_tc_init();

}  // namespace nst
