// C ABI of libb200nst.so (include/b200nst.h): handle, weight repacking, per-shape planning and
// the layer-by-layer forward of the three generator families.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <map>
#include <string>
#include <vector>
#include <algorithm>

#include "../../include/b200nst.h"
#include "conv_tc.cuh"
#include "api_util.cuh"

namespace nst {
cudaError_t launch_conv_direct(const ConvParams& p, const float* wdir, cudaStream_t stream);

constexpr int COMP_MAX_R = 16;
constexpr int COMP_MAX_J = 4;
constexpr int COMP_MAX_K = 8;
struct CompositeArgs {
  const void* srcs[COMP_MAX_K];
  const void* original;
  const float* masks;
  void* out;
  int K, R, J, H, W;
  int idx[COMP_MAX_R * COMP_MAX_J];
  float w[COMP_MAX_R * COMP_MAX_J];
};
cudaError_t launch_composite(const CompositeArgs& a, bool u8, cudaStream_t stream);
cudaError_t launch_stats_finalize(const double* part, int parts, int groups, int rr, int rot, int N, int C, int tiles_per_img, double inv_count,
                                  const float* gamma, const float* beta, float2* coeff, cudaStream_t stream);
// postproc.cu
struct ResizePackArgs {
  const float* src01; const uint8_t* orig; const float* mask; uint8_t* out;
  int N, Hi, Wi, H, W; float blend, blend_comp; int composite_mode;
};
cudaError_t launch_resize_post_pack(const ResizePackArgs& a, cudaStream_t stream);
constexpr int CB_MAX_R = 16, CB_MAX_J = 4, CB_MAX_K = 8;
struct CompositeBatchArgs {
  const void* srcs[CB_MAX_K]; int src_h[CB_MAX_K], src_w[CB_MAX_K];
  const uint8_t* original; const void* masks; uint8_t* out;
  int N, K, R, J, H, W; int mask_type; size_t mask_frame_stride;
  int idx[CB_MAX_R * CB_MAX_J]; float w[CB_MAX_R * CB_MAX_J];
};
cudaError_t launch_composite_batch(const CompositeBatchArgs& a, cudaStream_t stream);
struct CropTerm { const float* src; int sh, sw; float weight; };
struct CropRegion { int x1, y1, x2, y2; const float* mask; int mh, mw; int nterms; CropTerm term[CB_MAX_J]; };
struct CompositeCropsArgs {
  CropRegion reg[CB_MAX_R]; int R, H, W; const float* original; float* canvas; float* weight; uint8_t* out_u8; int finish;
};
cudaError_t launch_composite_crops(const CompositeCropsArgs& a, cudaStream_t stream);
struct CropsDilateArgs { const float* canvas_in; const float* weight_in; float* canvas_out; float* weight_out; uint8_t* out_u8; int H, W, k, last; };
cudaError_t launch_crops_dilate(const CropsDilateArgs& a, cudaStream_t stream);
struct FlowEmaArgs { const float* curr; const float* prev; const float* flow; float* out; int H, W; float alpha, alpha_comp; };
cudaError_t launch_flow_ema(const FlowEmaArgs& a, cudaStream_t stream);
constexpr int VOR_MAX_R = 16, BLUR_MAX_K = 255;
struct VoronoiArgs { float px[VOR_MAX_R], py[VOR_MAX_R], inv_w[VOR_MAX_R]; int R, H, W, weighted; uint8_t* label; };
cudaError_t launch_voronoi_label(const VoronoiArgs& a, cudaStream_t stream);
struct BlurArgs { const uint8_t* label; const float* src; float* dst; int R, H, W, k, vertical; float w[BLUR_MAX_K]; };
cudaError_t launch_blur_pass(const BlurArgs& a, cudaStream_t stream);
}  // namespace nst

using namespace nst;

// ---------------------------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";

static int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
// the same, for the entry points that live in masks.cu / temporal.cu (api_util.cuh)
int nst::api_fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
#define CUDA_TRY(expr)                                                                         \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess) return fail(NST_ERR_CUDA, "%s: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
  } while (0)

// ---------------------------------------------------------------------------------------------
// model description
// ---------------------------------------------------------------------------------------------
enum { BUF_IMAGE = -1, B_R1 = 0, B_R2, B_T1, B_T2, B_YA, B_YB, NBUF };

struct LayerSpec {
  std::string name;
  std::string wkey, bkey;        // conv weight / bias
  std::string in_gkey, in_bkey;  // InstanceNorm affine applied to the SOURCE tensor (previous layer's norm)
  int cin, cout, ks, stride;
  int map_mode, pad_mode, pad;
  int src, resid, side, out;
  int pre_relu, post_relu;
  int transposed, tanh_out;
};

static LayerSpec L(const std::string& name, const std::string& wkey, const std::string& bkey,
                   const std::string& gk, const std::string& bk, int cin, int cout, int ks, int stride,
                   int map_mode, int pad_mode, int src, int resid, int side, int out, int pre_relu, int post_relu,
                   int transposed = 0, int tanh_out = 0) {
  LayerSpec s;
  s.name = name; s.wkey = wkey; s.bkey = bkey; s.in_gkey = gk; s.in_bkey = bk;
  s.cin = cin; s.cout = cout; s.ks = ks; s.stride = stride;
  s.map_mode = map_mode; s.pad_mode = pad_mode; s.pad = (map_mode == MAP_ZERO_INSERT2) ? 1 : ks / 2;
  s.src = src; s.resid = resid; s.side = side; s.out = out;
  s.pre_relu = pre_relu; s.post_relu = post_relu; s.transposed = transposed; s.tanh_out = tanh_out;
  return s;
}

// transformer_net.TransformerNet (transformer_net.py:4-41)
static std::vector<LayerSpec> johnson_layers() {
  std::vector<LayerSpec> v;
  auto cw = [](const std::string& n) { return n + ".conv2d.weight"; };
  auto cb = [](const std::string& n) { return n + ".conv2d.bias"; };
  auto g = [](const std::string& n) { return n + ".weight"; };
  auto b = [](const std::string& n) { return n + ".bias"; };
  v.push_back(L("conv1", cw("conv1"), cb("conv1"), "", "", 3, 32, 9, 1, MAP_IDENTITY, PAD_REFLECT, BUF_IMAGE, -1, -1, B_R1, 0, 0));
  v.push_back(L("conv2", cw("conv2"), cb("conv2"), g("in1"), b("in1"), 32, 64, 3, 2, MAP_IDENTITY, PAD_REFLECT, B_R1, -1, -1, B_R2, 1, 0));
  v.push_back(L("conv3", cw("conv3"), cb("conv3"), g("in2"), b("in2"), 64, 128, 3, 2, MAP_IDENTITY, PAD_REFLECT, B_R2, -1, -1, B_T2, 1, 0));
  std::string prev_g = g("in3"), prev_b = b("in3");
  int ya = B_YA, yb = B_YB;
  for (int i = 1; i <= 5; ++i) {
    const std::string r = "res" + std::to_string(i);
    // block input y = [relu](IN(prev raw)) [+ previous block input]; materialised into `ya`
    v.push_back(L(r + ".conv1", cw(r + ".conv1"), cb(r + ".conv1"), prev_g, prev_b, 128, 128, 3, 1, MAP_IDENTITY, PAD_REFLECT,
                  B_T2, i == 1 ? -1 : yb, ya, B_T1, i == 1 ? 1 : 0, 0));
    v.push_back(L(r + ".conv2", cw(r + ".conv2"), cb(r + ".conv2"), g(r + ".in1"), b(r + ".in1"), 128, 128, 3, 1, MAP_IDENTITY, PAD_REFLECT,
                  B_T1, -1, -1, B_T2, 1, 0));
    prev_g = g(r + ".in2"); prev_b = b(r + ".in2");
    std::swap(ya, yb);
  }
  v.push_back(L("deconv1", cw("deconv1"), cb("deconv1"), prev_g, prev_b, 128, 64, 3, 1, MAP_UP2, PAD_REFLECT, B_T2, yb, -1, B_R2, 0, 0));
  v.push_back(L("deconv2", cw("deconv2"), cb("deconv2"), g("in4"), b("in4"), 64, 32, 3, 1, MAP_UP2, PAD_REFLECT, B_R2, -1, -1, B_R1, 1, 0));
  v.push_back(L("deconv3", cw("deconv3"), cb("deconv3"), g("in5"), b("in5"), 32, 3, 9, 1, MAP_IDENTITY, PAD_REFLECT, B_R1, -1, -1, BUF_IMAGE, 1, 0));
  return v;
}

// transformer_net_nst.TransformerNet (transformer_net_nst.py:62-127)
static std::vector<LayerSpec> nst_layers() {
  std::vector<LayerSpec> v;
  auto nw = [](const std::string& n) { return n + ".norm.weight"; };
  auto nb = [](const std::string& n) { return n + ".norm.bias"; };
  v.push_back(L("down1", "down1.conv.weight", "down1.conv.bias", "", "", 3, 32, 9, 1, MAP_PREPAD_REFLECT, PAD_ZERO, BUF_IMAGE, -1, -1, B_R1, 0, 0));
  v.push_back(L("down2", "down2.conv.weight", "down2.conv.bias", nw("down1"), nb("down1"), 32, 64, 3, 2, MAP_IDENTITY, PAD_ZERO, B_R1, -1, -1, B_R2, 1, 0));
  v.push_back(L("down3", "down3.conv.weight", "down3.conv.bias", nw("down2"), nb("down2"), 64, 128, 3, 2, MAP_IDENTITY, PAD_ZERO, B_R2, -1, -1, B_T2, 1, 0));
  std::string prev_g = nw("down3"), prev_b = nb("down3");
  int ya = B_YA, yb = B_YB;
  for (int i = 1; i <= 5; ++i) {
    const std::string r = "res" + std::to_string(i);
    v.push_back(L(r + ".conv1", r + ".conv1.weight", r + ".conv1.bias", prev_g, prev_b, 128, 128, 3, 1, MAP_IDENTITY, PAD_ZERO,
                  B_T2, i == 1 ? -1 : yb, ya, B_T1, i == 1 ? 1 : 0, 0));
    v.push_back(L(r + ".conv2", r + ".conv2.weight", r + ".conv2.bias", r + ".norm1.weight", r + ".norm1.bias", 128, 128, 3, 1,
                  MAP_IDENTITY, PAD_ZERO, B_T1, -1, -1, B_T2, 1, 0));
    prev_g = r + ".norm2.weight"; prev_b = r + ".norm2.bias";
    std::swap(ya, yb);
  }
  v.push_back(L("up1", "up1.conv.weight", "up1.conv.bias", prev_g, prev_b, 128, 64, 3, 1, MAP_ZERO_INSERT2, PAD_ZERO, B_T2, yb, -1, B_R2, 0, 0, 1));
  v.push_back(L("up2", "up2.conv.weight", "up2.conv.bias", nw("up1"), nb("up1"), 64, 32, 3, 1, MAP_ZERO_INSERT2, PAD_ZERO, B_R2, -1, -1, B_R1, 1, 0, 1));
  v.push_back(L("final", "final.weight", "final.bias", nw("up2"), nb("up2"), 32, 3, 9, 1, MAP_IDENTITY, PAD_ZERO, B_R1, -1, -1, BUF_IMAGE, 1, 0));
  return v;
}

// model.ReCoNet(frn=False) (model.py:75-116)
static std::vector<LayerSpec> reconet_layers() {
  std::vector<LayerSpec> v;
  auto cw = [](const std::string& p) { return p + ".layers.0.layers.1.weight"; };
  auto cb = [](const std::string& p) { return p + ".layers.0.layers.1.bias"; };
  auto g = [](const std::string& p) { return p + ".layers.1.weight"; };
  auto b = [](const std::string& p) { return p + ".layers.1.bias"; };
  const std::string e0 = "encoder.layers.0", e1 = "encoder.layers.1", e2 = "encoder.layers.2";
  v.push_back(L(e0, cw(e0), cb(e0), "", "", 3, 48, 9, 1, MAP_IDENTITY, PAD_REFLECT, BUF_IMAGE, -1, -1, B_R1, 0, 0));
  v.push_back(L(e1, cw(e1), cb(e1), g(e0), b(e0), 48, 96, 3, 2, MAP_IDENTITY, PAD_REFLECT, B_R1, -1, -1, B_R2, 1, 0));
  v.push_back(L(e2, cw(e2), cb(e2), g(e1), b(e1), 96, 192, 3, 2, MAP_IDENTITY, PAD_REFLECT, B_R2, -1, -1, B_T2, 1, 0));
  std::string prev_g = g(e2), prev_b = b(e2);
  int ya = B_YA, yb = B_YB;
  for (int i = 3; i <= 6; ++i) {
    const std::string r = "encoder.layers." + std::to_string(i);
    const std::string b0 = r + ".branch.0", b1 = r + ".branch.1";
    // ResLayer: x = relu(x + branch(x))  (model.py:57-60): the ReLU comes AFTER the add
    v.push_back(L(b0, cw(b0), cb(b0), prev_g, prev_b, 192, 192, 3, 1, MAP_IDENTITY, PAD_REFLECT,
                  B_T2, i == 3 ? -1 : yb, ya, B_T1, i == 3 ? 1 : 0, i == 3 ? 0 : 1));
    v.push_back(L(b1, cw(b1), cb(b1), g(b0), b(b0), 192, 192, 3, 1, MAP_IDENTITY, PAD_REFLECT, B_T1, -1, -1, B_T2, 1, 0));
    prev_g = g(b1); prev_b = b(b1);
    std::swap(ya, yb);
  }
  const std::string d1 = "decoder.layers.1", d3 = "decoder.layers.3", d4 = "decoder.layers.4";
  v.push_back(L(d1, cw(d1), cb(d1), prev_g, prev_b, 192, 96, 3, 1, MAP_UP2, PAD_REFLECT, B_T2, yb, -1, B_R2, 0, 1));
  v.push_back(L(d3, cw(d3), cb(d3), g(d1), b(d1), 96, 48, 3, 1, MAP_UP2, PAD_REFLECT, B_R2, -1, -1, B_R1, 1, 0));
  v.push_back(L(d4, d4 + ".layers.0.layers.1.weight", d4 + ".layers.0.layers.1.bias", g(d3), b(d3), 48, 3, 9, 1, MAP_IDENTITY,
                PAD_REFLECT, B_R1, -1, -1, BUF_IMAGE, 1, 0, 0, 1));
  return v;
}

// ---------------------------------------------------------------------------------------------
// handle
// ---------------------------------------------------------------------------------------------
struct LayerDev {
  float* wdir = nullptr;     // [ks*ks][cin][cout] fp32 (direct kernel)
  void* wpack = nullptr;     // tcgen05 operand tiles (16-bit)
  void* wpack2 = nullptr;    // second half of the output channels (phase layers split in two launches)
  int nsplit = 1;
  float* bias = nullptr;     // [cout]
  float* in_gamma = nullptr; // [cin]
  float* in_beta = nullptr;  // [cin]
  TcPlan tc_static;          // shape-independent part of the tiling
  bool tc_ok = false;
  int prec = 0;              // NST_PREC_* of THIS layer (a handle in NST_PREC_F16 may run single layers error-compensated, layer_precision())
};

struct LayerPlan {
  ConvParams p;              // pointers are filled per forward
  TcPlan t;
  bool use_tc = false;
  int in16 = 0, out16 = 0;   // element type of the activation tensors on the input side (src, resid, side) / of the output
  int producer = -1;         // layer that wrote `src` (its statistics feed this layer's InstanceNorm), or -1
  int consumer = -1;         // layer that reads `out` as its source (owns the gamma/beta of the norm in between), or -1
  size_t part_off = 0;       // bytes from the start of the statistics region: partial table [parts][nb][Cout][2] double
  size_t coeff_off = 0;      // bytes: [nb][Cout] float2 InstanceNorm (scale, shift) of this layer's output
  int out_H = 0, out_W = 0;  // extent of the output tensor (full resolution for the sub-pixel layers)
  int parts = 1;             // slots of the partial table (= SMs for the tcgen05 kernel, 1 for the direct kernel)
  double flops = 0, bytes = 0;
  // tensor maps of the TMA-fed operand path, encoded for `maps_ws` (re-encoded when the caller's workspace moves)
  CUtensorMap tmap[4], tmap_res;
  const void* maps_ws = nullptr;
  std::map<int, int> rot_cache;      // frames in the launch -> rotation of the round-robin tile sequence (pick_rotation)
};

struct BufDims { int C = 0, H = 0, W = 0; size_t off = 0; size_t bytes = 0; };   // off / bytes per frame

struct nst_handle {
  int device = 0, arch = 0, precision = 0, num_sms = 148;
  std::vector<LayerSpec> specs;
  std::vector<LayerDev> dev;
  bool weights_loaded = false;
  // plan cache
  int plan_N = 0, plan_H = 0, plan_W = 0, plan_nb = 0;
  int out_H = 0, out_W = 0;
  std::vector<LayerPlan> plans;
  BufDims bufs[NBUF];
  size_t act_bytes_per_frame = 0;
  int act16 = 0;                     // default element type of intermediate activations (layers next to an F16X3 layer differ)
  size_t stats_bytes = 0;            // partial tables + coefficient tables of one micro-batch
  size_t fit_bytes = 0;              // out01 planes of one micro-batch (uint8 "fit" surface at sizes that are not multiples of 4)
  size_t workspace_bytes = 0;
  unsigned int* err_host = nullptr;
  unsigned int* err_dev = nullptr;
  long long launches = 0;
  int microbatch = 8;
  int dbg = 0;
  unsigned long long* prof_dev = nullptr;     // NST_TC_PROF=1: per-role wait-cycle counters (debug)
  unsigned long long tc_layer_mask = ~0ull;   // debug: bit i clear -> layer i runs on the direct kernel
};

static int ceil_div(int a, int b) { return (a + b - 1) / b; }
static size_t round_up(size_t a, size_t b) { return (a + b - 1) / b * b; }

// ---------------------------------------------------------------------------------------------
// tcgen05 tiling
// ---------------------------------------------------------------------------------------------
static bool tc_supported(const LayerSpec& s) {
  if (!(s.ks == 3 || s.ks == 9)) return false;
  if (s.stride == 2 && s.ks != 3) return false;
  if (s.stride != 1 && s.stride != 2) return false;
  if (s.cin == 3) return s.ks == 9 && s.stride == 1;
  if (s.cin % 16 != 0) return false;
  if (s.cout != 3 && s.cout % 16 != 0) return false;
  if (s.cout > 256) return false;
  return true;
}

// Sub-pixel form of  nearest-upsample x2 -> ReflectionPad2d(1) -> Conv2d(k3)  (transformer_net.py:79-99, model.py:96-99):
// output pixel (2i+a, 2j+b) only sees the 2x2 low-res neighbourhood of (i,j), so the layer is a 3x3 conv over the LOW-RES
// tensor with replicate padding and N = 4 phases x Cout columns, where tap (r,s) feeds only some phases.  Entry table:
// low-res tap (r,s), first phase (= a*2+b), number of adjacent phases.  Entry 0 covers every column (it initialises the accumulator).
// kind 1 (upsample + reflect pad + conv3), kind 2 (ConvTranspose2d k3 s2 p1 op1, transformer_net_nst.py:46-59: output (2i+a, 2j+b)
// reads low-res rows i (kernel row 1 for a=0, row 0 for a=1 of the equivalent flipped kernel) and i+1 (row 2, a=1 only), zero padded).
static const int PH_NENT_K[3] = {0, 11, 5};
static const int PH_G1_K[3] = {0, 3, 2};                 // entries in the first weight stage
static const int PH_R_K[3][11]  = {{0}, {1, 0, 2, 1, 1, 1, 1, 0, 0, 2, 2}, {1, 2, 1, 1, 2}};
static const int PH_S_K[3][11]  = {{0}, {1, 1, 1, 0, 0, 2, 2, 0, 2, 0, 2}, {1, 1, 2, 2, 2}};
static const int PH_P0_K[3][11] = {{0}, {0, 0, 2, 0, 2, 1, 3, 0, 1, 2, 3}, {0, 2, 1, 3, 3}};
static const int PH_NP_K[3][11] = {{0}, {4, 2, 2, 1, 1, 1, 1, 1, 1, 1, 1}, {4, 2, 1, 1, 1}};
// original taps r' that land on low-res tap r for row/column phase a
static bool ph_hits(int kind, int a, int r, int rp) {
  if (kind == 1) {                                       // a=0: {0} {1,2} {};  a=1: {} {0,1} {2}
    if (a == 0) return (r == 0 && rp == 0) || (r == 1 && rp >= 1);
    return (r == 1 && rp <= 1) || (r == 2 && rp == 2);
  }
  if (a == 0) return r == 1 && rp == 1;                  // kind 2: a=0: {} {1} {};  a=1: {} {0} {2}
  return (r == 1 && rp == 0) || (r == 2 && rp == 2);
}
static int phase_kind(const LayerSpec& s, int precision) {
  if (getenv("NST_NO_PHASE") || precision == NST_PREC_F16X3) return 0;
  if (s.ks != 3 || s.stride != 1 || s.cout % 16 != 0 || 4 * s.cout > 256 || s.cin % 16 != 0) return 0;   // (callers pass the per-launch channel count)
  if (s.map_mode == MAP_UP2 && s.pad_mode == PAD_REFLECT) return 1;
  if (s.map_mode == MAP_ZERO_INSERT2 && s.pad_mode == PAD_ZERO) return 2;
  return 0;
}

// Operand-producer warps per layer.  16 warps x 128 registers fill the register file, so 10 producer warps are free; measured
// per layer (tools/ab.sh): the producer-bound forms gain 6-10 % with 10 (first layer, stride 2, residual operand, fold_s), the
// tensor- or epilogue-bound ones lose 3-5 % (more shared-memory traffic against the MMA) and keep 8.
static int tc_prod_warps(const LayerSpec& s, const TcPlan& t) {
  if (const char* e = getenv("NST_PROD_WARPS")) { const int v = atoi(e); if (v >= 1 && v <= TC_PROD_WARPS) return v; }
  return (s.cin == 3 || s.stride == 2 || s.resid >= 0 || t.fold_s) ? TC_PROD_WARPS : 8;
}

// Precision of one layer.  A handle in the default NST_PREC_F16 mode runs ReCoNet's three encoder convs error-compensated
// (fp16 hi+lo operands, three MMA terms, fp32 storage between them): the tanh output is x127.5 steep and the ablation in
// tools/precision_layers.py puts ~all of the |delta| >= 2 tail of the uint8 output on the rounding of these early layers
// (n(|d|>=2) 704 -> 4 of 783k values at 272x480; the 4-ppm tail at 3 disappears).  NST_X3_LAYERS=<bitmask> overrides.
static int layer_precision(int arch, int precision, int li) {
  if (precision != NST_PREC_F16) return precision;
  unsigned long long mask = arch == NST_ARCH_RECONET ? 0x7ull : 0ull;
  if (const char* e = getenv("NST_X3_LAYERS")) mask = strtoull(e, nullptr, 0);
  return ((mask >> li) & 1ull) ? NST_PREC_F16X3 : NST_PREC_F16;
}

static void tc_plan_static_(const LayerSpec& s, int precision, TcPlan& t);
static void tc_plan_static(const LayerSpec& s, int precision, TcPlan& t) {
  tc_plan_static_(s, precision, t);
  t.prod_warps = tc_prod_warps(s, t);
}
static void tc_plan_static_(const LayerSpec& s, int precision, TcPlan& t) {
  memset(&t, 0, sizeof(t));
  if (const int kind = phase_kind(s, precision)) {
    t.phase_mode = kind;
    const int PH_NENT = PH_NENT_K[kind];
    const int* PH_P0 = PH_P0_K[kind]; const int* PH_NP = PH_NP_K[kind];
    t.NPAD = 4 * s.cout;
    t.CK = (s.cin % 32 == 0) ? 32 : 16;
    t.planes = t.CK / 8; t.nchunks = s.cin / t.CK; t.nent = PH_NENT; t.k16_per_ent = t.CK / 16;
    t.is_bf16 = precision == NST_PREC_BF16;
    t.nterms = 1; t.a_tplanes = 1; t.b_tplanes = 1;
    t.ngroups = 2; t.tg = 8;
    t.grp_first[0] = 0; t.grp_first[1] = PH_G1_K[kind]; t.grp_first[2] = PH_NENT;
    uint32_t grp_bytes[2] = {0, 0};
    for (int e = 0; e < PH_NENT; ++e) {
      const int g = e < PH_G1_K[kind] ? 0 : 1;
      const int n = PH_NP[e] * s.cout;
      t.ent_info[e] = (uint32_t)(PH_P0[e] * s.cout) | ((uint32_t)(n / 16) << 9) | ((grp_bytes[g] / 16u) << 14);
      grp_bytes[g] += (uint32_t)(t.planes * n * 16);
    }
    t.b_ent_bytes = (uint32_t)(t.planes * t.NPAD * 16);
    t.b_stage_bytes = t.b_term_bytes = std::max(grp_bytes[0], grp_bytes[1]);
    t.b_stages = t.b_stage_bytes > 24576 ? 2 : 3;
    t.MB = std::max(1, 256 / t.NPAD);
    return;
  }
  t.NPAD = (int)round_up((size_t)s.cout, 16);
  // first layer (3 channels): a slot holds one pixel in 8 K positions (pair mode, K16 = 2 pixels, 3/8 useful) or two pixels
  // in 4 K positions each (quad mode, K16 = 4 pixels through LBO = 2 slots, 3/4 useful: 27 instead of 45 MMAs per M block)
  t.pair_mode = s.cin == 3 ? (getenv("NST_NO_QUAD") ? 1 : 2) : 0;
  // fold_s: 2 = tile pitch fixed at 32 so that a warp's 32 accumulator rows are one tile row and the horizontal-tap sum is 24 warp
  // shuffles (no shared-memory ring, no barriers); 1 = any pitch, rows exchanged through a ring (NST_FOLD_RING=1)
  t.fold_s = (s.cout == 3 && s.ks == 9 && s.cin != 3 && !getenv("NST_NO_FOLD")) ? (getenv("NST_FOLD_RING") ? 1 : 2) : 0;
  if (t.fold_s) t.NPAD = 32;
  if (t.pair_mode) {
    t.CK = 8; t.planes = 1; t.nchunks = 1;
    t.nent = t.pair_mode == 2 ? s.ks * ((s.ks + 3) / 4) : s.ks * ((s.ks + 1) / 2);
    t.k16_per_ent = 1;
  } else {
    t.CK = (s.cin % 32 == 0) ? 32 : 16;
    t.planes = t.CK / 8;
    t.nchunks = s.cin / t.CK;
    t.nent = t.fold_s ? s.ks : s.ks * s.ks;
    t.k16_per_ent = t.CK / 16;
  }
  t.is_bf16 = precision == NST_PREC_BF16;
  if (precision == NST_PREC_F16X3) { t.nterms = 3; t.a_tplanes = 2; t.b_tplanes = 2; }
  else { t.nterms = 1; t.a_tplanes = 1; t.b_tplanes = 1; }
  t.b_ent_bytes = (uint32_t)((t.pair_mode ? 2 : t.planes) * t.NPAD * 16);
  // weight stage: up to 24 KB (3 taps of a 128->128 layer) so one mbarrier wait feeds >= 12 MMAs
  int tg = (int)(24576 / (t.b_ent_bytes * t.b_tplanes));
  tg = std::max(1, std::min(tg, t.nent));
  if (s.ks == 9 && !t.pair_mode) tg = std::min(tg, 9);
  t.tg = tg;
  t.ngroups = ceil_div(t.nent, tg);
  t.b_term_bytes = (uint32_t)tg * t.b_ent_bytes;
  t.b_stage_bytes = t.b_term_bytes * (uint32_t)t.b_tplanes;
  t.b_stages = t.b_stage_bytes <= 8192 ? 6 : (t.b_stage_bytes <= 16384 ? 4 : 3);
  if (t.fold_s && t.nchunks * t.ngroups <= TC_MAX_B_STAGES && (size_t)t.nchunks * t.ngroups * t.b_stage_bytes <= 32768) {
    t.b_resident = 1;                    // 18-27 KB of weights stay in shared memory for the whole launch: room for larger tiles
    t.b_stages = t.nchunks * t.ngroups;
  }
  int mb = 256 / t.NPAD;                  // two accumulator buffers of MB*NPAD columns in 512
  mb = mb >= 8 ? 8 : (mb >= 4 ? 4 : (mb >= 2 ? 2 : 1));   // powers of two: the MMA issuer is instantiated for MB = 1, 2, 4, 8
  t.MB = mb;
  // Cout = 128: weights as the A operand (M = 128 channels), the halo tile as B (N = MB*128 positions) -- see TcPlan::swap_ab
  t.swap_ab = (s.cout == 128 && t.NPAD == 128 && !t.pair_mode && !t.fold_s && t.nterms == 1 && t.k16_per_ent == 2 && !getenv("NST_NO_SWAP")) ? 1 : 0;
}

// Shape-dependent part: tile extent, halo buffer, shared-memory budget.
// Lay out one candidate tile (TH x TW outputs, halo pitch P, MB M-blocks) in shared memory; false if it does not fit.
static bool tc_layout(const ConvParams& p, TcPlan& t, int MB, int bTH, int bTW, int bP) {
  const int e = p.ks - 1;
  const bool tma = t.tma_mode != 0;
  t.TH = bTH; t.TW = bTW; t.P = bP;
  const int Mused = ceil_div(bTH * bP, 128);
  const int MBu = std::min(MB, std::max(1, Mused));
  int max_off = 0;
  if (p.stride == 1) {
    t.halo_h = bTH + e; t.halo_w = bP; t.SUBPOS = 0;
    int k = 0;
    if (t.pair_mode) {
      if (t.pair_mode == 2) {
        for (int r = 0; r < p.ks; ++r) for (int sp = 0; sp < (p.ks + 3) / 4; ++sp) t.ent_off[k++] = r * bP + 4 * sp;
        max_off = e * bP + e + 4;
      } else {
        for (int r = 0; r < p.ks; ++r) for (int sp = 0; sp < (p.ks + 1) / 2; ++sp) t.ent_off[k++] = r * bP + 2 * sp;
        max_off = e * bP + e + 1;
      }
    } else {
      if (t.phase_mode) { for (int i = 0; i < PH_NENT_K[t.phase_mode]; ++i) t.ent_off[k++] = PH_R_K[t.phase_mode][i] * bP + PH_S_K[t.phase_mode][i]; max_off = e * bP + e; }
      else if (t.fold_s) { for (int r = 0; r < p.ks; ++r) t.ent_off[k++] = r * bP; max_off = e * bP; }
      else { for (int r = 0; r < p.ks; ++r) for (int s = 0; s < p.ks; ++s) t.ent_off[k++] = r * bP + s; max_off = e * bP + e; }
    }
    t.NPOS = std::max(MBu * 128 + max_off + 2, t.halo_h * bP + 2);
    t.tma_rows = t.halo_h; t.npos_used = t.halo_h * bP;
  } else {
    t.halo_h = 2 * bTH + 1; t.halo_w = 2 * bTW + 1;
    t.SUBPOS = (bTH + 1) * bP + 2;
    if (tma) t.SUBPOS = (int)round_up((size_t)t.SUBPOS, 8);            // every parity sub-plane is a TMA destination: 128-byte aligned
    int k = 0;
    for (int r = 0; r < 3; ++r) for (int s = 0; s < 3; ++s) t.ent_off[k++] = ((r & 1) * 2 + (s & 1)) * t.SUBPOS + (r >> 1) * bP + (s >> 1);
    max_off = 3 * t.SUBPOS + bP + 1;
    t.NPOS = std::max(MBu * 128 + max_off + 2, 4 * t.SUBPOS);
    t.tma_rows = bTH + 1; t.npos_used = 4 * t.SUBPOS;
  }
  if (tma && p.side_out != nullptr) {
    // the transform keeps the in-plane offsets of the slots a thread visits in registers: at most 2 iterations x 4 slots per chunk
    const int nblk = (t.npos_used + 7) / 8, blk_step = t.prod_warps * 32 / (8 * t.planes);
    if (blk_step < 1 || (nblk + blk_step - 1) / blk_step > 8) return false;
  }
  if (tma) {
    // TMA destinations (one box per plane) must be 128-byte aligned; the in-place transform touches a plane 128 bytes per
    // quarter-warp, so any plane stride is bank-conflict free.  Box extents are limited to 256 per dimension.
    if (bP > 256 || t.tma_rows > 256) return false;
    t.NPOS = (int)round_up((size_t)t.NPOS, 8);
  }
  // plane stride (in 16-byte units) chosen so the gather producers' 8-lane store phases cover all banks
  else if (t.planes == 4) { while (t.NPOS % 8 != 2) ++t.NPOS; }
  else if (t.planes == 2) { while (t.NPOS % 8 != 4) ++t.NPOS; }
  else { t.NPOS = (int)round_up((size_t)t.NPOS, 2); }
  t.a_plane_bytes = (uint32_t)t.NPOS * 16u;
  t.a_term_bytes = t.a_plane_bytes * (uint32_t)t.planes;
  t.a_res_off = (tma && p.resid != nullptr) ? t.a_term_bytes : 0u;      // (p.resid is only a flag at plan time)
  t.a_stage_bytes = (uint32_t)round_up((size_t)t.a_term_bytes * t.a_tplanes + t.a_res_off, 128);
  if (t.a_plane_bytes >= (1u << 18)) return false;
  const int b_total = t.b_stages * (int)t.b_stage_bytes;
  // barriers + TMEM slot + entry table (704) + the gather producers' two per-tile source tables [NPOS] (current / next tile)
  // + per-sample norm coefficients [Cin<=256] + epilogue scratch: fold ring, or per-warp double channel statistics
  t.tbl_ints = tma ? 0 : 2 * (int)round_up((size_t)t.NPOS, 4);
  t.stat_ch = (int)round_up((size_t)(t.phase_mode ? t.NPAD / 4 : t.NPAD), 16);
  t.epi_scratch_bytes = (uint32_t)(t.fold_s == 1 ? TC_FOLD_RING_BYTES : (t.fold_s == 2 ? 0 : 4 * t.stat_ch * 16));
  const int bar_bytes = 704 + t.tbl_ints * 4 + 256 * 8 + t.epi_groups * (int)t.epi_scratch_bytes;
  const int avail = TC_SMEM_LIMIT - b_total - bar_bytes;
  const int a_stages = avail / (int)t.a_stage_bytes;
  const int want = t.nchunks == 1 ? 2 : 3;
  if (a_stages < 2 && MB > 1) return false;
  if (a_stages < 1) return false;
  if (tma && a_stages < 2) return false;                                 // the loader must be able to run a stage ahead of the transform
  t.a_stages = std::min(a_stages, std::min(TC_MAX_A_STAGES, want + 1));
  int b_stages = t.b_stages;
  if (!t.b_resident && !getenv("NST_NO_DEEP_B")) {
    // leftover shared memory deepens the weight ring (up to 6 stages): a refill has L2 latency + the queueing of the TMA unit behind
    // the activation boxes, and the MMA issuer showed 10-30 % of its time waiting for weights with three 24 KB stages
    const int left = avail - t.a_stages * (int)t.a_stage_bytes;
    b_stages = std::min(6, t.b_stages + std::max(0, left / (int)t.b_stage_bytes));
  }
  t.b_stages = b_stages;
  t.MB = MBu;
  t.tiles_x = ceil_div(p.Wo, t.TW);
  t.tiles_y = ceil_div(p.Ho, t.TH);
  t.num_tiles = t.tiles_x * t.tiles_y * p.N;
  t.smem_bytes = (uint32_t)(t.a_stages * t.a_stage_bytes + t.b_stages * (int)t.b_stage_bytes + bar_bytes);
  return true;
}

// Shape-dependent part: tile extent, halo buffer, shared-memory budget.  Every (MB, TH) candidate is laid out and costed;
// the model is the measured one: an MMA costs its table value, an operand-producer item ~3.4 SM cycles, and the two overlap.
static bool tc_plan_geometry(const ConvParams& p, TcPlan& t, int force_TH = 0, int force_MB = 0) {
  if ((long long)p.N * p.Hs * p.Ws >= (1ll << 30)) return false;   // source-table entries are 30-bit pixel indices
  const int e = p.ks - 1;
  const TcPlan base = t;
  double best = 1e30;
  TcPlan best_plan = t;
  bool found = false;
  // measured tcgen05.mma cost (tools/umma_bench.cu): N=16 39, N=32 40, N=64 48, N=128 64, N=256 128 cycles
  const double mma_one = t.NPAD >= 128 ? t.NPAD / 2.0 : 32.0 + t.NPAD / 4.0;
  for (int MB = base.MB; MB >= 1; MB /= 2) {
    if (force_MB && MB != force_MB) continue;
    const int M = MB * 128;
    double mma_cyc = (double)MB * t.nchunks * t.nent * t.k16_per_ent * t.nterms * mma_one;
    if (t.phase_mode) mma_cyc = (double)MB * t.nchunks * t.k16_per_ent * (t.NPAD * (t.phase_mode == 1 ? 2.5 : 1.4));
    for (int TH = 1; TH <= std::min(64, M); ++TH) {
      if (force_TH && TH != force_TH && t.fold_s != 2) continue;      // (the pitch-32 image layer has one TH per MB)
      int P = M / TH;
      int TW = p.stride == 1 ? P - e : P - 1;
      if (TW < 1) continue;
      if (t.fold_s == 2) {                     // pitch 32 exactly (TH = M / 32); a narrow image just leaves columns unused
        if (P != 32 || TH * 32 != M) continue;
        TW = std::min(TW, p.Wo);
      } else
      if (TW > p.Wo) { TW = p.Wo; P = p.stride == 1 ? TW + e : TW + 1; }
      if (TW < 4 && p.Wo >= 4) continue;
      const int THe = std::min(TH, p.Ho);
      TcPlan cand = base;
      if (!tc_layout(p, cand, MB, THe, TW, P)) continue;
      const double tiles = (double)cand.tiles_x * cand.tiles_y;
      // short halo rows gather worse (tools/tile_sweep.py: wide, flat tiles win by 3-7 % at equal position count)
      const double row_factor = p.stride == 1 ? 1.0 + 8.0 / cand.P : 1.0;
      // operand cost per 16-byte slot: ~3.4 SM cycles for the gather producers; the TMA-fed in-place transform has no
      // table / address work and no row-length sensitivity (estimate, re-fit from profiles/r02_tcprof_*)
      const double prod_cyc = cand.tma_mode ? (double)cand.npos_used * cand.planes * cand.nchunks * (p.resid ? 1.6 : 1.2)
                                            : (double)cand.NPOS * cand.planes * cand.nchunks * 3.4 * row_factor;
      const double cost = tiles * (std::max(mma_cyc, prod_cyc) + 0.25 * std::min(mma_cyc, prod_cyc) + 1500.0);
      if (cost < best) { best = cost; best_plan = cand; found = true; }
    }
  }
  if (!found) return false;
  t = best_plan;
  return true;
}

// Host packing of one layer's weights into the weight-stage images the kernel bulk-copies:
// [chunk][group][term plane][entry in group][8-channel plane][NPAD][8] 16-bit.
static std::vector<uint16_t> tc_pack_weights(const LayerSpec& s, const TcPlan& t, const std::vector<float>& w_oihw) {
  const size_t stage_elems = t.b_stage_bytes / 2;
  std::vector<uint16_t> out((size_t)t.nchunks * t.ngroups * stage_elems, 0);
  if (t.phase_mode) {
    // entry e at [chunk][stage g][B offset]: [8-channel plane][n = (phase - phase0)*Cout + co][8] of the pre-summed phase weights
    const int kind = t.phase_mode, PH_NENT = PH_NENT_K[kind];
    const int *PH_R = PH_R_K[kind], *PH_S = PH_S_K[kind], *PH_P0 = PH_P0_K[kind], *PH_NP = PH_NP_K[kind];
    for (int ch = 0; ch < t.nchunks; ++ch) for (int e = 0; e < PH_NENT; ++e) {
      const int g = e < t.grp_first[1] ? 0 : 1;
      const int n_e = PH_NP[e] * s.cout;
      const size_t base = ((size_t)ch * t.ngroups + g) * stage_elems + (size_t)(t.ent_info[e] >> 14) * 8;
      for (int kp = 0; kp < t.planes; ++kp) for (int n = 0; n < n_e; ++n) for (int i = 0; i < 8; ++i) {
        const int ph = PH_P0[e] + n / s.cout, co = n % s.cout, ci = ch * t.CK + kp * 8 + i;
        const int a = ph >> 1, b = ph & 1;
        float v = 0.f;
        for (int rp = 0; rp < 3; ++rp) for (int sp = 0; sp < 3; ++sp)
          if (ph_hits(kind, a, PH_R[e], rp) && ph_hits(kind, b, PH_S[e], sp)) v += w_oihw[(((size_t)co * s.cin + ci) * 3 + rp) * 3 + sp];
        uint16_t u;
        if (t.is_bf16) { __nv_bfloat16 hv = __float2bfloat16_rn(v); memcpy(&u, &hv, 2); } else { __half hv = __float2half_rn(v); memcpy(&u, &hv, 2); }
        out[base + ((size_t)kp * n_e + n) * 8 + i] = u;
      }
    }
    return out;
  }
  auto conv16 = [&](float v) -> uint16_t {
    if (t.is_bf16) { __nv_bfloat16 h = __float2bfloat16_rn(v); uint16_t u; memcpy(&u, &h, 2); return u; }
    __half h = __float2half_rn(v); uint16_t u; memcpy(&u, &h, 2); return u;
  };
  auto back16 = [&](uint16_t u) -> float {
    if (t.is_bf16) { __nv_bfloat16 h; memcpy(&h, &u, 2); return __bfloat162float(h); }
    __half h; memcpy(&h, &u, 2); return __half2float(h);
  };
  const int ks = s.ks;
  const int kplanes = t.pair_mode ? 2 : t.planes;
  for (int ch = 0; ch < t.nchunks; ++ch) {
    for (int ent = 0; ent < t.nent; ++ent) {
      const int grp = ent / t.tg, e = ent % t.tg;
      for (int kp = 0; kp < kplanes; ++kp) {
        for (int n = 0; n < t.NPAD; ++n) {
          for (int i = 0; i < 8; ++i) {
            float v = 0.f;
            if (n < s.cout) {
              if (t.pair_mode == 2) {
                const int per_row = (ks + 3) / 4;
                const int r = ent / per_row, sp = ent % per_row;
                const int sx = 4 * sp + kp * 2 + i / 4, c = i % 4;
                if (sx < ks && c < s.cin) v = w_oihw[(((size_t)n * s.cin + c) * ks + r) * ks + sx];
              } else if (t.pair_mode) {
                const int per_row = (ks + 1) / 2;
                const int r = ent / per_row, sp = ent % per_row;
                const int sx = 2 * sp + kp;
                if (sx < ks && i < s.cin) v = w_oihw[(((size_t)n * s.cin + i) * ks + r) * ks + sx];
              } else {
                const int r = ent / ks, sx = ent % ks;
                const int ci = ch * t.CK + kp * 8 + i;
                v = w_oihw[(((size_t)n * s.cin + ci) * ks + r) * ks + sx];
              }
            }
            if (t.fold_s) {      // column n = s*3 + co of filter row r = ent
              v = 0.f;
              if (n < 3 * ks) v = w_oihw[(((size_t)(n % 3) * s.cin + (ch * t.CK + kp * 8 + i)) * ks + ent) * ks + n / 3];
            }
            const uint16_t hi = conv16(v);
            const size_t base = ((size_t)ch * t.ngroups + grp) * stage_elems;
            const size_t within = ((size_t)e * kplanes + kp) * t.NPAD * 8 + (size_t)n * 8 + i;
            out[base + within] = hi;
            if (t.b_tplanes == 2) out[base + t.b_term_bytes / 2 + within] = conv16(v - back16(hi));
          }
        }
      }
    }
  }
  return out;
}

// ---------------------------------------------------------------------------------------------
// lifetime
// ---------------------------------------------------------------------------------------------
extern "C" const char* nst_last_error(void) { return g_err; }
extern "C" const char* nst_version(void) { return "b200nst 0.2 (sm_100a, tcgen05 implicit-GEMM conv, TMA-fed operands)"; }

extern "C" int nst_create(int device, int arch, int precision, nst_handle** out) {
  if (out == nullptr) return fail(NST_ERR_INVALID, "nst_create: out is null");
  *out = nullptr;
  if (arch < 0 || arch > 2) return fail(NST_ERR_INVALID, "nst_create: unknown arch %d", arch);
  if (precision < 0 || precision > 3) return fail(NST_ERR_INVALID, "nst_create: unknown precision %d", precision);
  int ndev = 0;
  CUDA_TRY(cudaGetDeviceCount(&ndev));
  if (device < 0 || device >= ndev) return fail(NST_ERR_CUDA, "nst_create: CUDA device %d not present (%d visible); there is no CPU path", device, ndev);
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) return fail(NST_ERR_CUDA, "nst_create: device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
  CUDA_TRY(cudaSetDevice(device));
  nst_handle* h = new nst_handle();
  h->device = device; h->arch = arch; h->precision = precision;
  h->act16 = precision == NST_PREC_F16 ? 1 : 0;
  if (const char* a = getenv("NST_ACT16")) h->act16 = atoi(a) != 0 && precision != NST_PREC_F16X3;
  h->num_sms = prop.multiProcessorCount;
  h->specs = arch == NST_ARCH_JOHNSON ? johnson_layers() : (arch == NST_ARCH_NST ? nst_layers() : reconet_layers());
  h->dev.resize(h->specs.size());
  for (size_t li = 0; li < h->specs.size(); ++li) h->dev[li].prec = layer_precision(arch, precision, (int)li);
  if (const char* mbs = getenv("NST_MICROBATCH")) h->microbatch = std::max(1, atoi(mbs));
  if (const char* d = getenv("NST_TC_DBG")) h->dbg = atoi(d);
  if (const char* d = getenv("NST_TC_LAYER_MASK")) h->tc_layer_mask = strtoull(d, nullptr, 0);
  cudaError_t e = cudaHostAlloc((void**)&h->err_host, sizeof(unsigned int), cudaHostAllocMapped);
  if (e != cudaSuccess) { delete h; return fail(NST_ERR_CUDA, "cudaHostAlloc: %s", cudaGetErrorString(e)); }
  *h->err_host = 0;
  e = cudaHostGetDevicePointer((void**)&h->err_dev, h->err_host, 0);
  if (e == cudaSuccess) e = conv_tc_init();
  if (e == cudaSuccess && getenv("NST_TC_PROF") && atoi(getenv("NST_TC_PROF"))) e = cudaMalloc((void**)&h->prof_dev, (64 * 16 + 64 * 2 * 148) * sizeof(unsigned long long));
  if (e != cudaSuccess) { cudaFreeHost(h->err_host); delete h; return fail(NST_ERR_CUDA, "nst_create: %s", cudaGetErrorString(e)); }
  *out = h;
  return NST_OK;
}

static void free_layer(LayerDev& d) {
  cudaFree(d.wdir); cudaFree(d.wpack); cudaFree(d.wpack2); cudaFree(d.bias); cudaFree(d.in_gamma); cudaFree(d.in_beta);
  const int prec = d.prec;
  d = LayerDev();
  d.prec = prec;
}

// static (shape-independent) tcgen05 form of a layer: whether it has one, whether it runs as two channel halves, its tiling constants
static void layer_static_plan(const LayerSpec& s, LayerDev& d) {
  d.tc_ok = d.prec != NST_PREC_FP32_DIRECT && tc_supported(s);
  d.nsplit = 1;
  if (!d.tc_ok) return;
  // a sub-pixel layer whose 4*Cout accumulator columns do not fit TMEM (ReCoNet decoder.layers.1: 4*96) runs as two
  // launches over half of the output channels each
  LayerSpec half = s;
  half.cout = s.cout / 2;
  d.nsplit = (4 * s.cout > 256 && s.cout % 32 == 0 && phase_kind(half, d.prec) != 0 && !getenv("NST_NO_SPLIT")) ? 2 : 1;
  tc_plan_static(d.nsplit == 2 ? half : s, d.prec, d.tc_static);
}

extern "C" void nst_destroy(nst_handle* h) {
  if (h == nullptr) return;
  cudaSetDevice(h->device);
  for (auto& d : h->dev) free_layer(d);
  if (h->err_host) cudaFreeHost(h->err_host);
  cudaFree(h->prof_dev);
  delete h;
}

template <typename T>
static cudaError_t upload(T** dst, const std::vector<T>& src) {
  cudaError_t e = cudaMalloc((void**)dst, std::max<size_t>(src.size(), 1) * sizeof(T));
  if (e != cudaSuccess) return e;
  return cudaMemcpy(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice);
}

extern "C" int nst_load_weights(nst_handle* h, int n_tensors, const char* const* names, const float* const* ptrs,
                                const int64_t* shapes, int* n_missing, int* n_unexpected) {
  if (h == nullptr) return fail(NST_ERR_INVALID, "nst_load_weights: null handle");
  if (n_tensors < 0 || (n_tensors > 0 && (!names || !ptrs || !shapes))) return fail(NST_ERR_INVALID, "nst_load_weights: null argument");
  CUDA_TRY(cudaSetDevice(h->device));
  struct T { const float* p; int64_t s[4]; };
  std::map<std::string, T> m;
  for (int i = 0; i < n_tensors; ++i) {
    T t; t.p = ptrs[i];
    for (int k = 0; k < 4; ++k) t.s[k] = shapes[i * 4 + k];
    m[names[i]] = t;
  }
  // ---- pass 1: validate every tensor (names, shapes) before anything on the handle is touched ----
  std::map<std::string, bool> used;
  int missing = 0;
  std::string first_problem;
  auto want = [&](const std::string& key, int64_t s0, int64_t s1, int64_t s2, int64_t s3) {
    auto it = m.find(key);
    char msg[384];
    if (it == m.end() || it->second.p == nullptr) {
      ++missing;
      if (first_problem.empty()) { snprintf(msg, sizeof(msg), "missing tensor '%s'", key.c_str()); first_problem = msg; }
      return;
    }
    const T& t = it->second;
    used[key] = true;
    if ((t.s[0] != s0 || t.s[1] != s1 || t.s[2] != s2 || t.s[3] != s3) && first_problem.empty()) {
      snprintf(msg, sizeof(msg), "'%s' has shape [%lld,%lld,%lld,%lld], expected [%lld,%lld,%lld,%lld]", key.c_str(), (long long)t.s[0],
               (long long)t.s[1], (long long)t.s[2], (long long)t.s[3], (long long)s0, (long long)s1, (long long)s2, (long long)s3);
      first_problem = msg;
    }
  };
  for (const LayerSpec& s : h->specs) {
    if (s.transposed) want(s.wkey, s.cin, s.cout, s.ks, s.ks); else want(s.wkey, s.cout, s.cin, s.ks, s.ks);
    want(s.bkey, s.cout, 1, 1, 1);
    if (!s.in_gkey.empty()) { want(s.in_gkey, s.cin, 1, 1, 1); want(s.in_bkey, s.cin, 1, 1, 1); }
  }
  int unexpected = 0;
  for (auto& kv : m) if (!used.count(kv.first)) ++unexpected;
  if (n_missing) *n_missing = missing;
  if (n_unexpected) *n_unexpected = unexpected;
  if (!first_problem.empty())
    return fail(NST_ERR_WEIGHTS, "nst_load_weights: %s (%d missing, %d unexpected); the handle keeps its previous weights", first_problem.c_str(), missing, unexpected);
  // ---- pass 2: repack and upload into fresh buffers; the handle switches over only when every layer succeeded ----
  auto get = [&](const std::string& key, std::vector<float>& dst) {
    const T& t = m[key];
    dst.assign(t.p, t.p + (size_t)(t.s[0] * t.s[1] * t.s[2] * t.s[3]));
  };
  std::vector<LayerDev> fresh(h->specs.size());
  auto abandon = [&](cudaError_t e, const char* what) {
    for (auto& d : fresh) free_layer(d);
    return fail(NST_ERR_CUDA, "nst_load_weights: %s: %s; the handle keeps its previous weights", what, cudaGetErrorString(e));
  };
  for (size_t li = 0; li < h->specs.size(); ++li) {
    const LayerSpec& s = h->specs[li];
    LayerDev& d = fresh[li];
    d.prec = h->dev[li].prec;
    std::vector<float> w_raw, w, bias, gam, bet;
    get(s.wkey, w_raw);
    get(s.bkey, bias);
    // canonical OIHW of the equivalent direct convolution
    w.resize((size_t)s.cout * s.cin * s.ks * s.ks);
    if (s.transposed) {
      // ConvTranspose2d(k3,s2,p1,op1) == conv over the zero-inserted input with the flipped kernel:
      // W_eq[co][ci][r][s] = W_T[ci][co][k-1-r][k-1-s]
      for (int co = 0; co < s.cout; ++co) for (int ci = 0; ci < s.cin; ++ci) for (int r = 0; r < s.ks; ++r) for (int q = 0; q < s.ks; ++q)
        w[(((size_t)co * s.cin + ci) * s.ks + r) * s.ks + q] = w_raw[(((size_t)ci * s.cout + co) * s.ks + (s.ks - 1 - r)) * s.ks + (s.ks - 1 - q)];
    } else {
      w = w_raw;
    }
    std::vector<float> wdir((size_t)s.ks * s.ks * s.cin * s.cout);
    for (int co = 0; co < s.cout; ++co) for (int ci = 0; ci < s.cin; ++ci) for (int r = 0; r < s.ks; ++r) for (int q = 0; q < s.ks; ++q)
      wdir[((size_t)(r * s.ks + q) * s.cin + ci) * s.cout + co] = w[(((size_t)co * s.cin + ci) * s.ks + r) * s.ks + q];
    cudaError_t e;
    if ((e = upload(&d.wdir, wdir)) != cudaSuccess) return abandon(e, "upload weights");
    if ((e = upload(&d.bias, bias)) != cudaSuccess) return abandon(e, "upload bias");
    if (!s.in_gkey.empty()) {
      get(s.in_gkey, gam); get(s.in_bkey, bet);
      if ((e = upload(&d.in_gamma, gam)) != cudaSuccess) return abandon(e, "upload gamma");
      if ((e = upload(&d.in_beta, bet)) != cudaSuccess) return abandon(e, "upload beta");
    }
    layer_static_plan(s, d);
    if (d.tc_ok) {
      LayerSpec ps = s;
      ps.cout = s.cout / d.nsplit;
      d.tc_static.dbg = h->dbg;
      for (int part = 0; part < d.nsplit; ++part) {
        std::vector<float> wpart(w.begin() + (size_t)part * ps.cout * s.cin * s.ks * s.ks, w.begin() + (size_t)(part + 1) * ps.cout * s.cin * s.ks * s.ks);
        std::vector<uint16_t> pk = tc_pack_weights(ps, d.tc_static, wpart);
        if ((e = upload((uint16_t**)(part == 0 ? &d.wpack : &d.wpack2), pk)) != cudaSuccess) return abandon(e, "upload packed weights");
      }
    }
  }
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) return abandon(e, "synchronize");
  for (size_t li = 0; li < h->specs.size(); ++li) { free_layer(h->dev[li]); h->dev[li] = fresh[li]; }
  h->weights_loaded = true;
  h->plan_N = 0;
  h->plans.clear();
  return NST_OK;
}

// ---------------------------------------------------------------------------------------------
// planning
// ---------------------------------------------------------------------------------------------
static void out_shape(int arch, int H, int W, int* Ho, int* Wo) {
  if (arch == NST_ARCH_NST) { *Ho = H; *Wo = W; return; }
  *Ho = 4 * ceil_div(ceil_div(H, 2), 2);
  *Wo = 4 * ceil_div(ceil_div(W, 2), 2);
}

extern "C" int nst_output_shape(const nst_handle* h, int H, int W, int* Ho, int* Wo) {
  if (h == nullptr || Ho == nullptr || Wo == nullptr) return fail(NST_ERR_INVALID, "nst_output_shape: null argument");
  out_shape(h->arch, H, W, Ho, Wo);
  return NST_OK;
}

// A layer takes the TMA-fed operand path when its source is an fp16 planar-8 activation tensor that the conv reads through the
// identity map (the sub-pixel forms of the up-sampling layers are identity-mapped over the low-res tensor), with one MMA term.
static bool tma_eligible(const ConvParams& q, const TcPlan& t, int in16) {
  if (getenv("NST_NO_TMA")) return false;
  if (!in16 || t.nterms != 1 || t.pair_mode || t.is_bf16 || t.a_tplanes != 1) return false;
  if (q.map_mode != MAP_IDENTITY || (q.stride != 1 && q.stride != 2)) return false;
  if (q.stride == 2 && (q.Hs < 2 || q.Ws < 2 || q.resid != nullptr)) return false;
  if (q.Cin % 8 != 0) return false;
  return true;
}

// Epilogue groups of a TMA-fed layer.  tcprof (profiles/r02_tcprof_*): on the stride-2 32->64 layer, the Cin <= 96 sub-pixel layers and
// the folded image layer the MMA warp waits 20-50 % of its time for a free accumulator while the transform warps idle > 50 %, so
// four of them become a second epilogue group; everywhere else the transform needs its warps.  NST_EPI2_MASK=<layer bitmask> overrides.
static int epi_groups_for(const LayerSpec& s, const TcPlan& t, int li) {
  if (const char* e = getenv("NST_EPI2_MASK")) return ((strtoull(e, nullptr, 0) >> li) & 1ull) ? 2 : 1;
  if (t.swap_ab || s.resid >= 0) return 1;
  if (t.fold_s) return 2;
  if (t.phase_mode) return s.cin <= 96 ? 2 : 1;
  if (s.stride == 2) return 2;
  return 1;
}

static int make_plan(nst_handle* h, int N, int H, int W) {
  if (N < 1 || H < 1 || W < 1) return fail(NST_ERR_INVALID, "bad shape N=%d H=%d W=%d", N, H, W);
  if (h->arch == NST_ARCH_NST && (H <= 40 || W <= 40))
    return fail(NST_ERR_INVALID, "NST variant needs H,W > 40 (ReflectionPad2d(40), transformer_net_nst.py:74); got %dx%d", H, W);
  if (h->arch != NST_ARCH_NST && (H < 5 || W < 5))
    return fail(NST_ERR_INVALID, "input %dx%d too small for ReflectionPad2d(4)", H, W);
  if (h->plan_N == N && h->plan_H == H && h->plan_W == W && !h->plans.empty()) return NST_OK;
  const int nb = std::min(N, h->microbatch);
  const int nl = (int)h->specs.size();
  h->plans.assign(h->specs.size(), LayerPlan());
  for (int b = 0; b < NBUF; ++b) h->bufs[b] = BufDims();
  int producer[NBUF];
  for (int b = 0; b < NBUF; ++b) producer[b] = -1;
  // who reads what: producer / consumer links and the element type of every activation tensor.  A tensor is fp32 exactly when its
  // consumer is error-compensated (F16X3), so plain fp16 layers always read fp16.
  for (int li = 0; li < nl; ++li) {
    const LayerSpec& s = h->specs[li];
    LayerPlan& lp = h->plans[li];
    if (s.src != BUF_IMAGE) {
      lp.producer = producer[s.src];
      if (lp.producer < 0) return fail(NST_ERR_STATE, "internal: layer %s reads an unwritten buffer", s.name.c_str());
      h->plans[lp.producer].consumer = li;
    }
    if (s.out != BUF_IMAGE) producer[s.out] = li;
  }
  for (int li = 0; li < nl; ++li) {
    LayerPlan& lp = h->plans[li];
    const bool x3 = h->dev[li].prec == NST_PREC_F16X3;
    const bool cons_x3 = lp.consumer >= 0 && h->dev[lp.consumer].prec == NST_PREC_F16X3;
    (void)x3;
    lp.out16 = h->act16 && !cons_x3;          // an error-compensated consumer splits fp32 values into hi + lo: its source stays fp32
    if (h->precision == NST_PREC_F16X3 || h->precision == NST_PREC_FP32_DIRECT) lp.out16 = h->act16;
  }
  size_t stats_off = 0;
  for (size_t li = 0; li < h->specs.size(); ++li) {
    const LayerSpec& s = h->specs[li];
    LayerPlan& lp = h->plans[li];
    ConvParams& p = lp.p;
    memset(&p, 0, sizeof(p));
    p.N = nb; p.Cin = s.cin; p.Cout = s.cout; p.ks = s.ks; p.stride = s.stride;
    p.map_mode = s.map_mode; p.pad_mode = s.pad_mode; p.prepad = 40;
    p.pad_t = p.pad_l = s.pad;
    p.pre_relu = s.pre_relu; p.post_relu = s.post_relu; p.tanh_out = s.tanh_out;
    p.blend = 1.f;
    p.co_off = 0; p.co_pass = s.cout / h->dev[li].nsplit;
    lp.in16 = s.src == BUF_IMAGE ? h->act16 : h->plans[lp.producer].out16;
    if ((s.resid >= 0 || s.side >= 0) && lp.in16 != h->act16)
      return fail(NST_ERR_STATE, "internal: layer %s mixes fp32 source with fp16 residual tensors", s.name.c_str());
    p.act16 = lp.in16; p.out16 = lp.out16;
    if (s.src == BUF_IMAGE) { p.Hs = H; p.Ws = W; }
    else {
      const BufDims& bd = h->bufs[s.src];
      if (bd.C != s.cin) return fail(NST_ERR_STATE, "internal: layer %s expects %d channels, buffer has %d", s.name.c_str(), s.cin, bd.C);
      p.Hs = bd.H; p.Ws = bd.W;
    }
    switch (s.map_mode) {
      case MAP_UP2: case MAP_ZERO_INSERT2: p.Hc = 2 * p.Hs; p.Wc = 2 * p.Ws; break;
      case MAP_PREPAD_REFLECT: p.Hc = p.Hs + 80; p.Wc = p.Ws + 80; break;
      default: p.Hc = p.Hs; p.Wc = p.Ws;
    }
    int Hfull = (p.Hc + 2 * s.pad - s.ks) / s.stride + 1;
    int Wfull = (p.Wc + 2 * s.pad - s.ks) / s.stride + 1;
    if (s.pad_mode == PAD_REFLECT && (s.pad >= p.Hc || s.pad >= p.Wc))
      return fail(NST_ERR_INVALID, "layer %s: reflection pad %d needs input larger than %dx%d", s.name.c_str(), s.pad, p.Hc, p.Wc);
    p.Ho = Hfull; p.Wo = Wfull;
    if (s.out == BUF_IMAGE) {
      out_shape(h->arch, H, W, &h->out_H, &h->out_W);
      if (h->arch == NST_ARCH_NST) {       // centre crop (transformer_net_nst.py:122-125)
        p.oy_off = (Hfull - H) / 2; p.ox_off = (Wfull - W) / 2;
        p.Ho = H; p.Wo = W;
      } else if (Hfull != h->out_H || Wfull != h->out_W) {
        return fail(NST_ERR_STATE, "internal: output %dx%d != expected %dx%d", Hfull, Wfull, h->out_H, h->out_W);
      }
    } else {
      BufDims& bd = h->bufs[s.out];
      bd.C = s.cout; bd.H = p.Ho; bd.W = p.Wo;
      lp.out_H = p.Ho; lp.out_W = p.Wo;
      bd.bytes = std::max(bd.bytes, (size_t)s.cout * p.Ho * p.Wo * (lp.out16 ? 2 : 4));
    }
    if (s.side >= 0) {
      BufDims& bd = h->bufs[s.side];
      bd.C = s.cin; bd.H = p.Hs; bd.W = p.Ws;
      bd.bytes = std::max(bd.bytes, (size_t)s.cin * p.Hs * p.Ws * (lp.in16 ? 2 : 4));
    }
    lp.flops = 2.0 * p.Ho * p.Wo * (double)s.cout * s.cin * s.ks * s.ks;      // per frame
    {   // algorithmic HBM bytes per frame: each operand once, image tensors at their own element size
      const double in_esz = lp.in16 ? 2.0 : 4.0, out_esz = lp.out16 ? 2.0 : 4.0;
      const double in_b = s.src == BUF_IMAGE ? 3.0 * p.Hs * p.Ws : in_esz * p.Hs * p.Ws * s.cin;
      const double out_b = s.out == BUF_IMAGE ? 3.0 * p.Ho * p.Wo : out_esz * p.Ho * p.Wo * s.cout;
      lp.bytes = in_b + out_b + in_esz * (double)p.Hs * p.Ws * s.cin * ((s.resid >= 0 ? 1 : 0) + (s.side >= 0 ? 1 : 0));
    }
    lp.use_tc = h->dev[li].tc_ok && ((h->tc_layer_mask >> li) & 1ull);
    int fTH = 0, fMB = 0;
    if (const char* f = getenv("NST_TC_FORCE")) {      // tuning aid: "layer:TH:MB;layer:TH:MB" restricts the tile search (tools/tile_sweep.py)
      for (const char* q = f; q && *q; ) {
        int l = -1, th = 0, mb = 0;
        if (sscanf(q, "%d:%d:%d", &l, &th, &mb) == 3 && l == (int)li) { fTH = th; fMB = mb; }
        q = strchr(q, ';'); if (q) ++q;
      }
    }
    if (lp.use_tc) {
      lp.t = h->dev[li].tc_static;
      if (!lp.out16 || s.out == BUF_IMAGE) lp.t.swap_ab = 0;                 // the swapped epilogue stores fp16 planar-8
      ConvParams q = p;
      if (lp.t.phase_mode) {
        // tile over the low-res grid; the epilogue scatters the four phases into the [N, 2Hs, 2Ws, Cout] output
        q.phase_mode = lp.t.phase_mode; q.map_mode = MAP_IDENTITY; q.pad_mode = lp.t.phase_mode == 1 ? PAD_REPLICATE : PAD_ZERO;
        q.Hc = p.Hs; q.Wc = p.Ws; q.Ho = p.Hs; q.Wo = p.Ws;
      }
      q.resid = s.resid >= 0 ? (const void*)1 : nullptr;                   // plan-time flags: the layer has a residual operand / a side output
      q.side_out = s.side >= 0 ? (void*)1 : nullptr;
      lp.t.tma_mode = tma_eligible(q, lp.t, lp.in16) ? 1 : 0;
      lp.t.rr_tiles = getenv("NST_NO_RR") ? 0 : 1;
      lp.t.epi_groups = 1;
      if (lp.t.tma_mode) {
        lp.t.epi_groups = epi_groups_for(s, lp.t, (int)li);
        lp.t.prod_warps = std::min(lp.t.prod_warps, lp.t.epi_groups == 2 ? TC_PROD_WARPS - 5 : TC_PROD_WARPS - 1);
      }
      bool ok = tc_plan_geometry(q, lp.t, fTH, fMB);
      if (!ok && lp.t.tma_mode) {                                            // shapes the TMA boxes cannot express fall back to the gather producers
        const int swap = lp.t.swap_ab;
        lp.t = h->dev[li].tc_static;
        lp.t.swap_ab = swap;
        lp.t.epi_groups = 1;
        lp.t.rr_tiles = getenv("NST_NO_RR") ? 0 : 1;
        ok = tc_plan_geometry(q, lp.t, fTH, fMB);
      }
      q.resid = nullptr; q.side_out = nullptr;
      if (!ok)
        return fail(NST_ERR_INVALID, "layer %s (%d->%d k%d s%d, %dx%d): no tcgen05 tiling fits; only NST_PREC_FP32_DIRECT runs this shape",
                    s.name.c_str(), s.cin, s.cout, s.ks, s.stride, p.Hs, p.Ws);
      p = q;
    } else if (h->precision != NST_PREC_FP32_DIRECT && ((h->tc_layer_mask >> li) & 1ull)) {
      // never degrade silently to the fp32 FFMA kernel (~50x slower): the caller asked for a tensor-core precision
      return fail(NST_ERR_INVALID, "layer %s (%d->%d k%d s%d) has no tcgen05 form; only NST_PREC_FP32_DIRECT runs it", s.name.c_str(), s.cin,
                  s.cout, s.ks, s.stride);
    }
    if (s.out != BUF_IMAGE) {
      lp.parts = lp.use_tc ? h->num_sms * 2 : 1;                    // (CTA, epilogue group) slots
      lp.part_off = stats_off;
      stats_off += round_up((size_t)lp.parts * nb * s.cout * 2 * sizeof(double), 256);
      lp.coeff_off = stats_off;
      stats_off += round_up((size_t)nb * s.cout * sizeof(float2), 256);
    }
  }
  // A stride-2 TMA-fed layer reads its source through four (row, column) parity views.  When the producer stores every row as
  // [even columns | odd columns] ("column-split"), a view's columns are contiguous in memory and its boxes move whole 16*P-byte runs;
  // with the natural layout the TMA unit gathers 16-byte pieces 32 bytes apart and became the limiter of these layers (tcprof: the
  // MMA warp of conv3 waited 60 % of its time for operands).  Only tensors whose single reader is such a layer are split.
  if (!getenv("NST_NO_COLSPLIT")) {
    for (int li = 0; li < nl; ++li) {
      LayerPlan& lp = h->plans[li];
      const int c = lp.consumer;
      if (c < 0 || !lp.use_tc || lp.t.phase_mode || h->specs[li].out == BUF_IMAGE) continue;
      const LayerPlan& cp = h->plans[c];
      if (!cp.use_tc || !cp.t.tma_mode || h->specs[c].stride != 2 || h->specs[c].resid >= 0 || h->specs[c].side >= 0) continue;
      lp.p.out_colsplit = 1;
      h->plans[c].p.in_colsplit = 1;
    }
  }
  for (int li = 0; li < nl; ++li) {
    LayerPlan& lp = h->plans[li];
    lp.t.tma_wide = (lp.use_tc && lp.t.tma_mode && !getenv("NST_NO_TMA_WIDE") && (lp.p.stride == 1 || lp.p.in_colsplit) && lp.t.P <= 128) ? 1 : 0;
  }
  size_t off = 0;
  for (int b = 0; b < NBUF; ++b) {
    h->bufs[b].off = off;
    off += round_up(h->bufs[b].bytes, 256);
  }
  h->act_bytes_per_frame = off;
  h->stats_bytes = stats_off;
  // out01 planes of one micro-batch for the resize-back pass of the uint8 "fit" surface (sizes that are not multiples of 4)
  h->fit_bytes = (h->out_H != H || h->out_W != W) ? round_up((size_t)nb * 3 * h->out_H * h->out_W * sizeof(float), 256) : 0;
  h->workspace_bytes = round_up(off * nb, 256) + round_up(stats_off, 256) + h->fit_bytes + 256;
  h->plan_N = N; h->plan_H = H; h->plan_W = W; h->plan_nb = nb;
  return NST_OK;
}

extern "C" int nst_workspace_bytes(nst_handle* h, int N, int H, int W, size_t* bytes) {
  if (h == nullptr || bytes == nullptr) return fail(NST_ERR_INVALID, "nst_workspace_bytes: null argument");
  if (!h->weights_loaded) return fail(NST_ERR_STATE, "nst_workspace_bytes: call nst_load_weights first");
  int rc = make_plan(h, N, H, W);
  if (rc != NST_OK) return rc;
  *bytes = h->workspace_bytes;
  return NST_OK;
}

// ---------------------------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------------------------
struct IoSpec {
  int in_fmt, out_fmt;
  const void* x; void* y;
  float in_scale[3], in_shift[3], out_div[3], out_scale[3], out_shift[3];
  int swap_rb; double blend; const float* mask; int composite_mode;
  int fit;        // uint8 surface only: the result is H x W; frames whose size is not a multiple of 4 are resized back (pipeline.py:1512-1516)
};

// cuTensorMapEncodeTiled through the runtime's driver-entry-point lookup (libcuda is not linked)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess && qres == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)ptr;
  }
  return fn;
}

// Tensor map over a view of a planar-8 fp16 activation tensor.  `wide` (every contiguous view): 3-D {2*W 8-byte elements, H, N*C/8 planes},
// box {2*box_w, box_h, 1} -- a box row is one contiguous run of box_w slots.  (With the natural 4-D description {8 channels, W, H,
// planes} the inner box dimension is a single 16-byte slot and the TMA unit moves ~16 B per clock: measured, it was the limiter of
// the stride-2 layers and close to it on the residual ones.)  Otherwise 4-D fp16 {8, W, H, planes} with `sx` = 2 selecting every second
// column (parity views of a source stored in natural column order).  `sy` = 2 selects every second row starting at `base`.
static int encode_act_map(CUtensorMap* m, const void* base, int W, int H, size_t planes, int sx, int sy, size_t row_bytes, size_t plane_bytes,
                          int box_w, int box_h, bool wide) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (fn == nullptr) return fail(NST_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
  CUresult r;
  if (wide) {
    const cuuint64_t dims[3] = {(cuuint64_t)(2 * W), (cuuint64_t)H, (cuuint64_t)planes};
    const cuuint64_t strides[2] = {(cuuint64_t)(row_bytes * sy), (cuuint64_t)plane_bytes};
    const cuuint32_t box[3] = {(cuuint32_t)(2 * box_w), (cuuint32_t)box_h, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    r = fn(m, CU_TENSOR_MAP_DATA_TYPE_UINT64, 3, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  } else {
    const cuuint64_t dims[4] = {8, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)planes};
    const cuuint64_t strides[3] = {(cuuint64_t)(16 * sx), (cuuint64_t)(row_bytes * sy), (cuuint64_t)plane_bytes};
    const cuuint32_t box[4] = {8, (cuuint32_t)box_w, (cuuint32_t)box_h, 1};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  if (r != CUDA_SUCCESS)
    return fail(NST_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) for view %dx%d planes=%zu strides=%d/%zu/%zu box=%dx%d wide=%d", (int)r, W, H, planes,
                16 * sx, row_bytes * sy, plane_bytes, box_w, box_h, (int)wide);
  return NST_OK;
}

static int encode_layer_maps(LayerPlan& lp, const ConvParams& p, const void* src, const void* resid, int nb) {
  const TcPlan& t = lp.t;
  const size_t planes = (size_t)nb * (p.Cin >> 3);
  const size_t row_bytes = (size_t)p.Ws * 16, plane_bytes = (size_t)p.Hs * p.Ws * 16;
  int rc;
  const bool wide = t.tma_wide != 0;
  if (p.stride == 1) {
    if ((rc = encode_act_map(&lp.tmap[0], src, p.Ws, p.Hs, planes, 1, 1, row_bytes, plane_bytes, t.P, t.tma_rows, wide)) != NST_OK) return rc;
    if (resid != nullptr && (rc = encode_act_map(&lp.tmap_res, resid, p.Ws, p.Hs, planes, 1, 1, row_bytes, plane_bytes, t.P, t.tma_rows, wide)) != NST_OK) return rc;
  } else {
    for (int py = 0; py < 2; ++py) for (int px = 0; px < 2; ++px) {
      // column-split source: the odd columns of a row start (Ws + 1) / 2 slots in and are contiguous
      const char* base = (const char*)src + ((size_t)py * p.Ws + (p.in_colsplit ? (size_t)px * ((p.Ws + 1) / 2) : (size_t)px)) * 16;
      rc = encode_act_map(&lp.tmap[py * 2 + px], base, (p.Ws - px + 1) / 2, (p.Hs - py + 1) / 2, planes, p.in_colsplit ? 1 : 2, 2, row_bytes, plane_bytes,
                          t.P, t.tma_rows, wide);
      if (rc != NST_OK) return rc;
    }
  }
  return NST_OK;
}

// Rotation of the round-robin tile sequence (decode_tile in conv_tc.cu): simulate the per-CTA cost (border tiles run a second pass
// in the transform and store less) for a few rotations and keep the flattest.  Deterministic in (tiling, batch, grid).
static int pick_rotation(const TcPlan& t, int n_frames, int G) {
  if (getenv("NST_NO_ROT")) return 0;
  const int per_img = t.tiles_x * t.tiles_y, T = per_img * n_frames;
  if (G < 2 || T < 2 * G) return 0;
  static const int cand[] = {0, 1, 7, 13, 29, 41, 53, 61, 71, 89, 97, 113};
  int best = 0; long best_cost = -1;
  std::vector<int> cost(G);
  for (int rot : cand) {
    if (rot >= G) break;
    std::fill(cost.begin(), cost.end(), 0);
    for (int v = 0; v < T; ++v) {
      const int round = v / G, b = v - round * G;
      const int tile = (round + 1) * G <= T ? round * G + (b + round * rot) % G : v;
      const int r = tile % per_img, tx = r / t.tiles_y, ty = r - tx * t.tiles_y;
      const bool border = tx == 0 || ty == 0 || tx == t.tiles_x - 1 || ty == t.tiles_y - 1;
      cost[b] += border ? 3 : 2;
    }
    const long mx = *std::max_element(cost.begin(), cost.end());
    if (best_cost < 0 || mx < best_cost) { best_cost = mx; best = rot; }
  }
  return best;
}

static void print_tcprof(nst_handle* h, int n_frames) {
  std::vector<unsigned long long> all(64 * 16 + 64 * 2 * 148);
  if (cudaMemcpy(all.data(), h->prof_dev, all.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost) != cudaSuccess) return;
  for (size_t li = 0; li < h->specs.size() && li < 64; ++li) {
    const LayerPlan& lp = h->plans[li];
    if (!lp.use_tc) continue;
    const unsigned long long* pc = all.data() + li * 16;
    const int tiles = lp.t.tiles_x * lp.t.tiles_y * n_frames;
    const double g = std::min(tiles, h->num_sms) * (double)h->dev[li].nsplit;
    const double mhz = pc[15] ? 1e3 * (double)pc[14] / (double)pc[15] : 0.0;                  // CTA 0: SM cycles / ns
    fprintf(stderr, "[tcprof] %-12s %s%s EG=%d tiles=%d | %.0f MHz | per-CTA kcycles: loader total %.0f wait B_EMPTY %.0f | mma total %.0f wait ACC_EMPTY %.0f A_FULL %.0f B_FULL %.0f | "
                    "epi total %.0f wait ACC_FULL %.0f | prod total %.0f wait A_EMPTY %.0f | tma-issue total %.0f, transform waits TMA_FULL %.0f\n",
            h->specs[li].name.c_str(), lp.t.tma_mode ? "tma" : "gather", lp.t.swap_ab ? "/T" : "", lp.t.epi_groups, tiles, mhz,
            pc[8] / g / 1e3, pc[1] / g / 1e3, pc[9] / g / 1e3, pc[2] / g / 1e3, pc[3] / g / 1e3, pc[4] / g / 1e3, pc[10] / g / 1e3, pc[5] / g / 1e3,
            pc[11] / g / 1e3, pc[6] / g / 1e3, pc[12] / g / 1e3, pc[7] / g / 1e3);
    // wall-clock spans of the CTAs (last micro-batch, last split launch): start skew, mean busy span, tail
    const unsigned long long* ct = all.data() + 64 * 16 + li * 2 * 148;
    const int ctas = std::min(tiles, h->num_sms);
    unsigned long long s0 = ~0ull, s1 = 0, e0 = ~0ull, e1 = 0; double span = 0;
    for (int b = 0; b < ctas; ++b) { s0 = std::min(s0, ct[2 * b]); s1 = std::max(s1, ct[2 * b]); e0 = std::min(e0, ct[2 * b + 1]); e1 = std::max(e1, ct[2 * b + 1]); span += (double)(ct[2 * b + 1] - ct[2 * b]); }
    fprintf(stderr, "[tcspan] %-12s kernel %.1f us = first start .. last end | starts spread %.1f us | CTA span mean %.1f us | first end at %.1f us\n",
            h->specs[li].name.c_str(), (e1 - s0) / 1e3, (s1 - s0) / 1e3, span / ctas / 1e3, (e0 - s0) / 1e3);
  }
}

static int run_forward(nst_handle* h, const IoSpec& io, int N, int H, int W, void* workspace, size_t workspace_bytes,
                       cudaStream_t stream, float* layer_ms) {
  if (h == nullptr) return fail(NST_ERR_INVALID, "forward: null handle");
  if (!h->weights_loaded) return fail(NST_ERR_STATE, "forward: call nst_load_weights first");
  if (io.x == nullptr || io.y == nullptr) return fail(NST_ERR_INVALID, "forward: null tensor");
  CUDA_TRY(cudaSetDevice(h->device));
  int rc = make_plan(h, N, H, W);
  if (rc != NST_OK) return rc;
  if (workspace == nullptr || workspace_bytes < h->workspace_bytes)
    return fail(NST_ERR_WORKSPACE, "forward: workspace %zu bytes < required %zu", workspace_bytes, h->workspace_bytes);
  if (((uintptr_t)workspace & 255) != 0) return fail(NST_ERR_INVALID, "forward: workspace must be 256-byte aligned");
  if (*h->err_host != 0) return fail(NST_ERR_CUDA, "forward: a previous kernel reported pipeline error 0x%08x", *h->err_host);
  const int nb = h->plan_nb;
  char* act = (char*)workspace;
  char* stats = (char*)workspace + round_up(h->act_bytes_per_frame * nb, 256);
  const int Ho = h->out_H, Wo = h->out_W;
  const bool fit_resize = io.fit && (Ho != H || Wo != W);
  float* fit_buf = (float*)(stats + round_up(h->stats_bytes, 256));
  const size_t in_px = (size_t)H * W * 3, out_px = fit_resize ? in_px : (size_t)Ho * Wo * 3;
  const size_t in_elt = io.in_fmt == IN_U8_NHWC3 ? 1 : 4, out_elt = io.out_fmt == OUT_U8_NHWC3 ? 1 : 4;
  const bool blend_on = io.blend >= 0.0 && io.blend < 1.0;                 // pipeline.py:2089
  std::vector<cudaEvent_t> ev;
  if (layer_ms) {
    ev.resize(h->specs.size() + 1);
    for (auto& e : ev) CUDA_TRY(cudaEventCreate(&e));
    for (size_t i = 0; i < h->specs.size(); ++i) layer_ms[i] = 0.f;
  }
  auto buf = [&](int b) -> void* { return act + h->bufs[b].off * nb; };
  if (layer_ms && h->prof_dev) CUDA_TRY(cudaMemsetAsync(h->prof_dev, 0, (64 * 16 + 64 * 2 * 148) * sizeof(unsigned long long), stream));
  for (int n0 = 0; n0 < N; n0 += nb) {
    const int n = std::min(nb, N - n0);
    if (layer_ms) CUDA_TRY(cudaEventRecord(ev[0], stream));
    for (size_t li = 0; li < h->specs.size(); ++li) {
      const LayerSpec& s = h->specs[li];
      const LayerDev& d = h->dev[li];
      LayerPlan& lp = h->plans[li];
      ConvParams p = lp.p;
      p.N = n;
      if (s.src == BUF_IMAGE) {
        p.in_fmt = io.in_fmt;
        p.src = (const char*)io.x + (size_t)n0 * in_px * in_elt;
        for (int c = 0; c < 3; ++c) { p.in_scale[c] = io.in_scale[c]; p.in_shift[c] = io.in_shift[c]; }
        p.swap_rb = io.swap_rb;
      } else {
        p.in_fmt = IN_F32_NHWC;
        p.src = buf(s.src);
        p.in_coeff = (const float2*)(stats + h->plans[lp.producer].coeff_off);
      }
      p.resid = s.resid >= 0 ? buf(s.resid) : nullptr;
      p.side_out = s.side >= 0 ? buf(s.side) : nullptr;
      p.bias = d.bias;
      if (s.out == BUF_IMAGE) {
        p.out_fmt = fit_resize ? (int)OUT_F32_01_NCHW3 : io.out_fmt;
        p.out = fit_resize ? (void*)fit_buf : (void*)((char*)io.y + (size_t)n0 * out_px * out_elt);
        for (int c = 0; c < 3; ++c) { p.out_div[c] = io.out_div[c]; p.out_scale[c] = io.out_scale[c]; p.out_shift[c] = io.out_shift[c]; }
        p.swap_rb = io.swap_rb;
        // Python evaluates `blend * S + (1.0 - blend) * O` with double scalars that torch then narrows to fp32 (pipeline.py:2090)
        p.blend = blend_on ? (float)io.blend : 2.f;
        p.blend_comp = (float)(1.0 - io.blend);
        p.composite_mode = io.composite_mode;
        p.mask = (io.mask && !fit_resize) ? io.mask + (size_t)n0 * Ho * Wo : nullptr;
        p.orig = (io.in_fmt == IN_U8_NHWC3 && !fit_resize && (io.mask != nullptr || blend_on)) ? (const uint8_t*)io.x + (size_t)n0 * in_px : nullptr;
      } else {
        p.out_fmt = OUT_F32_NHWC_STATS;
        p.out = buf(s.out);
        p.out_part = (double*)(stats + lp.part_off);
      }
      int parts = 1, groups = 1, tiles_per_img = 0, rot = 0;
      if (lp.use_tc) {
        TcArgs a;
        a.p = p; a.t = lp.t;
        a.t.num_tiles = lp.t.tiles_x * lp.t.tiles_y * n;
        a.err_flag = h->err_dev;
        if (lp.t.tma_mode) {
          if (lp.maps_ws != workspace) {
            if ((rc = encode_layer_maps(lp, p, p.src, p.resid, nb)) != NST_OK) return rc;
            lp.maps_ws = workspace;
          }
          memcpy(a.tmap, lp.tmap, sizeof(a.tmap));
          a.tmap_res = lp.tmap_res;
        } else {
          memset(a.tmap, 0, sizeof(a.tmap)); memset(&a.tmap_res, 0, sizeof(a.tmap_res));
        }
        parts = conv_tc_grid(a.t, a.t.num_tiles, h->num_sms);
        if (a.t.rr_tiles) {                                    // cached per micro-batch size: the simulation is host work
          auto it = lp.rot_cache.find(n);
          if (it == lp.rot_cache.end()) it = lp.rot_cache.emplace(n, pick_rotation(lp.t, n, parts)).first;
          rot = it->second;
        }
        a.t.rr_rot = rot;
        groups = lp.t.epi_groups;
        tiles_per_img = lp.t.tiles_x * lp.t.tiles_y;
        for (int part = 0; part < d.nsplit; ++part) {
        a.wpack = part == 0 ? d.wpack : d.wpack2;
        a.p.co_off = part * a.p.co_pass;
        a.prof = (layer_ms && h->prof_dev && li < 64) ? h->prof_dev + li * 16 : nullptr;
        a.prof_cta = (a.prof && h->num_sms <= 148) ? h->prof_dev + 64 * 16 + li * 2 * 148 : nullptr;      // read back after the whole forward: no
        CUDA_TRY(launch_conv_tc(a, h->num_sms, stream));                                      // per-layer synchronisation, clocks stay in steady state
        if (part + 1 < d.nsplit) ++h->launches;
        }
      } else {
        if (s.out != BUF_IMAGE) CUDA_TRY(cudaMemsetAsync(p.out_part, 0, (size_t)n * s.cout * 2 * sizeof(double), stream));
        CUDA_TRY(launch_conv_direct(p, d.wdir, stream));
      }
      ++h->launches;
      if (s.out != BUF_IMAGE) {
        // fold the per-CTA partial statistics into the (scale, shift) the consumer applies (its gamma / beta belong to the norm in between)
        if (lp.consumer < 0) return fail(NST_ERR_STATE, "internal: layer %s has no consumer", s.name.c_str());
        const LayerDev& cd = h->dev[lp.consumer];
        CUDA_TRY(launch_stats_finalize(p.out_part, parts, groups, lp.use_tc ? lp.t.rr_tiles : 0, lp.use_tc ? rot : 0, n, s.cout, tiles_per_img, 1.0 / ((double)lp.out_H * lp.out_W), cd.in_gamma, cd.in_beta,
                                       (float2*)(stats + lp.coeff_off), stream));
        ++h->launches;
      }
      if (layer_ms) CUDA_TRY(cudaEventRecord(ev[li + 1], stream));
    }
    if (fit_resize) {
      // "lock stylised to content size" (pipeline.py:1512-1516), then the mask composite / blend / pack the image epilogue skipped
      ResizePackArgs ra;
      ra.src01 = fit_buf; ra.out = (uint8_t*)io.y + (size_t)n0 * in_px;
      ra.orig = (io.mask != nullptr || blend_on) ? (const uint8_t*)io.x + (size_t)n0 * in_px : nullptr;
      ra.mask = io.mask ? io.mask + (size_t)n0 * H * W : nullptr;
      ra.N = n; ra.Hi = Ho; ra.Wi = Wo; ra.H = H; ra.W = W;
      ra.blend = blend_on ? (float)io.blend : 2.f; ra.blend_comp = (float)(1.0 - io.blend); ra.composite_mode = io.composite_mode;
      CUDA_TRY(launch_resize_post_pack(ra, stream));
      ++h->launches;
    }
    if (layer_ms) {
      CUDA_TRY(cudaStreamSynchronize(stream));
      for (size_t li = 0; li < h->specs.size(); ++li) {
        float ms = 0.f;
        CUDA_TRY(cudaEventElapsedTime(&ms, ev[li], ev[li + 1]));
        layer_ms[li] += ms;
      }
    }
  }
  if (layer_ms) for (auto& e : ev) cudaEventDestroy(e);
  if (layer_ms && h->prof_dev) print_tcprof(h, std::min(N, nb));
  return NST_OK;
}

extern "C" int nst_forward_f32_nchw(nst_handle* h, const float* x, float* y, int N, int H, int W, void* workspace,
                                    size_t workspace_bytes, void* stream) {
  IoSpec io;
  memset(&io, 0, sizeof(io));
  io.in_fmt = IN_F32_NCHW3; io.out_fmt = OUT_F32_NCHW3; io.x = x; io.y = y;
  for (int c = 0; c < 3; ++c) { io.in_scale[c] = 1.f; io.out_div[c] = 1.f; io.out_scale[c] = 1.f; }
  io.blend = 1.0;
  return run_forward(h, io, N, H, W, workspace, workspace_bytes, (cudaStream_t)stream, nullptr);
}

static int fill_u8_io(nst_handle* h, IoSpec& io, const uint8_t* x, uint8_t* y, int H, int W, const float in_scale[3],
                      const float in_shift[3], const float out_div[3], const float out_scale[3], const float out_shift[3], int swap_rb,
                      double blend, const float* mask, int composite_mode, int fit = 0) {
  if (h == nullptr) return fail(NST_ERR_INVALID, "forward_u8: null handle");
  if (!in_scale || !in_shift || !out_div || !out_scale || !out_shift) return fail(NST_ERR_INVALID, "forward_u8: null affine");
  for (int c = 0; c < 3; ++c) if (out_div[c] == 0.f) return fail(NST_ERR_INVALID, "forward_u8: out_div[%d] is zero", c);
  int Ho, Wo;
  out_shape(h->arch, H, W, &Ho, &Wo);
  // H or W not a multiple of 4: the stride-2 convs round up, the output is the larger Ho x Wo frame (nst_output_shape) exactly like
  // the reference module's; the caller resizes it back (pipeline.py:1512-1516, frame_pipeline.py).  Blending with the H x W original
  // is only defined after that resize.
  if (!fit && (Ho != H || Wo != W) && (mask != nullptr || (blend >= 0.0 && blend < 1.0)))
    return fail(NST_ERR_INVALID, "forward_u8: blend / mask need H,W multiples of 4 for this architecture (got %dx%d -> %dx%d); resize first", H, W, Ho, Wo);
  memset(&io, 0, sizeof(io));
  io.in_fmt = IN_U8_NHWC3; io.out_fmt = OUT_U8_NHWC3; io.x = x; io.y = y;
  for (int c = 0; c < 3; ++c) {
    io.in_scale[c] = in_scale[c]; io.in_shift[c] = in_shift[c];
    io.out_div[c] = out_div[c]; io.out_scale[c] = out_scale[c]; io.out_shift[c] = out_shift[c];
  }
  io.swap_rb = swap_rb; io.blend = blend; io.mask = mask; io.composite_mode = composite_mode; io.fit = fit;
  return NST_OK;
}

extern "C" int nst_forward_u8_nhwc(nst_handle* h, const uint8_t* x, uint8_t* y, int N, int H, int W, const float in_scale[3],
                                   const float in_shift[3], const float out_div[3], const float out_scale[3], const float out_shift[3],
                                   int swap_rb, double blend, const float* mask, int composite_mode, void* workspace,
                                   size_t workspace_bytes, void* stream) {
  IoSpec io;
  int rc = fill_u8_io(h, io, x, y, H, W, in_scale, in_shift, out_div, out_scale, out_shift, swap_rb, blend, mask, composite_mode);
  if (rc != NST_OK) return rc;
  return run_forward(h, io, N, H, W, workspace, workspace_bytes, (cudaStream_t)stream, nullptr);
}

extern "C" int nst_forward_u8_nhwc_fit(nst_handle* h, const uint8_t* x, uint8_t* y, int N, int H, int W, const float in_scale[3],
                                       const float in_shift[3], const float out_div[3], const float out_scale[3], const float out_shift[3],
                                       int swap_rb, double blend, const float* mask, int composite_mode, void* workspace,
                                       size_t workspace_bytes, void* stream) {
  IoSpec io;
  int rc = fill_u8_io(h, io, x, y, H, W, in_scale, in_shift, out_div, out_scale, out_shift, swap_rb, blend, mask, composite_mode, 1);
  if (rc != NST_OK) return rc;
  return run_forward(h, io, N, H, W, workspace, workspace_bytes, (cudaStream_t)stream, nullptr);
}

extern "C" int nst_profile_layers_u8(nst_handle* h, const uint8_t* x, uint8_t* y, int N, int H, int W, void* workspace,
                                     size_t workspace_bytes, void* stream, float* ms) {
  if (ms == nullptr) return fail(NST_ERR_INVALID, "nst_profile_layers_u8: ms is null");
  const float one[3] = {1.f, 1.f, 1.f}, zero[3] = {0.f, 0.f, 0.f}, d255[3] = {255.f, 255.f, 255.f};
  IoSpec io;
  int rc = fill_u8_io(h, io, x, y, H, W, one, zero, d255, one, zero, 0, 1.0, nullptr, 0);
  if (rc != NST_OK) return rc;
  return run_forward(h, io, N, H, W, workspace, workspace_bytes, (cudaStream_t)stream, ms);
}

// ---------------------------------------------------------------------------------------------
// introspection
// ---------------------------------------------------------------------------------------------
extern "C" int nst_num_layers(const nst_handle* h) { return h ? (int)h->specs.size() : 0; }

extern "C" int nst_layer_info(const nst_handle* h, int layer, char* name, int name_cap, double* flops, double* bytes, int* kernel_kind) {
  if (h == nullptr || layer < 0 || layer >= (int)h->specs.size()) return fail(NST_ERR_INVALID, "nst_layer_info: bad layer %d", layer);
  if (h->plans.empty()) return fail(NST_ERR_STATE, "nst_layer_info: no plan yet (call nst_workspace_bytes or a forward)");
  if (name && name_cap > 0) { strncpy(name, h->specs[layer].name.c_str(), name_cap - 1); name[name_cap - 1] = 0; }
  if (flops) *flops = h->plans[layer].flops;
  if (bytes) *bytes = h->plans[layer].bytes;
  if (kernel_kind) *kernel_kind = h->plans[layer].use_tc ? (h->plans[layer].t.tma_mode ? 2 : 1) : 0;
  return NST_OK;
}

extern "C" long long nst_launch_count(const nst_handle* h) { return h ? h->launches : 0; }

// ---------------------------------------------------------------------------------------------
// compositing
// ---------------------------------------------------------------------------------------------
static int composite_common(const void* const* srcs, int K, const void* original, const float* masks, int R, int J,
                            const int* region_idx, const float* region_w, void* out, int H, int W, void* stream, bool u8) {
  if (K < 1 || K > COMP_MAX_K || R < 1 || R > COMP_MAX_R || J < 1 || J > COMP_MAX_J)
    return fail(NST_ERR_INVALID, "composite: need 1<=K<=%d, 1<=R<=%d, 1<=J<=%d (got K=%d R=%d J=%d)", COMP_MAX_K, COMP_MAX_R, COMP_MAX_J, K, R, J);
  if (!srcs || !masks || !region_idx || !region_w || !out || H < 1 || W < 1) return fail(NST_ERR_INVALID, "composite: null / empty argument");
  CompositeArgs a;
  memset(&a, 0, sizeof(a));
  for (int k = 0; k < K; ++k) { if (!srcs[k]) return fail(NST_ERR_INVALID, "composite: srcs[%d] is null", k); a.srcs[k] = srcs[k]; }
  a.original = original; a.masks = masks; a.out = out; a.K = K; a.R = R; a.J = J; a.H = H; a.W = W;
  for (int i = 0; i < R * J; ++i) {
    if (region_idx[i] >= K || region_idx[i] < -1) return fail(NST_ERR_INVALID, "composite: region index %d out of range", region_idx[i]);
    if (region_idx[i] == -1 && original == nullptr && region_w[i] != 0.f)
      return fail(NST_ERR_INVALID, "composite: assignment includes original (-1) but no original frame provided");
    a.idx[i] = region_idx[i]; a.w[i] = region_w[i];
  }
  CUDA_TRY(launch_composite(a, u8, (cudaStream_t)stream));
  return NST_OK;
}

extern "C" int nst_composite_regions_u8(const uint8_t* const* srcs, int K, const uint8_t* original, const float* masks, int R, int J,
                                        const int* region_idx, const float* region_w, uint8_t* out, int H, int W, void* stream) {
  return composite_common((const void* const*)srcs, K, original, masks, R, J, region_idx, region_w, out, H, W, stream, true);
}

extern "C" int nst_composite_regions_f32(const float* const* srcs, int K, const float* original, const float* masks, int R, int J,
                                         const int* region_idx, const float* region_w, float* out, int H, int W, void* stream) {
  return composite_common((const void* const*)srcs, K, original, masks, R, J, region_idx, region_w, out, H, W, stream, false);
}

extern "C" int nst_composite_regions_batch_u8(const uint8_t* const* srcs, const int* src_hw, int K, const uint8_t* original, const void* masks,
                                              int mask_type, int masks_per_frame, int R, int J, const int* region_idx, const float* region_w,
                                              uint8_t* out, int N, int H, int W, void* stream) {
  if (K < 1 || K > CB_MAX_K || R < 1 || R > CB_MAX_R || J < 1 || J > CB_MAX_J)
    return fail(NST_ERR_INVALID, "composite_batch: need 1<=K<=%d, 1<=R<=%d, 1<=J<=%d (got K=%d R=%d J=%d)", CB_MAX_K, CB_MAX_R, CB_MAX_J, K, R, J);
  if (!srcs || !masks || !region_idx || !region_w || !out || N < 1 || H < 1 || W < 1) return fail(NST_ERR_INVALID, "composite_batch: null / empty argument");
  if (mask_type < 0 || mask_type > 2) return fail(NST_ERR_INVALID, "composite_batch: mask_type must be 0 (fp32), 1 (fp16) or 2 (uint8)");
  CompositeBatchArgs a;
  memset(&a, 0, sizeof(a));
  for (int k = 0; k < K; ++k) {
    if (!srcs[k]) return fail(NST_ERR_INVALID, "composite_batch: srcs[%d] is null", k);
    a.srcs[k] = srcs[k];
    a.src_h[k] = src_hw ? src_hw[2 * k] : H; a.src_w[k] = src_hw ? src_hw[2 * k + 1] : W;
    if (a.src_h[k] < 1 || a.src_w[k] < 1) return fail(NST_ERR_INVALID, "composite_batch: source %d has extent %dx%d", k, a.src_h[k], a.src_w[k]);
  }
  a.original = original; a.masks = masks; a.out = out; a.N = N; a.K = K; a.R = R; a.J = J; a.H = H; a.W = W;
  a.mask_type = mask_type; a.mask_frame_stride = masks_per_frame ? (size_t)R * H * W : 0;
  for (int i = 0; i < R * J; ++i) {
    if (region_idx[i] >= K || region_idx[i] < -2) return fail(NST_ERR_INVALID, "composite_batch: region index %d out of range", region_idx[i]);
    if (region_idx[i] == -1 && original == nullptr)
      return fail(NST_ERR_INVALID, "composite_batch: assignment includes original (-1) but no original frame provided");
    a.idx[i] = region_idx[i]; a.w[i] = region_w[i];
  }
  CUDA_TRY(launch_composite_batch(a, (cudaStream_t)stream));
  return NST_OK;
}

extern "C" int nst_composite_from_crops_f32(int R, const int* bbox, const float* const* crop_masks, const int* mask_hw, const int* nterms,
                                            const float* const* term_src, const int* term_hw, const float* term_w, const float* original,
                                            float* out, uint8_t* out_u8, float* scratch, int H, int W, void* stream) {
  if (R < 1 || R > CB_MAX_R) return fail(NST_ERR_INVALID, "composite_from_crops: need 1 <= R <= %d (got %d)", CB_MAX_R, R);
  if (!bbox || !crop_masks || !mask_hw || !nterms || !term_src || !term_hw || !term_w || !out || H < 1 || W < 1)
    return fail(NST_ERR_INVALID, "composite_from_crops: null / empty argument");
  CompositeCropsArgs a;
  memset(&a, 0, sizeof(a));
  a.R = R; a.H = H; a.W = W; a.original = original;
  for (int r = 0; r < R; ++r) {
    CropRegion& g = a.reg[r];
    g.x1 = bbox[4 * r]; g.y1 = bbox[4 * r + 1]; g.x2 = bbox[4 * r + 2]; g.y2 = bbox[4 * r + 3];
    if (g.x1 < 0 || g.y1 < 0 || g.x2 > W || g.y2 > H || g.x2 <= g.x1 || g.y2 <= g.y1)
      return fail(NST_ERR_INVALID, "composite_from_crops: region %d bbox (%d,%d,%d,%d) outside the %dx%d frame", r, g.x1, g.y1, g.x2, g.y2, W, H);
    g.mask = crop_masks[r]; g.mh = mask_hw[2 * r]; g.mw = mask_hw[2 * r + 1];
    g.nterms = nterms[r];
    if (!g.mask || g.mh < 1 || g.mw < 1 || g.nterms < 1 || g.nterms > CB_MAX_J) return fail(NST_ERR_INVALID, "composite_from_crops: region %d is malformed", r);
    for (int j = 0; j < g.nterms; ++j) {
      CropTerm& tm = g.term[j];
      tm.src = term_src[r * CB_MAX_J + j]; tm.sh = term_hw[(r * CB_MAX_J + j) * 2]; tm.sw = term_hw[(r * CB_MAX_J + j) * 2 + 1];
      tm.weight = term_w[r * CB_MAX_J + j];
      if (tm.src == nullptr && original == nullptr) return fail(NST_ERR_INVALID, "Region uses original but no original provided");      // region_blend.py:2231
      if (tm.src != nullptr && (tm.sh < 1 || tm.sw < 1)) return fail(NST_ERR_INVALID, "composite_from_crops: region %d term %d has no extent", r, j);
    }
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (original != nullptr) {
    a.canvas = out; a.weight = nullptr; a.out_u8 = out_u8; a.finish = 1;
    CUDA_TRY(launch_composite_crops(a, st));
    return NST_OK;
  }
  // no original: accumulate, then the three max-pool gap-fill passes (region_blend.py:2271-2289); passes after the gaps are closed are
  // exact no-ops, so the reference's early `break` needs no host round trip.  scratch: 2 x 4 planes of H*W floats.
  if (scratch == nullptr) return fail(NST_ERR_INVALID, "composite_from_crops: without an original the gap fill needs 8*H*W floats of scratch");
  const size_t hw = (size_t)H * W;
  float* bufc[2] = {scratch, scratch + 4 * hw};
  a.canvas = bufc[0]; a.weight = bufc[0] + 3 * hw; a.out_u8 = nullptr; a.finish = 0;
  CUDA_TRY(launch_composite_crops(a, st));
  const int ks[3] = {5, 11, 21};
  for (int i = 0; i < 3; ++i) {
    CropsDilateArgs d;
    d.canvas_in = bufc[i & 1]; d.weight_in = bufc[i & 1] + 3 * hw;
    d.last = i == 2;
    d.canvas_out = d.last ? out : bufc[(i + 1) & 1]; d.weight_out = d.last ? nullptr : bufc[(i + 1) & 1] + 3 * hw;
    d.out_u8 = d.last ? out_u8 : nullptr; d.H = H; d.W = W; d.k = ks[i];
    CUDA_TRY(launch_crops_dilate(d, st));
  }
  return NST_OK;
}

// mask_r = feather(label == r): feather_mask / gaussian_blur_mask (region_blend.py:69-102) as two separable reflect-padded passes.
// label: [dev] H*W bytes (255 = no region); tmp: [dev] R*H*W floats.
int nst::feather_labels(const uint8_t* label, int R, int feather, float* masks, float* tmp, int H, int W, cudaStream_t st, const char* who) {
  BlurArgs b;
  memset(&b, 0, sizeof(b));
  b.R = R; b.H = H; b.W = W;
  if (feather <= 0) {                                                    // feather_mask returns the hard mask
    b.k = 1; b.w[0] = 1.f;
  } else {
    const float sigma = (float)feather / 3.0f;                           // 3-sigma rule (region_blend.py:101)
    int k = (int)(6 * sigma + 1);
    if (k % 2 == 0) ++k;
    k = std::max(3, k);
    if (k > BLUR_MAX_K) return fail(NST_ERR_INVALID, "%s: feather %d needs a %d-tap kernel (max %d)", who, feather, k, BLUR_MAX_K);
    if (k / 2 >= H || k / 2 >= W) return fail(NST_ERR_INVALID, "%s: reflect padding %d needs a frame larger than %dx%d", who, k / 2, W, H);
    float sum = 0.f;
    for (int t = 0; t < k; ++t) { const float x = (float)(t - k / 2); b.w[t] = expf(-(x * x) / (2 * sigma * sigma)); sum += b.w[t]; }
    for (int t = 0; t < k; ++t) b.w[t] /= sum;
    b.k = k;
  }
  b.label = label; b.src = label ? nullptr : masks; b.dst = tmp; b.vertical = 0;       // label == null: feather the soft masks in place
  CUDA_TRY(launch_blur_pass(b, st));
  b.label = nullptr; b.src = tmp; b.dst = masks; b.vertical = 1;
  CUDA_TRY(launch_blur_pass(b, st));
  return NST_OK;
}

extern "C" int nst_feather_masks_f32(float* masks, int R, int feather, float* scratch, int H, int W, void* stream) {
  if (!masks || !scratch || R < 1 || H < 2 || W < 2) return fail(NST_ERR_INVALID, "feather_masks: null / empty argument");
  if (feather <= 0) return NST_OK;                                       // feather_mask returns its argument
  return feather_labels(nullptr, R, feather, masks, scratch, H, W, (cudaStream_t)stream, "feather_masks");
}

extern "C" int nst_voronoi_masks_f32(const float* points_xy, const float* region_weights, int R, int feather, float* masks, float* scratch,
                                     int H, int W, void* stream) {
  if (!points_xy || !masks || !scratch || R < 1 || R > VOR_MAX_R || H < 2 || W < 2) return fail(NST_ERR_INVALID, "voronoi_masks: bad argument (1 <= R <= %d)", VOR_MAX_R);
  cudaStream_t st = (cudaStream_t)stream;
  VoronoiArgs v;
  memset(&v, 0, sizeof(v));
  v.R = R; v.H = H; v.W = W; v.weighted = region_weights != nullptr;
  v.label = reinterpret_cast<uint8_t*>(scratch);                        // scratch: H*W bytes of labels, then R*H*W floats
  float* tmp = scratch + ((size_t)H * W + 3) / 4;
  double wsum = 0.0;
  for (int i = 0; i < R && region_weights; ++i) wsum += region_weights[i];
  for (int i = 0; i < R; ++i) {
    v.px[i] = points_xy[2 * i]; v.py[i] = points_xy[2 * i + 1];
    // norm_weights = w * count / total; effective_dist = dist / (sqrt(weight) + 1e-6)   (region_blend.py:205-219)
    v.inv_w[i] = region_weights ? (float)(sqrt(region_weights[i] * R / wsum) + 1e-6) : 0.f;
  }
  CUDA_TRY(launch_voronoi_label(v, st));
  return feather_labels(v.label, R, feather, masks, tmp, H, W, st, "voronoi_masks");
}

extern "C" int nst_resize_post_pack_u8(const float* src01, int N, int Hi, int Wi, const uint8_t* original, const float* mask, double blend,
                                       int composite_mode, uint8_t* out, int H, int W, void* stream) {
  if (!src01 || !out || N < 1 || Hi < 1 || Wi < 1 || H < 1 || W < 1) return fail(NST_ERR_INVALID, "resize_post_pack: null / empty argument");
  const bool blend_on = blend >= 0.0 && blend < 1.0;
  if ((mask != nullptr || blend_on) && original == nullptr) return fail(NST_ERR_INVALID, "resize_post_pack: mask / blend need the original frames");
  ResizePackArgs ra;
  ra.src01 = src01; ra.out = out; ra.orig = (mask != nullptr || blend_on) ? original : nullptr; ra.mask = mask;
  ra.N = N; ra.Hi = Hi; ra.Wi = Wi; ra.H = H; ra.W = W;
  ra.blend = blend_on ? (float)blend : 2.f; ra.blend_comp = (float)(1.0 - blend); ra.composite_mode = composite_mode;
  CUDA_TRY(launch_resize_post_pack(ra, (cudaStream_t)stream));
  return NST_OK;
}

extern "C" int nst_flow_ema_f32(const float* curr, const float* prev, const float* flow, float* out, int H, int W, double alpha, void* stream) {
  if (!curr || !prev || !flow || !out || H < 1 || W < 1) return fail(NST_ERR_INVALID, "flow_ema: null / empty argument");
  if (prev == out) return fail(NST_ERR_INVALID, "flow_ema: out must not alias prev (the warp gathers from it)");
  FlowEmaArgs a;
  const double al = std::max(0.0, std::min(1.0, alpha));          // pipeline.py:1927
  a.curr = curr; a.prev = prev; a.flow = flow; a.out = out; a.H = H; a.W = W;
  a.alpha = (float)al; a.alpha_comp = (float)(1.0 - al);
  CUDA_TRY(launch_flow_ema(a, (cudaStream_t)stream));
  return NST_OK;
}

// ---------------------------------------------------------------------------------------------
// planner dump (no device needed): one line per layer with the tcgen05 tiling that would be used
// ---------------------------------------------------------------------------------------------
extern "C" int nst_describe_plan(int arch, int precision, int N, int H, int W, char* out, int cap) {
  if (out == nullptr || cap < 1) return fail(NST_ERR_INVALID, "nst_describe_plan: no output buffer");
  if (arch < 0 || arch > 2 || precision < 0 || precision > 3) return fail(NST_ERR_INVALID, "nst_describe_plan: bad arch/precision");
  nst_handle h;
  h.arch = arch; h.precision = precision;
  h.act16 = precision == NST_PREC_F16 ? 1 : 0;
  h.specs = arch == NST_ARCH_JOHNSON ? johnson_layers() : (arch == NST_ARCH_NST ? nst_layers() : reconet_layers());
  h.dev.resize(h.specs.size());
  if (const char* mbs = getenv("NST_MICROBATCH")) h.microbatch = std::max(1, atoi(mbs));
  for (size_t li = 0; li < h.specs.size(); ++li) {
    h.dev[li].prec = layer_precision(arch, precision, (int)li);
    layer_static_plan(h.specs[li], h.dev[li]);
  }
  int rc = make_plan(&h, N, H, W);
  if (rc != NST_OK) return rc;
  std::string s;
  char line[512];
  snprintf(line, sizeof(line), "arch=%d precision=%d N=%d (micro-batch %d) %dx%d -> %dx%d workspace=%.1f MB\n", arch, precision, N,
           h.plan_nb, H, W, h.out_H, h.out_W, h.workspace_bytes / 1048576.0);
  s += line;
  for (size_t li = 0; li < h.specs.size(); ++li) {
    const LayerSpec& sp = h.specs[li];
    const LayerPlan& lp = h.plans[li];
    const TcPlan& t = lp.t;
    if (lp.use_tc) {
      const double useful = (double)lp.p.Ho * lp.p.Wo / ((double)t.tiles_x * t.tiles_y * t.MB * 128);
      snprintf(line, sizeof(line),
               "%-28s %3d->%3d k%d s%d %4dx%-4d -> %4dx%-4d tc %s%s EG=%d PW=%d TH=%2d TW=%3d P=%3d MB=%d NPOS=%4d CK=%2d chunks=%d ent=%2d tg=%d "
               "A=%dx%uB B=%dx%uB smem=%uB tiles/img=%d rows_useful=%.2f GF=%.2f\n",
               sp.name.c_str(), sp.cin, sp.cout, sp.ks, sp.stride, lp.p.Hs, lp.p.Ws, lp.p.Ho, lp.p.Wo, t.tma_mode ? (t.swap_ab ? "tma/T " : "tma   ") : (t.swap_ab ? "gath/T" : "gather"),
               h.dev[li].prec == NST_PREC_F16X3 ? " x3" : (lp.out16 ? "" : " f32"), t.epi_groups, t.prod_warps, t.TH, t.TW, t.P, t.MB, t.NPOS,
               t.CK, t.nchunks, t.nent, t.tg, t.a_stages, t.a_stage_bytes, t.b_stages, t.b_stage_bytes, t.smem_bytes,
               t.tiles_x * t.tiles_y, useful, lp.flops * 1e-9);
    } else {
      snprintf(line, sizeof(line), "%-28s %3d->%3d k%d s%d %4dx%-4d -> %4dx%-4d direct GF=%.2f\n", sp.name.c_str(), sp.cin, sp.cout,
               sp.ks, sp.stride, lp.p.Hs, lp.p.Ws, lp.p.Ho, lp.p.Wo, lp.flops * 1e-9);
    }
    s += line;
  }
  strncpy(out, s.c_str(), cap - 1);
  out[cap - 1] = 0;
  return NST_OK;
}
