// conv_tc.h -- tcgen05/TMEM implicit-GEMM convolution fed by TMA (internal).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <vector>

#include "graph.h"
#include "kernels.h"

namespace dml {

struct TcOpPlan {
  bool valid = false;
  alignas(64) CUtensorMap map_a;  // im2col map over the NHWC input view
  alignas(64) CUtensorMap map_b;  // tiled map over packed weights [cout][taps*cin_pad]
  int bn = 0;        // N tile (output channels per CTA)
  int stages = 0;
  int group = 1;     // k-blocks per pipeline stage (v2 / v4)
  int tmem_cols = 0;
  int cpb = 0;       // 64-channel blocks per tap
  size_t smem = 0;
  // v2 (persistent kernel)
  int version = 2;
  alignas(64) CUtensorMap map_o;  // tiled map over the output view [M_capacity][cout], pitch = out_pitch
  int cw = 0;          // epilogue chunk width (columns per TMA store box)
  int acc_stride = 0;  // TMEM columns between the two accumulators
  int m_capacity = 0;  // rows the output map covers
  alignas(64) CUtensorMap map_r;  // residual view, same geometry as the output
  bool has_res = false;
  bool res_direct = false;  // v4: ONE residual tile buffer with per-chunk hand-back instead of two tile buffers
  bool a_tiled = false;  // A fetched with a tiled 2D map (1x1 stride-1 convs)
  alignas(64) CUtensorMap map_a2;  // second GEMM source (ConvParams::in2), pair kernel
  bool a2_tiled = false;
  // v3 (halo kernel)
  int wp = 0, rows_box = 0, tiles_per_img = 0, q_valid = 0, a_stage_bytes = 0, a_stages = 0, b_stages = 0;
  // v6 (halo kernel, resident filter)
  int a_cb_bytes = 0, row_bytes = 0;
  float best_ms = 0.f;  // autotune: time of the chosen plan
};

// The tail of a ResNet50 conv2 bottleneck as one kernel (conv_chain.inc): 3x3 -> 1x1 expand (+ residual, or + the folded
// projection shortcut as a second GEMM source) -> the next block's 1x1 reduce.
struct ChainArgs {
  const void* y1 = nullptr; int y1_pitch = 0;      // input of the 3x3: [n][h][w][64] bf16 (pitch in elements)
  const void* w3 = nullptr; const float* b3 = nullptr;   // packed [64][9][64], folded bias
  const void* we = nullptr; const float* be = nullptr;   // packed [256][64 (+ 64 of the projection)], folded bias (sum)
  const void* res = nullptr; int res_pitch = 0;    // residual [n*h*w][256], or nullptr
  const void* src2 = nullptr; int src2_pitch = 0;  // second GEMM source [n*h*w][64] (projection input), or nullptr
  void* out = nullptr; int out_pitch = 0;          // block output [n*h*w][256]
  const void* wr = nullptr; const float* br = nullptr;   // next block's reduce: packed [64][256], folded bias; or nullptr
  void* y1n = nullptr; int y1n_pitch = 0;          // its output [n*h*w][64]
  int n = 0, h = 0, w = 0;
  int rev = 0;  // ConvParams::rev
};
struct ChainPlan {
  bool valid = false;
  alignas(64) CUtensorMap map_halo, map_w3, map_we, map_wr, map_res, map_out, map_s2;
  int variant = 0;  // 0: residual + next reduce, 1: second source + next reduce, 2: residual, no next reduce
  int wp = 0, rpt = 0, rows = 0, ptiles_per_img = 0, box_bytes = 0, halo_bytes = 0, kb2 = 1;
  size_t smem = 0;
  int n_capacity = 0;
  float ms = 0.f, ms_separate = 0.f;  // autotune: this kernel against the three separate ones
};
bool chain_supported(const ChainArgs& a);
int chain_plan(ChainPlan* out, const ChainArgs& a, int n_capacity);
int launch_chain(const ChainPlan& plan, const ChainArgs& a, cudaStream_t s);

// A 1x1 expand (+ residual) and the next block's 1x1 reduce as one kernel (conv_duo.inc; ResNet50 conv3 / conv4).
struct DuoArgs {
  const void* y2 = nullptr; int y2_pitch = 0;      // input of the expand: [M][kmid] bf16 (pitch in elements)
  const void* we = nullptr; const float* be = nullptr;   // packed [nexp][kmid], folded bias
  const void* res = nullptr; int res_pitch = 0;    // residual [M][nexp], or nullptr (then src2 is given)
  const void* src2 = nullptr; int src2_pitch = 0;  // second GEMM source [M][kmid - k1]: the folded projection shortcut's input, or nullptr
  int s2_stride = 1;                               // ... read at this stride from an [n][h * stride][w * stride][kmid - k1] tensor (then n, h, w below are needed)
  int n = 0, h = 0, w = 0;                         // the output grid, M = n * h * w (only for s2_stride > 1)
  void* out = nullptr; int out_pitch = 0;          // block output [M][nexp]
  const void* wr = nullptr; const float* br = nullptr;   // next block's reduce: packed [nred][nexp], folded bias
  void* y1n = nullptr; int y1n_pitch = 0;          // its output [M][nred]
  int M = 0;                                       // pixels (n * h * w)
  int kmid = 0, nexp = 0, nred = 0;                // kmid: K of the expand, the second source's channels included
  int k1 = 0;                                      // channels that come from y2 (= kmid without a second source)
  int rev = 0;  // ConvParams::rev
};
struct DuoPlan {
  bool valid = false;
  alignas(64) CUtensorMap map_y2, map_we, map_wr, map_res, map_out, map_s2;
  int y2_bufs = 1, we_stages = 2, wr_stages = 2, slots = 4;
  bool no_slack = false;  // the shared-memory request leaves no room to align the base (it is 1 KiB-aligned in practice; the kernel checks)
  size_t smem = 0;
  int m_capacity = 0;
  float ms = 0.f, ms_separate = 0.f;  // autotune: this kernel against the two separate ones
};
bool duo_supported(const DuoArgs& a);
int duo_plan(DuoPlan* out, const DuoArgs& a, int m_capacity);
int launch_duo(const DuoPlan& plan, const DuoArgs& a, cudaStream_t s);

// A stride-1 convolution with 64 filters and the 3x3 stride-2 max-pool behind it as one kernel (conv_pool.inc).
struct ConvPoolPlan {
  bool valid = false;
  TcOpPlan conv;   // the resident-filter halo plan of the convolution (tensor maps, tile geometry)
  int hp = 0, wq = 0, pool_pad = 0, R = 0, bands = 0, T = 0, a_stages = 2;
  size_t smem = 0;
  float ms = 0.f, ms_separate = 0.f;  // autotune: this kernel against convolution + pool
};
// p: the convolution (p.out is not written); the pool's output is [n][hp][wq][pool_out_pitch] at launch time.
int conv_pool_plan(ConvPoolPlan* out, const ConvParams& p, int hp, int wq, int pool_pad, int n_capacity);
int launch_conv_pool(const ConvPoolPlan& plan, const ConvParams& p, void* pool_out, int pool_out_pitch, int pool_out_img0, cudaStream_t s);

struct TcPlan {
  std::vector<TcOpPlan> ops;  // indexed like Net::ops
  std::vector<DuoPlan> duos;  // indexed like Net::ops; valid at the expand that leads a duo the engine runs fused
  std::vector<ChainPlan> chains;  // indexed like Net::ops; valid at the 3x3 that leads a chain the engine runs fused
  std::vector<char> chained;      // op is computed inside a valid chain's launch
  int micro_batch = 0;
};

// Tile-shape overrides for tuning/tests (0 = heuristic).
struct TcTune {
  int bn = 0;
  int stages = 0;
  int group = 0;    // k-blocks per pipeline stage for v2 / v4 (0 = heuristic)
  int row_bytes = 0;   // halo kernels (v6 / v7): bytes of a shared-memory pixel row = 2 x channels per K block; 0 = by Cin (<= 16: 32, <= 32: 64, else 128)
  int res_direct = 0;  // v4 residual staging: 0 = heuristic, 1 = two tile buffers, 2 = one buffer, chunks handed back one by one
  int version = 0;  // 1 = one tile per CTA, 2 = persistent im2col, 3 = halo kernel, 4 = CTA-pair (cta_group::2) persistent, 5 = weights as A, 6 = halo kernel with the filter resident in shared memory, 7 = the same on CTA pairs; 0 = choose
};

bool tc_supported(const Op& op);
int tc_init();
size_t tc_scratch_bytes(const Net& net, int micro_batch);
int tc_plan(TcPlan* plan, const Net& net, uint8_t* ws, size_t off_arena, size_t off_input, size_t off_tc, uint8_t* wts,
            int micro_batch, int max_batch);
// Builds the plan of ONE conv over explicit pointers (tests, tuning).
int tc_plan_one(TcOpPlan* out, const ConvParams& p, int n_capacity, const TcTune& tune);
int launch_conv_tc_plan(const TcOpPlan& op, const ConvParams& p, cudaStream_t s);
int launch_conv_tc(const TcPlan& plan, int op_index, const ConvParams& p, int img0, cudaStream_t s, int* nlaunch);
// Runs the fused chain led by op `lead` (plan.chains[lead].valid) over the first n images of the arena.
int launch_chain_tc(const TcPlan& plan, const Net& net, uint8_t* wts, int lead, int n, uint8_t* ws, size_t off_arena, cudaStream_t s, int rev = 0);
// Runs the fused expand + next reduce led by op `lead` (plan.duos[lead].valid) over the first n images of the arena.
int launch_duo_tc(const TcPlan& plan, const Net& net, uint8_t* wts, int lead, int n, uint8_t* ws, size_t off_arena, cudaStream_t s, int rev = 0);
const char* tc_last_error();
// Tiled bf16 tensor map (rank <= 5) for kernels outside conv_tc.cu; swizzle_bytes in {0, 32, 64, 128}.
int tc_make_tiled_map(CUtensorMap* out, void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box, int swizzle_bytes);

}  // namespace dml
