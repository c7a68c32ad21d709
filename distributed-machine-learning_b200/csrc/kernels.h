// kernels.h -- launchers of the hand-written sm_100a kernels (internal).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace dml {

struct ConvParams {
  const void* in;     // NHWC activations, pointer already offset to the view's first channel
  const void* w;      // packed weights [cout][kh*kw][cin_pad]
  const float* bias;  // folded bias [cout]
  const void* res;    // residual (same dtype as out) or nullptr
  void* out;
  int n, hi, wi, cin, in_pitch;
  int ho, wo, cout, out_pitch, res_pitch;
  int kh, kw, stride, pad_h, pad_w, cin_pad;
  int relu;
  // second GEMM source (graph.h, Op::in2): a 1x1 convolution of stride stride2 over in2, accumulated into the same output
  const void* in2 = nullptr;
  int hi2 = 0, wi2 = 0, cin2 = 0, in2_pitch = 0, stride2 = 1, cin2_pad = 0;
  // Tile order.  Consecutive layers walk their tiles in OPPOSITE directions: what a layer wrote last is still in L2 (126 MB) when
  // the next layer starts, so the next layer begins with exactly those pixels instead of with the ones evicted long ago.
  int rev = 0;
  int norelu_cols = 0;  // output channels [0, norelu_cols) get no ReLU (graph.h, Op::norelu_cols); a multiple of 16
  // Resident-filter halo kernel, 16-channel input (the space-to-depth stems): the input rows are physically padded to
  // in_row_cells pixels (the left padding included) and pre-swizzled, graph.h Net::S2d::row_cells.  0 = plain NHWC.
  int in_row_cells = 0;
  int in_img0 = 0;   // tcgen05 path: first image / first output row of this launch inside the
  int out_row0 = 0;  // planned tensor maps (the pointers above are already offset for the CUDA-core path)
};

struct PoolParams {
  const void* in;
  void* out;
  int n, hi, wi, c, in_pitch;
  int ho, wo, out_pitch;
  int k, stride, pad;
  int rev = 0;                  // walk the outputs from the last to the first (ConvParams::rev)
  const float* bias = nullptr;  // average pool only: per-channel bias added after the average, then ReLU if relu != 0
  int relu = 0;                 // (the pool branch of an Inception block with the 1x1 conv moved in front, graph.h Op::post_pool)
};

struct PreParams {
  const uint8_t* pixels;
  const int64_t* img_off;
  const int32_t* hw;    // (H, W) per image
  const int32_t* xtab;  // n x S
  const int32_t* ytab;  // n x S
  void* out;            // base of the whole input tensor (image 0), 16-byte aligned
  int n, S;             // n = images in the tensor the tables cover
  int img_begin, img_end;  // images [img_begin, img_end) are produced by this launch
  int caffe;            // 1: output channel c reads source channel 2-c (RGB->BGR)
  int s2d;              // 1: bf16 space-to-depth output [n][S2][row_cells][16], S2 = (S+1)/2 (graph.h, Net::S2d)
  int s2d_row_cells;    // physical cells per row (row-padded, pre-swizzled form) and
  int s2d_first_cell;   // the physical column of data cell 0
};

// K1: NEAREST resize (host-computed index tables) + preprocess LUT; bf16 or fp32 out.
void launch_preprocess(const PreParams& p, const float* d_lut, bool fp32_out, cudaStream_t s);

// CUDA-core direct convolution (check mode, debug twin, C=3 stems).
void launch_conv_naive(const ConvParams& p, bool fp32_act, cudaStream_t s);

// K2: C_in = 3 stem convolution on the tensor cores (bf16 mode).
bool stem_tc_supported(const ConvParams& p);
int launch_stem_tc(const ConvParams& p, cudaStream_t s);

// K4 / K5: 3x3 max-pool (zero padding participates, as Keras' ZeroPadding2D does)
// and 3x3 s1 'same' average pool dividing by the in-bounds tap count.
void launch_maxpool(const PoolParams& p, bool fp32_act, cudaStream_t s);
void launch_avgpool(const PoolParams& p, bool fp32_act, cudaStream_t s);

// K6: global average pool -> fp32 [n][c]
void launch_gap(const void* in, float* out, int n, int hw, int c, int in_pitch, bool fp32_act, cudaStream_t s);

// K7: Dense(1000) -> logits scratch; softmax + top-5.
// tf32: TF32 tensor-core GEMM (bf16 engines); otherwise the exact fp32 CUDA-core kernel (check mode).
void launch_fc(const float* pooled, const float* w, const float* b, float* logits, int n, int cin, bool tf32, cudaStream_t s);
// Per-device one-time kernel attribute set-up (call with the device current).
void init_kernels();

void launch_softmax_top5(const float* logits, float* probs, int32_t* top_idx, float* top_prob, int n, cudaStream_t s);


// ---- programmatic dependent launch (PDL) ----
// Every kernel of a step is short (10-200 us) and fills the GPU by itself, so the launch latency and the prologue
// (barrier init, TMEM allocation, tensor-map prefetch) of kernel i+1 are exposed after the tail of kernel i.  Kernels
// launched through launch_pdl() may start while their predecessor drains; they call pdl_wait() before they touch any
// memory a previous kernel may have written or may still read, and pdl_launch_dependents() at their start.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

bool pdl_enabled();  // false with DML_NO_PDL=1
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

}  // namespace dml
