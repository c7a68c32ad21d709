// graph.h -- host-side description of the two networks the reference runs
// (keras.applications ResNet50 / InceptionV3, built at /root/reference/models.py:26,:51):
// tensors, ops, Keras-layout weight arrays and the activation-arena plan.
// Pure host C++, no CUDA.
#pragma once
#include <cstdint>
#include <string>
#include <utility>
#include <vector>

namespace dml {

struct WeightArr {
  std::string name;
  int rank = 0;
  int64_t shape[4] = {1, 1, 1, 1};
  int64_t offset = 0;  // floats into the blob
  int64_t count = 0;
};

struct Buf {
  int64_t elems = 0;      // per image
  int first_use = 1 << 30, last_use = -1;
  int64_t offset = 0;     // per-image element offset inside the arena
};

struct Tensor {
  std::string name;
  int h = 0, w = 0, c = 0;
  int buf = -1;     // -1: the engine's input tensor
  int c_off = 0;    // channel offset inside the buffer (concat views)
  int c_pitch = 0;  // channels per pixel of the buffer
};

enum OpKind { OP_CONV = 0, OP_MAXPOOL = 1, OP_AVGPOOL = 2, OP_GAP = 3 };

struct Op {
  OpKind kind = OP_CONV;
  std::string name;
  int in = -1, out = -1, res = -1;  // tensor ids (res: residual added before ReLU)
  int kh = 1, kw = 1, stride = 1, pad_h = 0, pad_w = 0;
  bool relu = false;
  // Keras weight arrays (indices into Net::warr; -1 = absent)
  int w_kernel = -1, w_bias = -1, w_gamma = -1, w_beta = -1, w_mean = -1, w_var = -1;
  float eps = 0.f;
  // packed weights
  int cin = 0, cout = 0, cin_pad = 0;
  int64_t wpack_off = 0;  // bytes into the weights buffer
  int64_t bias_off = 0;   // bytes into the weights buffer (fp32[cout])
  // Second GEMM source (ResNet50 projection shortcut folded into the block's last conv): out = relu(W x + W2 x2 + b + b2)
  // with x2 = tensor in2 read through a 1x1 convolution of stride stride2.  The packed weight rows are
  // [taps * cin_pad | cin2_pad] and the bias is the sum of both folded biases.  in2 < 0: none.
  // Sibling 1x1 convolutions of one input run as ONE GEMM (InceptionV3 branch heads): the leader executes with
  // merged_cout output channels into tensor merged_out (every member's output is a channel slice of it); the members'
  // packed weight / bias rows sit in one block in slice order (merge_row0); a follower (skip = true) only carries its
  // weight arrays.
  bool skip = false;
  std::vector<int> followers;            // leader: the ops whose outputs it also computes
  int merged_cout = 0, merged_out = -1;  // leader: summed Cout and the tensor that holds all slices
  int merge_row0 = 0;                    // leader and followers: first row of this op inside the merged weight / bias block
  int64_t merged_w_off = 0, merged_b_off = 0;  // leader: start of the merged block (bytes into the weights buffer)
  // Bottleneck-tail chain (ResNet50 conv2, conv_chain.inc): the 3x3 (the chain's lead) may run fused with the block's
  // 1x1 expand (chain_exp) and the next block's 1x1 reduce (chain_red, -1 for the stage's last block); chain_lead on
  // the members points back.  The engine decides at plan time whether the fused kernel or the separate ones run.
  int chain_exp = -1, chain_red = -1, chain_lead = -1;
  // Expand + next reduce duo (ResNet50 conv3 / conv4, conv_duo.inc): on a 1x1 expand with an identity residual, duo_red is
  // the next block's 1x1 reduce the engine may compute in the same launch; duo_lead on that reduce points back.
  int duo_red = -1, duo_lead = -1;
  // InceptionV3's pool branch is AveragePooling2D -> 1x1 conv -> BN -> ReLU.  The pool and the 1x1 conv are both linear and act
  // on different axes (space / channels), so they commute: relu(conv(avgpool(x)) + b) == relu(avgpool(conv(x)) + b).  In fused
  // engines the conv (post_pool = true) joins the block's merged head GEMM WITHOUT bias and ReLU (its output channels come first
  // in the merged range: Op::norelu_cols of the leader) and the pool runs afterwards on the conv's few output channels instead of
  // on the block's wide input, adding the folded bias (post_bias_off) and the ReLU.  pool_bias_of: on the pool op, the conv.
  bool post_pool = false;
  int64_t post_bias_off = 0;
  int norelu_cols = 0;
  int pool_bias_of = -1;
  int in2 = -1, cin2 = 0, cin2_pad = 0, stride2 = 1;
  // Branch streams (plan_streams): independent branches of an InceptionV3 block run on different CUDA streams.  stream 0 is
  // the caller's; before its launch the op waits for the events of the ops in `waits` (producers / memory hazards on other
  // streams); `signal`: some op on another stream waits for this one.  region: index into Net::regions (-1 outside the
  // blocks); solo: every other op of the region is an ancestor or a descendant (the block's head GEMM), so it runs alone.
  int stream = 0, region = -1;
  bool signal = false, solo = true;
  std::vector<int> waits;
  int w2_kernel = -1, w2_bias = -1, w2_gamma = -1, w2_beta = -1, w2_mean = -1, w2_var = -1;
  float eps2 = 0.f;
};

struct Net {
  int model = 0;
  int input_size = 0;  // S
  int mode = 0;        // 0 caffe, 1 tf
  std::vector<WeightArr> warr;
  int64_t total_floats = 0;
  std::vector<Tensor> tensors;
  std::vector<Buf> bufs;
  std::vector<Op> ops;
  int input_tensor = -1;
  int pooled_tensor = -1;  // GAP output (fp32, outside the arena)
  int fc_kernel = -1, fc_bias = -1;
  int64_t fc_w_off = 0, fc_b_off = 0;  // bytes into the weights buffer (fp32)
  int64_t arena_elems = 0;             // per image
  int64_t weights_bytes = 0;
  int64_t conv_macs = 0;               // per image
  // Blocks whose branches may run concurrently: [first op, last op].  Buffers first touched inside a region stay live to its
  // end, so the arena never makes one branch's scratch alias another's.  n_streams: streams plan_streams() used (1 = serial).
  std::vector<std::pair<int, int>> regions;
  int n_streams = 1;
  // Space-to-depth form of the C_in = 3 stride-2 stem (bf16 tensor-core mode only): the S x S x 3 input is stored as
  // S2 x S2 x 16 (channel (py*2+px)*3 + c = pixel (2y+py, 2x+px), channels 12..15 zero) and the k x k stride-2 filter
  // becomes a k2 x k2 stride-1 filter over it whose first tap sits pad2 cells before the output position.
  struct S2d {
    int op = -1;             // index of the stem op, -1 = none
    int size = 0;            // S2 = ceil(S / 2)
    int k2 = 0, pad2 = 0;
    // Physical row of the space-to-depth input: pad2 zero cells, the S2 data cells, zero cells up to row_cells (a multiple of 8,
    // >= S2 + k2 - 1).  With the horizontal padding in memory, a whole halo tile is ONE TMA box of 256-byte lines (8 cells) instead
    // of one 32-byte line per cell -- the TMA unit's cost is per line -- and since 8 | row_cells the 32-byte swizzle of a cell's
    // shared-memory row depends on its physical column only: K1 stores the cells already swizzled (the two 16-byte halves of the
    // cells with bit 2 of the physical column set are swapped) and the TMA copies them without a swizzle of its own.
    int row_cells = 0;
    int64_t w_off = 0;       // packed bf16 [cout][k2*k2][16], bytes into the weights buffer
  } s2d;
};

// Builds the graph.  keep_all: every tensor gets its own storage (parity taps).
// wbytes: bytes per packed conv weight element (2 = bf16, 4 = fp32 check mode).
// tc_pad: pad Cin of tensor-core-eligible convs to a multiple of 64.
// fuse_shortcut: ResNet50's projection shortcuts become a second GEMM source of the block's last conv (Op::in2) and
// InceptionV3's sibling 1x1 branch heads become one GEMM (Op::followers).
bool build_net(int model, bool keep_all, int wbytes, bool tc_pad, bool fuse_shortcut, Net* net, std::string* err, bool pool_commute = true);

// Assigns the ops of a net with regions to at most max_streams streams (Op::stream / waits / signal / solo) from the true
// data dependencies plus every memory hazard the arena plan creates.  Serial order on one stream is always a valid schedule;
// this only relaxes it.
void plan_streams(Net* net, int max_streams);

// Only the weight-array list (cheap; for dml_model_* queries).
bool build_weight_spec(int model, Net* net, std::string* err);

}  // namespace dml
