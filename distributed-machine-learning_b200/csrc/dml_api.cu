// dml_api.cu -- the C ABI of libdml_b200.so (include/dml_b200.h): engine
// object, weight folding/packing, the per-micro-batch layer walk.
//
// Replaces what /root/reference/models.py:23-71 does through Keras: model
// construction (:26, :51), load_img/preprocess_input (:34-38, :59-63),
// predict (:40, :65) and decode_predictions' top-5 selection (:42, :67).
#include "../../include/dml_b200.h"

#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "conv_tc.h"
#include "graph.h"
#include "kernels.h"

namespace {

thread_local std::string g_err;

int fail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}

#define CUDA_TRY(expr)                                                                        \
  do {                                                                                        \
    cudaError_t err__ = (expr);                                                               \
    if (err__ != cudaSuccess)                                                                 \
      return fail(DML_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(err__), __FILE__, __LINE__); \
  } while (0)

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

uint16_t f32_to_bf16_rn(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);  // NaN
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}

void make_lut(int mode, float out[768]) {
  // preprocess_input for uint8 input, per OUTPUT channel (models.py:38 / :63).
  const float mean_bgr[3] = {103.939f, 116.779f, 123.68f};
  for (int c = 0; c < 3; ++c)
    for (int v = 0; v < 256; ++v) {
      volatile float x = (float)v;
      if (mode == 0) {
        x = x - mean_bgr[c];  // caffe: BGR order, mean subtraction only
      } else {
        x = x / 127.5f;  // tf: true divide, then subtract (two rounded fp32 ops)
        x = x - 1.0f;
      }
      out[c * 256 + v] = x;
    }
}

}  // namespace

struct dml_engine {
  int device = 0, model = 0, max_batch = 0, micro_batch = 0, flags = 0;
  bool fp32 = false, use_tc = false, keep_all = false;
  int act_bytes = 2;
  dml::Net net;
  uint8_t* ws = nullptr;
  size_t ws_bytes = 0;
  uint8_t* wts = nullptr;
  size_t wts_bytes = 0;
  bool weights_loaded = false;
  size_t off_lut = 0, off_input = 0, off_arena = 0, off_pooled = 0, off_logits = 0, off_tc = 0, ws_need = 0;
  int64_t launches = 0;
  std::atomic<int> cancel{0};  // dml_cancel: checked between micro-batches
  dml::TcPlan tc;  // tensor maps etc. for the tcgen05 path
  bool s2d = false;          // K1 writes the space-to-depth input and the stem runs as a stride-1 halo conv (Net::S2d)
  size_t input_img_bytes = 0;
  dml::TcOpPlan stem_plan;
  std::vector<dml::ConvPoolPlan> conv_pool;  // indexed like Net::ops: valid where the conv and the max-pool behind it run as one kernel
  // Branch streams (graph.h, plan_streams): engine-owned side streams and one event per op another stream waits for.
  bool branch_streams = false;
  cudaStream_t side[7] = {};
  cudaEvent_t join_ev[7] = {};
  std::vector<cudaEvent_t> op_event;
  ~dml_engine() {
    for (cudaStream_t st : side) if (st) cudaStreamDestroy(st);
    for (cudaEvent_t ev : join_ev) if (ev) cudaEventDestroy(ev);
    for (cudaEvent_t ev : op_event) if (ev) cudaEventDestroy(ev);
  }
  // profiling (dml_profile_forward): one event pair per op launch
  bool profiling = false;
  std::vector<std::pair<int, std::pair<cudaEvent_t, cudaEvent_t>>> prof;

  void prof_begin(int op, cudaStream_t s) {
    if (!profiling) return;
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    cudaEventRecord(a, s);
    prof.push_back({op, {a, b}});
  }
  void prof_end(cudaStream_t s) {
    if (profiling) cudaEventRecord(prof.back().second.second, s);
  }

  void* tensor_ptr(int t, int img0) const {
    const dml::Tensor& tt = net.tensors[t];
    if (tt.buf == -1) return ws + off_input + input_img_bytes * img0;
    if (tt.buf == -2) return ws + off_pooled + (size_t)img0 * tt.c * 4;
    const dml::Buf& b = net.bufs[tt.buf];
    return ws + off_arena + ((size_t)b.offset * micro_batch + tt.c_off) * act_bytes;
  }
};

namespace {

int check_model(int model) {
  if (model != DML_MODEL_RESNET50 && model != DML_MODEL_INCEPTIONV3)
    return fail(DML_EINVAL, "unknown model id %d (0 = ResNet50, 1 = InceptionV3)", model);
  return 0;
}

// The stem as the stride-1 conv over the space-to-depth input that the halo kernel runs.
dml::ConvParams stem_s2d_params(const dml_engine* e, int img0, int cnt) {
  const dml::Net& net = e->net;
  const dml::Op& op = net.ops[net.s2d.op];
  const dml::Tensor& to = net.tensors[op.out];
  dml::ConvParams p{};
  p.in = e->ws + e->off_input;          // whole input tensor; in_img0 selects the image
  p.in_img0 = img0;
  p.w = e->wts + net.s2d.w_off;
  p.bias = reinterpret_cast<const float*>(e->wts + op.bias_off);
  p.out = e->tensor_ptr(op.out, 0);
  p.out_row0 = 0;
  p.n = cnt; p.hi = net.s2d.size; p.wi = net.s2d.size; p.cin = 16; p.in_pitch = 16; p.cin_pad = 16;
  p.in_row_cells = net.s2d.row_cells;
  p.ho = to.h; p.wo = to.w; p.cout = op.cout; p.out_pitch = to.c_pitch;
  p.kh = net.s2d.k2; p.kw = net.s2d.k2; p.stride = 1; p.pad_h = net.s2d.pad2; p.pad_w = net.s2d.pad2;
  p.relu = op.relu;
  return p;
}

int plan_stem(dml_engine* e) {
  if (!e->s2d) return 0;
  dml::TcTune tune;
  tune.version = 6;
  dml::ConvParams p = stem_s2d_params(e, 0, e->micro_batch);
  if (dml::tc_plan_one(&e->stem_plan, p, e->max_batch, tune) != 0)
    return fail(DML_ECUDA, "space-to-depth stem plan failed: %s", dml::tc_last_error());
  return 0;
}

// ConvParams of conv op oi over the engine's tensors (the stem in its space-to-depth form), as run_layers builds them.
dml::ConvParams conv_params_of(const dml_engine* e, int oi, int img0, int cnt) {
  const dml::Net& net = e->net;
  if (e->s2d && oi == net.s2d.op) return stem_s2d_params(e, img0, cnt);
  const dml::Op& op = net.ops[oi];
  const dml::Tensor& ti = net.tensors[op.in];
  const dml::Tensor& to = net.tensors[op.out];
  dml::ConvParams p{};
  p.in = e->tensor_ptr(op.in, img0);
  p.w = e->wts + op.merged_w_off;
  p.bias = reinterpret_cast<const float*>(e->wts + op.merged_b_off);
  p.out = e->tensor_ptr(op.out, img0);
  p.n = cnt; p.hi = ti.h; p.wi = ti.w; p.cin = op.cin; p.in_pitch = ti.c_pitch;
  p.ho = to.h; p.wo = to.w; p.cout = op.cout; p.out_pitch = to.c_pitch;
  p.kh = op.kh; p.kw = op.kw; p.stride = op.stride; p.pad_h = op.pad_h; p.pad_w = op.pad_w;
  p.cin_pad = op.cin_pad; p.relu = op.relu;
  return p;
}

// Convolution + max-pool pairs that may run as one kernel (conv_pool.inc): a stride-1 conv with 64 filters whose output is read by
// nothing but a 3x3 stride-2 max-pool (ResNet50 conv1 + pool1, InceptionV3 conv2d_2 + max_pooling2d).  Timed against the two kernels.
int plan_conv_pools(dml_engine* e) {
  const dml::Net& net = e->net;
  e->conv_pool.assign(net.ops.size(), dml::ConvPoolPlan());
  if (!e->use_tc || e->keep_all || getenv("DML_NO_CONV_POOL")) return 0;
  const bool tune_log = getenv("DML_TUNE_LOG") != nullptr;
  for (size_t i = 0; i + 1 < net.ops.size(); ++i) {
    const dml::Op& op = net.ops[i];
    const dml::Op& po = net.ops[i + 1];
    if (op.kind != dml::OP_CONV || op.skip || op.merged_cout > 0 || op.cout != 64 || op.res >= 0 || op.in2 >= 0 || !op.relu) continue;
    if (po.kind != dml::OP_MAXPOOL || po.in != op.out || po.kh != 3 || po.stride != 2 || po.pad_h > 1) continue;
    if (i < e->tc.chained.size() && e->tc.chained[i]) continue;
    const bool stem = e->s2d && (int)i == net.s2d.op;
    if (!stem && (op.stride != 1 || i >= e->tc.ops.size() || !e->tc.ops[i].valid)) continue;
    bool other_reader = false;
    for (size_t j = 0; j < net.ops.size(); ++j)
      if (j != i + 1 && (net.ops[j].in == op.out || net.ops[j].res == op.out || net.ops[j].in2 == op.out)) other_reader = true;
    if (other_reader) continue;
    const dml::Tensor& tp = net.tensors[po.out];
    dml::ConvParams p = conv_params_of(e, (int)i, 0, e->micro_batch);
    dml::ConvPoolPlan cp;
    if (dml::conv_pool_plan(&cp, p, tp.h, tp.w, po.pad_h, e->max_batch) != 0) {
      if (tune_log) fprintf(stderr, "[tune] conv + pool at %s not planned: %s\n", op.name.c_str(), dml::tc_last_error());
      continue;
    }
    void* pool_out = e->tensor_ptr(po.out, 0);
    bool use = true;
    if (!getenv("DML_NO_AUTOTUNE") && !getenv("DML_FORCE_CONV_POOL")) {
      cudaEvent_t e0, e1;
      if (cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess) return fail(DML_ECUDA, "cudaEventCreate failed");
      const dml::Tensor& to = net.tensors[op.out];
      dml::PoolParams pp{};
      pp.in = e->tensor_ptr(po.in, 0); pp.out = pool_out;
      pp.n = e->micro_batch; pp.hi = to.h; pp.wi = to.w; pp.c = to.c; pp.in_pitch = to.c_pitch;
      pp.ho = tp.h; pp.wo = tp.w; pp.out_pitch = tp.c_pitch; pp.k = 3; pp.stride = 2; pp.pad = po.pad_h;
      float best[2] = {1e30f, 1e30f};
      for (int v = 0; v < 2; ++v) {
        auto run = [&]() -> bool {
          if (v == 0) {
            if (dml::launch_conv_tc_plan(stem ? e->stem_plan : e->tc.ops[i], p, nullptr) != 0) return false;
            dml::launch_maxpool(pp, false, nullptr);
            return true;
          }
          return dml::launch_conv_pool(cp, p, pool_out, tp.c_pitch, 0, nullptr) == 0;
        };
        bool ok = run() && cudaDeviceSynchronize() == cudaSuccess;
        for (int rep = 0; ok && rep < 3; ++rep) {
          cudaEventRecord(e0, nullptr);
          for (int k = 0; k < 4; ++k) ok = ok && run();
          cudaEventRecord(e1, nullptr);
          ok = ok && cudaEventSynchronize(e1) == cudaSuccess;
          float ms = 0.f;
          if (ok) cudaEventElapsedTime(&ms, e0, e1);
          if (ok && ms / 4.f < best[v]) best[v] = ms / 4.f;
        }
        if (!ok) {
          cudaEventDestroy(e0); cudaEventDestroy(e1);
          return fail(DML_ECUDA, "conv + pool timing failed at %s: %s / %s", op.name.c_str(), dml::tc_last_error(), cudaGetErrorString(cudaGetLastError()));
        }
      }
      cudaEventDestroy(e0);
      cudaEventDestroy(e1);
      cp.ms = best[1]; cp.ms_separate = best[0];
      use = best[1] < best[0];
      if (tune_log)
        fprintf(stderr, "[tune] conv + pool %s + %s (bands of %d pooled rows, %d tiles): fused %.4f ms, separate %.4f ms -> %s\n", op.name.c_str(),
                po.name.c_str(), cp.R, cp.T, best[1], best[0], use ? "fused" : "separate");
    }
    if (use) e->conv_pool[i] = cp;
  }
  return 0;
}

// Streams and events of the branch schedule.  Measured on InceptionV3 batch 128 (B200): 47.2k -> 50.9k images/s.  The kernels
// keep their full grids: capping the SMs of the convolutions that run side by side (equal caps of 50 / 75 %, or shares in
// proportion to their measured times) lost 3 - 10 %, stream priorities changed nothing.
int plan_branch_streams(dml_engine* e) {
  const dml::Net& net = e->net;
  if (!e->branch_streams || net.n_streams < 2) return 0;
  for (int k = 0; k + 1 < net.n_streams; ++k) {
    if (!e->side[k]) CUDA_TRY(cudaStreamCreateWithFlags(&e->side[k], cudaStreamNonBlocking));
    if (!e->join_ev[k]) CUDA_TRY(cudaEventCreateWithFlags(&e->join_ev[k], cudaEventDisableTiming));
  }
  e->op_event.resize(net.ops.size(), nullptr);
  for (size_t i = 0; i < net.ops.size(); ++i)
    if (net.ops[i].signal && !e->op_event[i]) CUDA_TRY(cudaEventCreateWithFlags(&e->op_event[i], cudaEventDisableTiming));
  return 0;
}

int run_layers(dml_engine* e, int img0, int cnt, cudaStream_t s0) {
  const dml::Net& net = e->net;
  // Branch streams: per-op event timing needs the serial order, so profiling runs serially
  const bool multi = e->branch_streams && net.n_streams > 1 && !e->profiling && !e->op_event.empty();
  // consecutive launches walk their tiles in opposite directions (kernels.h, ConvParams::rev): K1 ascending, the stem descending, ...
  static const bool pingpong = getenv("DML_NO_PINGPONG") == nullptr;
  int rev = 0;
  for (size_t oi = 0; oi < net.ops.size(); ++oi) {
    const dml::Op& op = net.ops[oi];
    if (op.skip) continue;  // computed by its leader (Op::merge_with)
    if (e->use_tc && oi < e->tc.chained.size() && e->tc.chained[oi]) continue;  // computed inside its chain's launch
    const int out_t = op.merged_out >= 0 ? op.merged_out : op.out;
    const int cout_x = op.merged_cout > 0 ? op.merged_cout : op.cout;
    const dml::Tensor& ti = net.tensors[op.in];
    const dml::Tensor& to = net.tensors[out_t];
    cudaStream_t s = multi && op.stream > 0 ? e->side[op.stream - 1] : s0;
    if (multi)
      for (int w : op.waits) CUDA_TRY(cudaStreamWaitEvent(s, e->op_event[w], 0));
    e->prof_begin((int)oi, s);
    if (pingpong) rev ^= 1;
    if (op.kind == dml::OP_CONV) {
      dml::ConvParams p{};
      p.rev = rev;
      p.in = e->tensor_ptr(op.in, img0);
      p.w = e->wts + op.merged_w_off;  // = wpack_off unless the op leads a merged group (its rows need not come first)
      p.bias = reinterpret_cast<const float*>(e->wts + op.merged_b_off);
      p.res = op.res >= 0 ? e->tensor_ptr(op.res, img0) : nullptr;
      p.out = e->tensor_ptr(out_t, img0);
      p.n = cnt; p.hi = ti.h; p.wi = ti.w; p.cin = op.cin; p.in_pitch = ti.c_pitch;
      p.ho = to.h; p.wo = to.w; p.cout = cout_x; p.out_pitch = to.c_pitch;
      p.res_pitch = op.res >= 0 ? net.tensors[op.res].c_pitch : 0;
      p.kh = op.kh; p.kw = op.kw; p.stride = op.stride; p.pad_h = op.pad_h; p.pad_w = op.pad_w;
      p.cin_pad = op.cin_pad; p.relu = op.relu; p.norelu_cols = op.norelu_cols;
      if (op.in2 >= 0) {
        const dml::Tensor& t2 = net.tensors[op.in2];
        p.in2 = e->tensor_ptr(op.in2, img0);
        p.hi2 = t2.h; p.wi2 = t2.w; p.cin2 = op.cin2; p.in2_pitch = t2.c_pitch; p.stride2 = op.stride2; p.cin2_pad = op.cin2_pad;
      }
      int nl = 0;
      if (oi < e->conv_pool.size() && e->conv_pool[oi].valid) {
        // this convolution and the max-pool behind it as one kernel: the pool op (oi + 1) is done here
        const dml::Op& po = net.ops[oi + 1];
        if (multi)
          for (int w : po.waits) CUDA_TRY(cudaStreamWaitEvent(s, e->op_event[w], 0));
        dml::ConvParams pc = conv_params_of(e, (int)oi, img0, cnt);
        if (dml::launch_conv_pool(e->conv_pool[oi], pc, e->tensor_ptr(po.out, img0), net.tensors[po.out].c_pitch, 0, s) != 0)
          return fail(DML_ECUDA, "conv + pool kernel launch failed at %s: %s", op.name.c_str(), dml::tc_last_error());
        e->launches += 1;
        e->prof_end(s);
        if (multi && po.signal) CUDA_TRY(cudaEventRecord(e->op_event[oi + 1], s));
        if (multi && op.signal) CUDA_TRY(cudaEventRecord(e->op_event[oi], s));
        if (pingpong) rev ^= 1;  // (the pool's turn)
        ++oi;
        continue;
      }
      if (e->use_tc && oi < e->tc.chains.size() && e->tc.chains[oi].valid) {
        // 3x3 + expand (+ residual / folded shortcut) + the next block's reduce as one kernel
        if (dml::launch_chain_tc(e->tc, net, e->wts, (int)oi, cnt, e->ws, e->off_arena, s, rev) != 0)
          return fail(DML_ECUDA, "chain kernel launch failed at %s: %s", op.name.c_str(), dml::tc_last_error());
        nl = 1;
      } else if (e->use_tc && oi < e->tc.duos.size() && e->tc.duos[oi].valid) {
        // 1x1 expand + residual + the next block's reduce as one kernel
        if (dml::launch_duo_tc(e->tc, net, e->wts, (int)oi, cnt, e->ws, e->off_arena, s, rev) != 0)
          return fail(DML_ECUDA, "expand + reduce kernel launch failed at %s: %s", op.name.c_str(), dml::tc_last_error());
        nl = 1;
      } else if (e->s2d && (int)oi == net.s2d.op) {
        dml::ConvParams p2 = stem_s2d_params(e, img0, cnt);
        p2.rev = rev;
        if (dml::launch_conv_tc_plan(e->stem_plan, p2, s) != 0)
          return fail(DML_ECUDA, "stem conv launch failed for %s: %s", op.name.c_str(), dml::tc_last_error());
        nl = 1;
      } else if (e->use_tc && dml::tc_supported(op)) {
        int rc = dml::launch_conv_tc(e->tc, (int)oi, p, img0, s, &nl);
        if (rc != 0) return fail(DML_ECUDA, "tcgen05 conv launch failed for %s: %s", op.name.c_str(), dml::tc_last_error());
      } else if (e->use_tc && dml::stem_tc_supported(p)) {
        if (dml::launch_stem_tc(p, s) != 0) return fail(DML_ECUDA, "stem conv launch failed for %s", op.name.c_str());
        nl = 1;
      } else {
        dml::launch_conv_naive(p, e->fp32, s);
        nl = 1;
      }
      e->launches += nl;
    } else if (op.kind == dml::OP_MAXPOOL || op.kind == dml::OP_AVGPOOL) {
      dml::PoolParams p{};
      p.rev = rev;
      p.in = e->tensor_ptr(op.in, img0);
      p.out = e->tensor_ptr(op.out, img0);
      p.n = cnt; p.hi = ti.h; p.wi = ti.w; p.c = ti.c; p.in_pitch = ti.c_pitch;
      p.ho = to.h; p.wo = to.w; p.out_pitch = to.c_pitch;
      p.k = op.kh; p.stride = op.stride; p.pad = op.pad_h;
      if (op.pool_bias_of >= 0) {  // the pool branch's 1x1 conv ran first (inside the head GEMM): its folded bias and ReLU come here
        p.bias = reinterpret_cast<const float*>(e->wts + net.ops[op.pool_bias_of].post_bias_off);
        p.relu = 1;
      }
      if (op.kind == dml::OP_MAXPOOL) dml::launch_maxpool(p, e->fp32, s);
      else dml::launch_avgpool(p, e->fp32, s);
      e->launches += 1;
    } else if (op.kind == dml::OP_GAP) {
      dml::launch_gap(e->tensor_ptr(op.in, img0), reinterpret_cast<float*>(e->tensor_ptr(op.out, img0)), cnt,
                      ti.h * ti.w, ti.c, ti.c_pitch, e->fp32, s);
      e->launches += 1;
    }
    e->prof_end(s);
    if (multi && op.signal) CUDA_TRY(cudaEventRecord(e->op_event[oi], s));
  }
  if (multi)  // every side stream joins the caller's stream: whatever follows (and the next call) sees the whole forward pass done
    for (int k = 0; k + 1 < net.n_streams; ++k) {
      CUDA_TRY(cudaEventRecord(e->join_ev[k], e->side[k]));
      CUDA_TRY(cudaStreamWaitEvent(s0, e->join_ev[k], 0));
    }
  CUDA_TRY(cudaGetLastError());
  return 0;
}

int run_head(dml_engine* e, int n, float* d_probs, int32_t* d_idx, float* d_prob, cudaStream_t s) {
  const dml::Net& net = e->net;
  const float* pooled = reinterpret_cast<const float*>(e->ws + e->off_pooled);
  float* logits = reinterpret_cast<float*>(e->ws + e->off_logits);
  int cin = net.tensors[net.pooled_tensor].c;
  e->prof_begin((int)net.ops.size(), s);
  dml::launch_fc(pooled, reinterpret_cast<const float*>(e->wts + net.fc_w_off),
                 reinterpret_cast<const float*>(e->wts + net.fc_b_off), logits, n, cin, !e->fp32, s);
  e->prof_end(s);
  e->prof_begin((int)net.ops.size() + 1, s);
  dml::launch_softmax_top5(logits, d_probs, d_idx, d_prob, n, s);
  e->prof_end(s);
  e->launches += 2;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

int run_pre(dml_engine* e, const uint8_t* d_pixels, const int64_t* d_img_off, const int32_t* d_hw,
            const int32_t* d_xtab, const int32_t* d_ytab, int img0, int cnt, cudaStream_t s) {
  dml::PreParams p{};
  const int S = e->net.input_size;
  p.pixels = d_pixels;
  p.img_off = d_img_off;
  p.hw = d_hw;
  p.xtab = d_xtab;
  p.ytab = d_ytab;
  p.out = e->tensor_ptr(e->net.input_tensor, 0);
  p.n = img0 + cnt; p.S = S; p.caffe = e->net.mode == 0;
  p.img_begin = img0; p.img_end = img0 + cnt;
  p.s2d = e->s2d ? 1 : 0;
  p.s2d_row_cells = e->net.s2d.row_cells;
  p.s2d_first_cell = e->net.s2d.pad2;
  dml::launch_preprocess(p, reinterpret_cast<const float*>(e->ws + e->off_lut), e->fp32, s);
  e->launches += 1;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

int check_ready(dml_engine* e, int n) {
  if (!e) return fail(DML_EINVAL, "engine is NULL");
  if (!e->ws) return fail(DML_ESTATE, "no workspace bound: call dml_bind_workspace first");
  if (!e->wts || !e->weights_loaded) return fail(DML_ESTATE, "weights not loaded: call dml_bind_weights + dml_load_weights first");
  if (n < 0 || n > e->max_batch) return fail(DML_ESIZE, "n = %d is outside [0, max_batch = %d]", n, e->max_batch);
  return 0;
}

}  // namespace

extern "C" {

// Internal: lets the other translation units of the library report through dml_last_error().
int dml_set_error(int code, const char* msg) { return fail(code, "%s", msg); }

const char* dml_last_error(void) { return g_err.c_str(); }
const char* dml_version(void) {
#ifdef DML_LEGACY_VARIANTS
  return "dml_b200 0.1 (sm_100a) +legacy-variants";
#else
  return "dml_b200 0.1 (sm_100a)";
#endif
}

int dml_model_info(int model, int* input_size, int* preprocess_mode) {
  if (int rc = check_model(model)) return rc;
  if (input_size) *input_size = model == DML_MODEL_RESNET50 ? 224 : 299;
  if (preprocess_mode) *preprocess_mode = model == DML_MODEL_RESNET50 ? 0 : 1;
  return 0;
}

int dml_model_num_weights(int model, int* num_arrays, int64_t* total_floats) {
  if (int rc = check_model(model)) return rc;
  dml::Net net;
  std::string err;
  if (!dml::build_weight_spec(model, &net, &err)) return fail(DML_EINVAL, "%s", err.c_str());
  if (num_arrays) *num_arrays = (int)net.warr.size();
  if (total_floats) *total_floats = net.total_floats;
  return 0;
}

int dml_model_weight_info(int model, int i, char* name, int name_cap, int64_t shape[4], int* rank) {
  if (int rc = check_model(model)) return rc;
  dml::Net net;
  std::string err;
  if (!dml::build_weight_spec(model, &net, &err)) return fail(DML_EINVAL, "%s", err.c_str());
  if (i < 0 || i >= (int)net.warr.size()) return fail(DML_EINVAL, "weight index %d out of range [0, %zu)", i, net.warr.size());
  const dml::WeightArr& a = net.warr[i];
  if (name && name_cap > 0) snprintf(name, name_cap, "%s", a.name.c_str());
  if (shape) for (int d = 0; d < 4; ++d) shape[d] = a.shape[d];
  if (rank) *rank = a.rank;
  return 0;
}

int dml_preprocess_lut(int model, float out[768]) {
  if (int rc = check_model(model)) return rc;
  if (!out) return fail(DML_EINVAL, "out is NULL");
  make_lut(model == DML_MODEL_RESNET50 ? 0 : 1, out);
  return 0;
}

int dml_create(int device, int model, int max_batch, int micro_batch, int flags, dml_engine** out) {
  if (!out) return fail(DML_EINVAL, "out is NULL");
  *out = nullptr;
  if (int rc = check_model(model)) return rc;
  if (max_batch < 1 || max_batch > 4096) return fail(DML_ESIZE, "max_batch = %d is outside [1, 4096]", max_batch);
  if (micro_batch < 0) return fail(DML_EINVAL, "micro_batch = %d is negative", micro_batch);
  int ndev = 0;
  cudaError_t ce = cudaGetDeviceCount(&ndev);
  if (ce != cudaSuccess || ndev == 0)
    return fail(DML_ECUDA, "no CUDA device available (%s): this library has no CPU fallback", cudaGetErrorString(ce));
  if (device < 0 || device >= ndev) return fail(DML_EINVAL, "device %d out of range [0, %d)", device, ndev);
  CUDA_TRY(cudaSetDevice(device));
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return fail(DML_ECUDA, "device %d is sm_%d%d; this library is built for sm_100a (B200) only", device, prop.major, prop.minor);

  dml_engine* e = new dml_engine();
  e->device = device; e->model = model; e->max_batch = max_batch; e->flags = flags;
  e->fp32 = (flags & DML_FLAG_FP32_CHECK) != 0;
  e->use_tc = !e->fp32 && (flags & DML_FLAG_NO_TENSOR_CORE) == 0;
  e->keep_all = (flags & DML_FLAG_KEEP_ALL) != 0;
  e->act_bytes = e->fp32 ? 4 : 2;
  if (micro_batch == 0) micro_batch = max_batch;  // whole batch per layer: large launches beat L2-resident waves here
  e->micro_batch = micro_batch < max_batch ? micro_batch : max_batch;
  std::string err;
  // projection shortcuts folded into the block's last conv: tensor-core engines only (the CUDA-core kernel has one source)
  const bool fuse_sc = e->use_tc && !e->keep_all && getenv("DML_NO_FUSE_SHORTCUT") == nullptr;
  if (!dml::build_net(model, e->keep_all, e->fp32 ? 4 : 2, !e->fp32, fuse_sc, &e->net, &err, getenv("DML_NO_POOL_COMMUTE") == nullptr)) {
    delete e;
    return fail(DML_EINVAL, "%s", err.c_str());
  }
  // independent branches of a block on their own streams (bf16 tensor-core engines; DML_BRANCH_STREAMS=0: serial order)
  const char* bs = getenv("DML_BRANCH_STREAMS");
  e->branch_streams = e->use_tc && !e->keep_all && !(bs && atoi(bs) == 0);
  if (e->branch_streams) dml::plan_streams(&e->net, 4);
  const dml::Net& net = e->net;
  size_t off = 0;
  e->off_lut = off; off = align_up(off + 768 * 4, 256);
  e->s2d = e->use_tc && net.s2d.op >= 0 && getenv("DML_NO_S2D") == nullptr;
  e->input_img_bytes = e->s2d ? (size_t)net.s2d.size * net.s2d.row_cells * 16 * 2 : (size_t)net.input_size * net.input_size * 3 * e->act_bytes;
  e->off_input = off; off = align_up(off + (size_t)max_batch * e->input_img_bytes + 64, 256);
  e->off_arena = off; off = align_up(off + (size_t)net.arena_elems * e->micro_batch * e->act_bytes, 256);
  e->off_pooled = off; off = align_up(off + (size_t)max_batch * net.tensors[net.pooled_tensor].c * 4, 256);
  e->off_logits = off; off = align_up(off + (size_t)max_batch * 1000 * 4, 256);
  e->off_tc = off; off = align_up(off + dml::tc_scratch_bytes(net, e->micro_batch), 256);
  e->ws_need = off;
  dml::init_kernels();
  if (e->use_tc) {
    if (dml::tc_init() != 0) {
      std::string m = dml::tc_last_error();
      delete e;
      return fail(DML_ECUDA, "tcgen05 path init failed: %s", m.c_str());
    }
  }
  *out = e;
  return 0;
}

void dml_destroy(dml_engine* e) { delete e; }

int dml_workspace_bytes(dml_engine* e, size_t* bytes) {
  if (!e || !bytes) return fail(DML_EINVAL, "NULL argument");
  *bytes = e->ws_need;
  return 0;
}
int dml_weights_bytes(dml_engine* e, size_t* bytes) {
  if (!e || !bytes) return fail(DML_EINVAL, "NULL argument");
  *bytes = (size_t)e->net.weights_bytes;
  return 0;
}

int dml_bind_workspace(dml_engine* e, void* dev_ptr, size_t bytes) {
  if (!e || !dev_ptr) return fail(DML_EINVAL, "NULL argument");
  if (bytes < e->ws_need) return fail(DML_ESIZE, "workspace of %zu bytes is smaller than the %zu required", bytes, e->ws_need);
  if ((uintptr_t)dev_ptr % 256) return fail(DML_EINVAL, "workspace pointer must be 256-byte aligned");
  CUDA_TRY(cudaSetDevice(e->device));
  e->ws = reinterpret_cast<uint8_t*>(dev_ptr);
  e->ws_bytes = bytes;
  float lut[768];
  make_lut(e->net.mode, lut);
  CUDA_TRY(cudaMemcpy(e->ws + e->off_lut, lut, sizeof lut, cudaMemcpyHostToDevice));
  // Activation arena starts at zero so that padded/unused lanes are finite.
  CUDA_TRY(cudaMemset(e->ws + e->off_input, 0, e->ws_need - e->off_input));
  if (e->use_tc && e->wts) {
    if (dml::tc_plan(&e->tc, e->net, e->ws, e->off_arena, e->off_input, e->off_tc, e->wts, e->micro_batch, e->max_batch) != 0)
      return fail(DML_ECUDA, "tcgen05 plan failed: %s", dml::tc_last_error());
    if (plan_stem(e) != 0) return DML_ECUDA;
    if (int rc = plan_branch_streams(e)) return rc;
    if (int rc = plan_conv_pools(e)) return rc;
  }
  return 0;
}

int dml_bind_weights(dml_engine* e, void* dev_ptr, size_t bytes) {
  if (!e || !dev_ptr) return fail(DML_EINVAL, "NULL argument");
  if (bytes < (size_t)e->net.weights_bytes)
    return fail(DML_ESIZE, "weights buffer of %zu bytes is smaller than the %lld required", bytes, (long long)e->net.weights_bytes);
  if ((uintptr_t)dev_ptr % 256) return fail(DML_EINVAL, "weights pointer must be 256-byte aligned");
  e->wts = reinterpret_cast<uint8_t*>(dev_ptr);
  e->wts_bytes = bytes;
  e->weights_loaded = false;
  if (e->use_tc && e->ws) {
    CUDA_TRY(cudaSetDevice(e->device));
    if (dml::tc_plan(&e->tc, e->net, e->ws, e->off_arena, e->off_input, e->off_tc, e->wts, e->micro_batch, e->max_batch) != 0)
      return fail(DML_ECUDA, "tcgen05 plan failed: %s", dml::tc_last_error());
    if (plan_stem(e) != 0) return DML_ECUDA;
    if (int rc = plan_branch_streams(e)) return rc;
    if (int rc = plan_conv_pools(e)) return rc;
  }
  return 0;
}

int dml_load_weights(dml_engine* e, const float* blob, size_t num_floats) {
  if (!e || !blob) return fail(DML_EINVAL, "NULL argument");
  if (!e->wts) return fail(DML_ESTATE, "no weights buffer bound: call dml_bind_weights first");
  const dml::Net& net = e->net;
  if ((int64_t)num_floats != net.total_floats)
    return fail(DML_ESIZE, "weight blob has %zu floats, model needs %lld", num_floats, (long long)net.total_floats);
  CUDA_TRY(cudaSetDevice(e->device));
  std::vector<uint8_t> host((size_t)net.weights_bytes, 0);
  for (const dml::Op& op : net.ops) {
    if (op.kind != dml::OP_CONV) continue;
    const float* K = blob + net.warr[op.w_kernel].offset;  // HWIO
    const float* bias = op.w_bias >= 0 ? blob + net.warr[op.w_bias].offset : nullptr;
    const float* gamma = op.w_gamma >= 0 ? blob + net.warr[op.w_gamma].offset : nullptr;
    const float* beta = blob + net.warr[op.w_beta].offset;
    const float* mean = blob + net.warr[op.w_mean].offset;
    const float* var = blob + net.warr[op.w_var].offset;
    const int taps = op.kh * op.kw;
    float* bout = reinterpret_cast<float*>(host.data() + op.bias_off);
    for (int co = 0; co < op.cout; ++co) {
      // BN folded into the conv: s = gamma / sqrt(var + eps); W' = W*s; b' = beta + (b - mean)*s
      double s = (gamma ? (double)gamma[co] : 1.0) / std::sqrt((double)var[co] + (double)op.eps);
      bout[co] = (float)((double)beta[co] + ((bias ? (double)bias[co] : 0.0) - (double)mean[co]) * s);
      if (op.post_pool) {  // added by the pool op after averaging, not by the GEMM (graph.h, Op::post_pool)
        reinterpret_cast<float*>(host.data() + op.post_bias_off)[co] = bout[co];
        bout[co] = 0.f;
      }
      const size_t row_k = (size_t)taps * op.cin_pad + (size_t)op.cin2_pad;  // packed K of one output channel
      if (op.in2 >= 0) {
        // folded projection shortcut: its own BN scale on its columns [taps * cin_pad, + cin2), its folded bias added
        const float* K2 = blob + net.warr[op.w2_kernel].offset;  // [1][1][cin2][cout]
        const float* bias2 = op.w2_bias >= 0 ? blob + net.warr[op.w2_bias].offset : nullptr;
        const float* gamma2 = op.w2_gamma >= 0 ? blob + net.warr[op.w2_gamma].offset : nullptr;
        const double s2 = (gamma2 ? (double)gamma2[co] : 1.0) /
                          std::sqrt((double)(blob + net.warr[op.w2_var].offset)[co] + (double)op.eps2);
        const double b2 = (double)(blob + net.warr[op.w2_beta].offset)[co] +
                          ((bias2 ? (double)bias2[co] : 0.0) - (double)(blob + net.warr[op.w2_mean].offset)[co]) * s2;
        bout[co] = (float)((double)bout[co] + b2);
        for (int c = 0; c < op.cin2; ++c)
          reinterpret_cast<uint16_t*>(host.data() + op.wpack_off)[(size_t)co * row_k + (size_t)taps * op.cin_pad + c] =
              f32_to_bf16_rn((float)((double)K2[(size_t)c * op.cout + co] * s2));
      }
      for (int t = 0; t < taps; ++t)
        for (int c = 0; c < op.cin; ++c) {
          float v = (float)((double)K[((size_t)t * op.cin + c) * op.cout + co] * s);
          size_t idx = (size_t)co * row_k + (size_t)t * op.cin_pad + c;
          if (e->fp32) reinterpret_cast<float*>(host.data() + op.wpack_off)[idx] = v;
          else reinterpret_cast<uint16_t*>(host.data() + op.wpack_off)[idx] = f32_to_bf16_rn(v);
          if (!e->fp32 && net.s2d.op >= 0 && &op == &net.ops[net.s2d.op]) {
            // the same folded weight at its space-to-depth position: tap (r, s) reads pixel (2*oh + r - pad, ...)
            // = cell (oh + a, ow + b), sub-pixel (py, px) with r - pad = 2a + py
            const int r = t / op.kw, sx = t % op.kw;
            const int dr = r - op.pad_h, ds = sx - op.pad_w;
            const int a = dr >= 0 ? dr / 2 : -((-dr + 1) / 2), b = ds >= 0 ? ds / 2 : -((-ds + 1) / 2);
            const int py = dr - 2 * a, px = ds - 2 * b;
            const int tap2 = (a + net.s2d.pad2) * net.s2d.k2 + (b + net.s2d.pad2);
            const size_t idx2 = ((size_t)co * net.s2d.k2 * net.s2d.k2 + tap2) * 16 + (size_t)((py * 2 + px) * 3 + c);
            reinterpret_cast<uint16_t*>(host.data() + net.s2d.w_off)[idx2] = f32_to_bf16_rn(v);
          }
        }
    }
  }
  memcpy(host.data() + net.fc_w_off, blob + net.warr[net.fc_kernel].offset, (size_t)net.warr[net.fc_kernel].count * 4);
  memcpy(host.data() + net.fc_b_off, blob + net.warr[net.fc_bias].offset, 1000 * 4);
  CUDA_TRY(cudaMemcpy(e->wts, host.data(), host.size(), cudaMemcpyHostToDevice));
  e->weights_loaded = true;
  return 0;
}

int dml_weights_ready(dml_engine* e) {
  if (!e) return fail(DML_EINVAL, "engine is NULL");
  if (!e->wts) return fail(DML_ESTATE, "no weights buffer bound");
  e->weights_loaded = true;
  return 0;
}

int dml_preprocess(dml_engine* e, const uint8_t* d_pixels, const int64_t* d_img_off, const int32_t* d_hw,
                   const int32_t* d_xtab, const int32_t* d_ytab, int n, void* stream) {
  if (!e) return fail(DML_EINVAL, "engine is NULL");
  if (!e->ws) return fail(DML_ESTATE, "no workspace bound: call dml_bind_workspace first");
  if (n < 0 || n > e->max_batch) return fail(DML_ESIZE, "n = %d is outside [0, max_batch = %d]", n, e->max_batch);
  if (n == 0) return 0;
  if (!d_pixels || !d_img_off || !d_hw || !d_xtab || !d_ytab) return fail(DML_EINVAL, "NULL device pointer");
  CUDA_TRY(cudaSetDevice(e->device));
  return run_pre(e, d_pixels, d_img_off, d_hw, d_xtab, d_ytab, 0, n, (cudaStream_t)stream);
}

int dml_forward(dml_engine* e, int n, float* d_probs, int32_t* d_top5_idx, float* d_top5_prob, void* stream) {
  if (int rc = check_ready(e, n)) return rc;
  if (n == 0) return 0;
  if (!d_top5_idx || !d_top5_prob) return fail(DML_EINVAL, "NULL output pointer");
  CUDA_TRY(cudaSetDevice(e->device));
  cudaStream_t s = (cudaStream_t)stream;
  for (int i0 = 0; i0 < n; i0 += e->micro_batch) {
    if (e->cancel.exchange(0)) return fail(DML_ECANCELLED, "cancelled before image %d of %d", i0, n);
    int cnt = n - i0 < e->micro_batch ? n - i0 : e->micro_batch;
    if (int rc = run_layers(e, i0, cnt, s)) return rc;
  }
  return run_head(e, n, d_probs, d_top5_idx, d_top5_prob, s);
}

int dml_infer(dml_engine* e, const uint8_t* d_pixels, const int64_t* d_img_off, const int32_t* d_hw,
              const int32_t* d_xtab, const int32_t* d_ytab, int n, float* d_probs, int32_t* d_top5_idx,
              float* d_top5_prob, void* stream) {
  if (int rc = check_ready(e, n)) return rc;
  if (n == 0) return 0;
  if (!d_pixels || !d_img_off || !d_hw || !d_xtab || !d_ytab) return fail(DML_EINVAL, "NULL device pointer");
  if (!d_top5_idx || !d_top5_prob) return fail(DML_EINVAL, "NULL output pointer");
  CUDA_TRY(cudaSetDevice(e->device));
  cudaStream_t s = (cudaStream_t)stream;
  for (int i0 = 0; i0 < n; i0 += e->micro_batch) {
    if (e->cancel.exchange(0)) return fail(DML_ECANCELLED, "cancelled before image %d of %d", i0, n);
    int cnt = n - i0 < e->micro_batch ? n - i0 : e->micro_batch;
    if (int rc = run_pre(e, d_pixels, d_img_off, d_hw, d_xtab, d_ytab, i0, cnt, s)) return rc;
    if (int rc = run_layers(e, i0, cnt, s)) return rc;
  }
  return run_head(e, n, d_probs, d_top5_idx, d_top5_prob, s);
}

int dml_cancel(dml_engine* e) {
  if (!e) return fail(DML_EINVAL, "engine is NULL");
  e->cancel.store(1);
  return 0;
}

int dml_cancel_clear(dml_engine* e) {
  if (!e) return fail(DML_EINVAL, "engine is NULL");
  e->cancel.store(0);
  return 0;
}

int dml_num_tensors(dml_engine* e, int* n) {
  if (!e || !n) return fail(DML_EINVAL, "NULL argument");
  *n = (int)e->net.tensors.size();
  return 0;
}

int dml_tensor_info(dml_engine* e, int i, char* name, int name_cap, int* h, int* w, int* c, void** dev_ptr,
                    int* c_pitch, int* elem_bytes, int* keep) {
  if (!e) return fail(DML_EINVAL, "engine is NULL");
  if (i < 0 || i >= (int)e->net.tensors.size()) return fail(DML_EINVAL, "tensor index %d out of range", i);
  const dml::Tensor& t = e->net.tensors[i];
  if (name && name_cap > 0) snprintf(name, name_cap, "%s", t.name.c_str());
  if (h) *h = t.h;
  if (w) *w = t.w;
  if (c) *c = t.c;
  if (dev_ptr) *dev_ptr = e->ws ? e->tensor_ptr(i, 0) : nullptr;
  if (c_pitch) *c_pitch = t.c_pitch;
  if (elem_bytes) *elem_bytes = t.buf == -2 ? 4 : e->act_bytes;
  if (keep) *keep = e->keep_all || t.buf < 0;
  return 0;
}

int dml_input_tensor(dml_engine* e, void** dev_ptr, int* elem_bytes) {
  if (!e || !dev_ptr) return fail(DML_EINVAL, "NULL argument");
  if (!e->ws) return fail(DML_ESTATE, "no workspace bound");
  *dev_ptr = e->ws + e->off_input;
  if (elem_bytes) *elem_bytes = e->act_bytes;
  return 0;
}

int dml_input_layout(dml_engine* e, int* layout, int* cells, int* cell_channels) {
  if (!e || !layout) return fail(DML_EINVAL, "NULL argument");
  *layout = e->s2d ? 1 : 0;
  if (cells) *cells = e->s2d ? e->net.s2d.size : e->net.input_size;
  if (cell_channels) *cell_channels = e->s2d ? 16 : 3;
  return 0;
}

int dml_input_rows(dml_engine* e, int* row_cells, int* first_cell, int* swap_bit) {
  if (!e) return fail(DML_EINVAL, "engine is NULL");
  if (row_cells) *row_cells = e->s2d ? e->net.s2d.row_cells : e->net.input_size;
  if (first_cell) *first_cell = e->s2d ? e->net.s2d.pad2 : 0;
  if (swap_bit) *swap_bit = e->s2d ? 2 : -1;
  return 0;
}

int dml_logits(dml_engine* e, float** dev_ptr) {
  if (!e || !dev_ptr) return fail(DML_EINVAL, "NULL argument");
  if (!e->ws) return fail(DML_ESTATE, "no workspace bound");
  *dev_ptr = reinterpret_cast<float*>(e->ws + e->off_logits);
  return 0;
}

int dml_num_ops(dml_engine* e, int* n) {
  if (!e || !n) return fail(DML_EINVAL, "NULL argument");
  *n = (int)e->net.ops.size() + 2;
  return 0;
}

int dml_op_schedule(dml_engine* e, int i, int* stream, int* num_waits, int* num_streams) {
  if (!e) return fail(DML_EINVAL, "engine is NULL");
  const dml::Net& net = e->net;
  if (i < 0 || i >= (int)net.ops.size()) return fail(DML_EINVAL, "op index %d out of range", i);
  const bool on = e->branch_streams && net.n_streams > 1;
  if (stream) *stream = on ? net.ops[i].stream : 0;
  if (num_waits) *num_waits = on ? (int)net.ops[i].waits.size() : 0;
  if (num_streams) *num_streams = on ? net.n_streams : 1;
  return 0;
}

int dml_op_info(dml_engine* e, int i, char* name, int name_cap, int* kind, int* impl, int64_t* macs_per_img,
                int64_t* act_bytes_per_img, int64_t* weight_bytes) {
  if (!e) return fail(DML_EINVAL, "engine is NULL");
  const dml::Net& net = e->net;
  const int nops = (int)net.ops.size();
  if (i < 0 || i >= nops + 2) return fail(DML_EINVAL, "op index %d out of range", i);
  std::string nm;
  int k = 0, im = 0;
  int64_t macs = 0, ab = 0, wb = 0;
  const int cin_fc = net.tensors[net.pooled_tensor].c;
  if (i == nops) {
    nm = "predictions"; k = 4; macs = (int64_t)cin_fc * 1000; ab = (int64_t)cin_fc * 4 + 4000; wb = (int64_t)cin_fc * 4000 + 4000;
  } else if (i == nops + 1) {
    nm = "softmax_top5"; k = 5; ab = 4000 + 40;
  } else {
    const dml::Op& op = net.ops[i];
    const dml::Tensor& ti = net.tensors[op.in];
    const dml::Tensor& to = net.tensors[op.merged_out >= 0 ? op.merged_out : op.out];
    const int cout_x = op.merged_cout > 0 ? op.merged_cout : op.cout;
    nm = op.name; k = (int)op.kind;
    const bool chained = e->use_tc && (size_t)i < e->tc.chained.size() && e->tc.chained[i];
    const bool pooled_by_conv = i > 0 && (size_t)(i - 1) < e->conv_pool.size() && e->conv_pool[i - 1].valid;  // done inside the conv's launch
    if (op.skip || chained || pooled_by_conv) {  // a follower of a merged 1x1 pair / a member of a fused chain: reported by its leader
      if (name && name_cap > 0) snprintf(name, name_cap, "%s", nm.c_str());
      if (kind) *kind = k;
      if (impl) *impl = e->use_tc ? 1 : 0;
      if (macs_per_img) *macs_per_img = 0;
      if (act_bytes_per_img) *act_bytes_per_img = 0;
      if (weight_bytes) *weight_bytes = 0;
      return 0;
    }
    int64_t in_b = (int64_t)ti.h * ti.w * ti.c * e->act_bytes;
    int64_t out_b = op.kind == dml::OP_GAP ? (int64_t)to.c * 4 : (int64_t)to.h * to.w * to.c * e->act_bytes;
    ab = in_b + out_b + (op.res >= 0 ? out_b : 0);
    if (op.kind == dml::OP_CONV) {
      macs = (int64_t)to.h * to.w * cout_x * (op.kh * op.kw * op.cin + op.cin2);
      wb = (int64_t)cout_x * (op.kh * op.kw * op.cin + op.cin2) * (e->fp32 ? 4 : 2) + cout_x * 4;
      if (op.in2 >= 0) {  // the second source is read at stride2: only the pixels the 1x1 convolution touches
        const dml::Tensor& t2 = net.tensors[op.in2];
        ab += (int64_t)((t2.h + op.stride2 - 1) / op.stride2) * ((t2.w + op.stride2 - 1) / op.stride2) * t2.c * e->act_bytes;
      }
      im = (e->use_tc && (dml::tc_supported(op) || (op.cin == 3 && op.stride == 2))) ? 1 : 0;
      if (e->use_tc && (size_t)i < e->tc.chains.size() && e->tc.chains[i].valid) {
        // fused chain: this launch also does the expand (+ shortcut) and the next reduce; y2 and the re-read of `out` never reach HBM
        const dml::Op& oe = net.ops[op.chain_exp];
        const dml::Tensor& te = net.tensors[oe.out];
        const int64_t px = (int64_t)te.h * te.w;
        macs += px * oe.cout * (oe.cin + oe.cin2);
        wb += (int64_t)oe.cout * (oe.cin + oe.cin2) * 2 + oe.cout * 4;
        ab = in_b + px * oe.cout * 2 /*out*/ + (oe.res >= 0 ? px * oe.cout * 2 : px * oe.cin2 * 2) /*residual or second source*/;
        if (op.chain_red >= 0) {
          const dml::Op& orr = net.ops[op.chain_red];
          macs += px * orr.cout * orr.cin;
          wb += (int64_t)orr.cout * orr.cin * 2 + orr.cout * 4;
          ab += px * orr.cout * 2;
        }
      }
      if ((size_t)i < e->conv_pool.size() && e->conv_pool[i].valid) {
        // fused with the max-pool behind it: the convolution's own output never reaches HBM, the pooled rows do
        const dml::Tensor& tp = net.tensors[net.ops[i + 1].out];
        ab = in_b + (int64_t)tp.h * tp.w * tp.c * e->act_bytes;
      }
      if (e->use_tc && (size_t)i < e->tc.duos.size() && e->tc.duos[i].valid) {
        // fused duo: this launch also does the next block's reduce; the re-read of `out` never reaches HBM
        // (a folded projection shortcut's MACs and bytes are in the expand's own figures above)
        const dml::Op& orr = net.ops[op.duo_red];
        const int64_t px = (int64_t)to.h * to.w;
        macs += px * orr.cout * orr.cin;
        wb += (int64_t)orr.cout * orr.cin * 2 + orr.cout * 4;
        ab += px * orr.cout * 2;
      }
    }
  }
  if (name && name_cap > 0) snprintf(name, name_cap, "%s", nm.c_str());
  if (kind) *kind = k;
  if (impl) *impl = im;
  if (macs_per_img) *macs_per_img = macs;
  if (act_bytes_per_img) *act_bytes_per_img = ab;
  if (weight_bytes) *weight_bytes = wb;
  return 0;
}

int dml_profile_forward(dml_engine* e, int n, int32_t* d_top5_idx, float* d_top5_prob, float* op_ms, int cap,
                        void* stream) {
  if (int rc = check_ready(e, n)) return rc;
  if (!d_top5_idx || !d_top5_prob) return fail(DML_EINVAL, "NULL output pointer");
  const int nops = (int)e->net.ops.size() + 2;
  if (!op_ms || cap < nops) return fail(DML_EINVAL, "op_ms must hold at least %d floats", nops);
  CUDA_TRY(cudaSetDevice(e->device));
  cudaStream_t s = (cudaStream_t)stream;
  for (int i = 0; i < nops; ++i) op_ms[i] = 0.f;
  e->profiling = true;
  e->prof.clear();
  int rc = 0;
  for (int i0 = 0; i0 < n && rc == 0; i0 += e->micro_batch) {
    int cnt = n - i0 < e->micro_batch ? n - i0 : e->micro_batch;
    rc = run_layers(e, i0, cnt, s);
  }
  if (rc == 0) rc = run_head(e, n, nullptr, d_top5_idx, d_top5_prob, s);
  e->profiling = false;
  cudaError_t ce = cudaStreamSynchronize(s);
  for (auto& pr : e->prof) {
    float ms = 0.f;
    if (ce == cudaSuccess && rc == 0 && cudaEventElapsedTime(&ms, pr.second.first, pr.second.second) == cudaSuccess)
      op_ms[pr.first] += ms;
    cudaEventDestroy(pr.second.first);
    cudaEventDestroy(pr.second.second);
  }
  e->prof.clear();
  if (rc) return rc;
  CUDA_TRY(ce);
  return 0;
}

int dml_launch_count(dml_engine* e, int64_t* launches) {
  if (!e || !launches) return fail(DML_EINVAL, "NULL argument");
  *launches = e->launches;
  return 0;
}

int dml_conv2d_bf16(int impl, const void* d_in, const void* d_w, const float* d_bias, const void* d_res, void* d_out,
                    int n, int hi, int wi, int cin, int in_pitch, int cout, int out_pitch, int res_pitch, int kh,
                    int kw, int stride, int pad_h, int pad_w, int cin_pad, int relu, int tile_n, int stages,
                    void* stream) {
  if (!d_in || !d_w || !d_bias || !d_out) return fail(DML_EINVAL, "NULL device pointer");
  if (n < 1 || stride < 1 || kh < 1 || kw < 1) return fail(DML_EINVAL, "bad conv geometry");
  dml::ConvParams p{};
  p.in = d_in; p.w = d_w; p.bias = d_bias; p.res = d_res; p.out = d_out;
  p.n = n; p.hi = hi; p.wi = wi; p.cin = cin; p.in_pitch = in_pitch;
  p.ho = (hi + 2 * pad_h - kh) / stride + 1;
  p.wo = (wi + 2 * pad_w - kw) / stride + 1;
  p.cout = cout; p.out_pitch = out_pitch; p.res_pitch = res_pitch;
  p.kh = kh; p.kw = kw; p.stride = stride; p.pad_h = pad_h; p.pad_w = pad_w; p.cin_pad = cin_pad; p.relu = relu;
  if (impl == 0) {
    dml::launch_conv_naive(p, false, (cudaStream_t)stream);
    CUDA_TRY(cudaGetLastError());
    return 0;
  }
  if (impl == 3) {
#ifndef DML_LEGACY_VARIANTS
    return fail(DML_EINVAL, "impl 3 (the patch-expanding C=3 stem kernel) is a superseded A/B twin: not in this build (compile with -DDML_LEGACY_VARIANTS)");
#endif
    if (!dml::stem_tc_supported(p)) return fail(DML_EINVAL, "shape is not a C_in = 3 stride-2 stem convolution");
    if (dml::launch_stem_tc(p, (cudaStream_t)stream) != 0) return fail(DML_ECUDA, "stem conv launch failed");
    return 0;
  }
  if (impl < 1 || impl > 8) return fail(DML_EINVAL, "impl must be 0 (CUDA cores), 1 (tcgen05, tile per CTA), 2 (tcgen05, persistent im2col), 3 (C=3 stem), 4 (tcgen05 halo kernel), 5 (CTA pairs), 6 (weights as A), 7 (halo kernel, resident filter) or 8 (the same on CTA pairs)");
  if (dml::tc_init() != 0) return fail(DML_ECUDA, "%s", dml::tc_last_error());
  dml::TcOpPlan plan;
  dml::TcTune tune;
  tune.bn = tile_n; tune.stages = stages & 0xff; tune.group = (stages >> 8) & 0xff; tune.res_direct = (stages >> 16) & 3; tune.row_bytes = ((stages >> 18) & 1) ? 64 : 0; tune.version = impl == 4 ? 3 : (impl == 5 ? 4 : (impl == 6 ? 5 : (impl == 7 ? 6 : (impl == 8 ? 7 : impl))));  // 5 = CTA-pair kernel, 6 = weights-as-A kernel, 7 = resident-filter halo kernel, 8 = the same on CTA pairs
  if (dml::tc_plan_one(&plan, p, n, tune) != 0) return fail(DML_EINVAL, "%s", dml::tc_last_error());
  if (dml::launch_conv_tc_plan(plan, p, (cudaStream_t)stream) != 0) return fail(DML_ECUDA, "%s", dml::tc_last_error());
  return 0;
}

int dml_chain_bf16(const void* d_y1, const void* d_w3, const float* d_b3, const void* d_we, const float* d_be,
                   const void* d_res, const void* d_src2, void* d_out, const void* d_wr, const float* d_br, void* d_y1n,
                   int n, int h, int w, void* stream) {
  if (n < 1) return fail(DML_EINVAL, "n must be positive");
  if (dml::tc_init() != 0) return fail(DML_ECUDA, "%s", dml::tc_last_error());
  dml::ChainArgs a;
  a.y1 = d_y1; a.y1_pitch = 64;
  a.w3 = d_w3; a.b3 = d_b3; a.we = d_we; a.be = d_be;
  a.res = d_res; a.res_pitch = 256;
  a.src2 = d_src2; a.src2_pitch = 64;
  a.out = d_out; a.out_pitch = 256;
  a.wr = d_wr; a.br = d_br; a.y1n = d_y1n; a.y1n_pitch = 64;
  a.n = n; a.h = h; a.w = w;
  dml::ChainPlan plan;
  if (dml::chain_plan(&plan, a, n) != 0) return fail(DML_EINVAL, "%s", dml::tc_last_error());
  if (dml::launch_chain(plan, a, (cudaStream_t)stream) != 0) return fail(DML_ECUDA, "%s", dml::tc_last_error());
  return 0;
}

int dml_duo_bf16(const void* d_y2, const void* d_we, const float* d_be, const void* d_res, const void* d_src2, void* d_out,
                 const void* d_wr, const float* d_br, void* d_y1n, int n, int h, int w, int kmid, int k2, int s2_stride, int nexp,
                 int nred, void* stream) {
  const int m = n * h * w;
  if (n < 1 || h < 1 || w < 1) return fail(DML_EINVAL, "n, h, w must be positive");
  if (dml::tc_init() != 0) return fail(DML_ECUDA, "%s", dml::tc_last_error());
  dml::DuoArgs a;
  a.y2 = d_y2; a.y2_pitch = kmid;
  a.we = d_we; a.be = d_be;
  a.res = d_res; a.res_pitch = nexp;
  a.src2 = d_src2; a.src2_pitch = k2; a.s2_stride = d_src2 ? s2_stride : 1;
  a.n = n; a.h = h; a.w = w;
  a.out = d_out; a.out_pitch = nexp;
  a.wr = d_wr; a.br = d_br; a.y1n = d_y1n; a.y1n_pitch = nred;
  a.M = m; a.k1 = kmid; a.kmid = kmid + (d_src2 ? k2 : 0); a.nexp = nexp; a.nred = nred;
  dml::DuoPlan plan;
  if (dml::duo_plan(&plan, a, m) != 0) return fail(DML_EINVAL, "%s", dml::tc_last_error());
  if (dml::launch_duo(plan, a, (cudaStream_t)stream) != 0) return fail(DML_ECUDA, "%s", dml::tc_last_error());
  return 0;
}

int dml_conv_pool_bf16(const void* d_in, const void* d_w, const float* d_bias, void* d_pool_out, int n, int hi, int wi, int cin,
                       int cin_pad, int kh, int kw, int pad, int relu, int pool_pad, void* stream) {
  if (n < 1) return fail(DML_EINVAL, "n must be positive");
  if (dml::tc_init() != 0) return fail(DML_ECUDA, "%s", dml::tc_last_error());
  dml::ConvParams p{};
  p.in = d_in; p.w = d_w; p.bias = d_bias; p.res = nullptr; p.out = d_pool_out;  // (the convolution's own output is never written)
  p.n = n; p.hi = hi; p.wi = wi; p.cin = cin; p.in_pitch = cin;
  p.ho = hi + 2 * pad - kh + 1; p.wo = wi + 2 * pad - kw + 1;
  p.cout = 64; p.out_pitch = 64; p.kh = kh; p.kw = kw; p.stride = 1; p.pad_h = pad; p.pad_w = pad; p.cin_pad = cin_pad; p.relu = relu;
  const int hp = (p.ho + 2 * pool_pad - 3) / 2 + 1, wq = (p.wo + 2 * pool_pad - 3) / 2 + 1;
  dml::ConvPoolPlan plan;
  if (dml::conv_pool_plan(&plan, p, hp, wq, pool_pad, n) != 0) return fail(DML_EINVAL, "%s", dml::tc_last_error());
  if (dml::launch_conv_pool(plan, p, d_pool_out, 64, 0, (cudaStream_t)stream) != 0) return fail(DML_ECUDA, "%s", dml::tc_last_error());
  return 0;
}

int dml_sync(dml_engine* e) {
  if (!e) return fail(DML_EINVAL, "engine is NULL");
  CUDA_TRY(cudaSetDevice(e->device));
  CUDA_TRY(cudaDeviceSynchronize());
  return 0;
}

}  // extern "C"
