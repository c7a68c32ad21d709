// graph.cc -- ResNet50 / InceptionV3 as the reference's Keras builds them
// (/root/reference/models.py:26 InceptionV3(...), :51 ResNet50(...)); layer
// semantics restated from the published keras.applications definitions
// (SURVEY.md Appendix A/B).  Weight arrays are listed in layer creation order,
// each layer in get_weights() order.
#include "graph.h"

#include <algorithm>
#include <cstdio>

namespace dml {
namespace {

struct Builder {
  Net* net;
  bool keep_all;
  int nconv = 0;  // InceptionV3 auto-naming counter
  bool fuse_shortcut = false;
  bool pool_commute = false;  // InceptionV3 pool branches: 1x1 conv first (inside the merged head GEMM), average pool after

  int add_warr(const std::string& name, std::initializer_list<int64_t> shape) {
    WeightArr a;
    a.name = name;
    a.rank = (int)shape.size();
    int i = 0;
    a.count = 1;
    for (int64_t s : shape) { a.shape[i++] = s; a.count *= s; }
    a.offset = net->total_floats;
    net->total_floats += a.count;
    net->warr.push_back(a);
    return (int)net->warr.size() - 1;
  }

  int new_buf(int h, int w, int c) {
    Buf b;
    b.elems = (int64_t)h * w * c;
    net->bufs.push_back(b);
    return (int)net->bufs.size() - 1;
  }
  int new_tensor(const std::string& name, int h, int w, int c) {
    Tensor t;
    t.name = name; t.h = h; t.w = w; t.c = c;
    t.buf = new_buf(h, w, c);
    t.c_off = 0; t.c_pitch = c;
    net->tensors.push_back(t);
    return (int)net->tensors.size() - 1;
  }
  int view(int parent, const std::string& name, int c_off, int c) {
    Tensor t = net->tensors[parent];
    t.name = name; t.c_off = c_off; t.c = c;
    net->tensors.push_back(t);
    return (int)net->tensors.size() - 1;
  }
  void touch(int tensor, int op_index) {
    int b = net->tensors[tensor].buf;
    if (b < 0) return;
    net->bufs[b].first_use = std::min(net->bufs[b].first_use, op_index);
    net->bufs[b].last_use = std::max(net->bufs[b].last_use, op_index);
  }
  int push(Op op) {
    int idx = (int)net->ops.size();
    touch(op.in, idx);
    touch(op.out, idx);
    if (op.res >= 0) touch(op.res, idx);
    if (op.in2 >= 0) touch(op.in2, idx);
    net->ops.push_back(op);
    return idx;
  }

  // Conv2D (+bias) + BatchNormalization (+ReLU) (+residual add before the ReLU).
  // out < 0: a fresh tensor is created.  Returns the output tensor id.
  int conv(int in, int out, const std::string& cname, const std::string& bname, int cout, int kh, int kw,
           int stride, bool same, bool bias, bool bn_scale, float eps, bool relu, int res) {
    const Tensor ti = net->tensors[in];
    Op op;
    op.kind = OP_CONV; op.name = cname;
    op.kh = kh; op.kw = kw; op.stride = stride;
    op.pad_h = same ? (kh - 1) / 2 : 0;
    op.pad_w = same ? (kw - 1) / 2 : 0;
    int ho = (ti.h + 2 * op.pad_h - kh) / stride + 1;
    int wo = (ti.w + 2 * op.pad_w - kw) / stride + 1;
    op.cin = ti.c; op.cout = cout;
    op.w_kernel = add_warr(cname + "/kernel", {kh, kw, ti.c, cout});
    if (bias) op.w_bias = add_warr(cname + "/bias", {cout});
    if (bn_scale) op.w_gamma = add_warr(bname + "/gamma", {cout});
    op.w_beta = add_warr(bname + "/beta", {cout});
    op.w_mean = add_warr(bname + "/moving_mean", {cout});
    op.w_var = add_warr(bname + "/moving_variance", {cout});
    op.eps = eps; op.relu = relu; op.res = res;
    if (out < 0) out = new_tensor(cname + ":out", ho, wo, cout);
    op.in = in; op.out = out;
    net->conv_macs += (int64_t)ho * wo * cout * kh * kw * ti.c;
    push(op);
    return out;
  }
  int pool(OpKind kind, int in, int out, const std::string& name, int k, int stride, int pad) {
    const Tensor ti = net->tensors[in];
    Op op;
    op.kind = kind; op.name = name;
    op.kh = op.kw = k; op.stride = stride; op.pad_h = op.pad_w = pad;
    int ho = (ti.h + 2 * pad - k) / stride + 1, wo = (ti.w + 2 * pad - k) / stride + 1;
    if (out < 0) out = new_tensor(name + ":out", ho, wo, ti.c);
    op.in = in; op.out = out; op.cin = op.cout = ti.c;
    push(op);
    return out;
  }
  // Two sibling 1x1 reductions of tensor x (f1 and f2 filters) whose outputs are the channel slices [0, f1) and [f1, f1+f2)
  // of one tensor.  red1 / red2 create them at their place in Keras' layer order; with fuse_shortcut the second becomes a
  // follower of the first (Op::merge_with) and the first computes both.
  struct Pair { int t = -1, v1 = -1, v2 = -1, op1 = -1, f1 = 0, f2 = 0; };
  Pair pair_begin(int x, const std::string& name, int f1, int f2) {
    Pair p;
    p.f1 = f1; p.f2 = f2;
    if (!fuse_shortcut) return p;
    const Tensor tx = net->tensors[x];
    p.t = new_tensor(name, tx.h, tx.w, f1 + f2);
    p.v1 = view(p.t, name + ":a", 0, f1);
    p.v2 = view(p.t, name + ":b", f1, f2);
    return p;
  }
  int red1(int x, Pair& p) {
    int out = cb(x, p.v1, p.f1, 1, 1);
    p.op1 = (int)net->ops.size() - 1;
    return out;
  }
  int red2(int x, Pair& p) {
    int out = cb(x, p.v2, p.f2, 1, 1);
    if (p.t >= 0) {
      const int i2 = (int)net->ops.size() - 1;
      Op& lead = net->ops[p.op1];
      lead.followers.push_back(i2); lead.merged_cout = p.f1 + p.f2; lead.merged_out = p.t;
      net->ops[i2].skip = true;
      net->ops[i2].merge_row0 = p.f1;
    }
    return out;
  }
  // Block head with a 1x1 branch that goes straight into the concat: [reduction 1 | reduction 2 | concat] share ONE buffer,
  // the two reductions in R = f1 + f2 scratch channels in FRONT of the concat, so that (red1, red2, branch1x1) are one
  // contiguous channel range = one GEMM.  The block's output tensor is the view behind the scratch channels.
  struct Head { int cat = -1, b1 = -1, v1 = -1, v2 = -1, merged = -1, lead = -1, R = 0, f1 = 0, f2 = 0, fb1 = 0, P = 0, vp = -1; };
  // fp: filters of the pool branch's 1x1 conv (0 = none / not commuted): its pre-pool output takes the first fp scratch channels
  Head head_begin(int x, const std::string& n, int fb1, int f1, int f2, int ccat, int fp = 0) {
    Head h;
    h.f1 = f1; h.f2 = f2; h.fb1 = fb1;
    const Tensor tx = net->tensors[x];
    if (!fuse_shortcut) {
      h.cat = new_tensor(n, tx.h, tx.w, ccat);
      h.b1 = view(h.cat, n + ":b1", 0, fb1);
      return h;
    }
    h.P = pool_commute ? fp : 0;
    h.R = h.P + f1 + f2;
    const int buf = new_tensor(n + ":buf", tx.h, tx.w, h.R + ccat);
    h.cat = view(buf, n, h.R, ccat);
    h.b1 = view(buf, n + ":b1", h.R, fb1);
    if (h.P) h.vp = view(buf, n + ":prepool", 0, h.P);
    h.v1 = view(buf, n + ":red:a", h.P, f1);
    h.v2 = view(buf, n + ":red:b", h.P + f1, f2);
    h.merged = view(buf, n + ":head", 0, h.R + fb1);
    return h;
  }
  // The pool branch of a block: AveragePooling2D(3, 1, same) -> conv2d_bn(f, 1, 1) into the concat slice at c_off.
  void pool_branch(int x, Head& h, const std::string& n, int c_off, int f) {
    if (h.P == 0) {
      int t = pool(OP_AVGPOOL, x, -1, n + ":avgpool", 3, 1, 1);
      cb(t, cat_view(h, n + ":bp", c_off, f), f, 1, 1);
      return;
    }
    // commuted: the 1x1 conv (no bias, no ReLU) is one more member of the head GEMM, the pool follows on its f channels
    cb(x, h.vp, f, 1, 1);
    const int ci = (int)net->ops.size() - 1;
    net->ops[ci].skip = true;
    net->ops[ci].merge_row0 = 0;
    net->ops[ci].post_pool = true;
    net->ops[h.lead].followers.push_back(ci);
    net->ops[h.lead].norelu_cols = h.P;
    const int pi_out = pool(OP_AVGPOOL, h.vp, cat_view(h, n + ":bp", c_off, f), n + ":avgpool", 3, 1, 1);
    (void)pi_out;
    Op& po = net->ops.back();
    po.pool_bias_of = ci;
    po.relu = true;
    touch(h.merged, h.lead);  // (the leader writes the scratch channels)
  }
  int cat_view(const Head& h, const std::string& name, int c_off, int c) { return view(h.cat, name, h.R + c_off, c); }
  void head_b1(int x, Head& h) {  // branch1x1: first in Keras' layer order, so it leads the group
    cb(x, h.b1, h.fb1, 1, 1);
    h.lead = (int)net->ops.size() - 1;
    if (h.merged >= 0) {
      Op& o = net->ops[h.lead];
      o.merged_cout = h.R + h.fb1; o.merged_out = h.merged; o.merge_row0 = h.R;
    }
  }
  int head_red(int x, Head& h, int which) {
    int out = cb(x, which == 0 ? h.v1 : h.v2, which == 0 ? h.f1 : h.f2, 1, 1);
    if (h.merged >= 0) {
      const int i = (int)net->ops.size() - 1;
      net->ops[i].skip = true;
      net->ops[i].merge_row0 = h.P + (which == 0 ? 0 : h.f1);
      net->ops[h.lead].followers.push_back(i);
    }
    return out;
  }
  // InceptionV3 conv2d_bn: Conv2D(no bias) -> BN(scale=False, eps 1e-3) -> ReLU
  int cb(int in, int out, int f, int kh, int kw, bool same = true, int stride = 1) {
    int i = nconv++;
    std::string cn = i == 0 ? "conv2d" : "conv2d_" + std::to_string(i);
    std::string bn = i == 0 ? "batch_normalization" : "batch_normalization_" + std::to_string(i);
    return conv(in, out, cn, bn, f, kh, kw, stride, same, false, false, 1e-3f, true, -1);
  }
  void head(int x) {
    const Tensor t = net->tensors[x];
    Op op;
    op.kind = OP_GAP; op.name = "avg_pool";
    op.in = x; op.cin = op.cout = t.c;
    Tensor p;
    p.name = "avg_pool"; p.h = 1; p.w = 1; p.c = t.c; p.buf = -2; p.c_pitch = t.c;
    net->tensors.push_back(p);
    op.out = net->pooled_tensor = (int)net->tensors.size() - 1;
    push(op);
    net->fc_kernel = add_warr("predictions/kernel", {t.c, 1000});
    net->fc_bias = add_warr("predictions/bias", {1000});
  }
};

void build_resnet50(Builder& b) {
  Net* net = b.net;
  const float eps = 1.001e-5f;
  int x = net->input_tensor;
  // conv1_pad(3) + conv1_conv 7x7 s2 valid == 7x7 s2 with zero padding 3 -> 112x112
  {
    const Tensor ti = net->tensors[x];
    Op op;
    op.kind = OP_CONV; op.name = "conv1_conv";
    op.kh = op.kw = 7; op.stride = 2; op.pad_h = op.pad_w = 3;
    op.cin = 3; op.cout = 64;
    op.w_kernel = b.add_warr("conv1_conv/kernel", {7, 7, 3, 64});
    op.w_bias = b.add_warr("conv1_conv/bias", {64});
    op.w_gamma = b.add_warr("conv1_bn/gamma", {64});
    op.w_beta = b.add_warr("conv1_bn/beta", {64});
    op.w_mean = b.add_warr("conv1_bn/moving_mean", {64});
    op.w_var = b.add_warr("conv1_bn/moving_variance", {64});
    op.eps = eps; op.relu = true;
    int ho = (ti.h + 6 - 7) / 2 + 1;
    op.in = x; op.out = b.new_tensor("conv1_relu", ho, ho, 64);
    net->conv_macs += (int64_t)ho * ho * 64 * 49 * 3;
    b.push(op);
    x = op.out;
  }
  // pool1_pad(1, zeros) + MaxPooling2D(3, 2, valid) -> 56x56
  x = b.pool(OP_MAXPOOL, x, -1, "pool1_pool", 3, 2, 1);
  net->tensors[x].name = "pool1_pool";
  const int filters[4] = {64, 128, 256, 512}, blocks[4] = {3, 4, 6, 3}, stride1[4] = {1, 2, 2, 2};
  for (int s = 0; s < 4; ++s) {
    int f = filters[s];
    for (int blk = 1; blk <= blocks[s]; ++blk) {
      std::string n = "conv" + std::to_string(s + 2) + "_block" + std::to_string(blk);
      int stride = blk == 1 ? stride1[s] : 1;
      int sc = x;
      const int x0 = x;
      Op proj;  // the projection shortcut when it is folded into _3_conv: only its weight arrays are registered here
      const bool fold = blk == 1 && b.fuse_shortcut;
      if (fold) {
        const std::string cn = n + "_0_conv", bn = n + "_0_bn";
        const int cin0 = net->tensors[x].c;
        proj.w_kernel = b.add_warr(cn + "/kernel", {1, 1, cin0, 4 * f});
        proj.w_bias = b.add_warr(cn + "/bias", {4 * f});
        proj.w_gamma = b.add_warr(bn + "/gamma", {4 * f});
        proj.w_beta = b.add_warr(bn + "/beta", {4 * f});
        proj.w_mean = b.add_warr(bn + "/moving_mean", {4 * f});
        proj.w_var = b.add_warr(bn + "/moving_variance", {4 * f});
        sc = -1;
      } else if (blk == 1) {
        sc = b.conv(x, -1, n + "_0_conv", n + "_0_bn", 4 * f, 1, 1, stride, false, true, true, eps, false, -1);
      }
      int y = b.conv(x, -1, n + "_1_conv", n + "_1_bn", f, 1, 1, stride, false, true, true, eps, true, -1);
      y = b.conv(y, -1, n + "_2_conv", n + "_2_bn", f, 3, 3, 1, true, true, true, eps, true, -1);
      x = b.conv(y, -1, n + "_3_conv", n + "_3_bn", 4 * f, 1, 1, 1, false, true, true, eps, true, sc);
      if (fold) {
        const int oi = (int)net->ops.size() - 1;
        Op& o3 = net->ops[oi];
        o3.in2 = x0; o3.cin2 = net->tensors[x0].c; o3.stride2 = stride; o3.eps2 = eps;
        o3.w2_kernel = proj.w_kernel; o3.w2_bias = proj.w_bias; o3.w2_gamma = proj.w_gamma;
        o3.w2_beta = proj.w_beta; o3.w2_mean = proj.w_mean; o3.w2_var = proj.w_var;
        b.touch(x0, oi);  // the block input stays alive until the folded conv has read it
        const Tensor& to = net->tensors[x];
        net->conv_macs += (int64_t)to.h * to.w * to.c * o3.cin2;
      }
      net->tensors[x].name = n + "_out";
    }
  }
  if (b.fuse_shortcut) {
    // conv2: every block's (3x3, expand) and the following block's reduce form a chain the engine may run as one kernel
    for (int oi = 0; oi + 1 < (int)net->ops.size(); ++oi) {
      Op& o3 = net->ops[oi];
      Op& oe = net->ops[oi + 1];
      if (o3.kind != OP_CONV || o3.kh != 3 || o3.kw != 3 || o3.stride != 1 || o3.cin != 64 || o3.cout != 64) continue;
      if (oe.kind != OP_CONV || oe.kh != 1 || oe.in != o3.out || oe.cin != 64 || oe.cout != 256 || !oe.relu) continue;
      if (!((oe.res >= 0) != (oe.in2 >= 0)) || (oe.in2 >= 0 && (oe.stride2 != 1 || oe.cin2 != 64))) continue;
      o3.chain_exp = oi + 1;
      oe.chain_lead = oi;
      int last = oi + 1;
      if (oi + 2 < (int)net->ops.size()) {
        Op& orr = net->ops[oi + 2];
        if (orr.kind == OP_CONV && orr.kh == 1 && orr.kw == 1 && orr.stride == 1 && orr.in == oe.out && orr.cin == 256 && orr.cout == 64 &&
            orr.relu && orr.res < 0 && orr.in2 < 0) {
          o3.chain_red = oi + 2;
          orr.chain_lead = oi;
          last = oi + 2;
        }
      }
      if (oe.in2 >= 0 && o3.chain_red < 0) {  // the fused kernel has no (second source, no next reduce) variant
        o3.chain_exp = -1; oe.chain_lead = -1;
        continue;
      }
      // the fused kernel reads and writes every tensor of the chain during ONE launch: all of them are live from the
      // lead to the last member
      for (int k = oi; k <= last; ++k) {
        const Op& m = net->ops[k];
        for (int t : {m.in, m.out, m.res, m.in2})
          if (t >= 0) { b.touch(t, oi); b.touch(t, last); }
      }
    }
  }
  if (b.fuse_shortcut) {
    // conv3 / conv4: an identity-residual block's 1x1 expand and the following block's 1x1 reduce may run as one kernel
    for (int oi = 0; oi + 1 < (int)net->ops.size(); ++oi) {
      Op& oe = net->ops[oi];
      Op& orr = net->ops[oi + 1];
      if (oe.kind != OP_CONV || oe.kh != 1 || oe.kw != 1 || oe.stride != 1 || !oe.relu || oe.cout != 4 * oe.cin) continue;
      if ((oe.res >= 0) == (oe.in2 >= 0)) continue;
      // with an identity residual: conv3 / conv4 (conv2's identity blocks run as chains); with the folded projection shortcut as a
      // second GEMM source on the same pixel grid: conv2_block1 (where the chain kernel loses against the separate layers)
      if (oe.res >= 0 && (oe.chain_lead >= 0 || (oe.cin != 128 && oe.cin != 256))) continue;
      // (at stride 2 -- conv3_block1, K = 128 + 256 -- the kernel reads the shortcut's input through an im2col-mode tensor map; conv4 / conv5's
      // first blocks have K = 768 / 1536: their 128-pixel tile of A alone is 192 / 384 KB)
      if (oe.in2 >= 0 && (oe.stride2 > 2 || oe.cin2 % 64 || oe.cin % 64 || oe.cin + oe.cin2 > 384)) continue;
      if (orr.kind != OP_CONV || orr.kh != 1 || orr.kw != 1 || orr.stride != 1 || orr.in != oe.out || orr.cin != oe.cout ||
          orr.cout != oe.cin || !orr.relu || orr.res >= 0 || orr.in2 >= 0)
        continue;
      oe.duo_red = oi + 1;
      orr.duo_lead = oi;
      for (int k = oi; k <= oi + 1; ++k) {  // one launch reads and writes all of these
        const Op& m = net->ops[k];
        for (int t : {m.in, m.out, m.res, m.in2})
          if (t >= 0) { b.touch(t, oi); b.touch(t, oi + 1); }
      }
    }
  }
  b.head(x);
}

void build_inception_v3(Builder& b) {
  Net* net = b.net;
  int x = net->input_tensor;
  x = b.cb(x, -1, 32, 3, 3, false, 2);
  net->tensors[x].name = "stem_conv0";
  x = b.cb(x, -1, 32, 3, 3, false);
  x = b.cb(x, -1, 64, 3, 3);
  x = b.pool(OP_MAXPOOL, x, -1, "max_pooling2d", 3, 2, 0);
  x = b.cb(x, -1, 80, 1, 1, false);
  x = b.cb(x, -1, 192, 3, 3, false);
  x = b.pool(OP_MAXPOOL, x, -1, "max_pooling2d_1", 3, 2, 0);
  net->tensors[x].name = "stem_pool2";

  const int pool_f[3] = {32, 64, 64};
  for (int i = 0; i < 3; ++i) {  // mixed0..2, 35x35
    std::string n = "mixed" + std::to_string(i);
    const Tensor tx = net->tensors[x];
    const int r0 = (int)net->ops.size();
    Builder::Head hd = b.head_begin(x, n, 64, 48, 64, 64 + 64 + 96 + pool_f[i], pool_f[i]);
    b.head_b1(x, hd);
    int t = b.head_red(x, hd, 0);
    b.cb(t, b.cat_view(hd, n + ":b5", 64, 64), 64, 5, 5);
    t = b.head_red(x, hd, 1);
    t = b.cb(t, -1, 96, 3, 3);
    b.cb(t, b.cat_view(hd, n + ":b3d", 128, 96), 96, 3, 3);
    b.pool_branch(x, hd, n, 224, pool_f[i]);
    net->regions.push_back({r0, (int)net->ops.size() - 1});
    x = hd.cat;
  }
  {  // mixed3, 17x17
    const Tensor tx = net->tensors[x];
    int ho = (tx.h - 3) / 2 + 1;
    const int r0 = (int)net->ops.size();
    int cat = b.new_tensor("mixed3", ho, ho, 384 + 96 + tx.c);
    b.cb(x, b.view(cat, "mixed3:b3", 0, 384), 384, 3, 3, false, 2);
    int t = b.cb(x, -1, 64, 1, 1);
    t = b.cb(t, -1, 96, 3, 3);
    b.cb(t, b.view(cat, "mixed3:b3d", 384, 96), 96, 3, 3, false, 2);
    b.pool(OP_MAXPOOL, x, b.view(cat, "mixed3:bp", 480, tx.c), "max_pooling2d_2", 3, 2, 0);
    net->regions.push_back({r0, (int)net->ops.size() - 1});
    x = cat;
  }
  const int c7[4] = {128, 160, 160, 192};
  for (int i = 0; i < 4; ++i) {  // mixed4..7, 17x17
    std::string n = "mixed" + std::to_string(i + 4);
    const Tensor tx = net->tensors[x];
    int c = c7[i];
    const int r0 = (int)net->ops.size();
    Builder::Head hd = b.head_begin(x, n, 192, c, c, 768, 192);
    b.head_b1(x, hd);
    int t = b.head_red(x, hd, 0);
    t = b.cb(t, -1, c, 1, 7);
    b.cb(t, b.cat_view(hd, n + ":b7", 192, 192), 192, 7, 1);
    t = b.head_red(x, hd, 1);
    t = b.cb(t, -1, c, 7, 1);
    t = b.cb(t, -1, c, 1, 7);
    t = b.cb(t, -1, c, 7, 1);
    b.cb(t, b.cat_view(hd, n + ":b7d", 384, 192), 192, 1, 7);
    b.pool_branch(x, hd, n, 576, 192);
    net->regions.push_back({r0, (int)net->ops.size() - 1});
    x = hd.cat;
  }
  {  // mixed8, 8x8
    const Tensor tx = net->tensors[x];
    int ho = (tx.h - 3) / 2 + 1;
    const int r0 = (int)net->ops.size();
    int cat = b.new_tensor("mixed8", ho, ho, 320 + 192 + tx.c);
    Builder::Pair pr = b.pair_begin(x, "mixed8:red", 192, 192);
    int t = b.red1(x, pr);
    b.cb(t, b.view(cat, "mixed8:b3", 0, 320), 320, 3, 3, false, 2);
    t = b.red2(x, pr);
    t = b.cb(t, -1, 192, 1, 7);
    t = b.cb(t, -1, 192, 7, 1);
    b.cb(t, b.view(cat, "mixed8:b7x3", 320, 192), 192, 3, 3, false, 2);
    b.pool(OP_MAXPOOL, x, b.view(cat, "mixed8:bp", 512, tx.c), "max_pooling2d_3", 3, 2, 0);
    net->regions.push_back({r0, (int)net->ops.size() - 1});
    x = cat;
  }
  for (int i = 9; i <= 10; ++i) {  // mixed9, mixed10, 8x8
    std::string n = "mixed" + std::to_string(i);
    const Tensor tx = net->tensors[x];
    const int r0 = (int)net->ops.size();
    Builder::Head hd = b.head_begin(x, n, 320, 384, 448, 2048, 192);
    b.head_b1(x, hd);
    int t = b.head_red(x, hd, 0);
    b.cb(t, b.cat_view(hd, n + ":b3a", 320, 384), 384, 1, 3);
    b.cb(t, b.cat_view(hd, n + ":b3b", 704, 384), 384, 3, 1);
    t = b.head_red(x, hd, 1);
    t = b.cb(t, -1, 384, 3, 3);
    b.cb(t, b.cat_view(hd, n + ":bda", 1088, 384), 384, 1, 3);
    b.cb(t, b.cat_view(hd, n + ":bdb", 1472, 384), 384, 3, 1);
    b.pool_branch(x, hd, n, 1856, 192);
    net->regions.push_back({r0, (int)net->ops.size() - 1});
    x = hd.cat;
  }
  b.head(x);
}

int64_t align_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

// Greedy first-fit arena plan: two buffers may overlap in address only if
// their [first_use, last_use] op ranges are disjoint.
void plan_arena(Net* net, bool keep_all) {
  const int nb = (int)net->bufs.size();
  std::vector<int> order(nb);
  for (int i = 0; i < nb; ++i) order[i] = i;
  std::stable_sort(order.begin(), order.end(),
                   [&](int a, int b) { return net->bufs[a].first_use < net->bufs[b].first_use; });
  std::vector<int> placed;
  int64_t top = 0;
  for (int bi : order) {
    Buf& b = net->bufs[bi];
    int64_t sz = align_up(b.elems, 64);
    int64_t off = 0;
    if (keep_all) {
      off = top;
    } else {
      // candidate offsets: 0 and the end of every conflicting placed buffer
      std::vector<std::pair<int64_t, int64_t>> busy;  // [start, end)
      for (int pj : placed) {
        const Buf& p = net->bufs[pj];
        bool live_overlap = !(p.last_use < b.first_use || b.last_use < p.first_use);
        if (live_overlap) busy.push_back({p.offset, p.offset + align_up(p.elems, 64)});
      }
      std::sort(busy.begin(), busy.end());
      for (auto& iv : busy) {
        if (off + sz <= iv.first) break;
        off = std::max(off, iv.second);
      }
    }
    b.offset = off;
    top = std::max(top, off + sz);
    placed.push_back(bi);
  }
  net->arena_elems = top;
}

void plan_weights(Net* net, int wbytes, bool tc_pad) {
  int64_t off = 0;
  for (Op& op : net->ops) {
    if (op.kind != OP_CONV) continue;
    op.cin_pad = (tc_pad && op.cin >= 16) ? (int)align_up(op.cin, 64) : op.cin;
    op.cin2_pad = op.in2 >= 0 ? (int)align_up(op.cin2, 64) : 0;
    if (op.skip) continue;  // a follower's rows were placed right behind its leader's
    const int rows = op.merged_cout > 0 ? op.merged_cout : op.cout;
    const int64_t row_bytes = (int64_t)(op.kh * op.kw * op.cin_pad + op.cin2_pad) * wbytes;
    op.merged_w_off = off;
    off = align_up(off + rows * row_bytes, 256);
    op.merged_b_off = off;
    off = align_up(off + (int64_t)rows * 4, 256);
    op.wpack_off = op.merged_w_off + op.merge_row0 * row_bytes;
    op.bias_off = op.merged_b_off + (int64_t)op.merge_row0 * 4;
    for (int fi : op.followers) {
      Op& f = net->ops[fi];
      f.cin_pad = op.cin_pad;
      f.wpack_off = op.merged_w_off + f.merge_row0 * row_bytes;
      f.bias_off = op.merged_b_off + (int64_t)f.merge_row0 * 4;
    }
  }
  for (Op& op : net->ops)
    if (op.kind == OP_CONV && op.post_pool) {  // the bias the pool op adds after averaging
      op.post_bias_off = off;
      off = align_up(off + (int64_t)op.cout * 4, 256);
    }
  // floor division by 2 of (tap - pad): cells a in [a_min, a_max] of the space-to-depth grid
  auto fdiv2 = [](int v) { return v >= 0 ? v / 2 : -((-v + 1) / 2); };
  if (tc_pad && !net->ops.empty()) {
    const Op& st = net->ops[0];
    if (st.kind == OP_CONV && st.cin == 3 && st.stride == 2 && st.kh == st.kw && st.pad_h == st.pad_w) {
      const int a_min = fdiv2(-st.pad_h), a_max = fdiv2(st.kh - 1 - st.pad_h);
      net->s2d.op = 0;
      net->s2d.size = (net->input_size + 1) / 2;
      net->s2d.k2 = a_max - a_min + 1;
      net->s2d.pad2 = -a_min;
      net->s2d.row_cells = (int)align_up(net->s2d.size + net->s2d.k2 - 1, 8);
      net->s2d.w_off = off;
      off = align_up(off + (int64_t)st.cout * net->s2d.k2 * net->s2d.k2 * 16 * 2, 256);
    }
  }
  const WeightArr& fk = net->warr[net->fc_kernel];
  net->fc_w_off = off;
  off = align_up(off + fk.count * 4, 256);
  net->fc_b_off = off;
  off = align_up(off + 1000 * 4, 256);
  net->weights_bytes = off;
}

bool build(int model, bool keep_all, int wbytes, bool tc_pad, bool fuse_shortcut, bool spec_only, Net* net, std::string* err, bool pool_commute) {
  *net = Net();
  net->model = model;
  Builder b{net, keep_all};
  b.fuse_shortcut = fuse_shortcut && !keep_all && tc_pad;
  b.pool_commute = b.fuse_shortcut && pool_commute;
  Tensor in;
  in.name = "input";
  in.buf = -1;
  in.c = 3; in.c_pitch = 3;
  if (model == 0) {
    net->input_size = 224; net->mode = 0;
  } else if (model == 1) {
    net->input_size = 299; net->mode = 1;
  } else {
    if (err) *err = "unknown model id " + std::to_string(model) + " (0 = ResNet50, 1 = InceptionV3)";
    return false;
  }
  in.h = in.w = net->input_size;
  net->tensors.push_back(in);
  net->input_tensor = 0;
  if (model == 0) build_resnet50(b); else build_inception_v3(b);
  if (spec_only) return true;
  for (size_t r = 0; r < net->regions.size(); ++r) {
    const int f = net->regions[r].first, l = net->regions[r].second;
    for (int i = f; i <= l; ++i) net->ops[i].region = (int)r;
    for (Buf& bf : net->bufs)  // a branch's scratch must not alias another branch's: everything born in the block lives to its end
      if (bf.first_use >= f && bf.first_use <= l && bf.last_use < l) bf.last_use = l;
  }
  plan_arena(net, keep_all);
  plan_weights(net, wbytes, tc_pad);
  return true;
}

}  // namespace

namespace {

// One memory access of an op: a channel range of a buffer (arena address range for the aliasing test).
struct Access {
  int buf;          // >= 0: arena buffer; -1 input tensor; -2 pooled vector
  int c0, c1;       // channels [c0, c1) of the buffer's pixel
  int64_t a0, a1;   // per-image element range of the buffer inside the arena
};

Access access_of(const Net& net, int t, bool whole_blocks) {
  const Tensor& tt = net.tensors[t];
  Access a{tt.buf, tt.c_off, tt.c_off + tt.c, 0, 0};
  if (whole_blocks) {  // tensor-core loaders fetch whole 64-channel blocks: the tail block may reach into the next slice
    a.c1 = tt.c_off + (int)align_up(tt.c, 64);
    if (a.c1 > tt.c_pitch) a.c1 = tt.c_pitch;
  }
  if (tt.buf >= 0) {
    a.a0 = net.bufs[tt.buf].offset;
    a.a1 = a.a0 + align_up(net.bufs[tt.buf].elems, 64);
  }
  return a;
}

bool conflicts(const Access& x, const Access& y) {
  if (x.buf < 0 || y.buf < 0) return x.buf == y.buf;
  if (x.buf == y.buf) return x.c0 < y.c1 && y.c0 < x.c1;
  return x.a0 < y.a1 && y.a0 < x.a1;  // two buffers the arena placed on the same addresses
}

}  // namespace

void plan_streams(Net* net, int max_streams) {
  const int n = (int)net->ops.size();
  for (Op& op : net->ops) { op.stream = 0; op.signal = false; op.solo = true; op.waits.clear(); }
  net->n_streams = 1;
  if (net->regions.empty() || max_streams < 2) return;
  if (max_streams > 8) max_streams = 8;
  std::vector<std::vector<Access>> rd(n), wr(n);
  std::vector<int> exec;  // ops that launch something, in serial order
  for (int i = 0; i < n; ++i) {
    const Op& op = net->ops[i];
    if (op.skip) continue;  // computed by its leader, whose accesses cover it
    exec.push_back(i);
    rd[i].push_back(access_of(*net, op.in, op.kind == OP_CONV));
    if (op.res >= 0) rd[i].push_back(access_of(*net, op.res, false));
    if (op.in2 >= 0) rd[i].push_back(access_of(*net, op.in2, true));
    wr[i].push_back(access_of(*net, op.merged_out >= 0 ? op.merged_out : op.out, false));
  }
  // hazards (read-after-write, write-after-read, write-after-write) and their transitive closure
  std::vector<std::vector<int>> deps(n);
  std::vector<std::vector<char>> anc(n, std::vector<char>(n, 0));
  for (size_t xi = 0; xi < exec.size(); ++xi) {
    const int i = exec[xi];
    for (size_t xj = 0; xj < xi; ++xj) {
      const int j = exec[xj];
      bool hz = false;
      for (const Access& w : wr[j]) {
        for (const Access& r : rd[i]) hz = hz || conflicts(w, r);
        for (const Access& w2 : wr[i]) hz = hz || conflicts(w, w2);
      }
      for (const Access& r : rd[j])
        for (const Access& w2 : wr[i]) hz = hz || conflicts(r, w2);
      if (!hz) continue;
      deps[i].push_back(j);
      anc[i][j] = 1;
      for (int k = 0; k < n; ++k) anc[i][k] = anc[i][k] | anc[j][k];
    }
  }
  std::vector<int> tail(max_streams, -1);
  int used = 1;
  for (int i : exec) {
    Op& op = net->ops[i];
    int k = -1;
    if (tail[0] < 0) k = 0;
    for (int d : deps[i])  // continue the chain of a direct producer (the caller's stream first)
      if (k < 0 && tail[0] == d) k = 0;
    if (k < 0) {
      int best = -1;
      for (int q = 1; q < max_streams; ++q)
        for (int d : deps[i])
          if (tail[q] == d && d > best) { best = d; k = q; }
    }
    if (k < 0)  // a stream that has nothing this op could overlap with: empty, or its last op is an ancestor
      for (int q = 0; q < max_streams && k < 0; ++q)
        if (tail[q] < 0 || anc[i][tail[q]]) k = q;
    if (k < 0) {  // all streams busy with independent work: queue behind the one whose last op is the oldest
      k = 0;
      for (int q = 1; q < max_streams; ++q)
        if (tail[q] < tail[k]) k = q;
    }
    op.stream = k;
    if (k + 1 > used) used = k + 1;
    for (int d : deps[i]) {
      if (net->ops[d].stream == k) continue;  // stream order
      bool covered = tail[k] >= 0 && anc[tail[k]][d];  // already ordered before this stream's previous op
      for (int d2 : deps[i]) covered = covered || (d2 != d && anc[d2][d]);
      if (covered) continue;
      op.waits.push_back(d);
      net->ops[d].signal = true;
    }
    tail[k] = i;
  }
  net->n_streams = used;
  for (int i : exec) {
    Op& op = net->ops[i];
    if (op.region < 0) continue;
    for (int j : exec)
      if (j != i && net->ops[j].region == op.region && !anc[i][j] && !anc[j][i]) op.solo = false;
  }
}

bool build_net(int model, bool keep_all, int wbytes, bool tc_pad, bool fuse_shortcut, Net* net, std::string* err, bool pool_commute) {
  return build(model, keep_all, wbytes, tc_pad, fuse_shortcut, false, net, err, pool_commute);
}
bool build_weight_spec(int model, Net* net, std::string* err) {
  return build(model, true, 4, false, false, true, net, err, false);
}

}  // namespace dml
