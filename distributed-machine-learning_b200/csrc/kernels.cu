// kernels.cu -- the HBM-bound kernels of the predict path and the CUDA-core
// convolution used by the fp32 check mode.  sm_100a.
//
//  K1 preprocess : load_img NEAREST resize + img_to_array + preprocess_input
//                  (/root/reference/models.py:34-38, :59-63)
//  K4 maxpool / K5 avgpool / K6 global average pool / K7 dense+softmax+top-5:
//                  pieces of modelObj.predict and decode_predictions
//                  (/root/reference/models.py:40-42, :65-67)
#include "kernels.h"

#include <cstdlib>

#include <cuda_bf16.h>

namespace dml {
namespace {

__device__ __forceinline__ float to_f(float v) { return v; }
__device__ __forceinline__ float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// ---------------------------------------------------------------------------
// K1: each thread produces 8 consecutive output elements (16 B of bf16).
// ---------------------------------------------------------------------------
template <typename TO>
__global__ void __launch_bounds__(256) preprocess_kernel(PreParams p, const float* __restrict__ lut_g, int caffe) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float lut[768];
  for (int i = threadIdx.x; i < 768; i += blockDim.x) lut[i] = lut_g[i];
  __syncthreads();
  const int S = p.S;
  const int64_t per_img = (int64_t)S * S * 3;
  // Element range of this launch inside the whole tensor; threads own 8-element
  // groups aligned in the GLOBAL element index so that full groups are 16-byte stores
  // even when an image starts at an odd element (299*299*3 is odd).
  const int64_t lo = per_img * p.img_begin, hi = per_img * p.img_end;
  const int64_t g0 = (lo / 8) * 8 + ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (g0 >= hi) return;
  const int64_t e0 = g0 < lo ? lo : g0;
  const int64_t e1 = g0 + 8 < hi ? g0 + 8 : hi;
  int img = (int)(e0 / per_img);
  int rem = (int)(e0 - (int64_t)img * per_img);
  int pix = rem / 3;
  int ch = rem - pix * 3;
  int y = pix / S;
  int x = pix - y * S;

  const uint8_t* src = p.pixels + p.img_off[img];
  int W = p.hw[2 * img + 1];
  int sy = p.ytab[img * S + y];
  int sx = p.xtab[img * S + x];
  TO vals[8];
  const int first = (int)(e0 - g0), cnt = (int)(e1 - e0);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    vals[i] = from_f<TO>(0.f);
    if (i >= first && i < first + cnt) {
      int sc = caffe ? 2 - ch : ch;
      uint8_t v = src[((int64_t)sy * W + sx) * 3 + sc];
      vals[i] = from_f<TO>(lut[ch * 256 + v]);
      if (++ch == 3) {
        ch = 0;
        if (++x == S) {
          x = 0;
          if (++y == S) {
            y = 0;
            ++img;
            if (img < p.img_end) {
              src = p.pixels + p.img_off[img];
              W = p.hw[2 * img + 1];
            }
          }
          if (img < p.img_end) sy = p.ytab[img * S + y];
        }
        if (img < p.img_end) sx = p.xtab[img * S + x];
      }
    }
  }
  TO* out = reinterpret_cast<TO*>(p.out) + g0;
  if (cnt == 8) {
    if (sizeof(TO) == 2) {
      *reinterpret_cast<uint4*>(out) = *reinterpret_cast<uint4*>(vals);
    } else {
      reinterpret_cast<uint4*>(out)[0] = reinterpret_cast<uint4*>(vals)[0];
      reinterpret_cast<uint4*>(out)[1] = reinterpret_cast<uint4*>(vals)[1];
    }
  } else {
    for (int i = first; i < first + cnt; ++i) out[i] = vals[i];
  }
}

// K1, space-to-depth flavour: one thread per output cell (y2, x2) = the 2x2 pixel block at (2*y2, 2*x2); 12 values
// (sub-pixel-major, then channel) + 4 zero channels = 32 bytes.  Same gather + LUT as above, so the values are the
// bf16 roundings of exactly the same fp32 numbers; pixels past the image edge (odd S) are zeros.
__global__ void __launch_bounds__(256) preprocess_s2d_kernel(PreParams p, const float* __restrict__ lut_g, int caffe) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float lut[768];
  for (int i = threadIdx.x; i < 768; i += blockDim.x) lut[i] = lut_g[i];
  __syncthreads();
  // grid: x walks the cells of one image, y the images of this launch -- no 64-bit division per thread (those divisions were most
  // of this kernel's ~400 instructions per thread; ncu round 2: 75 % issue activity at 20 % of DRAM throughput)
  const int S = p.S, S2 = (S + 1) >> 1;
  const int rem = (int)(blockIdx.x * blockDim.x + threadIdx.x);
  if (rem >= S2 * S2) return;
  const int img = p.img_begin + (int)blockIdx.y;
  // rem / S2 for rem < 2^16: (rem + 0.5) / S2 is at least 0.5 / S2 away from every integer, far more than the float rounding error
  const int y2 = __float2int_rz(((float)rem + 0.5f) * (1.0f / (float)S2)), x2 = rem - y2 * S2;
  const uint8_t* src = p.pixels + p.img_off[img];
  const int W = p.hw[2 * img + 1];
  __nv_bfloat16 vals[16];
#pragma unroll
  for (int i = 12; i < 16; ++i) vals[i] = __float2bfloat16_rn(0.f);
#pragma unroll
  for (int py = 0; py < 2; ++py) {
    const int y = 2 * y2 + py;
    const int sy = y < S ? p.ytab[img * S + y] : 0;
#pragma unroll
    for (int px = 0; px < 2; ++px) {
      const int x = 2 * x2 + px;
      const bool in = y < S && x < S;
      const int sx = x < S ? p.xtab[img * S + x] : 0;
      const uint8_t* px_src = src + ((int64_t)sy * W + sx) * 3;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const int sc = caffe ? 2 - c : c;
        vals[(py * 2 + px) * 3 + c] = __float2bfloat16_rn(in ? lut[c * 256 + px_src[sc]] : 0.f);
      }
    }
  }
  // physical position: rows of s2d_row_cells cells, data cell x2 at column s2d_first_cell + x2; the 16-byte halves of a cell are
  // swapped where bit 2 of the physical column is set (= the 32-byte swizzle of the shared-memory row the cell lands in)
  const int xp = x2 + p.s2d_first_cell;
  uint4* out = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.out) + (((int64_t)img * S2 + y2) * p.s2d_row_cells + xp) * 16);
  const int sw = (xp >> 2) & 1;
  out[sw] = reinterpret_cast<uint4*>(vals)[0];
  out[sw ^ 1] = reinterpret_cast<uint4*>(vals)[1];
}

// ---------------------------------------------------------------------------
// CUDA-core direct convolution: 64 pixels x 64 output channels per block,
// 4x4 outputs per thread, K = taps*Cin walked in chunks of 16.
// ---------------------------------------------------------------------------
template <typename TA, typename TW>
__global__ void __launch_bounds__(256) conv_naive_kernel(ConvParams p) {
  __shared__ float As[16][65];
  __shared__ float Bs[16][65];
  const TA* __restrict__ in = reinterpret_cast<const TA*>(p.in);
  const TW* __restrict__ w = reinterpret_cast<const TW*>(p.w);
  const int tid = threadIdx.x;
  const int64_t M = (int64_t)p.n * p.ho * p.wo;
  const int K = p.kh * p.kw * p.cin;
  const int64_t m0 = (int64_t)blockIdx.x * 64;
  const int n0 = blockIdx.y * 64;
  const int tx = tid & 15, ty = tid >> 4;
  const int lp = tid >> 2, lk = (tid & 3) * 4;

  // pixel this thread loads for the A tile
  const int64_t lm = m0 + lp;
  const bool lm_ok = lm < M;
  int ln = 0, lho = 0, lwo = 0;
  if (lm_ok) {
    ln = (int)(lm / (p.ho * p.wo));
    int r = (int)(lm - (int64_t)ln * p.ho * p.wo);
    lho = r / p.wo;
    lwo = r - lho * p.wo;
  }
  const int hbase = lho * p.stride - p.pad_h, wbase = lwo * p.stride - p.pad_w;
  const int lco = n0 + lp;
  const int64_t wrow = (int64_t)lco * p.kh * p.kw * p.cin_pad;

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < K; k0 += 16) {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      int k = k0 + lk + e;
      float av = 0.f, bv = 0.f;
      if (k < K) {
        int tap = k / p.cin;
        int c = k - tap * p.cin;
        int r = tap / p.kw;
        int s = tap - r * p.kw;
        int hi = hbase + r, wi = wbase + s;
        if (lm_ok && hi >= 0 && hi < p.hi && wi >= 0 && wi < p.wi)
          av = to_f(in[(((int64_t)ln * p.hi + hi) * p.wi + wi) * p.in_pitch + c]);
        if (lco < p.cout) bv = to_f(w[wrow + (int64_t)tap * p.cin_pad + c]);
      }
      As[lk + e][lp] = av;
      Bs[lk + e][lp] = bv;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

  TA* out = reinterpret_cast<TA*>(p.out);
  const TA* res = reinterpret_cast<const TA*>(p.res);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int64_t m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int co = n0 + tx * 4 + j;
      if (co >= p.cout) continue;
      float v = acc[i][j] + p.bias[co];
      if (res) v += to_f(res[m * p.res_pitch + co]);
      if (p.relu) v = fmaxf(v, 0.f);
      out[m * p.out_pitch + co] = from_f<TA>(v);
    }
  }
}

// ---------------------------------------------------------------------------
// K4 / K5 pools: one thread per (output pixel, 8-channel group) for bf16,
// (output pixel, 4-channel group) for fp32 -- 16-byte accesses either way.
// ---------------------------------------------------------------------------
template <typename T> struct Vec16;
template <> struct Vec16<float> { static constexpr int N = 4; };
template <> struct Vec16<__nv_bfloat16> { static constexpr int N = 8; };

template <typename T, bool AVG>
__global__ void __launch_bounds__(256) pool_kernel(PoolParams p) {
  constexpr int V = Vec16<T>::N;
  const int cg = p.c / V;
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t total = (int64_t)p.n * p.ho * p.wo * cg;
  if (idx >= total) return;
  int g = (int)(idx % cg);
  int64_t pix = idx / cg;
  int wo = (int)(pix % p.wo);
  int64_t t = pix / p.wo;
  int ho = (int)(t % p.ho);
  int n = (int)(t / p.ho);
  const T* in = reinterpret_cast<const T*>(p.in);
  float acc[V];
  // max-pool: Keras pads with ZEROS (ZeroPadding2D) where pad > 0, so an
  // out-of-bounds tap contributes 0, not -inf.
#pragma unroll
  for (int i = 0; i < V; ++i) acc[i] = AVG ? 0.f : -INFINITY;
  int valid = 0;
  for (int r = 0; r < p.k; ++r) {
    int hi = ho * p.stride - p.pad + r;
    for (int s = 0; s < p.k; ++s) {
      int wi = wo * p.stride - p.pad + s;
      bool ok = hi >= 0 && hi < p.hi && wi >= 0 && wi < p.wi;
      if (ok) {
        ++valid;
        uint4 raw = *reinterpret_cast<const uint4*>(in + (((int64_t)n * p.hi + hi) * p.wi + wi) * p.in_pitch + g * V);
        const T* v = reinterpret_cast<const T*>(&raw);
#pragma unroll
        for (int i = 0; i < V; ++i) {
          float f = to_f(v[i]);
          acc[i] = AVG ? acc[i] + f : fmaxf(acc[i], f);
        }
      } else if (!AVG) {
#pragma unroll
        for (int i = 0; i < V; ++i) acc[i] = fmaxf(acc[i], 0.f);
      }
    }
  }
  T o[V];
#pragma unroll
  for (int i = 0; i < V; ++i) {
    float v = AVG ? acc[i] / (float)valid : acc[i];
    if (AVG && p.bias) {  // pool branch with the 1x1 conv in front (PoolParams::bias)
      v += __ldg(p.bias + g * V + i);
      if (p.relu) v = fmaxf(v, 0.f);
    }
    o[i] = from_f<T>(v);
  }
  T* out = reinterpret_cast<T*>(p.out) + (((int64_t)n * p.ho + ho) * p.wo + wo) * p.out_pitch + g * V;
  *reinterpret_cast<uint4*>(out) = *reinterpret_cast<uint4*>(o);
}

// K4 fast path (bf16, 3x3): the nine 16-byte loads are issued back to back and reduced with packed
// bf16x2 max (exact: max of bf16 values is a bf16 value); an out-of-bounds tap contributes 0.
__global__ void __launch_bounds__(256) maxpool3_bf16_kernel(PoolParams p) {
  pdl_launch_dependents();
  pdl_wait();
  // 32-bit index arithmetic (the launcher checks the thread count fits): 64-bit divisions were most of this kernel's
  // instructions (ncu: 67 % issue activity at 62 % of DRAM throughput)
  const uint32_t cg = (uint32_t)p.c >> 3;
  const uint32_t idx = (p.rev ? gridDim.x - 1u - blockIdx.x : blockIdx.x) * blockDim.x + threadIdx.x;  // PoolParams::rev: last pixels first
  if (idx >= (uint32_t)p.n * (uint32_t)p.ho * (uint32_t)p.wo * cg) return;
  const uint32_t pix = idx / cg;
  const int g = (int)(idx - pix * cg);
  const uint32_t t = pix / (uint32_t)p.wo;
  const int wo = (int)(pix - t * (uint32_t)p.wo);
  const int n = (int)(t / (uint32_t)p.ho);
  const int ho = (int)(t - (uint32_t)n * (uint32_t)p.ho);
  const __nv_bfloat16* in = reinterpret_cast<const __nv_bfloat16*>(p.in) + (int64_t)n * p.hi * p.wi * p.in_pitch + g * 8;
  const int h0 = ho * p.stride - p.pad, w0 = wo * p.stride - p.pad;
  uint4 raw[9];
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int s = 0; s < 3; ++s) {
      const int hi = h0 + r, wi = w0 + s;
      raw[r * 3 + s] = make_uint4(0, 0, 0, 0);  // bf16 zeros = what ZeroPadding2D contributes
      if (hi >= 0 && hi < p.hi && wi >= 0 && wi < p.wi)
        raw[r * 3 + s] = *reinterpret_cast<const uint4*>(in + (hi * p.wi + wi) * p.in_pitch);  // one image < 2^31 elements
    }
  // pad == 0: every tap is in bounds, so starting from tap 0 is exact; pad > 0: zeros take part like any tap
  __nv_bfloat162 m[4];
  {
    const __nv_bfloat162* v = reinterpret_cast<const __nv_bfloat162*>(&raw[0]);
#pragma unroll
    for (int i = 0; i < 4; ++i) m[i] = v[i];
  }
#pragma unroll
  for (int k = 1; k < 9; ++k) {
    const __nv_bfloat162* v = reinterpret_cast<const __nv_bfloat162*>(&raw[k]);
#pragma unroll
    for (int i = 0; i < 4; ++i) m[i] = __hmax2(m[i], v[i]);
  }
  __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(p.out) + (((int64_t)n * p.ho + ho) * p.wo + wo) * p.out_pitch + g * 8;
  *reinterpret_cast<uint4*>(out) = *reinterpret_cast<const uint4*>(m);
}

// K5 fast path (3x3, stride 1, pad 1): one thread per (image, output column, channel
// group) walks down the rows keeping the last three horizontal 3-tap sums, so each
// output costs 3 loads instead of 9; divisor = valid rows x valid columns.
constexpr int AVG_ROWS = 6;  // output rows per thread: more threads in flight than one thread per column
template <typename T>
__global__ void __launch_bounds__(256) avgpool3_kernel(PoolParams p) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int V = Vec16<T>::N;
  const int cg = p.c / V;
  const int segs = (p.hi + AVG_ROWS - 1) / AVG_ROWS;
  // 32-bit index arithmetic (the launcher checks the thread count fits)
  const uint32_t idx = (p.rev ? gridDim.x - 1u - blockIdx.x : blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= (uint32_t)p.n * (uint32_t)segs * (uint32_t)p.wi * (uint32_t)cg) return;
  const uint32_t q1 = idx / (uint32_t)cg;
  const int g = (int)(idx - q1 * (uint32_t)cg);
  const uint32_t q2 = q1 / (uint32_t)p.wi;
  const int wo = (int)(q1 - q2 * (uint32_t)p.wi);
  const int n = (int)(q2 / (uint32_t)segs);
  const int seg = (int)(q2 - (uint32_t)n * (uint32_t)segs);
  const T* in = reinterpret_cast<const T*>(p.in) + (int64_t)n * p.hi * p.wi * p.in_pitch + g * V;
  T* out = reinterpret_cast<T*>(p.out) + (int64_t)n * p.hi * p.wi * p.out_pitch + g * V;
  const int w0 = wo > 0 ? wo - 1 : 0, w1 = wo < p.wi - 1 ? wo + 1 : p.wi - 1;
  const float ncols = (float)(w1 - w0 + 1);
  const int r_begin = seg * AVG_ROWS, r_end = r_begin + AVG_ROWS < p.hi ? r_begin + AVG_ROWS : p.hi;
  float h0[V], h1[V], h2[V];  // horizontal sums of rows r-1, r, r+1
  float bv[V];
#pragma unroll
  for (int i = 0; i < V; ++i) {
    h0[i] = h1[i] = h2[i] = 0.f;
    bv[i] = p.bias ? __ldg(p.bias + g * V + i) : 0.f;
  }
  auto hsum = [&](int r, float (&h)[V]) {
    uint4 raw[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const int wi = w0 + j;
      raw[j] = make_uint4(0, 0, 0, 0);
      if (wi <= w1) raw[j] = *reinterpret_cast<const uint4*>(in + (r * p.wi + wi) * p.in_pitch);  // one image < 2^31 elements
    }
    const T* va = reinterpret_cast<const T*>(&raw[0]);
    const T* vb = reinterpret_cast<const T*>(&raw[1]);
    const T* vc = reinterpret_cast<const T*>(&raw[2]);
#pragma unroll
    for (int i = 0; i < V; ++i) h[i] = (to_f(va[i]) + to_f(vb[i])) + to_f(vc[i]);  // left to right: w0, w0+1, w0+2 (an absent column adds an exact 0)
  };
  if (r_begin > 0) hsum(r_begin - 1, h0);
  hsum(r_begin, h1);
#pragma unroll
  for (int k = 0; k < AVG_ROWS; ++k) {
    const int r = r_begin + k;
    if (r >= r_end) break;
    const bool has_next = r + 1 < p.hi;
    if (has_next) {
      hsum(r + 1, h2);
    } else {
#pragma unroll
      for (int i = 0; i < V; ++i) h2[i] = 0.f;
    }
    const float nrows = (float)((r > 0) + 1 + has_next);
    const float inv = 1.f / (nrows * ncols);
    T o[V];
#pragma unroll
    for (int i = 0; i < V; ++i) {
      // h0 is all zeros above the first row and h2 below the last one: adding them is exact
      float sum = h1[i] + h0[i] + h2[i];
      float v = sum * inv;
      if (p.bias) {  // pool branch with the 1x1 conv in front: folded bias and ReLU come after the average
        v += bv[i];
        if (p.relu) v = fmaxf(v, 0.f);
      }
      o[i] = from_f<T>(v);
    }
    *reinterpret_cast<uint4*>(out + (r * p.wi + wo) * p.out_pitch) = *reinterpret_cast<uint4*>(o);
#pragma unroll
    for (int i = 0; i < V; ++i) { h0[i] = h1[i]; h1[i] = h2[i]; }
  }
}

// K6: one thread per (image, V-channel group), walks the HxW pixels.
template <typename T>
__global__ void __launch_bounds__(256) gap_kernel(const T* __restrict__ in, float* __restrict__ out, int n, int hw, int c,
                                                  int in_pitch) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int V = Vec16<T>::N;
  const int cg = c / V;
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n * cg) return;
  int g = idx % cg, img = idx / cg;
  float acc[V];
#pragma unroll
  for (int i = 0; i < V; ++i) acc[i] = 0.f;
  const T* base = in + (int64_t)img * hw * in_pitch + g * V;
  // eight independent loads in flight per thread, summed in pixel order (same result as a plain loop: one thread's
  // 49-64 dependent-latency loads were the whole cost of this kernel)
  for (int px0 = 0; px0 < hw; px0 += 8) {
    uint4 raw[8];
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (px0 + j < hw) raw[j] = *reinterpret_cast<const uint4*>(base + (int64_t)(px0 + j) * in_pitch);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (px0 + j >= hw) break;
      const T* v = reinterpret_cast<const T*>(&raw[j]);
#pragma unroll
      for (int i = 0; i < V; ++i) acc[i] += to_f(v[i]);
    }
  }
  float inv = 1.f / (float)hw;
#pragma unroll
  for (int i = 0; i < V; ++i) out[(int64_t)img * c + g * V + i] = acc[i] * inv;
}

// K7a: logits[img][j] = pooled[img] . W[:, j] + b[j].  Block = 8 images x 128
// columns, 8 warps each owning a K slice (W read as float4, 4 columns per lane);
// the 8 partial sums are reduced through shared memory in a fixed order.
constexpr int FC_IMGS = 8;
constexpr int FC_COLS = 128;
constexpr int FC_WARPS = 8;
__global__ void __launch_bounds__(FC_WARPS * 32) fc_kernel(const float* __restrict__ pooled, const float* __restrict__ w,
                                                          const float* __restrict__ b, float* __restrict__ logits, int n,
                                                          int cin) {
  extern __shared__ float xs[];  // FC_IMGS x cin, later reused for the partial sums
  const int img0 = blockIdx.x * FC_IMGS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int col = blockIdx.y * FC_COLS + lane * 4;
  for (int i = threadIdx.x; i < FC_IMGS * cin; i += FC_WARPS * 32) {
    int im = img0 + i / cin;
    xs[i] = im < n ? pooled[(int64_t)im * cin + (i % cin)] : 0.f;
  }
  __syncthreads();
  float acc[FC_IMGS][4];
#pragma unroll
  for (int i = 0; i < FC_IMGS; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  const int kslice = cin / FC_WARPS;
  const int k0 = warp * kslice;
  if (col < 1000) {
#pragma unroll 4
    for (int k = k0; k < k0 + kslice; ++k) {
      const float4 wv = *reinterpret_cast<const float4*>(w + (int64_t)k * 1000 + col);
#pragma unroll
      for (int i = 0; i < FC_IMGS; ++i) {
        const float x = xs[i * cin + k];
        acc[i][0] = fmaf(x, wv.x, acc[i][0]);
        acc[i][1] = fmaf(x, wv.y, acc[i][1]);
        acc[i][2] = fmaf(x, wv.z, acc[i][2]);
        acc[i][3] = fmaf(x, wv.w, acc[i][3]);
      }
    }
  }
  __syncthreads();
  float* part = xs;  // [warp][img][128]
#pragma unroll
  for (int i = 0; i < FC_IMGS; ++i)
    *reinterpret_cast<float4*>(part + ((warp * FC_IMGS + i) * FC_COLS) + lane * 4) =
        make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
  __syncthreads();
  for (int o = threadIdx.x; o < FC_IMGS * FC_COLS; o += FC_WARPS * 32) {
    const int i = o / FC_COLS, c = o % FC_COLS;
    const int cc = blockIdx.y * FC_COLS + c;
    if (cc >= 1000 || img0 + i >= n) continue;
    float sum = b[cc];
#pragma unroll
    for (int wv = 0; wv < FC_WARPS; ++wv) sum += part[(wv * FC_IMGS + i) * FC_COLS + c];
    logits[(int64_t)(img0 + i) * 1000 + cc] = sum;
  }
}

// K7b: softmax over 1000 logits + top-5 (descending prob; ties -> larger index).
__global__ void __launch_bounds__(256) softmax_top5_kernel(const float* __restrict__ logits, float* __restrict__ probs,
                                                           int32_t* __restrict__ top_idx, float* __restrict__ top_prob,
                                                           int n) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float sp[1000];
  __shared__ float red[8];
  __shared__ int redi[8];
  const int img = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const float* lg = logits + (int64_t)img * 1000;
  float mx = -INFINITY;
  for (int j = tid; j < 1000; j += 256) mx = fmaxf(mx, lg[j]);
  for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if (lane == 0) red[wid] = mx;
  __syncthreads();
  mx = red[0];
  for (int i = 1; i < 8; ++i) mx = fmaxf(mx, red[i]);
  __syncthreads();
  float sum = 0.f;
  for (int j = tid; j < 1000; j += 256) {
    float e = expf(lg[j] - mx);
    sp[j] = e;
    sum += e;
  }
  for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if (lane == 0) red[wid] = sum;
  __syncthreads();
  sum = 0.f;
  for (int i = 0; i < 8; ++i) sum += red[i];
  __syncthreads();
  for (int j = tid; j < 1000; j += 256) {
    float pr = sp[j] / sum;
    sp[j] = pr;
    if (probs) probs[(int64_t)img * 1000 + j] = pr;
  }
  __syncthreads();
  for (int t = 0; t < 5; ++t) {
    float bv = -1.f;
    int bi = -1;
    for (int j = tid; j < 1000; j += 256) {
      float v = sp[j];
      if (v > bv || (v == bv && j > bi)) { bv = v; bi = j; }
    }
    for (int o = 16; o; o >>= 1) {
      float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > bv || (ov == bv && oi > bi)) { bv = ov; bi = oi; }
    }
    if (lane == 0) { red[wid] = bv; redi[wid] = bi; }
    __syncthreads();
    if (tid == 0) {
      for (int i = 1; i < 8; ++i)
        if (red[i] > bv || (red[i] == bv && redi[i] > bi)) { bv = red[i]; bi = redi[i]; }
      top_idx[img * 5 + t] = bi;
      top_prob[img * 5 + t] = bv;
      sp[bi] = -2.f;  // remove from the next rounds
    }
    __syncthreads();
  }
}

}  // namespace

void launch_preprocess(const PreParams& p, const float* d_lut, bool fp32_out, cudaStream_t s) {
  if (p.s2d) {
    const int S2 = (p.S + 1) / 2;
    const int imgs = p.img_end - p.img_begin;
    if (imgs <= 0) return;
    for (int i0 = 0; i0 < imgs; i0 += 65535) {  // grid.y limit
      PreParams q = p;
      q.img_begin = p.img_begin + i0;
      q.img_end = q.img_begin + (imgs - i0 < 65535 ? imgs - i0 : 65535);
      launch_pdl(preprocess_s2d_kernel, dim3((unsigned)((S2 * S2 + 255) / 256), (unsigned)(q.img_end - q.img_begin)), dim3(256), 0, s, q, d_lut, p.caffe);
    }
    return;
  }
  int64_t per_img = (int64_t)p.S * p.S * 3;
  int64_t lo = per_img * p.img_begin, hi = per_img * p.img_end;
  int64_t threads = (hi - (lo / 8) * 8 + 7) / 8;
  if (threads <= 0) return;
  unsigned blocks = (unsigned)((threads + 255) / 256);
  if (fp32_out) launch_pdl(preprocess_kernel<float>, dim3(blocks), dim3(256), 0, s, p, d_lut, p.caffe);
  else launch_pdl(preprocess_kernel<__nv_bfloat16>, dim3(blocks), dim3(256), 0, s, p, d_lut, p.caffe);
}

void launch_conv_naive(const ConvParams& p, bool fp32_act, cudaStream_t s) {
  int64_t M = (int64_t)p.n * p.ho * p.wo;
  dim3 grid((unsigned)((M + 63) / 64), (unsigned)((p.cout + 63) / 64));
  if (fp32_act) conv_naive_kernel<float, float><<<grid, 256, 0, s>>>(p);
  else conv_naive_kernel<__nv_bfloat16, __nv_bfloat16><<<grid, 256, 0, s>>>(p);
}

template <bool AVG>
static void launch_pool(const PoolParams& p, bool fp32_act, cudaStream_t s) {
  int V = fp32_act ? 4 : 8;
  int64_t total = (int64_t)p.n * p.ho * p.wo * (p.c / V);
  unsigned blocks = (unsigned)((total + 255) / 256);
  if (fp32_act) pool_kernel<float, AVG><<<blocks, 256, 0, s>>>(p);
  else pool_kernel<__nv_bfloat16, AVG><<<blocks, 256, 0, s>>>(p);
}
void launch_maxpool(const PoolParams& p, bool fp32_act, cudaStream_t s) {
  if (!fp32_act && p.k == 3 && (int64_t)p.n * p.ho * p.wo * (p.c / 8) < (1ll << 31) && (int64_t)p.hi * p.wi * p.in_pitch < (1ll << 31)) {
    int64_t total = (int64_t)p.n * p.ho * p.wo * (p.c / 8);
    launch_pdl(maxpool3_bf16_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, s, p);
    return;
  }
  launch_pool<false>(p, fp32_act, s);
}
void launch_avgpool(const PoolParams& p, bool fp32_act, cudaStream_t s) {
  if (p.k == 3 && p.stride == 1 && p.pad == 1 && p.ho == p.hi && p.wo == p.wi &&
      (int64_t)p.n * ((p.hi + AVG_ROWS - 1) / AVG_ROWS) * p.wi * (p.c / (fp32_act ? 4 : 8)) < (1ll << 31) &&
      (int64_t)p.hi * p.wi * (p.in_pitch > p.out_pitch ? p.in_pitch : p.out_pitch) < (1ll << 31)) {
    int V = fp32_act ? 4 : 8;
    int64_t total = (int64_t)p.n * ((p.hi + AVG_ROWS - 1) / AVG_ROWS) * p.wi * (p.c / V);
    unsigned blocks = (unsigned)((total + 255) / 256);
    if (fp32_act) launch_pdl(avgpool3_kernel<float>, dim3(blocks), dim3(256), 0, s, p);
    else launch_pdl(avgpool3_kernel<__nv_bfloat16>, dim3(blocks), dim3(256), 0, s, p);
    return;
  }
  launch_pool<true>(p, fp32_act, s);
}

void launch_gap(const void* in, float* out, int n, int hw, int c, int in_pitch, bool fp32_act, cudaStream_t s) {
  int V = fp32_act ? 4 : 8;
  int total = n * (c / V);
  unsigned blocks = (unsigned)((total + 255) / 256);
  if (fp32_act) launch_pdl(gap_kernel<float>, dim3(blocks), dim3(256), 0, s, reinterpret_cast<const float*>(in), out, n, hw, c, in_pitch);
  else launch_pdl(gap_kernel<__nv_bfloat16>, dim3(blocks), dim3(256), 0, s, reinterpret_cast<const __nv_bfloat16*>(in), out, n, hw, c, in_pitch);
}

// Dense layer of the bf16 engines: logits = pooled x W + b as a TF32 tensor-core GEMM (mma.sync m16n8k8, fp32
// accumulate).  A 256 x 2048 x 1000 product is too small for a tcgen05 pipeline to pay off and the fp32 CUDA-core
// kernel above re-reads W once per 8 images; here a CTA owns 32 images x 64 classes and streams K in chunks of 32
// through a cp.async double buffer.  The fp32 check mode keeps the exact fp32 kernel.
constexpr int FT_BM = 32, FT_BN = 64, FT_BK = 32, FT_APAD = 36, FT_BPAD = 72, FT_STAGES = 4;  // 4 stages = 55 KB: four CTAs per SM, so the 512 CTAs of a 256-image batch are one wave
constexpr size_t FT_SMEM = (size_t)FT_STAGES * (FT_BM * FT_APAD + FT_BK * FT_BPAD) * sizeof(float);
__device__ __forceinline__ uint32_t to_tf32(float v) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
  return r;
}
__device__ __forceinline__ void cp_async16(float* smem_dst, const float* gsrc, bool pred) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  const int bytes = pred ? 16 : 0;  // 0 source bytes: the 16 destination bytes are zero filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(gsrc), "r"(bytes) : "memory");
}
// Split K: the gridDim.z CTAs of a thread-block cluster (1 x 1 x S) each take cin / S of the K axis; the partial 32 x 64 tiles meet in the
// shared memory of the cluster's first CTA, which adds them in rank order (deterministic), adds the bias and writes the logits.
// 128 CTAs walking 64 K chunks one after the other were bound by the latency of each chunk; 512 CTAs walk 16.
__global__ void __launch_bounds__(128) fc_tf32_kernel(const float* __restrict__ pooled, const float* __restrict__ w,
                                                      const float* __restrict__ b, float* __restrict__ logits, int n, int cin) {
  pdl_launch_dependents();
  pdl_wait();
  const int ksplit = (int)gridDim.z, krank = (int)blockIdx.z;
  const int kspan = cin / ksplit, kbase = krank * kspan;
  extern __shared__ __align__(16) float ft_smem[];
  float (*As)[FT_BM][FT_APAD] = reinterpret_cast<float (*)[FT_BM][FT_APAD]>(ft_smem);
  float (*Bs)[FT_BK][FT_BPAD] = reinterpret_cast<float (*)[FT_BK][FT_BPAD]>(ft_smem + FT_STAGES * FT_BM * FT_APAD);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, t = lane & 3;
  const int m0 = blockIdx.x * FT_BM, n0 = blockIdx.y * FT_BN;
  const int wm = (warp & 1) * 16, wn = (warp >> 1) * 32;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  auto load_stage = [&](int st, int k0) {
#pragma unroll
    for (int i = 0; i < 2; ++i) {  // A: 32 rows x 8 float4
      const int idx = tid + i * 128, r = idx >> 3, c4 = (idx & 7) * 4;
      const bool ok = m0 + r < n;
      cp_async16(&As[st][r][c4], pooled + (int64_t)(ok ? m0 + r : 0) * cin + kbase + k0 + c4, ok);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {  // B: 32 k rows x 16 float4
      const int idx = tid + i * 128, r = idx >> 4, c4 = (idx & 15) * 4;
      const bool ok = n0 + c4 < 1000;
      cp_async16(&Bs[st][r][c4], w + (int64_t)(kbase + k0 + r) * 1000 + (ok ? n0 + c4 : 0), ok);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  const int nk = kspan / FT_BK;
  // FT_STAGES - 1 chunks in flight (L2 latency, not bandwidth, bounds a CTA that waits for every chunk)
  for (int i = 0; i < FT_STAGES - 1; ++i) {
    if (i < nk) load_stage(i, i * FT_BK);
    else asm volatile("cp.async.commit_group;" ::: "memory");
  }
  for (int kc = 0; kc < nk; ++kc) {
    const int st = kc % FT_STAGES;
    asm volatile("cp.async.wait_group %0;" ::"n"(FT_STAGES - 2) : "memory");
    __syncthreads();  // chunk kc has landed for everyone, and everyone is done with the stage refilled below
    if (kc + FT_STAGES - 1 < nk) load_stage((kc + FT_STAGES - 1) % FT_STAGES, (kc + FT_STAGES - 1) * FT_BK);
    else asm volatile("cp.async.commit_group;" ::: "memory");
#pragma unroll
    for (int k8 = 0; k8 < FT_BK; k8 += 8) {
      const uint32_t a0 = to_tf32(As[st][wm + g][k8 + t]), a1 = to_tf32(As[st][wm + g + 8][k8 + t]);
      const uint32_t a2 = to_tf32(As[st][wm + g][k8 + t + 4]), a3 = to_tf32(As[st][wm + g + 8][k8 + t + 4]);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint32_t b0 = to_tf32(Bs[st][k8 + t][wn + j * 8 + g]), b1 = to_tf32(Bs[st][k8 + t + 4][wn + j * 8 + g]);
        asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                     : "+f"(acc[j][0]), "+f"(acc[j][1]), "+f"(acc[j][2]), "+f"(acc[j][3])
                     : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
      }
    }
  }
  if (ksplit > 1) {
    // every CTA parks its partial tile [32][64] in its own shared memory (the pipeline buffers are done with), the cluster meets,
    // rank 0 reads the others' tiles through distributed shared memory and adds them in rank order
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    float* part = ft_smem;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = wn + j * 8 + 2 * t;
      *reinterpret_cast<float2*>(part + (wm + g) * FT_BN + c) = make_float2(acc[j][0], acc[j][1]);
      *reinterpret_cast<float2*>(part + (wm + g + 8) * FT_BN + c) = make_float2(acc[j][2], acc[j][3]);
    }
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (krank == 0) {
      const uint32_t local = (uint32_t)__cvta_generic_to_shared(part);
      for (int r = 1; r < ksplit; ++r) {
        uint32_t remote;
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(r));
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int c = wn + j * 8 + 2 * t;
          float2 v0, v1;
          asm volatile("ld.shared::cluster.v2.f32 {%0, %1}, [%2];" : "=f"(v0.x), "=f"(v0.y) : "r"(remote + (uint32_t)((wm + g) * FT_BN + c) * 4u));
          asm volatile("ld.shared::cluster.v2.f32 {%0, %1}, [%2];" : "=f"(v1.x), "=f"(v1.y) : "r"(remote + (uint32_t)((wm + g + 8) * FT_BN + c) * 4u));
          acc[j][0] += v0.x; acc[j][1] += v0.y; acc[j][2] += v1.x; acc[j][3] += v1.y;
        }
      }
    }
    // nobody leaves while rank 0 may still be reading its shared memory
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (krank != 0) return;
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int col = n0 + wn + j * 8 + 2 * t;
    if (col >= 1000) continue;
    const float b0 = b[col], b1 = b[col + 1];
    const int r0 = m0 + wm + g, r1 = r0 + 8;
    if (r0 < n) *reinterpret_cast<float2*>(logits + (int64_t)r0 * 1000 + col) = make_float2(acc[j][0] + b0, acc[j][1] + b1);
    if (r1 < n) *reinterpret_cast<float2*>(logits + (int64_t)r1 * 1000 + col) = make_float2(acc[j][2] + b0, acc[j][3] + b1);
  }
}

bool pdl_enabled() {
  static const bool on = getenv("DML_NO_PDL") == nullptr;
  return on;
}

void init_kernels() {
  cudaFuncSetAttribute(fc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
  cudaFuncSetAttribute(fc_tf32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FT_SMEM);
}

void launch_fc(const float* pooled, const float* w, const float* b, float* logits, int n, int cin, bool tf32, cudaStream_t s) {
  if (tf32 && cin % FT_BK == 0 && n > 0) {
    // K split over a cluster of 4 (2) CTAs when cin allows whole 32-element chunks per CTA
    const unsigned ks = cin % (4 * FT_BK) == 0 ? 4u : (cin % (2 * FT_BK) == 0 ? 2u : 1u);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)((n + FT_BM - 1) / FT_BM), (1000 + FT_BN - 1) / FT_BN, ks);
    cfg.blockDim = dim3(128);
    cfg.dynamicSmemBytes = FT_SMEM;
    cfg.stream = s;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = ks;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    cudaLaunchKernelEx(&cfg, fc_tf32_kernel, pooled, w, b, logits, n, cin);
    return;
  }
  dim3 grid((unsigned)((n + FC_IMGS - 1) / FC_IMGS), (1000 + FC_COLS - 1) / FC_COLS);
  // the activations buffer is reused for the per-warp partial sums [FC_WARPS][FC_IMGS][FC_COLS]: size it for both
  size_t smem = (size_t)FC_IMGS * cin * sizeof(float);
  const size_t part = (size_t)FC_WARPS * FC_IMGS * FC_COLS * sizeof(float);
  if (smem < part) smem = part;
  fc_kernel<<<grid, FC_WARPS * 32, smem, s>>>(pooled, w, b, logits, n, cin);
}

void launch_softmax_top5(const float* logits, float* probs, int32_t* top_idx, float* top_prob, int n, cudaStream_t s) {
  launch_pdl(softmax_top5_kernel, dim3((unsigned)n), dim3(256), 0, s, logits, probs, top_idx, top_prob, n);
}

}  // namespace dml
