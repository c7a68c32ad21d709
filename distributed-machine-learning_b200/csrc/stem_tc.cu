// stem_tc.cu -- K2: the C_in = 3 stem convolutions of predict
// (/root/reference/models.py:40, :65): ResNet50 conv1_conv 7x7 s2 on the
// zero-padded image (+bias, BN, ReLU) and InceptionV3 conv2d 3x3 s2 valid
// (BN no-scale, ReLU).  K = kh*kw*3 is too small and ragged for TMA im2col, so
// each CTA stages the input patch of an 8 x 16 output tile in shared memory,
// expands it there into the K-major 128-byte-swizzled A tile the tensor core
// reads (128 pixels x Kpad), and issues tcgen05.mma (M=128, N=cout, K=16 x Kpad/16)
// with the accumulator in TMEM.  Epilogue: +bias, ReLU, bf16, one full pixel row
// (cout*2 bytes, contiguous in NHWC) per thread.
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdlib>

#include "kernels.h"

namespace dml {
namespace {

constexpr int TILE_H = 8, TILE_W = 16;  // 128 output pixels
constexpr int STEM_THREADS = 256;

struct StemParams {
  const __nv_bfloat16* in;   // [n][hi][wi][3]
  const __nv_bfloat16* w;    // [cout][kh][kw*3], contiguous
  const float* bias;
  __nv_bfloat16* out;        // [n][ho][wo][out_pitch]
  int n, hi, wi, ho, wo, cout, out_pitch;
  int kh, kw, pad;
  int seg, segpad;           // kw*3 real elements per filter row, padded to a multiple of 8 in the GEMM K axis
  int kpad, kblocks;         // K = kh*segpad rounded up to 16; 64-element blocks
  int tiles_x, tiles_y, total_tiles;
  int patch_h, patch_w, ppitch;  // input patch of one tile, row pitch in elements (even)
  int tmem_cols;
  int debug;  // bit0: no prefetch, bit1: zero whole patch buffer every tile
};

__device__ __forceinline__ uint32_t s_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t sw128_desc(uint32_t addr) {
  return (uint64_t)((addr >> 4) & 0x3FFF) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) |
         ((uint64_t)2 << 61);
}

// byte offset of the 16-byte chunk holding elements [k, k+8) (k % 8 == 0) of `row`
// in a [kblocks][rows][64] bf16 tile with 128-byte swizzle
__device__ __forceinline__ uint32_t sw_chunk(int row, int k, int rows) {
  const int kb = k >> 6, j = (k & 63) >> 3;
  return (uint32_t)(kb * rows * 128 + row * 128 + ((j ^ (row & 7)) << 4));
}

// Persistent: each CTA stages the weights once, then walks tiles
// tile = blockIdx.x + i*gridDim.x: patch -> smem, expand to the A tile, MMA, epilogue.
__global__ void __launch_bounds__(STEM_THREADS) stem_tc_kernel(const StemParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (s_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* sbase = smem_raw + (base - s_u32(smem_raw));
  const uint32_t a_bytes = (uint32_t)p.kblocks * 128u * 128u;
  const uint32_t b_bytes = (uint32_t)p.kblocks * (uint32_t)p.cout * 128u;
  uint8_t* sA = sbase;
  uint8_t* sB = sA + a_bytes;
  __nv_bfloat16* patch = reinterpret_cast<__nv_bfloat16*>(sB + b_bytes);
  const int patch_elems = p.patch_h * p.ppitch;
  uint8_t* tail = reinterpret_cast<uint8_t*>(patch) + (((patch_elems + 64) * 2 + 15) & ~15);  // +64: over-read slack
  const uint32_t bar = s_u32(tail);
  const uint32_t tmem_slot = bar + 8u;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(p.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // weights -> swizzled B tile, once per CTA: B[co][r*segpad + e] = w[co][r][e] (0 for e >= seg or k >= kh*segpad)
  for (int i = tid; i < p.cout * p.kblocks * 64; i += STEM_THREADS) {
    const int co = i / (p.kblocks * 64), k = i - co * (p.kblocks * 64);
    const int r = k / p.segpad, e = k - r * p.segpad;
    __nv_bfloat16 v = (r < p.kh && e < p.seg) ? p.w[((size_t)co * p.kh + r) * p.seg + e] : __float2bfloat16(0.f);
    *reinterpret_cast<__nv_bfloat16*>(sB + sw_chunk(co, k & ~7, p.cout) + ((k & 7) << 1)) = v;
  }
  // zero the A columns beyond kh*segpad (never written by the expansion), and the patch slack
  for (int i = tid; i < 128 * ((p.kblocks * 64 - p.kh * p.segpad) >> 3); i += STEM_THREADS) {
    const int m = i & 127, k = p.kh * p.segpad + ((i >> 7) << 3);
    *reinterpret_cast<uint4*>(sA + sw_chunk(m, k, 128)) = make_uint4(0, 0, 0, 0);
  }
  // the pad columns [patch_w*3, ppitch) and the slack are over-read by the expansion (they meet zero
  // weights, but a stale NaN bit pattern times zero is still NaN): clear the whole buffer once
  for (int i = tid; i < patch_elems + 64; i += STEM_THREADS) patch[i] = __float2bfloat16(0.f);
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

  const int q = warp & 3, half = warp >> 2;
  const int ncol = p.cout / 2;
  const int tiles_per_img = p.tiles_x * p.tiles_y;
  const int prow = p.patch_w * 3;
  uint32_t phase = 0;

  // Patch rows are fetched into registers one tile ahead, so the global-load latency
  // hides behind the previous tile's MMA and epilogue.
  constexpr int MAX_ROWS = 11;  // ceil(patch_h / 2) with patch_h <= 22
  __nv_bfloat16 pre[MAX_ROWS];
  auto fetch = [&](int tile) {
    const int img = tile / tiles_per_img;
    const int tr = tile - img * tiles_per_img;
    const int ph0 = 2 * ((tr / p.tiles_x) * TILE_H) - p.pad, pw0 = 2 * ((tr % p.tiles_x) * TILE_W) - p.pad;
    const __nv_bfloat16* img_in = p.in + (size_t)img * p.hi * p.wi * 3;
    const int c = tid & 127;
    const int wi = pw0 + c / 3;
    const bool col_ok = c < prow && wi >= 0 && wi < p.wi;
#pragma unroll
    for (int j = 0; j < MAX_ROWS; ++j) {
      const int r = (tid >> 7) + 2 * j;
      const int hi = ph0 + r;
      pre[j] = __float2bfloat16(0.f);
      if (col_ok && r < p.patch_h && hi >= 0 && hi < p.hi) pre[j] = img_in[((size_t)hi * p.wi + pw0) * 3 + c];
    }
  };
  if ((int)blockIdx.x < p.total_tiles) fetch(blockIdx.x);

  for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
    const int img = tile / tiles_per_img;
    const int tr = tile - img * tiles_per_img;
    const int ty0 = (tr / p.tiles_x) * TILE_H, tx0 = (tr % p.tiles_x) * TILE_W;

    if (p.debug & 1) fetch(tile);
    // 1. prefetched input patch -> smem (zeros outside the image = the conv's zero padding)
    {
      const int c = tid & 127;
      if (c < prow) {
#pragma unroll
        for (int j = 0; j < MAX_ROWS; ++j) {
          const int r = (tid >> 7) + 2 * j;
          if (r < p.patch_h) patch[r * p.ppitch + c] = pre[j];
        }
      }
    }
    __syncthreads();
    // 2. expand: A[m][r*segpad .. +segpad) = patch[2ty + r][6tx .. +segpad)  (the tail past seg meets zero weights)
    for (int i = tid; i < 128 * p.kh; i += STEM_THREADS) {
      const int m = i & 127, r = i >> 7;
      const int ty = m >> 4, tx = m & 15;
      const uint32_t* src = reinterpret_cast<const uint32_t*>(patch + (2 * ty + r) * p.ppitch + 6 * tx);
      const int k0 = r * p.segpad;
      for (int e = 0; e < p.segpad; e += 8) {
        uint4 v = make_uint4(src[(e >> 1)], src[(e >> 1) + 1], src[(e >> 1) + 2], src[(e >> 1) + 3]);
        *reinterpret_cast<uint4*>(sA + sw_chunk(m, k0 + e, 128)) = v;
      }
    }
    // generic-proxy writes must be visible to the tensor core (async proxy)
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

    // 3. MMA: D[128 x cout] = A[128 x kpad] * B[cout x kpad]^T
    if (tid == 0) {
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.cout >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      const uint32_t a0 = base, b0 = base + a_bytes;
      for (int i = 0; i < p.kpad / 16; ++i) {
        const int kb = i >> 2, k16 = i & 3;
        const uint64_t da = sw128_desc(a0 + (uint32_t)kb * 128u * 128u) + (uint64_t)(2 * k16);
        const uint64_t db = sw128_desc(b0 + (uint32_t)kb * (uint32_t)p.cout * 128u) + (uint64_t)(2 * k16);
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_base),
            "l"(da), "l"(db), "r"(idesc), "r"((uint32_t)(i != 0))
            : "memory");
      }
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
    }
    if (!(p.debug & 1) && tile + (int)gridDim.x < p.total_tiles) fetch(tile + gridDim.x);
    // 4. epilogue: warps 0-3 take columns [0, cout/2), warps 4-7 the other half; TMEM lane quarter = warp % 4
    {
      uint32_t done = 0;
      long long t0 = clock64();
      while (!done) {
        asm volatile(
            "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(phase)
            : "memory");
        if (!done && clock64() - t0 > 4000000000LL) __trap();
      }
    }
    phase ^= 1u;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const int m = q * 32 + lane;
    const int oh = ty0 + (m >> 4), ow = tx0 + (m & 15);
    const bool ok = oh < p.ho && ow < p.wo;
    __nv_bfloat16* orow = p.out + (((size_t)img * p.ho + oh) * p.wo + ow) * p.out_pitch + half * ncol;
    for (int c0 = 0; c0 < ncol; c0 += 16) {
      uint32_t v[16];
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
          : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
            "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
          : "r"(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(half * ncol + c0))
          : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      if (ok) {
        uint32_t pk[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float f0 = fmaxf(__uint_as_float(v[2 * i]) + __ldg(p.bias + half * ncol + c0 + 2 * i), 0.f);
          float f1 = fmaxf(__uint_as_float(v[2 * i + 1]) + __ldg(p.bias + half * ncol + c0 + 2 * i + 1), 0.f);
          __nv_bfloat162 h2 = __floats2bfloat162_rn(f0, f1);
          pk[i] = *reinterpret_cast<uint32_t*>(&h2);
        }
        *reinterpret_cast<uint4*>(orow + c0) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        *reinterpret_cast<uint4*>(orow + c0 + 8) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
      }
    }
    // the next tile rewrites the patch/A tile and re-targets the accumulator
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
  }
}

}  // namespace

bool stem_tc_supported(const ConvParams& p) {
  return p.cin == 3 && p.stride == 2 && p.cin_pad == 3 && p.pad_h == p.pad_w && (p.cout == 64 || p.cout == 32) &&
         p.relu && p.res == nullptr && p.kh * ((p.kw * 3 + 7) / 8 * 8) <= 192 && (30 + p.kw) * 3 <= 128 && p.kh <= 8;
}

int launch_stem_tc(const ConvParams& c, cudaStream_t s) {
  StemParams p{};
  p.in = reinterpret_cast<const __nv_bfloat16*>(c.in);
  p.w = reinterpret_cast<const __nv_bfloat16*>(c.w);
  p.bias = c.bias;
  p.out = reinterpret_cast<__nv_bfloat16*>(c.out);
  p.n = c.n; p.hi = c.hi; p.wi = c.wi; p.ho = c.ho; p.wo = c.wo; p.cout = c.cout; p.out_pitch = c.out_pitch;
  p.kh = c.kh; p.kw = c.kw; p.pad = c.pad_h;
  p.seg = c.kw * 3;
  p.segpad = (p.seg + 7) / 8 * 8;
  p.kpad = (c.kh * p.segpad + 15) / 16 * 16;
  p.kblocks = (p.kpad + 63) / 64;
  p.tiles_x = (c.wo + TILE_W - 1) / TILE_W;
  p.tiles_y = (c.ho + TILE_H - 1) / TILE_H;
  p.total_tiles = c.n * p.tiles_x * p.tiles_y;
  p.patch_h = 2 * TILE_H + c.kh - 2;
  p.patch_w = 2 * TILE_W + c.kw - 2;
  p.ppitch = (p.patch_w * 3 + 1) / 2 * 2;
  p.tmem_cols = c.cout <= 32 ? 32 : 64;
  size_t smem = 1024 + (size_t)p.kblocks * 128 * 128 + (size_t)p.kblocks * c.cout * 128 +
                ((((size_t)p.patch_h * p.ppitch + 64) * 2 + 15) & ~(size_t)15) + 64;
  static bool attr = false;
  static int num_sms = 148;
  if (!attr) {
    if (cudaFuncSetAttribute(stem_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024) != cudaSuccess) return -1;
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
    attr = true;
  }
  int per_sm = (int)((220 * 1024) / (smem + 1024));
  if (per_sm > 4) per_sm = 4;
  if (per_sm < 1) per_sm = 1;
  if (const char* e = getenv("DML_STEM_DEBUG")) p.debug = atoi(e);
  if (const char* e = getenv("DML_STEM_PER_SM")) per_sm = atoi(e);
  int grid = num_sms * per_sm;
  if (grid > p.total_tiles) grid = p.total_tiles;
  stem_tc_kernel<<<grid, STEM_THREADS, smem, s>>>(p);
  return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

}  // namespace dml
