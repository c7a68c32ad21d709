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
#include <cstdio>
#include <cstdlib>

#include "conv_tc.h"
#include "kernels.h"

// Superseded by the space-to-depth stem (the stem as a stride-1 halo convolution, conv_tc6_kernel): kept as an A/B twin in
// test builds (-DDML_LEGACY_VARIANTS); the product library carries stubs only.
#ifndef DML_LEGACY_VARIANTS
namespace dml {
bool stem_tc_supported(const ConvParams&) { return false; }
int launch_stem_tc(const ConvParams&, cudaStream_t) { return -1; }
}  // namespace dml
#else
namespace dml {
namespace {

constexpr int TILE_H = 8, TILE_W = 16;  // 128 output pixels

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

constexpr int STEM_BUILDERS = 256;                  // warps 0..7 build A tiles
constexpr int STEM_THREADS_TOTAL = STEM_BUILDERS + 32 + 128;  // + MMA warp (8) + epilogue warps (9..12)

__device__ __forceinline__ void s_mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void s_mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void s_mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  long long t0 = clock64();
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) break;
    if (clock64() - t0 > 4000000000LL) __trap();  // a protocol bug aborts the kernel instead of hanging the GPU
  }
}
__device__ __forceinline__ void s_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool s_elect() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
  return pred != 0;
}

// Persistent and warp-specialised.  Builder warps (0..7) stage the weights once, then per 8x16 tile
// fetch the input patch (one tile ahead, in registers), store it to shared memory and expand it into
// one of TWO A buffers; the MMA warp (8) issues the K/16 tcgen05.mma of a tile as soon as its A buffer is
// full and an accumulator is free; the epilogue warps (9..12) drain the other accumulator meanwhile.
// So patch expansion, tensor-core work and the output stores of three consecutive tiles overlap.
__global__ void __launch_bounds__(STEM_THREADS_TOTAL, 1) stem_tc_kernel(const __grid_constant__ CUtensorMap map_o,
                                                                       const StemParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (s_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* sbase = smem_raw + (base - s_u32(smem_raw));
  const uint32_t a_bytes = (uint32_t)p.kblocks * 128u * 128u;
  const uint32_t b_bytes = (uint32_t)p.kblocks * (uint32_t)p.cout * 128u;
  uint8_t* sA0 = sbase;                        // A buffers 0 and 1
  uint8_t* sB = sA0 + 2u * a_bytes;
  const int patch_elems = p.patch_h * p.ppitch;
  const uint32_t patch_bytes = (uint32_t)(((patch_elems + 64) * 2 + 15) & ~15);  // +64 elements of over-read slack
  __nv_bfloat16* patch0 = reinterpret_cast<__nv_bfloat16*>(sB + b_bytes);
  // output staging: 2 x [128 pixels][cout] bf16, 1024-byte aligned (swizzled rows for the TMA store)
  const uint32_t out_bytes = 128u * (uint32_t)p.cout * 2u;
  const uint32_t sO = (s_u32(reinterpret_cast<uint8_t*>(patch0) + 2u * patch_bytes) + 1023u) & ~1023u;
  const uint32_t bars = sO + 2u * out_bytes;
  const uint32_t afull0 = bars, aempty0 = bars + 16u, tfull0 = bars + 32u, tempty0 = bars + 48u;
  const uint32_t tmem_slot = bars + 64u;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tiles_per_img = p.tiles_x * p.tiles_y;
  const int acc_cols = p.cout <= 32 ? 32 : 64;

  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      s_mbar_init(afull0 + 8u * i, STEM_BUILDERS / 32);  // one arrive per builder warp
      s_mbar_init(aempty0 + 8u * i, 1);
      s_mbar_init(tfull0 + 8u * i, 1);
      s_mbar_init(tempty0 + 8u * i, 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 8) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(p.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (warp < 8) {
    // weights -> swizzled B tile, once per CTA: B[co][r*segpad + e] = w[co][r][e] (0 for e >= seg or k >= kh*segpad)
    for (int i = tid; i < p.cout * p.kblocks * 64; i += STEM_BUILDERS) {
      const int co = i / (p.kblocks * 64), k = i - co * (p.kblocks * 64);
      const int r = k / p.segpad, e = k - r * p.segpad;
      __nv_bfloat16 v = (r < p.kh && e < p.seg) ? p.w[((size_t)co * p.kh + r) * p.seg + e] : __float2bfloat16(0.f);
      *reinterpret_cast<__nv_bfloat16*>(sB + sw_chunk(co, k & ~7, p.cout) + ((k & 7) << 1)) = v;
    }
    // A columns beyond kh*segpad are never written by the expansion: clear them in both buffers
    for (int i = tid; i < 2 * 128 * ((p.kblocks * 64 - p.kh * p.segpad) >> 3); i += STEM_BUILDERS) {
      const int buf = i & 1, j = i >> 1;
      const int m = j & 127, k = p.kh * p.segpad + ((j >> 7) << 3);
      *reinterpret_cast<uint4*>(sA0 + buf * a_bytes + sw_chunk(m, k, 128)) = make_uint4(0, 0, 0, 0);
    }
    // the pad columns [patch_w*3, ppitch) and the slack are over-read by the expansion (they meet zero
    // weights, but a stale NaN bit pattern times zero is still NaN): clear both patch buffers once
    for (int i = tid; i < (int)(2u * patch_bytes / 2u); i += STEM_BUILDERS) patch0[i] = __float2bfloat16(0.f);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

  if (warp < 8) {
    // ---------------- builders ----------------
    const int prow = p.patch_w * 3;
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
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
      const uint32_t buf = it & 1u, ph = (it >> 1) & 1u;
      __nv_bfloat16* patch = patch0 + buf * (patch_bytes / 2u);
      uint8_t* sA = sA0 + buf * a_bytes;
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
      asm volatile("bar.sync 1, 256;" ::: "memory");  // builders only
      if (tile + (int)gridDim.x < p.total_tiles) fetch(tile + gridDim.x);  // loads fly during the expansion
      s_mbar_wait(aempty0 + 8u * buf, ph ^ 1u);  // the MMAs that read this A buffer two tiles ago have retired
      // 2. expand: A[m][r*segpad .. +segpad) = patch[2ty + r][6tx .. +segpad)  (the tail past seg meets zero weights)
      for (int i = tid; i < 128 * p.kh; i += STEM_BUILDERS) {
        const int m = i & 127, r = i >> 7;
        const int ty = m >> 4, tx = m & 15;
        const uint32_t* src = reinterpret_cast<const uint32_t*>(patch + (2 * ty + r) * p.ppitch + 6 * tx);
        const int k0 = r * p.segpad;
        for (int e = 0; e < p.segpad; e += 8) {
          uint4 v = make_uint4(src[(e >> 1)], src[(e >> 1) + 1], src[(e >> 1) + 2], src[(e >> 1) + 3]);
          *reinterpret_cast<uint4*>(sA + sw_chunk(m, k0 + e, 128)) = v;
        }
      }
      // generic-proxy writes must be visible to the tensor core (async proxy) before the warp reports in
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) s_mbar_arrive(afull0 + 8u * buf);
    }
  } else if (warp == 8) {
    // ---------------- MMA issuer: D[128 x cout] = A[128 x kpad] * B[cout x kpad]^T ----------------
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.cout >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint32_t b0 = base + 2u * a_bytes;
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
      const uint32_t buf = it & 1u, ph = (it >> 1) & 1u;
      s_mbar_wait(tempty0 + 8u * buf, ph ^ 1u);  // the epilogue has drained this accumulator
      s_mbar_wait(afull0 + 8u * buf, ph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t a0 = base + buf * a_bytes;
      const uint32_t d_tmem = tmem_base + buf * (uint32_t)acc_cols;
      if (s_elect()) {
        for (int i = 0; i < p.kpad / 16; ++i) {
          const int kb = i >> 2, k16 = i & 3;
          const uint64_t da = sw128_desc(a0 + (uint32_t)kb * 128u * 128u) + (uint64_t)(2 * k16);
          const uint64_t db = sw128_desc(b0 + (uint32_t)kb * (uint32_t)p.cout * 128u) + (uint64_t)(2 * k16);
          asm volatile(
              "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
              "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
              "l"(da), "l"(db), "r"(idesc), "r"((uint32_t)(i != 0))
              : "memory");
        }
        s_commit(aempty0 + 8u * buf);
        s_commit(tfull0 + 8u * buf);
      }
      __syncwarp();
    }
  } else {
    // ---------------- epilogue: 4 warps = 128 tile rows, TMEM lane quarter = warp % 4 ----------------
    const int q = warp & 3;
    const int m = q * 32 + lane;
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
      const uint32_t buf = it & 1u, ph = (it >> 1) & 1u;
      const int img = tile / tiles_per_img;
      const int tr = tile - img * tiles_per_img;
      const int ty0 = (tr / p.tiles_x) * TILE_H, tx0 = (tr % p.tiles_x) * TILE_W;
      s_mbar_wait(tfull0 + 8u * buf, ph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + buf * (uint32_t)acc_cols;
      // the TMA store that read this staging buffer two tiles ago must be done reading it
      if (warp == 9 && lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
      asm volatile("bar.sync 2, 128;" ::: "memory");
      const uint32_t row_bytes = (uint32_t)p.cout * 2u;  // 128 B (cout 64) or 64 B (cout 32)
      const uint32_t rowaddr = sO + buf * out_bytes + (uint32_t)m * row_bytes;
      const uint32_t swz = p.cout == 64 ? (uint32_t)(m & 7) : (uint32_t)((m >> 1) & 3);
      for (int c0 = 0; c0 < p.cout; c0 += 16) {
        uint32_t v[16];
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
              "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
            : "r"(t_row + (uint32_t)c0)
            : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        uint32_t pk[8];
        const float4* bp = reinterpret_cast<const float4*>(p.bias + c0);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 b4 = __ldg(bp + i);
          const float f0 = fmaxf(__uint_as_float(v[4 * i + 0]) + b4.x, 0.f), f1 = fmaxf(__uint_as_float(v[4 * i + 1]) + b4.y, 0.f);
          const float f2 = fmaxf(__uint_as_float(v[4 * i + 2]) + b4.z, 0.f), f3 = fmaxf(__uint_as_float(v[4 * i + 3]) + b4.w, 0.f);
          __nv_bfloat162 lo = __floats2bfloat162_rn(f0, f1), hi = __floats2bfloat162_rn(f2, f3);
          pk[2 * i] = *reinterpret_cast<uint32_t*>(&lo);
          pk[2 * i + 1] = *reinterpret_cast<uint32_t*>(&hi);
        }
        const uint32_t j0 = (uint32_t)(c0 >> 3);
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(rowaddr + ((j0 ^ swz) << 4)), "r"(pk[0]), "r"(pk[1]),
                     "r"(pk[2]), "r"(pk[3])
                     : "memory");
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(rowaddr + (((j0 + 1u) ^ swz) << 4)), "r"(pk[4]),
                     "r"(pk[5]), "r"(pk[6]), "r"(pk[7])
                     : "memory");
      }
      // one 4-D TMA store writes the 8 x 16 pixel block (rows/columns past the image edge are clipped)
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      asm volatile("bar.sync 2, 128;" ::: "memory");
      if (warp == 9 && lane == 0) {
        asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(&map_o),
                     "r"(sO + buf * out_bytes), "r"(0), "r"(tx0), "r"(ty0), "r"(img)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) s_mbar_arrive(tempty0 + 8u * buf);
    }
    if (warp == 9 && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 8) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
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
  p.tmem_cols = c.cout <= 32 ? 64 : 128;  // two accumulators
  size_t smem = 1024 + 2 * (size_t)p.kblocks * 128 * 128 + (size_t)p.kblocks * c.cout * 128 +
                2 * ((((size_t)p.patch_h * p.ppitch + 64) * 2 + 15) & ~(size_t)15) + 1024 + 2 * 128 * (size_t)c.cout * 2 + 128;
  // output as (C, Wo, Ho, N) with the view's channel pitch: one box = the 8 x 16 pixel block of a tile
  alignas(64) CUtensorMap map_o;
  {
    const uint64_t dims[4] = {(uint64_t)c.cout, (uint64_t)c.wo, (uint64_t)c.ho, (uint64_t)c.n};
    const uint64_t strides[3] = {(uint64_t)c.out_pitch * 2, (uint64_t)c.wo * c.out_pitch * 2, (uint64_t)c.ho * c.wo * c.out_pitch * 2};
    const uint32_t box[4] = {(uint32_t)c.cout, (uint32_t)TILE_W, (uint32_t)TILE_H, 1};
    if (tc_make_tiled_map(&map_o, c.out, 4, dims, strides, box, c.cout == 64 ? 128 : 64) != 0) return -1;
  }
  // the opt-in shared-memory size is a per-device attribute: set it once on every device that launches the kernel
  static bool attr[64] = {};
  static int sms[64] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return -1;
  if (!attr[dev]) {
    if (cudaFuncSetAttribute(stem_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024) != cudaSuccess) return -1;
    if (cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms[dev] <= 0) sms[dev] = 148;
    attr[dev] = true;
  }
  const int num_sms = sms[dev];
  int per_sm = (int)((220 * 1024) / (smem + 1024));
  if (per_sm > 2) per_sm = 2;  // 416 threads per CTA
  if (per_sm < 1) per_sm = 1;
  int grid = num_sms * per_sm;
  if (grid > p.total_tiles) grid = p.total_tiles;
  stem_tc_kernel<<<grid, STEM_THREADS_TOTAL, smem, s>>>(map_o, p);
  return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

}  // namespace dml
#endif  // DML_LEGACY_VARIANTS
