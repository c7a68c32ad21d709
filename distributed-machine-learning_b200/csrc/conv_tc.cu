// conv_tc.cu -- K3: every Conv2D(+BN+bias+ReLU+residual) of modelObj.predict
// (/root/reference/models.py:40, :65) with Cin >= 16 as an implicit GEMM on the
// 5th-gen tensor cores: D[pixels, cout] = im2col(A)[pixels, taps*cin] * W^T.
//
//   A tile : 128 output pixels x 64 channels of one filter tap, fetched by ONE
//            im2col-mode TMA (cp.async.bulk.tensor.4d...im2col) with hardware
//            zero fill for 'same' padding and element strides for stride 2.
//   B tile : bn output channels x 64 of the packed bf16 weights, tiled TMA.
//   MMA    : tcgen05.mma.cta_group::1.kind::f16, M=128, N=bn, K=16, bf16 in,
//            fp32 accumulators in TMEM.
//   Epilogue: tcgen05.ld -> +folded bias (+residual) -> ReLU -> bf16 -> strided
//            NHWC store (channel slice of a concat buffer when pitch > cout).
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM alloc + MMA
// issue, warps 2..5 = epilogue (TMEM lane quarter = warp_id % 4).
#include "conv_tc.h"

#include <cuda_bf16.h>
#include <cudaTypedefs.h>

#include <array>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <chrono>
#include <fstream>
#include <map>
#include <sstream>
#include <string>
#include <dlfcn.h>
#include <sys/stat.h>
#include <unistd.h>

namespace dml {
namespace {

thread_local std::string g_tc_err;
int tc_fail(const char* fmt, ...) {
  char buf[768];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_tc_err = buf;
  return -1;
}

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;                       // bf16 elements = one 128-byte swizzle row
constexpr int A_STAGE_BYTES = BLOCK_M * BLOCK_K * 2;  // 16 KiB
constexpr int TC_THREADS = 192;
constexpr int SMEM_LIMIT = 227 * 1024;

struct TcKernelParams {
  const float* bias;
  const __nv_bfloat16* res;
  __nv_bfloat16* out;
  int M;         // valid output pixels (n * ho * wo)
  int ho, wo;
  int stride, pad_h, pad_w;
  int kw, num_kb, cpb;
  int bn, out_pitch, res_pitch, relu;
  int stages, tmem_cols;
};

// ------------------------------- PTX helpers -------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// Bounded wait: a protocol bug traps (kernel error) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(done)
      : "r"(bar), "r"(parity)
      : "memory");
  if (done) return;  // already complete: no clock read on the fast path
  long long t0 = clock64();
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) break;
    if (clock64() - t0 > 4000000000LL) __trap();
  }
}
// Same, polling with the non-suspending test_wait: for the one-warp producer / MMA roles, whose wake-up
// latency after a failed (suspended) try_wait is on the critical path of every pipeline stage.
__device__ __forceinline__ void mbar_wait_spin(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  uint32_t spins = 0;
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) break;
    if (++spins > 400000000u) __trap();
  }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int x, int y) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
      "l"(map), "r"(bar), "r"(x), "r"(y)
      : "memory");
}
__device__ __forceinline__ void tma_load_im2col_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c, int w, int h,
                                                   int n, uint16_t off_w, uint16_t off_h) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};" ::"r"(dst),
      "l"(map), "r"(bar), "r"(c), "r"(w), "r"(h), "r"(n), "h"(off_w), "h"(off_h)
      : "memory");
}
// One lane of a converged warp; the rest of the loop stays warp-uniform so that addresses and
// descriptors live in uniform registers (no per-instruction R2UR / serialisation loop).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accum)
      : "memory");
}
// K-major, 128-byte swizzle: rows 128 B apart, 8-row groups 1024 B apart (SBO),
// LBO unused (1), descriptor version 1 (sm_100), layout type 2 (SWIZZLE_128B).
__device__ __forceinline__ uint64_t make_sw128_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void tmem_ld16_nowait(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}

#ifdef DML_LEGACY_VARIANTS  // superseded A/B twins (one tile per CTA; halo kernel with streamed weights): test builds only
// ------------------------------- the kernel --------------------------------
__global__ void __launch_bounds__(TC_THREADS) conv_tc_kernel(const __grid_constant__ CUtensorMap map_a,
                                                             const __grid_constant__ CUtensorMap map_b,
                                                             const TcKernelParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B tiles need 1024-B alignment
  const int stages = p.stages;
  const uint32_t b_stage_bytes = (uint32_t)p.bn * 128u;
  const uint32_t sA = base;
  const uint32_t sB = sA + (uint32_t)stages * A_STAGE_BYTES;
  const uint32_t bars = sB + (uint32_t)stages * b_stage_bytes;
  const uint32_t full0 = bars, empty0 = bars + 8u * stages, tmem_full = bars + 16u * stages;
  const uint32_t tmem_slot = tmem_full + 8u;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * BLOCK_M;
  const int n0 = blockIdx.y * p.bn;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    for (int i = 0; i < stages; ++i) {
      mbar_init(full0 + 8u * i, 1);
      mbar_init(empty0 + 8u * i, 1);
    }
    mbar_init(tmem_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(p.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

  if (warp == 0) {
    if (lane == 0) {
      // ---- TMA producer: one (tap, 64-channel block) per pipeline stage ----
      const int howo = p.ho * p.wo;
      const int img = m0 / howo;
      const int rem = m0 - img * howo;
      const int oh = rem / p.wo, ow = rem - oh * p.wo;
      const int w_base = ow * p.stride - p.pad_w;  // im2col base pixel = top-left of the filter window
      const int h_base = oh * p.stride - p.pad_h;
      const uint32_t tx_bytes = A_STAGE_BYTES + b_stage_bytes;
      for (int j = 0; j < p.num_kb; ++j) {
        const int s = j % stages;
        const uint32_t ph = (uint32_t)(j / stages) & 1u;
        mbar_wait(empty0 + 8u * s, ph ^ 1u);
        mbar_arrive_expect_tx(full0 + 8u * s, tx_bytes);
        const int tap = j / p.cpb, cb = j - tap * p.cpb;
        const int r = tap / p.kw, sx = tap - r * p.kw;
        tma_load_im2col_4d(sA + (uint32_t)s * A_STAGE_BYTES, &map_a, full0 + 8u * s, cb * BLOCK_K, w_base, h_base, img,
                           (uint16_t)sx, (uint16_t)r);
        tma_load_2d(sB + (uint32_t)s * b_stage_bytes, &map_b, full0 + 8u * s, j * BLOCK_K, n0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ---- MMA issuer ----
      // idesc: D=f32 (bit 4), A=bf16 (bit 7), B=bf16 (bit 10), K-major A and B,
      //        N>>3 at bit 17, M>>4 at bit 24.
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.bn >> 3) << 17) | ((uint32_t)(BLOCK_M >> 4) << 24);
      for (int j = 0; j < p.num_kb; ++j) {
        const int s = j % stages;
        const uint32_t ph = (uint32_t)(j / stages) & 1u;
        mbar_wait(full0 + 8u * s, ph);
        tc_fence_after();
        const uint64_t da = make_sw128_desc(sA + (uint32_t)s * A_STAGE_BYTES);
        const uint64_t db = make_sw128_desc(sB + (uint32_t)s * b_stage_bytes);
#pragma unroll
        for (int k = 0; k < BLOCK_K / 16; ++k) {
          // advancing K by 16 bf16 = 32 bytes inside the swizzle row: +2 in the (addr >> 4) field
          umma_bf16(tmem_base, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)((j | k) != 0));
        }
        umma_commit(empty0 + 8u * s);  // frees the smem stage when these MMAs retire
      }
      umma_commit(tmem_full);  // accumulator complete
    }
  } else {
    // ---- epilogue: warp q owns TMEM lanes [32q, 32q+32) = tile rows ----
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const long long m = (long long)m0 + row;
    const bool row_ok = m < p.M;
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    __nv_bfloat16* orow = p.out + m * p.out_pitch + n0;
    const __nv_bfloat16* rrow = p.res ? p.res + m * p.res_pitch + n0 : nullptr;
    for (int c0 = 0; c0 < p.bn; c0 += 16) {
      uint32_t v[16];
      tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
      if (row_ok) {
        float f[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) f[i] = __uint_as_float(v[i]) + __ldg(p.bias + n0 + c0 + i);
        if (rrow) {
          uint4 r0 = *reinterpret_cast<const uint4*>(rrow + c0);
          uint4 r1 = *reinterpret_cast<const uint4*>(rrow + c0 + 8);
          const __nv_bfloat16* rb0 = reinterpret_cast<const __nv_bfloat16*>(&r0);
          const __nv_bfloat16* rb1 = reinterpret_cast<const __nv_bfloat16*>(&r1);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            f[i] += __bfloat162float(rb0[i]);
            f[8 + i] += __bfloat162float(rb1[i]);
          }
        }
        if (p.relu) {
#pragma unroll
          for (int i = 0; i < 16; ++i) f[i] = fmaxf(f[i], 0.f);
        }
        __align__(16) __nv_bfloat16 o[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) o[i] = __float2bfloat16_rn(f[i]);
        *reinterpret_cast<uint4*>(orow + c0) = *reinterpret_cast<const uint4*>(o);
        *reinterpret_cast<uint4*>(orow + c0 + 8) = *reinterpret_cast<const uint4*>(o + 8);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
  }
}


#endif  // DML_LEGACY_VARIANTS
// ------------------- v2: persistent, warp-specialised kernel -------------------
// One CTA per SM walks tiles tile = blockIdx.x + i*gridDim.x (N tiles of one M tile
// adjacent, so concurrent CTAs share the A rows in L2).  The smem ring and the
// TMA producer run ahead across tile boundaries; two TMEM accumulators let the
// MMA warp start tile i+1 while the epilogue warps drain tile i.  The epilogue
// stages bf16 rows in swizzled shared memory and writes them with TMA stores
// (full 128-byte lines; rows past the tensor end are clipped by the TMA unit).
struct TcV2Params {
  const float* bias;
  const __nv_bfloat16* res;
  int M, ho, wo;
  int stride, pad_h, pad_w;
  int kw, num_kb, cpb;
  int bn, res_pitch, relu;
  int stages, tmem_cols, acc_stride;
  int m_tiles, n_tiles;
  int cw;  // epilogue chunk width in columns: 64 / 32 / 16 (TMA store swizzle 128B / 64B / 32B)
  int has_res;      // residual tiles are TMA-loaded into shared memory by the producer warp
  int out_row0;     // first output row of this launch inside the output/residual maps
  int in_img0;      // first image of this launch inside the input map
  int a_tiled;      // 1x1 stride-1 convs: A is a plain [M][cin] matrix, fetched with tiled TMA
  int in_row0;      // first input row (pixel) of this launch for the tiled A map
  int last_kmma;    // K=16 steps that carry real channels in the last 64-channel block of a tap (1..4)
  int group;        // 64-channel k-blocks per pipeline stage: one barrier round trip per `group` x 4 MMAs
  int prof;         // DML_TC_PROF: per-role cycle counters into g_tc2_prof (v2 only)
  int M_cap;        // rows of the output / residual tensors
  int kb1;          // k-blocks of the first GEMM source; [kb1, num_kb) come from the second one (pair kernel)
  int stride2, a2_tiled, in2_row0;
  int norelu_cols;  // output columns [0, norelu_cols) keep their sign (no ReLU): multiple of 16
  int rev;          // walk the tiles from the last to the first (see ConvParams::rev)
};

__device__ long long g_tc2_prof[160 * 8];

constexpr int OUT_BUF_BYTES = 4096;  // 32 rows x 128 B
constexpr int OUT_STAGING_BYTES = 8 * 2 * OUT_BUF_BYTES;  // 8 epilogue warps x 2 buffers
constexpr int TC2_THREADS = 320;                       // TMA warp, MMA warp, 8 epilogue warps
constexpr int TC2X_THREADS = 352;                      // pair kernel: + one warp that loads the residual tiles

__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, uint32_t src, int x, int y) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(map), "r"(src), "r"(x),
               "r"(y)
               : "memory");
}

__global__ void __launch_bounds__(TC2_THREADS, 1) conv_tc2_kernel(const __grid_constant__ CUtensorMap map_a,
                                                                 const __grid_constant__ CUtensorMap map_b,
                                                                 const __grid_constant__ CUtensorMap map_o,
                                                                 const __grid_constant__ CUtensorMap map_r,
                                                                 const TcV2Params p) {
  pdl_launch_dependents();
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int stages = p.stages;
  const uint32_t b_stage_bytes = (uint32_t)p.bn * 128u;
  const int G = p.group;
  const uint32_t sA = base;
  const uint32_t sB = sA + (uint32_t)(stages * G) * A_STAGE_BYTES;
  const uint32_t sO = sB + (uint32_t)(stages * G) * b_stage_bytes;
  const uint32_t res_tile_bytes = p.has_res ? (uint32_t)BLOCK_M * (uint32_t)p.bn * 2u : 0u;
  const uint32_t sR = sO + OUT_STAGING_BYTES;  // 2 residual tiles: [chunk][128 rows][cw*2 bytes], swizzled like the output
  const uint32_t bars = sR + 2u * res_tile_bytes;
  const uint32_t full0 = bars, empty0 = bars + 8u * stages;
  const uint32_t tfull0 = bars + 16u * stages, tempty0 = tfull0 + 16u;
  const uint32_t rfull0 = tempty0 + 16u, rempty0 = rfull0 + 16u;
  const uint32_t tmem_slot = rempty0 + 16u;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_tiles = p.m_tiles * p.n_tiles;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_o) : "memory");
    for (int i = 0; i < stages; ++i) {
      mbar_init(full0 + 8u * i, 1);
      mbar_init(empty0 + 8u * i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(tfull0 + 8u * i, 1);
      mbar_init(tempty0 + 8u * i, 8);  // one arrive per epilogue warp
      mbar_init(rfull0 + 8u * i, 1);
      mbar_init(rempty0 + 8u * i, 8);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(p.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");
  pdl_wait();  // everything above overlapped the previous kernel's tail; its results are visible from here on

  if (warp == 0) {
    // ---------------- TMA producer (whole warp walks the loop, one elected lane issues) ----------------
    const int howo = p.ho * p.wo;
    const uint32_t tx_bytes = A_STAGE_BYTES + b_stage_bytes;
    uint32_t it = 0, lt = 0;
    long long prof_wait = 0;
    const long long prof_t0 = p.prof ? clock64() : 0;
    const uint32_t chunk_bytes = (uint32_t)BLOCK_M * (uint32_t)p.cw * 2u;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++lt) {
      const int tix = p.rev ? total_tiles - 1 - tile : tile;
      const int mt = tix / p.n_tiles, nt = tix - mt * p.n_tiles;
      const int m0 = mt * BLOCK_M, n0 = nt * p.bn;
      if (p.has_res) {
        // whole residual tile of this output tile, one box per column chunk
        const uint32_t rb = lt & 1u, rph = (lt >> 1) & 1u;
        mbar_wait_spin(rempty0 + 8u * rb, rph ^ 1u);
        if (elect_one()) {
          mbar_arrive_expect_tx(rfull0 + 8u * rb, res_tile_bytes);
          for (int c = 0, ci = 0; c < p.bn; c += p.cw, ++ci)
            tma_load_2d(sR + rb * res_tile_bytes + (uint32_t)ci * chunk_bytes, &map_r, rfull0 + 8u * rb, n0 + c,
                        p.out_row0 + m0);
        }
        __syncwarp();
      }
      const int img = m0 / howo;
      const int rem = m0 - img * howo;
      const int oh = rem / p.wo, ow = rem - oh * p.wo;
      const int w_base = ow * p.stride - p.pad_w, h_base = oh * p.stride - p.pad_h;
      int cb = 0, r = 0, sx = 0;
      for (int j0 = 0; j0 < p.num_kb; j0 += G, ++it) {
        const int cnt = p.num_kb - j0 < G ? p.num_kb - j0 : G;  // k-blocks in this stage
        const uint32_t s = it % (uint32_t)stages;
        const uint32_t ph = (it / (uint32_t)stages) & 1u;
        const long long w0 = p.prof ? clock64() : 0;
        mbar_wait_spin(empty0 + 8u * s, ph ^ 1u);
        if (p.prof) prof_wait += clock64() - w0;
        if (elect_one()) mbar_arrive_expect_tx(full0 + 8u * s, (uint32_t)cnt * tx_bytes);
        __syncwarp();
        for (int g = 0; g < cnt; ++g) {
          if (elect_one()) {
            const uint32_t sub = s * (uint32_t)G + (uint32_t)g;
            if (p.a_tiled)
              tma_load_2d(sA + sub * A_STAGE_BYTES, &map_a, full0 + 8u * s, cb * BLOCK_K, p.in_row0 + m0);
            else
              tma_load_im2col_4d(sA + sub * A_STAGE_BYTES, &map_a, full0 + 8u * s, cb * BLOCK_K, w_base, h_base,
                                 img + p.in_img0, (uint16_t)sx, (uint16_t)r);
            tma_load_2d(sB + sub * b_stage_bytes, &map_b, full0 + 8u * s, (j0 + g) * BLOCK_K, n0);
          }
          __syncwarp();
          if (++cb == p.cpb) {
            cb = 0;
            if (++sx == p.kw) { sx = 0; ++r; }
          }
        }
      }
    }
    if (p.prof && lane == 0) {
      g_tc2_prof[blockIdx.x * 8 + 0] = prof_wait;
      g_tc2_prof[blockIdx.x * 8 + 1] = clock64() - prof_t0;
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer (whole warp walks the loop, one elected lane issues) ----------------
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.bn >> 3) << 17) | ((uint32_t)(BLOCK_M >> 4) << 24);
    uint32_t it = 0, lt = 0;
    long long prof_te = 0, prof_full = 0;
    const long long prof_t0 = p.prof ? clock64() : 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++lt) {
      const uint32_t a = lt & 1u, aph = (lt >> 1) & 1u;
      const long long w1 = p.prof ? clock64() : 0;
      mbar_wait_spin(tempty0 + 8u * a, aph ^ 1u);  // epilogue has drained this accumulator
      if (p.prof) prof_te += clock64() - w1;
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + a * (uint32_t)p.acc_stride;
      int cb = 0;
      for (int j0 = 0; j0 < p.num_kb; j0 += G, ++it) {
        const int cnt = p.num_kb - j0 < G ? p.num_kb - j0 : G;
        const uint32_t s = it % (uint32_t)stages;
        const uint32_t ph = (it / (uint32_t)stages) & 1u;
        const long long w2 = p.prof ? clock64() : 0;
        mbar_wait_spin(full0 + 8u * s, ph);
        if (p.prof) prof_full += clock64() - w2;
        tc_fence_after();
        for (int g = 0; g < cnt; ++g) {
          const uint32_t sub = s * (uint32_t)G + (uint32_t)g;
          const uint64_t da = make_sw128_desc(sA + sub * A_STAGE_BYTES);
          const uint64_t db = make_sw128_desc(sB + sub * b_stage_bytes);
          // channels past Cin in the last block of a tap are zero padding: skip their K steps
          const int nk = (cb == p.cpb - 1) ? p.last_kmma : BLOCK_K / 16;
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BLOCK_K / 16; ++k)
              if (k < nk) umma_bf16(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)((j0 | g | k) != 0));
          }
          __syncwarp();
          if (++cb == p.cpb) cb = 0;
        }
        if (elect_one()) umma_commit(empty0 + 8u * s);
        __syncwarp();
      }
      if (elect_one()) umma_commit(tfull0 + 8u * a);
      __syncwarp();
    }
    if (p.prof && lane == 0) {
      g_tc2_prof[blockIdx.x * 8 + 2] = prof_te;
      g_tc2_prof[blockIdx.x * 8 + 3] = prof_full;
      g_tc2_prof[blockIdx.x * 8 + 4] = clock64() - prof_t0;
      g_tc2_prof[blockIdx.x * 8 + 7] = lt;
    }
  } else {
    // ---------------- epilogue warps ----------------
    // 8 warps: TMEM lane quarter q = warp % 4 (hardware rule), and the two warps of a quarter split the column chunks
    const int q = warp & 3;
    const int ew = warp - 2, half = ew >> 2;
    const uint32_t my_out = sO + (uint32_t)ew * 2u * OUT_BUF_BYTES;
    const int cw = p.cw;
    const uint32_t row_bytes = (uint32_t)cw * 2u;
    const uint32_t chunk_bytes = (uint32_t)BLOCK_M * row_bytes;
    const uint32_t swz = cw == 64 ? (uint32_t)(lane & 7) : (cw == 32 ? (uint32_t)((lane >> 1) & 3) : (uint32_t)((lane >> 2) & 1));
    uint32_t lt = 0, nstore = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++lt) {
      const int tix = p.rev ? total_tiles - 1 - tile : tile;
      const int mt = tix / p.n_tiles, nt = tix - mt * p.n_tiles;
      const int m0 = mt * BLOCK_M, n0 = nt * p.bn;
      const uint32_t a = lt & 1u, aph = (lt >> 1) & 1u;
      if (p.has_res) mbar_wait(rfull0 + 8u * a, aph);
      mbar_wait(tfull0 + 8u * a, aph);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + a * (uint32_t)p.acc_stride;
      const uint32_t r_row = sR + a * res_tile_bytes + (uint32_t)(q * 32 + lane) * row_bytes;
      int ci = half;
      for (int c0 = half * cw; c0 < p.bn; c0 += 2 * cw, ++nstore, ci += 2) {
        const uint32_t sbuf = my_out + (nstore & 1u) * OUT_BUF_BYTES;
        // the TMA store that last read this buffer (two stores ago) must be done reading
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        __syncwarp();
        const uint32_t rowaddr = sbuf + (uint32_t)lane * row_bytes;
        for (int g = 0; g < cw; g += 32) {
          // two 16-column TMEM loads in flight per wait
          uint32_t v[2][16];
          const bool two = g + 16 < cw;
          tmem_ld16_nowait(t_row + (uint32_t)(c0 + g), v[0]);
          if (two) tmem_ld16_nowait(t_row + (uint32_t)(c0 + g + 16), v[1]);
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            if (h == 1 && !two) break;
            const int gg = g + 16 * h;
            const float4* bp = reinterpret_cast<const float4*>(p.bias + n0 + c0 + gg);
            float f[16];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              float4 b4 = __ldg(bp + i);
              f[4 * i + 0] = __uint_as_float(v[h][4 * i + 0]) + b4.x;
              f[4 * i + 1] = __uint_as_float(v[h][4 * i + 1]) + b4.y;
              f[4 * i + 2] = __uint_as_float(v[h][4 * i + 2]) + b4.z;
              f[4 * i + 3] = __uint_as_float(v[h][4 * i + 3]) + b4.w;
            }
            const uint32_t j0 = (uint32_t)(gg >> 3);  // 16-byte chunk index of this group inside the row
            if (p.has_res) {
              uint32_t rr[8];
              const uint32_t ra = r_row + (uint32_t)ci * chunk_bytes;
              asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                           : "=r"(rr[0]), "=r"(rr[1]), "=r"(rr[2]), "=r"(rr[3])
                           : "r"(ra + ((j0 ^ swz) << 4)));
              asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                           : "=r"(rr[4]), "=r"(rr[5]), "=r"(rr[6]), "=r"(rr[7])
                           : "r"(ra + (((j0 + 1u) ^ swz) << 4)));
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                f[2 * i] += __uint_as_float(rr[i] << 16);             // low bf16 of the pair
                f[2 * i + 1] += __uint_as_float(rr[i] & 0xffff0000u);  // high bf16
              }
            }
            if (p.relu && n0 + c0 + gg >= p.norelu_cols) {
#pragma unroll
              for (int i = 0; i < 16; ++i) f[i] = fmaxf(f[i], 0.f);
            }
            uint32_t pk[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              __nv_bfloat162 h2 = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
              pk[i] = *reinterpret_cast<uint32_t*>(&h2);
            }
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(rowaddr + ((j0 ^ swz) << 4)), "r"(pk[0]),
                         "r"(pk[1]), "r"(pk[2]), "r"(pk[3])
                         : "memory");
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(rowaddr + (((j0 + 1u) ^ swz) << 4)), "r"(pk[4]),
                         "r"(pk[5]), "r"(pk[6]), "r"(pk[7])
                         : "memory");
          }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(&map_o, sbuf, n0 + c0, p.out_row0 + m0 + q * 32);
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
      // every tcgen05.ld / residual read of this warp is done: hand both buffers back
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(tempty0 + 8u * a);
        if (p.has_res) mbar_arrive(rempty0 + 8u * a);
      }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
  }
}


__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

#ifdef DML_LEGACY_VARIANTS
// ------------- v3: stride-1 multi-tap convs with the input halo staged ONCE -------------
// im2col through L2 re-reads every input pixel once per filter tap (9x for 3x3) and the
// L2 -> SM path (~6 TB/s, no faster than HBM) is what bounds those layers.  Here a tile is
// 128 consecutive positions q of one image's PADDED raster (q = hp*Wp + wp, Wp = W + 2*pad_w).
// The producer loads the rows that cover [q, q + 128 + (kh-1)*Wp + kw-1) with one tiled TMA box
// per 64-channel block (zero fill outside the image = the conv's padding); the A operand of tap
// (r, s) is then the SAME shared memory seen through a descriptor that starts (r*Wp + s) pixels
// (128 bytes each) later.  Positions with wp >= Wo or hp >= Ho are computed and dropped.
struct TcV3Params {
  const float* bias;
  __nv_bfloat16* out;
  int n_img, ho, wo, wp, out_pitch;
  int kh, kw, pad_h, pad_w, cblocks;
  int tiles_per_img, q_valid;   // positions [0, q_valid) of an image hold all its outputs
  int rows_box;                 // input rows per TMA box
  int a_stage_bytes, a_stages, b_stages;
  int bn, n_tiles, relu;
  int tmem_cols, acc_stride, cw;
  int in_img0, out_img0;
};


__global__ void __launch_bounds__(TC_THREADS, 1) conv_tc3_kernel(const __grid_constant__ CUtensorMap map_a,
                                                                 const __grid_constant__ CUtensorMap map_b,
                                                                 const TcV3Params p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t b_stage_bytes = (uint32_t)p.bn * 128u;
  const uint32_t sA = base;
  const uint32_t sB = sA + (uint32_t)p.a_stages * (uint32_t)p.a_stage_bytes;
  const uint32_t sO = sB + (uint32_t)p.b_stages * b_stage_bytes;      // 2 x [128 rows][128 B] staging
  const uint32_t sOff = sO + 2u * 16384u;                              // 2 x 128 int32 output offsets
  const uint32_t bars = sOff + 1024u;
  const uint32_t afull0 = bars, aempty0 = afull0 + 8u * p.a_stages;
  const uint32_t bfull0 = aempty0 + 8u * p.a_stages, bempty0 = bfull0 + 8u * p.b_stages;
  const uint32_t tfull0 = bempty0 + 8u * p.b_stages, tempty0 = tfull0 + 16u;
  const uint32_t tmem_slot = tempty0 + 16u;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int taps = p.kh * p.kw;
  const int total_tiles = p.n_img * p.tiles_per_img * p.n_tiles;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    for (int i = 0; i < p.a_stages; ++i) { mbar_init(afull0 + 8u * i, 1); mbar_init(aempty0 + 8u * i, 1); }
    for (int i = 0; i < p.b_stages; ++i) { mbar_init(bfull0 + 8u * i, 1); mbar_init(bempty0 + 8u * i, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(tfull0 + 8u * i, 1); mbar_init(tempty0 + 8u * i, 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(p.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

  // tile -> (image, first position, n tile); N tiles of one position range are adjacent
  auto decode = [&](int tile, int& img, int& l0, int& n0) {
    const int nt = tile % p.n_tiles;
    const int mt = tile / p.n_tiles;
    img = mt / p.tiles_per_img;
    l0 = (mt - img * p.tiles_per_img) * BLOCK_M;
    n0 = nt * p.bn;
  };

  if (warp == 0) {
    // ---------------- producer: one halo box per channel block, one weight tile per (block, tap) ----------------
    uint32_t ita = 0, itb = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      int img, l0, n0;
      decode(tile, img, l0, n0);
      const int row0 = l0 / p.wp;  // first padded row the tile touches
      for (int cb = 0; cb < p.cblocks; ++cb, ++ita) {
        const uint32_t sa = ita % (uint32_t)p.a_stages, pha = (ita / (uint32_t)p.a_stages) & 1u;
        mbar_wait(aempty0 + 8u * sa, pha ^ 1u);
        if (elect_one()) {
          mbar_arrive_expect_tx(afull0 + 8u * sa, (uint32_t)p.rows_box * (uint32_t)p.wp * 128u);
          // 4D tiled box {64 ch, Wp, rows_box, 1} at (channel, -pad_w, row0 - pad_h, image): OOB -> zeros
          asm volatile(
              "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
                  sA + sa * (uint32_t)p.a_stage_bytes),
              "l"(&map_a), "r"(afull0 + 8u * sa), "r"(cb * BLOCK_K), "r"(-p.pad_w), "r"(row0 - p.pad_h), "r"(img + p.in_img0)
              : "memory");
        }
        __syncwarp();
        for (int t = 0; t < taps; ++t, ++itb) {
          const uint32_t sb = itb % (uint32_t)p.b_stages, phb = (itb / (uint32_t)p.b_stages) & 1u;
          mbar_wait(bempty0 + 8u * sb, phb ^ 1u);
          if (elect_one()) {
            mbar_arrive_expect_tx(bfull0 + 8u * sb, b_stage_bytes);
            tma_load_2d(sB + sb * b_stage_bytes, &map_b, bfull0 + 8u * sb, (t * p.cblocks + cb) * BLOCK_K, n0);
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer (warp-uniform loop, one elected lane issues) ----------------
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.bn >> 3) << 17) | ((uint32_t)(BLOCK_M >> 4) << 24);
    uint32_t ita = 0, itb = 0, lt = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++lt) {
      int img, l0, n0;
      decode(tile, img, l0, n0);
      const int loff = l0 - (l0 / p.wp) * p.wp;  // tile's first position inside the loaded rows
      const uint32_t a = lt & 1u, aph = (lt >> 1) & 1u;
      mbar_wait(tempty0 + 8u * a, aph ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + a * (uint32_t)p.acc_stride;
      uint32_t first = 1;
      for (int cb = 0; cb < p.cblocks; ++cb, ++ita) {
        const uint32_t sa = ita % (uint32_t)p.a_stages, pha = (ita / (uint32_t)p.a_stages) & 1u;
        mbar_wait(afull0 + 8u * sa, pha);
        tc_fence_after();
        const uint32_t a_base = sA + sa * (uint32_t)p.a_stage_bytes;
        int r = 0, sx = 0;
        for (int t = 0; t < taps; ++t, ++itb) {
          const uint32_t sb = itb % (uint32_t)p.b_stages, phb = (itb / (uint32_t)p.b_stages) & 1u;
          mbar_wait(bfull0 + 8u * sb, phb);
          tc_fence_after();
          const uint32_t a_addr = a_base + (uint32_t)(loff + r * p.wp + sx) * 128u;
          // the swizzle is a function of the absolute shared-memory address, so a start that is only
          // 128-byte aligned needs no base_offset (checked on B200: base_offset = (addr >> 7) & 7 is wrong)
          const uint64_t da = make_sw128_desc(a_addr);
          const uint64_t db = make_sw128_desc(sB + sb * b_stage_bytes);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BLOCK_K / 16; ++k)
              umma_bf16(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (first && k == 0) ? 0u : 1u);
            umma_commit(bempty0 + 8u * sb);
          }
          __syncwarp();
          first = 0;
          if (++sx == p.kw) { sx = 0; ++r; }
        }
        if (elect_one()) umma_commit(aempty0 + 8u * sa);
        __syncwarp();
      }
      if (elect_one()) umma_commit(tfull0 + 8u * a);
      __syncwarp();
    }
  } else {
    // ---------------- epilogue: 4 warps = 128 positions; staged in smem, written with coalesced 16-byte stores ----------------
    const int q = warp & 3;
    const int et = q * 32 + lane;  // epilogue thread id = tile row
    const int cw = p.cw;
    const int chunks16 = cw >> 3;  // 16-byte pieces per staged row
    uint32_t lt = 0, nchunk = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++lt) {
      int img, l0, n0;
      decode(tile, img, l0, n0);
      const uint32_t a = lt & 1u, aph = (lt >> 1) & 1u;
      // output offset of this thread's position (elements), -1 if the position is padding
      {
        const int pos = l0 + et;
        const int hp = pos / p.wp, wq = pos - hp * p.wp;
        int off = -1;
        if (hp < p.ho && wq < p.wo) off = (((img + p.out_img0) * p.ho + hp) * p.wo + wq) * p.out_pitch + n0;
        asm volatile("st.shared.b32 [%0], %1;" ::"r"(sOff + (lt & 1u) * 512u + (uint32_t)et * 4u), "r"(off) : "memory");
      }
      mbar_wait(tfull0 + 8u * a, aph);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + a * (uint32_t)p.acc_stride;
      for (int c0 = 0; c0 < p.bn; c0 += cw, ++nchunk) {
        const uint32_t sbuf = sO + (nchunk & 1u) * 16384u;
        const uint32_t rowaddr = sbuf + (uint32_t)et * 128u;
        const uint32_t swz = (uint32_t)(et & 7);
        for (int g = 0; g < cw; g += 32) {
          uint32_t v[2][16];
          const bool two = g + 16 < cw;
          tmem_ld16_nowait(t_row + (uint32_t)(c0 + g), v[0]);
          if (two) tmem_ld16_nowait(t_row + (uint32_t)(c0 + g + 16), v[1]);
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            if (h == 1 && !two) break;
            const int gg = g + 16 * h;
            const float4* bp = reinterpret_cast<const float4*>(p.bias + n0 + c0 + gg);
            uint32_t pk[8];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              float4 b4 = __ldg(bp + i);
              float f0 = __uint_as_float(v[h][4 * i + 0]) + b4.x, f1 = __uint_as_float(v[h][4 * i + 1]) + b4.y;
              float f2 = __uint_as_float(v[h][4 * i + 2]) + b4.z, f3 = __uint_as_float(v[h][4 * i + 3]) + b4.w;
              if (p.relu) { f0 = fmaxf(f0, 0.f); f1 = fmaxf(f1, 0.f); f2 = fmaxf(f2, 0.f); f3 = fmaxf(f3, 0.f); }
              __nv_bfloat162 lo = __floats2bfloat162_rn(f0, f1), hi = __floats2bfloat162_rn(f2, f3);
              pk[2 * i] = *reinterpret_cast<uint32_t*>(&lo);
              pk[2 * i + 1] = *reinterpret_cast<uint32_t*>(&hi);
            }
            const uint32_t j0 = (uint32_t)(gg >> 3);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(rowaddr + ((j0 ^ swz) << 4)), "r"(pk[0]),
                         "r"(pk[1]), "r"(pk[2]), "r"(pk[3])
                         : "memory");
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(rowaddr + (((j0 + 1u) ^ swz) << 4)), "r"(pk[4]),
                         "r"(pk[5]), "r"(pk[6]), "r"(pk[7])
                         : "memory");
          }
        }
        named_bar_sync(1, 128);  // the whole 128 x cw chunk (and the offset table) is staged
        // cooperative copy-out: consecutive threads take consecutive 16-byte pieces of a row
        for (int idx = et; idx < BLOCK_M * chunks16; idx += 128) {
          const int m = idx / chunks16, j = idx - m * chunks16;
          int off;
          asm volatile("ld.shared.b32 %0, [%1];" : "=r"(off) : "r"(sOff + (lt & 1u) * 512u + (uint32_t)m * 4u) : "memory");
          if (off >= 0) {
            uint4 val;
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                         : "=r"(val.x), "=r"(val.y), "=r"(val.z), "=r"(val.w)
                         : "r"(sbuf + (uint32_t)m * 128u + (((uint32_t)j ^ (uint32_t)(m & 7)) << 4))
                         : "memory");
            *reinterpret_cast<uint4*>(p.out + off + c0 + j * 8) = val;
          }
        }
        // the other staging buffer is free again once everybody left the copy-out two chunks ago:
        // the bar.sync of the NEXT chunk orders that (each buffer is reused every second chunk)
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty0 + 8u * a);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
  }
}


#endif  // DML_LEGACY_VARIANTS
// ------------- v6: halo kernel with the WHOLE filter resident in shared memory -------------
// For stride-1 multi-tap layers with few channels (ResNet50 conv2 3x3, the InceptionV3 stem 3x3s) one
// tcgen05.mma is short (N <= 96) and the issuing warp cannot hide a barrier round trip behind it
// (tools/mma_bench.cu: +106 cycles per try_wait at N = 64).  Here nothing is waited for inside a tile:
// every tap's weight tile is loaded ONCE per CTA, the input halo of a tile (all channel blocks) arrives
// behind one barrier, and the MMA warp issues taps x blocks x K-steps instructions back to back.
// Channel blocks are 64 wide (128-byte rows, SWIZZLE_128B), for Cin <= 32 32 wide (64-byte rows, SWIZZLE_64B)
// and for Cin <= 16 16 wide (32-byte rows, SWIZZLE_32B: one K = 16 step per tap), which shrinks the staged
// bytes of those layers accordingly.  The padded raster is Wp = Wo + kw - 1 wide and starts pad_w columns /
// pad_h rows before the image (the pads on the far side may differ: the space-to-depth stems use that).
constexpr int TC6_THREADS = 320;  // producer warp, MMA warp, 8 epilogue warps

// (image, tile inside the image) of a persistent CTA's tiles without a division per tile
struct TileWalk {
  int img, mt, tpi, dimg, dmt, rev;
  // rev: walk from tile total - 1 - first downwards (ConvParams::rev)
  __device__ TileWalk(int tiles_per_img, int first, int stride, int total = 0, int reverse = 0) : tpi(tiles_per_img), rev(reverse) {
    if (rev) first = total - 1 - first;
    if (first < 0) first = 0;  // (a CTA without tiles never uses the walk)
    img = first / tpi; mt = first - img * tpi;
    dimg = stride / tpi; dmt = stride - dimg * tpi;
  }
  __device__ void next() {
    if (!rev) {
      img += dimg; mt += dmt;
      if (mt >= tpi) { mt -= tpi; ++img; }
    } else {
      img -= dimg; mt -= dmt;
      if (mt < 0) { mt += tpi; --img; }
    }
  }
};

struct TcV6Params {
  const float* bias;
  __nv_bfloat16* out;
  int n_img, ho, wo, wp, out_pitch;
  int kh, kw, pad_h, pad_w, cblocks;
  int tiles_per_img, rows_box;
  int a_cb_bytes, a_stage_bytes, a_stages;
  int bn, relu, tmem_cols, acc_stride, cw;
  int in_img0, out_img0;
  int row_bytes;   // 128 or 64
  int last_kmma;   // K = 16 steps that carry real channels in the last channel block
  int cin_pad;     // weight row layout: [tap][cin_pad]
  int prof;
  int wide_store;  // output rows are 32-byte aligned: 256-bit stores
  int n_acc;       // TMEM accumulators (2 or 4)
  int n_tiles;     // Cout / bn
  float inv_wp;    // 1 / wp
  int rev;         // walk the tiles from the last to the first (ConvParams::rev)
};

// q / wp for 0 <= q < 2^16 and wp <= 256 without an integer division: (q + 0.5) / wp is at least 0.5 / 256 away from
// every integer, far more than the float rounding error of the product
__device__ __forceinline__ int div_wp(int q, float inv_wp) { return __float2int_rz(((float)q + 0.5f) * inv_wp); }

__device__ __forceinline__ uint64_t make_sw_desc(uint32_t smem_addr, uint32_t row_bytes) {
  // rows `row_bytes` apart, 8-row groups 8 * row_bytes apart (SBO); layout 2 = SWIZZLE_128B, 4 = SWIZZLE_64B, 6 = SWIZZLE_32B
  const uint64_t layout = row_bytes == 128u ? 2ull : (row_bytes == 64u ? 4ull : 6ull);
  return (uint64_t)((smem_addr >> 4) & 0x3FFF) | ((uint64_t)1 << 16) | ((uint64_t)((8u * row_bytes) >> 4) << 32) |
         ((uint64_t)1 << 46) | (layout << 61);
}

// KH x KW, CBLOCKS and ROW_BYTES are compile-time for the hot shapes (0 = take them from the parameters): the MMA
// warp's loop then unrolls into bare tcgen05.mma instructions whose descriptors differ by immediates.  A single
// warp runs ~50 dependent (uniform-datapath) instructions per k-block in the generic loop, which costs more
// than the four short MMAs it issues (DML_TC_PROF counters).
template <int KH, int KW, int CBLOCKS, int ROW_BYTES>
__global__ void __launch_bounds__(TC6_THREADS, 1) conv_tc6_kernel(const __grid_constant__ CUtensorMap map_a,
                                                                 const __grid_constant__ CUtensorMap map_b,
                                                                 const TcV6Params p) {
  pdl_launch_dependents();
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int taps = p.kh * p.kw;
  const uint32_t row_bytes = (uint32_t)p.row_bytes;
  const uint32_t b_tile_bytes = (uint32_t)p.bn * row_bytes;
  const uint32_t sB = base;                                               // taps x cblocks weight tiles, resident
  const uint32_t sA = sB + (((uint32_t)(taps * p.cblocks) * b_tile_bytes + 1023u) & ~1023u);
  const uint32_t sO = sA + (uint32_t)p.a_stages * (uint32_t)p.a_stage_bytes;  // bn biases (fp32)
  const uint32_t bars = (sO + (uint32_t)p.bn * 4u + 15u) & ~15u;
  const uint32_t afull0 = bars, aempty0 = afull0 + 8u * p.a_stages;
  const uint32_t bfull = aempty0 + 8u * p.a_stages;
  const uint32_t tfull0 = bfull + 8u, tempty0 = tfull0 + 32u;  // up to 4 accumulators
  const uint32_t tmem_slot = tempty0 + 32u;
  const uint32_t n_acc = (uint32_t)p.n_acc;  // 2 or 4
  const uint32_t acc_shift = n_acc == 4u ? 2u : 1u;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_tiles = p.n_img * p.tiles_per_img;
  // Cout may be split into n_tiles slices when the whole filter does not fit: CTA c keeps slice c % n_tiles resident
  // and walks the position tiles as virtual CTA c / n_tiles of gridDim.x / n_tiles
  const int vcta = (int)blockIdx.x / p.n_tiles, vgrid = (int)gridDim.x / p.n_tiles;
  const int n0 = ((int)blockIdx.x - vcta * p.n_tiles) * p.bn;
  long long prof[4] = {0, 0, 0, 0};  // DML_TC_PROF: producer wait empty | mma wait tmem-empty, wait A | epilogue wait accumulator
  const long long prof_t0 = p.prof ? clock64() : 0;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    for (int i = 0; i < p.a_stages; ++i) { mbar_init(afull0 + 8u * i, 1); mbar_init(aempty0 + 8u * i, 1); }
    mbar_init(bfull, 1);
    for (int i = 0; i < 4; ++i) { mbar_init(tfull0 + 8u * i, 1); mbar_init(tempty0 + 8u * i, 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(p.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");
  pdl_wait();  // everything above overlapped the previous kernel's tail; its results are visible from here on

  if (warp == 0) {
    // ---------------- producer: the filter once, then one halo (all channel blocks) per tile ----------------
    const int kb = p.row_bytes >> 1;  // channels per block
    if (elect_one()) {
      mbar_arrive_expect_tx(bfull, (uint32_t)(taps * p.cblocks) * b_tile_bytes);
      for (int t = 0; t < taps; ++t)
        for (int cb = 0; cb < p.cblocks; ++cb)
          tma_load_2d(sB + (uint32_t)(t * p.cblocks + cb) * b_tile_bytes, &map_b, bfull, t * p.cin_pad + cb * kb, n0);
    }
    __syncwarp();
    uint32_t ita = 0;
    const uint32_t box_bytes = (uint32_t)p.rows_box * (uint32_t)p.wp * row_bytes;
    TileWalk tw(p.tiles_per_img, vcta, vgrid, total_tiles, p.rev);
    uint32_t sa = 0, pha = 0;
    for (int tile = vcta; tile < total_tiles; tile += vgrid, ++ita, tw.next(), pha ^= (++sa == (uint32_t)p.a_stages), sa = sa == (uint32_t)p.a_stages ? 0u : sa) {
      const int img = tw.img;
      const int l0 = tw.mt * BLOCK_M;
      const int row0 = div_wp(l0, p.inv_wp);  // first padded row the tile touches
      const long long w0 = p.prof ? clock64() : 0;
      mbar_wait(aempty0 + 8u * sa, pha ^ 1u);
      if (p.prof) prof[0] += clock64() - w0;
      if (elect_one()) {
        mbar_arrive_expect_tx(afull0 + 8u * sa, (uint32_t)p.cblocks * box_bytes);
        for (int cb = 0; cb < p.cblocks; ++cb)
          // 4D tiled box {channels, Wp, rows_box, 1} at (channel, -pad_w, row0 - pad_h, image): out of bounds -> zeros
          asm volatile(
              "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
                  sA + sa * (uint32_t)p.a_stage_bytes + (uint32_t)cb * (uint32_t)p.a_cb_bytes),
              "l"(&map_a), "r"(afull0 + 8u * sa), "r"(cb * kb), "r"(-p.pad_w), "r"(row0 - p.pad_h), "r"(img + p.in_img0)
              : "memory");
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer: no waits inside a tile ----------------
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.bn >> 3) << 17) | ((uint32_t)(BLOCK_M >> 4) << 24);
    const int ksteps = p.row_bytes >> 5;
    mbar_wait(bfull, 0);
    uint32_t ita = 0;
    TileWalk tw(p.tiles_per_img, vcta, vgrid, total_tiles, p.rev);
    uint32_t sa = 0, pha = 0;
    for (int tile = vcta; tile < total_tiles; tile += vgrid, ++ita, tw.next(), pha ^= (++sa == (uint32_t)p.a_stages), sa = sa == (uint32_t)p.a_stages ? 0u : sa) {
      const int l0 = tw.mt * BLOCK_M;
      const int loff = l0 - div_wp(l0, p.inv_wp) * p.wp;  // tile's first position inside the loaded rows
      const uint32_t a = ita & (n_acc - 1u), aph = (ita >> acc_shift) & 1u;
      const long long w1 = p.prof ? clock64() : 0;
      mbar_wait(tempty0 + 8u * a, aph ^ 1u);
      const long long w2 = p.prof ? clock64() : 0;
      mbar_wait(afull0 + 8u * sa, pha);
      if (p.prof) { const long long w3 = clock64(); prof[1] += w2 - w1; prof[2] += w3 - w2; }
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + a * (uint32_t)p.acc_stride;
      const uint32_t a_base = sA + sa * (uint32_t)p.a_stage_bytes + (uint32_t)loff * row_bytes;
      if (KH > 0) {
        // compile-time shape: one election per tile, descriptors = tile base + per-tap constants
        const uint64_t hi = make_sw_desc(0, (uint32_t)ROW_BYTES);
        const uint32_t a_lo = a_base >> 4, b_lo = sB >> 4;
        const uint32_t wp16 = ((uint32_t)p.wp * (uint32_t)ROW_BYTES) >> 4;
        const uint32_t acb16 = (uint32_t)p.a_cb_bytes >> 4, bt16 = b_tile_bytes >> 4;
        const int nk_last = p.last_kmma;
        if (elect_one()) {
#pragma unroll
          for (int t = 0; t < KH * KW; ++t) {
#pragma unroll
            for (int cb = 0; cb < CBLOCKS; ++cb) {
              const uint32_t ao = a_lo + (uint32_t)(t / KW) * wp16 + (uint32_t)((t % KW) * (ROW_BYTES >> 4)) + (uint32_t)cb * acb16;
              const uint32_t bo = b_lo + (uint32_t)(t * CBLOCKS + cb) * bt16;
#pragma unroll
              for (int k = 0; k < ROW_BYTES / 32; ++k)
                if (cb < CBLOCKS - 1 || k < nk_last)
                  umma_bf16(d_tmem, hi | (uint64_t)((ao + 2u * k) & 0x3FFFu), hi | (uint64_t)((bo + 2u * k) & 0x3FFFu), idesc,
                            (uint32_t)((t | cb | k) != 0));
            }
          }
          umma_commit(aempty0 + 8u * sa);
          umma_commit(tfull0 + 8u * a);
        }
        __syncwarp();
        continue;
      }
      int r = 0, sx = 0;
      for (int t = 0; t < taps; ++t) {
        for (int cb = 0; cb < p.cblocks; ++cb) {
          // the swizzle is a function of the absolute shared-memory address: a start that is only row-aligned
          // needs no base_offset (tests/test_gpu_kernels.py covers both row widths)
          const uint64_t da = make_sw_desc(a_base + (uint32_t)cb * (uint32_t)p.a_cb_bytes + (uint32_t)(r * p.wp + sx) * row_bytes, row_bytes);
          const uint64_t db = make_sw_desc(sB + (uint32_t)(t * p.cblocks + cb) * b_tile_bytes, row_bytes);
          const int nk = (cb == p.cblocks - 1) ? p.last_kmma : ksteps;
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
              if (k < nk) umma_bf16(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)((t | cb | k) != 0));
          }
          __syncwarp();
        }
        if (++sx == p.kw) { sx = 0; ++r; }
      }
      if (elect_one()) {
        umma_commit(aempty0 + 8u * sa);
        umma_commit(tfull0 + 8u * a);
      }
      __syncwarp();
    }
  } else {
    // ---------------- epilogue: 4 warps = 128 positions; every thread owns one position ----------------
    // The thread converts its accumulator row 16 channels at a time and writes the two 16-byte pieces straight
    // to its pixel in global memory.  Measured against a swizzled staging buffer + cooperative copy-out (2200-3500
    // cycles per tile: one warp per scheduler cannot hide that instruction stream) and against one 1-D bulk copy
    // per thread (~22 cycles of TMA issue per copy): the strided 16-byte stores cost ~900 cycles per tile and L2
    // merges them into full lines before they reach HBM.
    // 8 warps in two groups of 4 (TMEM lane quarter q = warp % 4, hardware rule).  Group g drains the tiles with
    // lt % 2 == g, so two tiles are in the epilogue at once: one tile's chain (TMEM load -> bias/ReLU/pack -> store)
    // is latency-bound, and with up to four accumulators the MMA warp keeps issuing meanwhile.
    const int q = warp & 3;
    const int grp = (warp - 2) >> 2;
    const int et = q * 32 + lane;  // tile row
    const uint32_t sbias = sO;
    for (int c = (int)threadIdx.x - 64; c < p.bn; c += 256) {
      const float bv = __ldg(p.bias + n0 + c);
      asm volatile("st.shared.f32 [%0], %1;" ::"r"(sbias + (uint32_t)c * 4u), "f"(bv) : "memory");
    }
    named_bar_sync(1, 256);
    const int g_begin = 0, g_end = p.bn;
    uint32_t lt = (uint32_t)grp;
    TileWalk tw(p.tiles_per_img, vcta + grp * vgrid, 2 * vgrid, total_tiles, p.rev);
    for (int tile = vcta + grp * vgrid; tile < total_tiles; tile += 2 * vgrid, lt += 2, tw.next()) {
      const int img = tw.img;
      const int l0 = tw.mt * BLOCK_M;
      const uint32_t a = lt & (n_acc - 1u), aph = (lt >> acc_shift) & 1u;
      const int pos = l0 + et;
      const int hp = div_wp(pos, p.inv_wp), wq = pos - hp * p.wp;
      const bool valid = hp < p.ho && wq < p.wo;
      const long long off = ((long long)((img + p.out_img0) * p.ho + hp) * p.wo + wq) * p.out_pitch + n0;
      const long long w4 = p.prof ? clock64() : 0;
      mbar_wait(tfull0 + 8u * a, aph);
      if (p.prof) prof[3] += clock64() - w4;
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + a * (uint32_t)p.acc_stride;
      for (int g = g_begin; g < g_end; g += 32) {
        uint32_t v[2][16];
        const bool two = g + 16 < g_end;
        tmem_ld16_nowait(t_row + (uint32_t)g, v[0]);
        if (two) tmem_ld16_nowait(t_row + (uint32_t)(g + 16), v[1]);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          if (h == 1 && !two) break;
          const int gg = g + 16 * h;
          uint32_t pk[8];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float4 b4;
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                         : "=f"(b4.x), "=f"(b4.y), "=f"(b4.z), "=f"(b4.w)
                         : "r"(sbias + (uint32_t)(gg + 4 * i) * 4u));
            float f0 = __uint_as_float(v[h][4 * i + 0]) + b4.x, f1 = __uint_as_float(v[h][4 * i + 1]) + b4.y;
            float f2 = __uint_as_float(v[h][4 * i + 2]) + b4.z, f3 = __uint_as_float(v[h][4 * i + 3]) + b4.w;
            if (p.relu) { f0 = fmaxf(f0, 0.f); f1 = fmaxf(f1, 0.f); f2 = fmaxf(f2, 0.f); f3 = fmaxf(f3, 0.f); }
            __nv_bfloat162 lo = __floats2bfloat162_rn(f0, f1), hi = __floats2bfloat162_rn(f2, f3);
            pk[2 * i] = *reinterpret_cast<uint32_t*>(&lo);
            pk[2 * i + 1] = *reinterpret_cast<uint32_t*>(&hi);
          }
          if (valid && p.wide_store != 2) {
            if (p.wide_store) {  // one 32-byte store: a whole sector per thread
              asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p.out + off + gg), "r"(pk[0]), "r"(pk[1]),
                           "r"(pk[2]), "r"(pk[3]), "r"(pk[4]), "r"(pk[5]), "r"(pk[6]), "r"(pk[7])
                           : "memory");
            } else {
              uint4* dst = reinterpret_cast<uint4*>(p.out + off + gg);
              dst[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
              dst[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
            }
          }
        }
      }
      // accumulator fully read: hand it back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty0 + 8u * a);
    }
  }
  if (p.prof && lane == 0 && warp <= 2) {
    if (warp == 0) g_tc2_prof[blockIdx.x * 8 + 0] = prof[0];
    if (warp == 1) { g_tc2_prof[blockIdx.x * 8 + 1] = prof[1]; g_tc2_prof[blockIdx.x * 8 + 2] = prof[2]; }
    if (warp == 2) { g_tc2_prof[blockIdx.x * 8 + 3] = prof[3]; g_tc2_prof[blockIdx.x * 8 + 4] = clock64() - prof_t0; }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
  }
}


// ------------- v4: the persistent kernel on CTA PAIRS (tcgen05 cta_group::2, M = 256) -------------
// One tcgen05.mma costs ~150-250 cycles of tensor-pipe time on B200 however little it computes
// (measured: MMA-only runs), so a 128 x N x 16 instruction cannot reach peak.  Two CTAs of a
// cluster (one SM pair) therefore share each instruction: every CTA loads its own 128 rows of A and
// HALF of the weight tile, the leader CTA issues tcgen05.mma.cta_group::2 (M = 256, N = bn) and both
// CTAs receive their 128 accumulator rows in their own TMEM.  Twice the work per instruction, and a
// third less operand traffic per output.
constexpr uint32_t PEER_BIT_MASK = 0xFEFFFFFFu;  // shared::cluster address -> same offset in the pair's even CTA

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma2_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int x, int y) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
      "l"(map), "r"(bar & PEER_BIT_MASK), "r"(x), "r"(y)
      : "memory");
}
__device__ __forceinline__ void tma2_load_im2col_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c, int w, int h,
                                                    int n, uint16_t off_w, uint16_t off_h) {
  asm volatile(
      "cp.async.bulk.tensor.4d.im2col.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};" ::"r"(dst),
      "l"(map), "r"(bar & PEER_BIT_MASK), "r"(c), "r"(w), "r"(h), "r"(n), "h"(off_w), "h"(off_h)
      : "memory");
}
__device__ __forceinline__ void umma2_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accum)
      : "memory");
}
__device__ __forceinline__ void umma2_commit_both(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
               "h"((uint16_t)3)
               : "memory");
}
// Branch-free forms for the lean loops: every lane runs the instruction stream, elect.sync picks the one lane whose
// predicate lets the TMA / MMA / commit through.  Without a C++ branch around them the operands stay warp-uniform
// for the compiler (uniform registers, no R2UR / BSSY / BSYNC per k-block), which is what the single issuing warp's
// time goes into otherwise (DML_TC_PROF: ~400-650 busy cycles per k-block against 256-512 cycles of MMA work).
__device__ __forceinline__ void tma2_load_2d_elect(uint32_t dst, const CUtensorMap* map, uint32_t bar, int x, int y) {
  asm volatile(
      "{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\t"
      "@P cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n\t}" ::"r"(dst),
      "l"(map), "r"(bar & PEER_BIT_MASK), "r"(x), "r"(y)
      : "memory");
}
__device__ __forceinline__ void tma2_load_im2col_4d_elect(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c, int w, int h,
                                                          int n, uint16_t off_w, uint16_t off_h) {
  asm volatile(
      "{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\t"
      "@P cp.async.bulk.tensor.4d.im2col.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};\n\t}" ::"r"(dst),
      "l"(map), "r"(bar & PEER_BIT_MASK), "r"(c), "r"(w), "r"(h), "r"(n), "h"(off_w), "h"(off_h)
      : "memory");
}
// four K = 16 steps of one 64-channel k-block; lo words of the descriptors advance by 2 (32 bytes) per step
__device__ __forceinline__ void umma2_bf16_x4_elect(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t desc_hi, uint32_t idesc,
                                                    uint32_t accum_first) {
  asm volatile(
      "{\n\t.reg .pred P, Q;\n\t.reg .b64 da, db;\n\t.reg .b32 al, bl;\n\t"
      "elect.sync _|P, 0xffffffff;\n\t"
      "setp.ne.b32 Q, %5, 0;\n\t"
      "mov.b64 da, {%1, %3};\n\tmov.b64 db, {%2, %3};\n\t"
      "@P tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %4, Q;\n\t"
      "add.u32 al, %1, 2;\n\tadd.u32 bl, %2, 2;\n\tmov.b64 da, {al, %3};\n\tmov.b64 db, {bl, %3};\n\t"
      "@P tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %4, 1;\n\t"
      "add.u32 al, %1, 4;\n\tadd.u32 bl, %2, 4;\n\tmov.b64 da, {al, %3};\n\tmov.b64 db, {bl, %3};\n\t"
      "@P tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %4, 1;\n\t"
      "add.u32 al, %1, 6;\n\tadd.u32 bl, %2, 6;\n\tmov.b64 da, {al, %3};\n\tmov.b64 db, {bl, %3};\n\t"
      "@P tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %4, 1;\n\t}" ::"r"(tmem_d),
      "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accum_first)
      : "memory");
}
__device__ __forceinline__ void umma2_commit_both_elect(uint32_t bar) {
  asm volatile(
      "{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\t"
      "@P tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n\t}" ::"r"(bar),
      "h"((uint16_t)3)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive_leader(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(bar & PEER_BIT_MASK) : "memory");
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(TC2X_THREADS, 1)
    conv_tc2x_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                     const __grid_constant__ CUtensorMap map_o, const __grid_constant__ CUtensorMap map_r,
                     const __grid_constant__ CUtensorMap map_a2, const TcV2Params p) {
  pdl_launch_dependents();
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int stages = p.stages;
  const uint32_t b_half_bytes = (uint32_t)p.bn * 64u;  // bn/2 weight rows of 128 bytes
  const int G = p.group;
  const uint32_t sA = base;
  const uint32_t sB = sA + (uint32_t)(stages * G) * A_STAGE_BYTES;
  const uint32_t sO = sB + (uint32_t)(stages * G) * b_half_bytes;
  const uint32_t res_tile_bytes = p.has_res ? (uint32_t)BLOCK_M * (uint32_t)p.bn * 2u : 0u;
  const uint32_t res_bufs = p.has_res == 2 ? 1u : 2u;  // has_res == 2: ONE tile buffer whose column chunks are released one by one
  const uint32_t sR = sO + OUT_STAGING_BYTES;
  const uint32_t bars = sR + res_bufs * res_tile_bytes;
  const uint32_t full0 = bars, empty0 = bars + 8u * stages;
  const uint32_t tfull0 = bars + 16u * stages, tempty0 = tfull0 + 16u;
  const uint32_t rfull0 = tempty0 + 16u, rempty0 = rfull0 + 32u;  // 2 tile buffers (mode 1) or up to 4 chunk slots (mode 2)
  const uint32_t tmem_slot = rempty0 + 32u;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
  const int total_tiles = p.m_tiles * p.n_tiles;  // m_tiles counts 256-row tiles here

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_o) : "memory");
    if (p.kb1 < p.num_kb) asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a2) : "memory");
    for (int i = 0; i < stages; ++i) {
      mbar_init(full0 + 8u * i, 1);   // leader: its own arrive.expect_tx covers the bytes of both CTAs
      mbar_init(empty0 + 8u * i, 1);  // one multicast commit per stage
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(tfull0 + 8u * i, 1);
      mbar_init(tempty0 + 8u * i, 16);  // leader: 8 epilogue warps of each CTA
    }
    for (int i = 0; i < 4; ++i) {
      mbar_init(rfull0 + 8u * i, 1);
      mbar_init(rempty0 + 8u * i, p.has_res == 2 ? 4 : 8);  // chunk slot: the 4 warps that read it; tile buffer: all 8
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(p.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // both CTAs' barriers exist before anything remote touches them
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");
  pdl_wait();  // everything above overlapped the previous kernel's tail; its results are visible from here on

  if (warp == 0) {
    // ---------------- TMA producer (both CTAs): own 128 A rows, own half of the weight tile ----------------
    const int howo = p.ho * p.wo;
    const uint32_t tx_bytes = 2u * A_STAGE_BYTES + 2u * b_half_bytes;
    uint32_t it = 0, lt = 0;
    long long prof_pw = 0;
    const long long prof_p0 = p.prof ? clock64() : 0;
    const uint32_t chunk_bytes = (uint32_t)BLOCK_M * (uint32_t)p.cw * 2u;
    for (int tile = cluster_id; tile < total_tiles; tile += num_clusters, ++lt) {
      const int tix = p.rev ? total_tiles - 1 - tile : tile;
      const int mt = tix / p.n_tiles, nt = tix - mt * p.n_tiles;
      const int m0 = mt * 2 * BLOCK_M + (int)rank * BLOCK_M, n0 = nt * p.bn;
      const int img = m0 / howo;
      const int rem = m0 - img * howo;
      const int oh = rem / p.wo, ow = rem - oh * p.wo;
      const int w_base = ow * p.stride - p.pad_w, h_base = oh * p.stride - p.pad_h;
      int cb = 0, r = 0, sx = 0;
      for (int j0 = 0; j0 < p.num_kb; j0 += G, ++it) {
        const int cnt = p.num_kb - j0 < G ? p.num_kb - j0 : G;
        const uint32_t s = it % (uint32_t)stages;
        const uint32_t ph = (it / (uint32_t)stages) & 1u;
        const long long w0 = p.prof ? clock64() : 0;
        mbar_wait_spin(empty0 + 8u * s, ph ^ 1u);
        if (p.prof) prof_pw += clock64() - w0;
        if (leader && elect_one()) mbar_arrive_expect_tx(full0 + 8u * s, (uint32_t)cnt * tx_bytes);
        __syncwarp();
        for (int g = 0; g < cnt; ++g) {
          {
            const uint32_t sub = s * (uint32_t)G + (uint32_t)g;
            const int j = j0 + g;
            if (j >= p.kb1) {
              // second GEMM source (the folded projection shortcut): 1x1 convolution of stride stride2 over another tensor
              if (p.a2_tiled)
                tma2_load_2d_elect(sA + sub * A_STAGE_BYTES, &map_a2, full0 + 8u * s, (j - p.kb1) * BLOCK_K, p.in2_row0 + m0);
              else
                tma2_load_im2col_4d_elect(sA + sub * A_STAGE_BYTES, &map_a2, full0 + 8u * s, (j - p.kb1) * BLOCK_K, ow * p.stride2,
                                          oh * p.stride2, img + p.in_img0, (uint16_t)0, (uint16_t)0);
            } else if (p.a_tiled)
              tma2_load_2d_elect(sA + sub * A_STAGE_BYTES, &map_a, full0 + 8u * s, cb * BLOCK_K, p.in_row0 + m0);
            else
              tma2_load_im2col_4d_elect(sA + sub * A_STAGE_BYTES, &map_a, full0 + 8u * s, cb * BLOCK_K, w_base, h_base,
                                        img + p.in_img0, (uint16_t)sx, (uint16_t)r);
            tma2_load_2d_elect(sB + sub * b_half_bytes, &map_b, full0 + 8u * s, (j0 + g) * BLOCK_K, n0 + (int)rank * (p.bn >> 1));
          }
          if (++cb == p.cpb) {
            cb = 0;
            if (++sx == p.kw) { sx = 0; ++r; }
          }
        }
      }
    }
    if (p.prof && leader && lane == 0) {
      g_tc2_prof[cluster_id * 8 + 0] = prof_pw;
      g_tc2_prof[cluster_id * 8 + 1] = clock64() - prof_p0;
    }
  } else if (warp == 1) {
    if (leader) {
      // ---------------- MMA issuer (leader CTA only; warp-uniform loop, one elected lane issues) ----------------
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.bn >> 3) << 17) | ((uint32_t)((2 * BLOCK_M) >> 4) << 24);
      uint32_t it = 0, lt = 0;
      const uint32_t desc_hi = (uint32_t)(make_sw128_desc(0) >> 32);
      const uint32_t a_lo0 = (sA >> 4) | (1u << 16), b_lo0 = (sB >> 4) | (1u << 16);  // low word: address >> 4, LBO field = 1
      long long prof_te = 0, prof_full = 0;
      const long long prof_m0 = p.prof ? clock64() : 0;
      for (int tile = cluster_id; tile < total_tiles; tile += num_clusters, ++lt) {
        const uint32_t a = lt & 1u, aph = (lt >> 1) & 1u;
        const long long w1 = p.prof ? clock64() : 0;
        mbar_wait_spin(tempty0 + 8u * a, aph ^ 1u);  // both CTAs' epilogues have drained this accumulator
        if (p.prof) prof_te += clock64() - w1;
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + a * (uint32_t)p.acc_stride;
        int cb = 0;
        for (int j0 = 0; j0 < p.num_kb; j0 += G, ++it) {
          const int cnt = p.num_kb - j0 < G ? p.num_kb - j0 : G;
          const uint32_t s = it % (uint32_t)stages;
          const uint32_t ph = (it / (uint32_t)stages) & 1u;
          const long long w2 = p.prof ? clock64() : 0;
          mbar_wait_spin(full0 + 8u * s, ph);
          if (p.prof) prof_full += clock64() - w2;
          tc_fence_after();
          if (p.last_kmma == BLOCK_K / 16) {
            // every k-block carries four K steps (Cin is a multiple of 64): lean, branch-free issue
            for (int g = 0; g < cnt; ++g) {
              const uint32_t sub = s * (uint32_t)G + (uint32_t)g;
              umma2_bf16_x4_elect(d_tmem, a_lo0 + sub * (A_STAGE_BYTES >> 4), b_lo0 + sub * (b_half_bytes >> 4), desc_hi, idesc,
                                  (uint32_t)((j0 | g) != 0));
            }
            umma2_commit_both_elect(empty0 + 8u * s);
            continue;
          }
          for (int g = 0; g < cnt; ++g) {
            const uint32_t sub = s * (uint32_t)G + (uint32_t)g;
            const uint64_t da = make_sw128_desc(sA + sub * A_STAGE_BYTES);
            const uint64_t db = make_sw128_desc(sB + sub * b_half_bytes);
            const int nk = (cb == p.cpb - 1) ? p.last_kmma : BLOCK_K / 16;
            if (elect_one()) {
#pragma unroll
              for (int k = 0; k < BLOCK_K / 16; ++k)
                if (k < nk) umma2_bf16(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)((j0 | g | k) != 0));
            }
            __syncwarp();
            if (++cb == p.cpb) cb = 0;
          }
          if (elect_one()) umma2_commit_both(empty0 + 8u * s);  // frees the stage in BOTH CTAs
          __syncwarp();
        }
        if (elect_one()) umma2_commit_both(tfull0 + 8u * a);
        __syncwarp();
      }
      if (p.prof && lane == 0) {
        g_tc2_prof[cluster_id * 8 + 2] = prof_te;
        g_tc2_prof[cluster_id * 8 + 3] = prof_full;
        g_tc2_prof[cluster_id * 8 + 4] = clock64() - prof_m0;
        g_tc2_prof[cluster_id * 8 + 7] = lt;
      }
    }
  } else if (warp == 10) {
    // ---------------- residual producer (both CTAs): its own warp, so that waiting for the epilogue to hand a residual slot
    // back never holds up the operand stream of the next tile (with few k-blocks per tile the operand ring cannot run a whole
    // tile ahead, and the MMA warp starved behind these waits) ----------------
    if (p.has_res) {
      const uint32_t chunk_bytes = (uint32_t)BLOCK_M * (uint32_t)p.cw * 2u;
      uint32_t lt = 0;
      for (int tile = cluster_id; tile < total_tiles; tile += num_clusters, ++lt) {
        const int tix = p.rev ? total_tiles - 1 - tile : tile;
        const int mt = tix / p.n_tiles, nt = tix - mt * p.n_tiles;
        const int m0 = mt * 2 * BLOCK_M + (int)rank * BLOCK_M, n0 = nt * p.bn;
        if (p.has_res == 1) {
          const uint32_t rb = lt & 1u, rph = (lt >> 1) & 1u;
          mbar_wait_spin(rempty0 + 8u * rb, rph ^ 1u);
          if (elect_one()) {
            mbar_arrive_expect_tx(rfull0 + 8u * rb, res_tile_bytes);
            for (int c = 0, ci = 0; c < p.bn; c += p.cw, ++ci)
              tma_load_2d(sR + rb * res_tile_bytes + (uint32_t)ci * chunk_bytes, &map_r, rfull0 + 8u * rb, n0 + c, p.out_row0 + m0);
          }
          __syncwarp();
        } else {
          // single residual buffer: chunk slot c is refilled as soon as the previous tile's epilogue has read it
          const uint32_t rph = lt & 1u;
          for (int c = 0, ci = 0; c < p.bn; c += p.cw, ++ci) {
            mbar_wait_spin(rempty0 + 8u * ci, rph ^ 1u);
            if (elect_one()) {
              mbar_arrive_expect_tx(rfull0 + 8u * ci, chunk_bytes);
              tma_load_2d(sR + (uint32_t)ci * chunk_bytes, &map_r, rfull0 + 8u * ci, n0 + c, p.out_row0 + m0);
            }
            __syncwarp();
          }
        }
      }
    }
  } else {
    // ---------------- epilogue warps (both CTAs, own 128 rows) ----------------
    // 8 warps: TMEM lane quarter q = warp % 4 (hardware rule), and the two warps of a quarter split the column chunks
    const int q = warp & 3;
    const int ew = warp - 2, half = ew >> 2;
    const uint32_t my_out = sO + (uint32_t)ew * 2u * OUT_BUF_BYTES;
    const int cw = p.cw;
    const uint32_t row_bytes = (uint32_t)cw * 2u;
    const uint32_t chunk_bytes = (uint32_t)BLOCK_M * row_bytes;
    const uint32_t swz = cw == 64 ? (uint32_t)(lane & 7) : (cw == 32 ? (uint32_t)((lane >> 1) & 3) : (uint32_t)((lane >> 2) & 1));
    uint32_t lt = 0, nstore = 0;
    for (int tile = cluster_id; tile < total_tiles; tile += num_clusters, ++lt) {
      const int tix = p.rev ? total_tiles - 1 - tile : tile;
      const int mt = tix / p.n_tiles, nt = tix - mt * p.n_tiles;
      const int m0 = mt * 2 * BLOCK_M + (int)rank * BLOCK_M, n0 = nt * p.bn;
      const uint32_t a = lt & 1u, aph = (lt >> 1) & 1u;
      if (p.has_res == 1) mbar_wait(rfull0 + 8u * a, aph);
      mbar_wait(tfull0 + 8u * a, aph);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + a * (uint32_t)p.acc_stride;
      const uint32_t r_row = sR + (p.has_res == 2 ? 0u : a * res_tile_bytes) + (uint32_t)(q * 32 + lane) * row_bytes;
      int ci = half;
      for (int c0 = half * cw; c0 < p.bn; c0 += 2 * cw, ++nstore, ci += 2) {
        const uint32_t sbuf = my_out + (nstore & 1u) * OUT_BUF_BYTES;
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        __syncwarp();
        const uint32_t rowaddr = sbuf + (uint32_t)lane * row_bytes;
        if (p.has_res == 2) mbar_wait(rfull0 + 8u * (uint32_t)ci, lt & 1u);  // this chunk of the residual tile has landed
        for (int g = 0; g < cw; g += 32) {
          uint32_t v[2][16];
          const bool two = g + 16 < cw;
          tmem_ld16_nowait(t_row + (uint32_t)(c0 + g), v[0]);
          if (two) tmem_ld16_nowait(t_row + (uint32_t)(c0 + g + 16), v[1]);
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            if (h == 1 && !two) break;
            const int gg = g + 16 * h;
            const float4* bp = reinterpret_cast<const float4*>(p.bias + n0 + c0 + gg);
            float f[16];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              float4 b4 = __ldg(bp + i);
              f[4 * i + 0] = __uint_as_float(v[h][4 * i + 0]) + b4.x;
              f[4 * i + 1] = __uint_as_float(v[h][4 * i + 1]) + b4.y;
              f[4 * i + 2] = __uint_as_float(v[h][4 * i + 2]) + b4.z;
              f[4 * i + 3] = __uint_as_float(v[h][4 * i + 3]) + b4.w;
            }
            const uint32_t j0 = (uint32_t)(gg >> 3);
            if (p.has_res) {
              uint32_t rr[8];
              const uint32_t ra = r_row + (uint32_t)ci * chunk_bytes;
              asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                           : "=r"(rr[0]), "=r"(rr[1]), "=r"(rr[2]), "=r"(rr[3])
                           : "r"(ra + ((j0 ^ swz) << 4)));
              asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                           : "=r"(rr[4]), "=r"(rr[5]), "=r"(rr[6]), "=r"(rr[7])
                           : "r"(ra + (((j0 + 1u) ^ swz) << 4)));
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                f[2 * i] += __uint_as_float(rr[i] << 16);
                f[2 * i + 1] += __uint_as_float(rr[i] & 0xffff0000u);
              }
            }
            if (p.relu && n0 + c0 + gg >= p.norelu_cols) {
#pragma unroll
              for (int i = 0; i < 16; ++i) f[i] = fmaxf(f[i], 0.f);
            }
            uint32_t pk[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              __nv_bfloat162 h2 = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
              pk[i] = *reinterpret_cast<uint32_t*>(&h2);
            }
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(rowaddr + ((j0 ^ swz) << 4)), "r"(pk[0]),
                         "r"(pk[1]), "r"(pk[2]), "r"(pk[3])
                         : "memory");
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(rowaddr + (((j0 + 1u) ^ swz) << 4)), "r"(pk[4]),
                         "r"(pk[5]), "r"(pk[6]), "r"(pk[7])
                         : "memory");
          }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(&map_o, sbuf, n0 + c0, p.out_row0 + m0 + q * 32);
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          if (p.has_res == 2) mbar_arrive(rempty0 + 8u * (uint32_t)ci);  // this warp is done with its rows of chunk ci
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive_leader(tempty0 + 8u * a);  // the leader's MMA warp waits for both CTAs
        if (p.has_res == 1) mbar_arrive(rempty0 + 8u * a);
      }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // the peer may still be reading/arriving on this CTA's memory
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
  }
}


// ------------- v6x: the resident-filter halo kernel on CTA PAIRS (cta_group::2, M = 256) -------------
// For layers whose filter fits in the shared memory of TWO SMs: each CTA of a pair keeps HALF of the filter's
// output channels resident and stages the halo of its own 128-position tile; the leader issues M = 256 MMAs over
// both.  The two tiles of a pair must start at the same shared-memory offset (one descriptor serves both CTAs), so
// tiles are ROW ALIGNED here: rpt = floor(128 / Wp) whole padded rows per CTA, first position at the start of the
// box; the MMA rows past rpt * Wp compute garbage that is never stored.
struct TcV6xParams {
  const float* bias;
  __nv_bfloat16* out;
  int n_img, ho, wo, wp, out_pitch;
  int kh, kw, pad_h, pad_w, cblocks;
  int rpt, ptiles_per_img;       // output rows per CTA tile; pair tiles (2 * rpt rows) per image
  int rows_box, a_cb_bytes, a_stage_bytes, a_stages;
  int bn, n_tiles, relu, tmem_cols, acc_stride, n_acc;
  int in_img0, out_img0;
  int row_bytes, last_kmma, cin_pad, wide_store;
  float inv_wp;
  int rev;  // walk the tiles from the last to the first (ConvParams::rev)
};

__device__ __forceinline__ void tma2_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst),
      "l"(map), "r"(bar & PEER_BIT_MASK), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

template <int KH, int KW, int CBLOCKS, int ROW_BYTES>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(TC6_THREADS, 1)
    conv_tc6x_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, const TcV6xParams p) {
  pdl_launch_dependents();
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int taps = p.kh * p.kw;
  const uint32_t row_bytes = (uint32_t)p.row_bytes;
  const uint32_t b_tile_bytes = (uint32_t)(p.bn >> 1) * row_bytes;       // this CTA's half of the N tile
  const uint32_t sB = base;
  const uint32_t sA = sB + (((uint32_t)(taps * p.cblocks) * b_tile_bytes + 1023u) & ~1023u);
  const uint32_t sO = sA + (uint32_t)p.a_stages * (uint32_t)p.a_stage_bytes;  // bn biases (fp32)
  const uint32_t bars = (sO + (uint32_t)p.bn * 4u + 15u) & ~15u;
  const uint32_t afull0 = bars, aempty0 = afull0 + 8u * p.a_stages;
  const uint32_t bfull = aempty0 + 8u * p.a_stages;
  const uint32_t tfull0 = bfull + 8u, tempty0 = tfull0 + 32u;
  const uint32_t tmem_slot = tempty0 + 32u;
  const uint32_t n_acc = (uint32_t)p.n_acc;
  const uint32_t acc_shift = n_acc == 4u ? 2u : 1u;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cl = (int)blockIdx.x >> 1, ncl = (int)gridDim.x >> 1;
  const int vcl = cl / p.n_tiles, vn = ncl / p.n_tiles;        // Cout slice cl % n_tiles stays resident in this pair
  const int n0 = (cl - vcl * p.n_tiles) * p.bn;
  const int total_tiles = p.n_img * p.ptiles_per_img;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    for (int i = 0; i < p.a_stages; ++i) { mbar_init(afull0 + 8u * i, 1); mbar_init(aempty0 + 8u * i, 1); }
    mbar_init(bfull, 1);
    for (int i = 0; i < 4; ++i) { mbar_init(tfull0 + 8u * i, 1); mbar_init(tempty0 + 8u * i, 8); }  // leader: 4 warps of each CTA
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(p.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // both CTAs' barriers exist before anything remote touches them
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");
  pdl_wait();

  if (warp == 0) {
    // ---------------- producer (both CTAs): own filter half once, then the halo of the own tile ----------------
    const int kb = p.row_bytes >> 1;
    if (elect_one()) {
      if (leader) mbar_arrive_expect_tx(bfull, 2u * (uint32_t)(taps * p.cblocks) * b_tile_bytes);
      for (int t = 0; t < taps; ++t)
        for (int cb = 0; cb < p.cblocks; ++cb)
          tma2_load_2d(sB + (uint32_t)(t * p.cblocks + cb) * b_tile_bytes, &map_b, bfull, t * p.cin_pad + cb * kb,
                       n0 + (int)rank * (p.bn >> 1));
    }
    __syncwarp();
    uint32_t ita = 0, sa = 0, pha = 0;
    const uint32_t box_bytes = (uint32_t)p.rows_box * (uint32_t)p.wp * row_bytes;
    TileWalk tw(p.ptiles_per_img, vcl, vn, total_tiles, p.rev);
    for (int tile = vcl; tile < total_tiles; tile += vn, ++ita, tw.next(), pha ^= (++sa == (uint32_t)p.a_stages), sa = sa == (uint32_t)p.a_stages ? 0u : sa) {
      const int row0 = (tw.mt * 2 + (int)rank) * p.rpt;  // first output row of this CTA's tile (may be past the image: zero fill)
      mbar_wait(aempty0 + 8u * sa, pha ^ 1u);
      if (elect_one()) {
        if (leader) mbar_arrive_expect_tx(afull0 + 8u * sa, 2u * (uint32_t)p.cblocks * box_bytes);
        for (int cb = 0; cb < p.cblocks; ++cb)
          tma2_load_4d(sA + sa * (uint32_t)p.a_stage_bytes + (uint32_t)cb * (uint32_t)p.a_cb_bytes, &map_a, afull0 + 8u * sa, cb * kb,
                       -p.pad_w, row0 - p.pad_h, tw.img + p.in_img0);
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    if (leader) {
      // ---------------- MMA issuer (leader only): no waits inside a tile ----------------
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.bn >> 3) << 17) | ((uint32_t)((2 * BLOCK_M) >> 4) << 24);
      const int ksteps = p.row_bytes >> 5;
      mbar_wait(bfull, 0);
      uint32_t ita = 0, sa = 0, pha = 0;
      for (int tile = vcl; tile < total_tiles; tile += vn, ++ita, pha ^= (++sa == (uint32_t)p.a_stages), sa = sa == (uint32_t)p.a_stages ? 0u : sa) {
        const uint32_t a = ita & (n_acc - 1u), aph = (ita >> acc_shift) & 1u;
        mbar_wait(tempty0 + 8u * a, aph ^ 1u);
        mbar_wait(afull0 + 8u * sa, pha);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + a * (uint32_t)p.acc_stride;
        const uint32_t a_base = sA + sa * (uint32_t)p.a_stage_bytes;
        if (KH > 0) {
          const uint64_t hi = make_sw_desc(0, (uint32_t)ROW_BYTES);
          const uint32_t a_lo = a_base >> 4, b_lo = sB >> 4;
          const uint32_t wp16 = ((uint32_t)p.wp * (uint32_t)ROW_BYTES) >> 4;
          const uint32_t acb16 = (uint32_t)p.a_cb_bytes >> 4, bt16 = b_tile_bytes >> 4;
          const int nk_last = p.last_kmma;
          if (elect_one()) {
#pragma unroll
            for (int t = 0; t < KH * KW; ++t) {
#pragma unroll
              for (int cb = 0; cb < CBLOCKS; ++cb) {
                const uint32_t ao = a_lo + (uint32_t)(t / KW) * wp16 + (uint32_t)((t % KW) * (ROW_BYTES >> 4)) + (uint32_t)cb * acb16;
                const uint32_t bo = b_lo + (uint32_t)(t * CBLOCKS + cb) * bt16;
#pragma unroll
                for (int k = 0; k < ROW_BYTES / 32; ++k)
                  if (cb < CBLOCKS - 1 || k < nk_last)
                    umma2_bf16(d_tmem, hi | (uint64_t)((ao + 2u * k) & 0x3FFFu), hi | (uint64_t)((bo + 2u * k) & 0x3FFFu), idesc,
                               (uint32_t)((t | cb | k) != 0));
              }
            }
            umma2_commit_both(aempty0 + 8u * sa);
            umma2_commit_both(tfull0 + 8u * a);
          }
          __syncwarp();
          continue;
        }
        int r = 0, sx = 0;
        for (int t = 0; t < taps; ++t) {
          for (int cb = 0; cb < p.cblocks; ++cb) {
            const uint64_t da = make_sw_desc(a_base + (uint32_t)cb * (uint32_t)p.a_cb_bytes + (uint32_t)(r * p.wp + sx) * row_bytes, row_bytes);
            const uint64_t db = make_sw_desc(sB + (uint32_t)(t * p.cblocks + cb) * b_tile_bytes, row_bytes);
            const int nk = (cb == p.cblocks - 1) ? p.last_kmma : ksteps;
            if (elect_one()) {
#pragma unroll
              for (int k = 0; k < 4; ++k)
                if (k < nk) umma2_bf16(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)((t | cb | k) != 0));
            }
            __syncwarp();
          }
          if (++sx == p.kw) { sx = 0; ++r; }
        }
        if (elect_one()) {
          umma2_commit_both(aempty0 + 8u * sa);
          umma2_commit_both(tfull0 + 8u * a);
        }
        __syncwarp();
      }
    }
  } else {
    // ---------------- epilogue (both CTAs, own 128 rows): two groups of 4 warps drain alternate tiles ----------------
    const int q = warp & 3;
    const int grp = (warp - 2) >> 2;
    const int et = q * 32 + lane;  // tile row = position et of the row-aligned tile
    const uint32_t sbias = sO;
    for (int c = (int)threadIdx.x - 64; c < p.bn; c += 256) {
      const float bv = __ldg(p.bias + n0 + c);
      asm volatile("st.shared.f32 [%0], %1;" ::"r"(sbias + (uint32_t)c * 4u), "f"(bv) : "memory");
    }
    named_bar_sync(1, 256);
    const int trow = div_wp(et, p.inv_wp), tcol = et - trow * p.wp;  // the same for every tile
    uint32_t lt = (uint32_t)grp;
    TileWalk tw(p.ptiles_per_img, vcl + grp * vn, 2 * vn, total_tiles, p.rev);
    for (int tile = vcl + grp * vn; tile < total_tiles; tile += 2 * vn, lt += 2, tw.next()) {
      const uint32_t a = lt & (n_acc - 1u), aph = (lt >> acc_shift) & 1u;
      const int orow = (tw.mt * 2 + (int)rank) * p.rpt + trow;
      const bool valid = trow < p.rpt && orow < p.ho && tcol < p.wo;
      const long long off = ((long long)((tw.img + p.out_img0) * p.ho + orow) * p.wo + tcol) * p.out_pitch + n0;
      mbar_wait(tfull0 + 8u * a, aph);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + a * (uint32_t)p.acc_stride;
      for (int g = 0; g < p.bn; g += 32) {
        uint32_t v[2][16];
        const bool two = g + 16 < p.bn;
        tmem_ld16_nowait(t_row + (uint32_t)g, v[0]);
        if (two) tmem_ld16_nowait(t_row + (uint32_t)(g + 16), v[1]);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          if (h == 1 && !two) break;
          const int gg = g + 16 * h;
          uint32_t pk[8];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float4 b4;
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                         : "=f"(b4.x), "=f"(b4.y), "=f"(b4.z), "=f"(b4.w)
                         : "r"(sbias + (uint32_t)(gg + 4 * i) * 4u));
            float f0 = __uint_as_float(v[h][4 * i + 0]) + b4.x, f1 = __uint_as_float(v[h][4 * i + 1]) + b4.y;
            float f2 = __uint_as_float(v[h][4 * i + 2]) + b4.z, f3 = __uint_as_float(v[h][4 * i + 3]) + b4.w;
            if (p.relu) { f0 = fmaxf(f0, 0.f); f1 = fmaxf(f1, 0.f); f2 = fmaxf(f2, 0.f); f3 = fmaxf(f3, 0.f); }
            __nv_bfloat162 lo = __floats2bfloat162_rn(f0, f1), hi = __floats2bfloat162_rn(f2, f3);
            pk[2 * i] = *reinterpret_cast<uint32_t*>(&lo);
            pk[2 * i + 1] = *reinterpret_cast<uint32_t*>(&hi);
          }
          if (valid) {
            if (p.wide_store) {
              asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p.out + off + gg), "r"(pk[0]), "r"(pk[1]),
                           "r"(pk[2]), "r"(pk[3]), "r"(pk[4]), "r"(pk[5]), "r"(pk[6]), "r"(pk[7])
                           : "memory");
            } else {
              uint4* dst = reinterpret_cast<uint4*>(p.out + off + gg);
              dst[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
              dst[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_leader(tempty0 + 8u * a);  // the leader's MMA warp waits for both CTAs
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // the peer may still be reading this CTA's operands / arriving on its barriers
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
  }
}


#define DML_V6X_SHAPES(X) X(3, 3, 1, 128) X(3, 3, 2, 128) X(5, 5, 1, 128) X(1, 7, 2, 128) X(7, 1, 2, 128) X(1, 7, 3, 128) X(7, 1, 3, 128) \
    X(3, 3, 3, 64) X(1, 7, 5, 64) X(7, 1, 5, 64)

// ------------- v5: "weights as A" for layers with few output channels -------------
// A tcgen05.mma takes ~190 cycles on B200 whatever its N (measured), so with pixels on M (128 per
// instruction) a Cout <= 128 layer can use at most N/256 of the tensor pipe.  Here the roles are
// swapped: A = the weight tile (Cout rows, zero-filled up to 128 by TMA), B = 256 pixels of the
// im2col matrix, D[cout][pixel] in TMEM: twice the pixels per instruction.  The epilogue
// transposes through shared memory (thread = one channel, 2-byte stores side by side) and one TMA
// store writes the [256 pixels][Cout] tile.  No residual input (the layers it targets have none).
struct TcV5Params {
  const float* bias;
  int M, ho, wo;
  int stride, pad_h, pad_w;
  int kw, num_kb, cpb, last_kmma;
  int cout, box_cols, relu;
  int stages;
  int p_tiles, c_tiles;
  int out_row0, in_img0, a_tiled, in_row0;
};

constexpr int V5_PIX = 256;
constexpr int V5_B_STAGE_BYTES = V5_PIX * 128;

__global__ void __launch_bounds__(TC_THREADS, 1) conv_tc5_kernel(const __grid_constant__ CUtensorMap map_w,
                                                                 const __grid_constant__ CUtensorMap map_x,
                                                                 const __grid_constant__ CUtensorMap map_o,
                                                                 const TcV5Params p) {
  pdl_launch_dependents();
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int stages = p.stages;
  const uint32_t sW = base;                                         // weight tiles: 128 rows x 128 B
  const uint32_t sX = sW + (uint32_t)stages * A_STAGE_BYTES;        // pixel tiles: 256 rows x 128 B
  const uint32_t sO = sX + (uint32_t)stages * V5_B_STAGE_BYTES;     // [256 pixels][box_cols] bf16
  const uint32_t out_row_bytes = (uint32_t)p.box_cols * 2u;
  const uint32_t bars = (sO + (uint32_t)V5_PIX * out_row_bytes + 15u) & ~15u;
  const uint32_t full0 = bars, empty0 = bars + 8u * stages;
  const uint32_t tfull0 = bars + 16u * stages, tempty0 = tfull0 + 16u;
  const uint32_t tmem_slot = tempty0 + 16u;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_tiles = p.p_tiles * p.c_tiles;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_w) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_x) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_o) : "memory");
    for (int i = 0; i < stages; ++i) {
      mbar_init(full0 + 8u * i, 1);
      mbar_init(empty0 + 8u * i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(tfull0 + 8u * i, 1);
      mbar_init(tempty0 + 8u * i, 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");
  pdl_wait();  // everything above overlapped the previous kernel's tail; its results are visible from here on

  if (warp == 0) {
    // ---------------- TMA producer ----------------
    const int howo = p.ho * p.wo;
    const uint32_t tx_bytes = A_STAGE_BYTES + V5_B_STAGE_BYTES;
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int pt = tile / p.c_tiles, ct = tile - pt * p.c_tiles;
      const int m0 = pt * V5_PIX, c0 = ct * BLOCK_M;
      const int img = m0 / howo;
      const int rem = m0 - img * howo;
      const int oh = rem / p.wo, ow = rem - oh * p.wo;
      const int w_base = ow * p.stride - p.pad_w, h_base = oh * p.stride - p.pad_h;
      int cb = 0, r = 0, sx = 0;
      for (int j = 0; j < p.num_kb; ++j, ++it) {
        const uint32_t s = it % (uint32_t)stages;
        const uint32_t ph = (it / (uint32_t)stages) & 1u;
        mbar_wait_spin(empty0 + 8u * s, ph ^ 1u);
        if (elect_one()) {
          mbar_arrive_expect_tx(full0 + 8u * s, tx_bytes);
          tma_load_2d(sW + s * A_STAGE_BYTES, &map_w, full0 + 8u * s, j * BLOCK_K, c0);  // rows past Cout: zeros
          if (p.a_tiled)
            tma_load_2d(sX + s * V5_B_STAGE_BYTES, &map_x, full0 + 8u * s, cb * BLOCK_K, p.in_row0 + m0);
          else
            tma_load_im2col_4d(sX + s * V5_B_STAGE_BYTES, &map_x, full0 + 8u * s, cb * BLOCK_K, w_base, h_base,
                               img + p.in_img0, (uint16_t)sx, (uint16_t)r);
        }
        __syncwarp();
        if (++cb == p.cpb) {
          cb = 0;
          if (++sx == p.kw) { sx = 0; ++r; }
        }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer: D[cout 128][pixel 256] += W[128 x 16] * X[256 x 16]^T ----------------
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(V5_PIX >> 3) << 17) | ((uint32_t)(BLOCK_M >> 4) << 24);
    uint32_t it = 0, lt = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++lt) {
      const uint32_t a = lt & 1u, aph = (lt >> 1) & 1u;
      mbar_wait_spin(tempty0 + 8u * a, aph ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + a * (uint32_t)V5_PIX;
      int cb = 0;
      for (int j = 0; j < p.num_kb; ++j, ++it) {
        const uint32_t s = it % (uint32_t)stages;
        const uint32_t ph = (it / (uint32_t)stages) & 1u;
        mbar_wait_spin(full0 + 8u * s, ph);
        tc_fence_after();
        const uint64_t dw = make_sw128_desc(sW + s * A_STAGE_BYTES);
        const uint64_t dx = make_sw128_desc(sX + s * V5_B_STAGE_BYTES);
        const int nk = (cb == p.cpb - 1) ? p.last_kmma : BLOCK_K / 16;
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < BLOCK_K / 16; ++k)
            if (k < nk) umma_bf16(d_tmem, dw + (uint64_t)(2 * k), dx + (uint64_t)(2 * k), idesc, (uint32_t)((j | k) != 0));
          umma_commit(empty0 + 8u * s);
        }
        __syncwarp();
        if (++cb == p.cpb) cb = 0;
      }
      if (elect_one()) umma_commit(tfull0 + 8u * a);
      __syncwarp();
    }
  } else {
    // ---------------- epilogue: thread = one output channel (TMEM lane), 256 pixels in its columns ----------------
    const int q = warp & 3;
    const int et = q * 32 + lane;
    uint32_t lt = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++lt) {
      const int pt = tile / p.c_tiles, ct = tile - pt * p.c_tiles;
      const int m0 = pt * V5_PIX, c0 = ct * BLOCK_M;
      const uint32_t a = lt & 1u, aph = (lt >> 1) & 1u;
      const int c = c0 + et;
      const bool c_ok = et < p.box_cols && c < p.cout;
      const float bias = c_ok ? __ldg(p.bias + c) : 0.f;
      mbar_wait(tfull0 + 8u * a, aph);
      tc_fence_after();
      // the previous tile's TMA store must have finished reading the staging buffer
      if (et == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      named_bar_sync(2, 128);
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + a * (uint32_t)V5_PIX;
      const uint32_t my_col = sO + (uint32_t)et * 2u;
      if (q * 32 < p.box_cols) {  // warp-uniform: warps whose 32 channels are all padding only release the accumulator
        for (int g = 0; g < V5_PIX; g += 32) {
          uint32_t v[2][16];
          tmem_ld16_nowait(t_row + (uint32_t)g, v[0]);
          tmem_ld16_nowait(t_row + (uint32_t)(g + 16), v[1]);
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
          if (c_ok) {
#pragma unroll
            for (int h = 0; h < 2; ++h)
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                float f = __uint_as_float(v[h][i]) + bias;
                if (p.relu) f = fmaxf(f, 0.f);
                const __nv_bfloat16 b = __float2bfloat16_rn(f);
                asm volatile("st.shared.b16 [%0], %1;" ::"r"(my_col + (uint32_t)(g + 16 * h + i) * out_row_bytes),
                             "h"(*reinterpret_cast<const uint16_t*>(&b))
                             : "memory");
              }
          }
        }
      }
      tc_fence_before();
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      named_bar_sync(2, 128);
      if (et == 0) {
        tma_store_2d(&map_o, sO, c0, p.out_row0 + m0);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
      if (lane == 0) mbar_arrive(tempty0 + 8u * a);
    }
    if (et == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

#include "conv_chain.inc"
#include "conv_duo.inc"
#include "conv_pool.inc"

}  // namespace

// ------------------------------- host side ---------------------------------
namespace {

PFN_cuTensorMapEncodeTiled g_encode_tiled = nullptr;
PFN_cuTensorMapEncodeIm2col g_encode_im2col = nullptr;
int g_driver_version = 0;

int g_num_sms = 148;
int g_smem_aligned = -1;  // 1: dynamic shared memory starts at a multiple of 1024 bytes (probed in tc_init)

int pick_bn(int cout, int override_bn) {
  if (override_bn > 0) return (cout % override_bn == 0 && override_bn % 16 == 0 && override_bn <= 256) ? override_bn : 0;
  if (cout <= 128) return cout;
  for (int bn = 128; bn >= 16; bn -= 16)
    if (cout % bn == 0) return bn;
  return 0;
}

// Persistent kernel: the widest N tile that still gives every SM a tile.
int pick_bn_v2(int cout, long long m_tiles, int override_bn) {
  if (override_bn > 0) return pick_bn(cout, override_bn);
  int best = 0;
  long long best_tiles = -1;
  for (int bn = cout < 256 ? cout : 256; bn >= 16; bn -= 16) {
    if (cout % bn) continue;
    long long tiles = m_tiles * (cout / bn);
    if (tiles >= g_num_sms) return bn;
    if (bn >= 64 || best == 0) {
      if (tiles > best_tiles) { best_tiles = tiles; best = bn; }
    }
  }
  return best;
}

}  // namespace

const char* tc_last_error() { return g_tc_err.c_str(); }

bool tc_supported(const Op& op) {
  return op.kind == OP_CONV && op.cin >= 16 && op.cin % 8 == 0 && op.cout % 16 == 0 && op.cin_pad % BLOCK_K == 0 &&
         pick_bn(op.cout, 0) > 0;
}

int tc_init() {
  if (!g_encode_tiled) {
    cudaDriverEntryPointQueryResult qres;
    void* fn = nullptr;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn)
      return tc_fail("cuTensorMapEncodeTiled not available from the driver (%s)", cudaGetErrorString(e));
    g_encode_tiled = reinterpret_cast<PFN_cuTensorMapEncodeTiled>(fn);
    fn = nullptr;
    e = cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn)
      return tc_fail("cuTensorMapEncodeIm2col not available from the driver (%s)", cudaGetErrorString(e));
    g_encode_im2col = reinterpret_cast<PFN_cuTensorMapEncodeIm2col>(fn);
    cudaDriverGetVersion(&g_driver_version);
  }
  cudaError_t e = cudaSuccess;
#ifdef DML_LEGACY_VARIANTS
  e = cudaFuncSetAttribute(conv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  if (e != cudaSuccess) return tc_fail("cudaFuncSetAttribute(conv_tc_kernel, smem %d): %s", SMEM_LIMIT, cudaGetErrorString(e));
  e = cudaFuncSetAttribute(conv_tc3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  if (e != cudaSuccess) return tc_fail("cudaFuncSetAttribute(conv_tc3_kernel, smem %d): %s", SMEM_LIMIT, cudaGetErrorString(e));
#endif
  e = cudaFuncSetAttribute(conv_tc5_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  if (e != cudaSuccess) return tc_fail("cudaFuncSetAttribute(conv_tc5_kernel, smem %d): %s", SMEM_LIMIT, cudaGetErrorString(e));
  e = cudaSuccess;
#define DML_V6X_ATTR(KH_, KW_, CB_, RB_) \
  if (e == cudaSuccess) e = cudaFuncSetAttribute(conv_tc6x_kernel<KH_, KW_, CB_, RB_>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  DML_V6X_SHAPES(DML_V6X_ATTR) DML_V6X_ATTR(0, 0, 0, 0)
#undef DML_V6X_ATTR
  if (e != cudaSuccess) return tc_fail("cudaFuncSetAttribute(conv_tc6x_kernel, smem %d): %s", SMEM_LIMIT, cudaGetErrorString(e));
#define DML_V6_SHAPES_INIT(X) X(3, 3, 1, 128) X(3, 3, 1, 64) X(3, 3, 2, 128) X(4, 4, 1, 32) X(2, 2, 1, 32) X(5, 5, 1, 128) \
    X(1, 7, 2, 128) X(7, 1, 2, 128) X(1, 7, 3, 128) X(7, 1, 3, 128) X(1, 3, 6, 128) X(3, 1, 6, 128) X(3, 3, 3, 64) X(1, 7, 5, 64) X(7, 1, 5, 64) \
    X(0, 0, 0, 0)
  e = cudaSuccess;
#define DML_V6_ATTR(KH_, KW_, CB_, RB_) \
  if (e == cudaSuccess) e = cudaFuncSetAttribute(conv_tc6_kernel<KH_, KW_, CB_, RB_>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  DML_V6_SHAPES_INIT(DML_V6_ATTR)
#undef DML_V6_ATTR
  if (e != cudaSuccess) return tc_fail("cudaFuncSetAttribute(conv_tc6_kernel, smem %d): %s", SMEM_LIMIT, cudaGetErrorString(e));
  e = cudaFuncSetAttribute(conv_chain_kernel<true, false, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(conv_chain_kernel<false, true, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(conv_chain_kernel<true, false, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(conv_chain_kernel<true, false, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(conv_chain_kernel<false, true, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(conv_chain_kernel<true, false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  if (e != cudaSuccess) return tc_fail("cudaFuncSetAttribute(conv_chain_kernel, smem %d): %s", SMEM_LIMIT, cudaGetErrorString(e));
  if (e == cudaSuccess) e = cudaFuncSetAttribute(conv_tc6p_kernel<4, 4, 1, 32>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(conv_tc6p_kernel<3, 3, 1, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(conv_duo_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(conv_duo_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(conv_duo_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(conv_duo_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  if (e != cudaSuccess) return tc_fail("cudaFuncSetAttribute(conv_duo_kernel) failed: %s", cudaGetErrorString(e));
  e = cudaFuncSetAttribute(conv_tc2x_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  if (e != cudaSuccess) return tc_fail("cudaFuncSetAttribute(conv_tc2x_kernel, smem %d): %s", SMEM_LIMIT, cudaGetErrorString(e));
  e = cudaFuncSetAttribute(conv_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
  if (e != cudaSuccess) return tc_fail("cudaFuncSetAttribute(conv_tc2_kernel, smem %d): %s", SMEM_LIMIT, cudaGetErrorString(e));
  int dev = 0, sms = 0;
  if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && sms > 0)
    g_num_sms = sms;
  if (g_smem_aligned < 0) {  // does dynamic shared memory start 1 KiB-aligned?  (duo_plan: the conv3_block1 shape fits only without slack)
    int mod = -1;
    smem_base_probe_kernel<<<1, 32, 2048>>>();
    if (cudaMemcpyFromSymbol(&mod, g_smem_base_mod, sizeof mod) == cudaSuccess) g_smem_aligned = mod == 0 ? 1 : 0;
    else (void)cudaGetLastError();
  }
  return 0;
}

size_t tc_scratch_bytes(const Net&, int) { return 0; }

int tc_make_tiled_map(CUtensorMap* out, void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box, int swizzle_bytes) {
  if (!g_encode_tiled && tc_init() != 0) return -1;
  cuuint64_t d[5], st[4];
  cuuint32_t b[5], es[5];
  for (int i = 0; i < rank; ++i) { d[i] = dims[i]; b[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) st[i] = strides_bytes[i];
  CUtensorMapSwizzle sw = swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                          : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                          : swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE;
  CUresult r = g_encode_tiled(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, base, d, st, b, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                              CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return tc_fail("cuTensorMapEncodeTiled failed (%d)", (int)r);
  return 0;
}

int tc_plan_one(TcOpPlan* out, const ConvParams& p, int n_capacity, const TcTune& tune) {
  if (!g_encode_tiled) return tc_fail("tc_init was not called");
#ifndef DML_LEGACY_VARIANTS
  if (tune.version == 1 || tune.version == 3)
    return tc_fail("kernel variant %d is a superseded A/B twin: not in this build (compile with -DDML_LEGACY_VARIANTS)", tune.version);
#endif
  TcOpPlan op;
  // ---- v5: weights as the A operand, 256 pixels per instruction (Cout <= 128, no residual) ----
  if (tune.version == 5) {
    if (p.cout > 128 || p.cout % 8 != 0 || p.res != nullptr || p.out == nullptr || p.cin_pad % BLOCK_K)
      return tc_fail("shape is not eligible for the weights-as-A kernel");
    op.version = 5;
    op.bn = p.cout;
    op.cpb = p.cin_pad / BLOCK_K;
    op.cw = p.cout < 128 ? p.cout : 128;  // TMA store box columns
    const int stage_bytes = A_STAGE_BYTES + V5_B_STAGE_BYTES;
    const int fixed = 1024 + V5_PIX * op.cw * 2 + 256;
    op.stages = (SMEM_LIMIT - fixed) / stage_bytes;
    if (op.stages > 6) op.stages = 6;
    if (tune.stages > 0 && tune.stages < op.stages) op.stages = tune.stages;
    if (op.stages < 2) return tc_fail("not enough shared memory for the weights-as-A kernel");
    op.smem = (size_t)op.stages * stage_bytes + fixed;
    const long long M_cap5 = (long long)n_capacity * p.ho * p.wo;
    op.m_capacity = (int)M_cap5;
    op.a_tiled = p.kh == 1 && p.kw == 1 && p.stride == 1 && p.pad_h == 0 && p.pad_w == 0;
    CUresult r5;
    if (op.a_tiled) {
      cuuint64_t adim[2] = {(cuuint64_t)p.cin, (cuuint64_t)n_capacity * p.hi * p.wi};
      cuuint64_t astr[1] = {(cuuint64_t)p.in_pitch * 2};
      cuuint32_t abox[2] = {BLOCK_K, V5_PIX};
      cuuint32_t aes[2] = {1, 1};
      r5 = g_encode_tiled(&op.map_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(p.in), adim, astr, abox, aes,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    } else {
      cuuint64_t gdim[4] = {(cuuint64_t)p.cin, (cuuint64_t)p.wi, (cuuint64_t)p.hi, (cuuint64_t)n_capacity};
      cuuint64_t gstr[3] = {(cuuint64_t)p.in_pitch * 2, (cuuint64_t)p.wi * p.in_pitch * 2, (cuuint64_t)p.hi * p.wi * p.in_pitch * 2};
      int lower[2] = {-p.pad_w, -p.pad_h};
      int upper[2] = {p.pad_w - (p.kw - 1), p.pad_h - (p.kh - 1)};
      cuuint32_t estr[4] = {1, (cuuint32_t)p.stride, (cuuint32_t)p.stride, 1};
      r5 = g_encode_im2col(&op.map_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(p.in), gdim, gstr, lower, upper,
                           BLOCK_K, V5_PIX, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r5 == CUDA_SUCCESS && g_driver_version <= 13010 && (size_t)n_capacity * p.hi * p.wi * p.in_pitch * 2 < 131072)
        reinterpret_cast<uint64_t*>(&op.map_a)[1] &= ~(1ull << 21);
    }
    if (r5 != CUDA_SUCCESS) return tc_fail("tensor map for the pixel operand failed (%d)", (int)r5);
    const cuuint64_t ktot = (cuuint64_t)p.kh * p.kw * p.cin_pad;
    cuuint64_t bdim[2] = {ktot, (cuuint64_t)p.cout};
    cuuint64_t bstr[1] = {ktot * 2};
    cuuint32_t bbox[2] = {BLOCK_K, (cuuint32_t)BLOCK_M};  // rows past Cout are out of bounds -> zero filled
    cuuint32_t bes[2] = {1, 1};
    r5 = g_encode_tiled(&op.map_b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(p.w), bdim, bstr, bbox, bes,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r5 != CUDA_SUCCESS) return tc_fail("cuTensorMapEncodeTiled(weights as A) failed (%d)", (int)r5);
    cuuint64_t odim[2] = {(cuuint64_t)p.cout, (cuuint64_t)M_cap5};
    cuuint64_t ostr[1] = {(cuuint64_t)p.out_pitch * 2};
    cuuint32_t obox[2] = {(cuuint32_t)op.cw, V5_PIX};
    cuuint32_t oes[2] = {1, 1};
    r5 = g_encode_tiled(&op.map_o, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, p.out, odim, ostr, obox, oes, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r5 != CUDA_SUCCESS) return tc_fail("cuTensorMapEncodeTiled(output, weights-as-A) failed (%d): cout=%d cw=%d", (int)r5, p.cout, op.cw);
    op.valid = true;
    *out = op;
    return 0;
  }
  // ---- v6x (version 7): the resident-filter halo kernel on CTA pairs, row-aligned tiles ----
  if (tune.version == 7) {
    const int wp = p.wo + p.kw - 1;
    // K blocks of 32 channels (64-byte rows) on request: Cin = 80 / 96 / 160 pad to 96 / 96 / 160 channels instead of 128 / 128 / 192
    const int row_bytes = p.cin <= 16 ? 32 : ((p.cin <= 32 || tune.row_bytes == 64) ? 64 : 128);
    const int kb = row_bytes / 2;
    const int cblocks = (p.cin + kb - 1) / kb;
    const int taps = p.kh * p.kw;
    const bool eligible = p.stride == 1 && taps > 1 && p.res == nullptr && wp <= BLOCK_M && p.out != nullptr && p.cout % 16 == 0 &&
                          p.cin % 8 == 0 && (p.in_pitch * 2) % 16 == 0 && (p.out_pitch * 2) % 16 == 0;
    if (!eligible) return tc_fail("shape is not eligible for the pair resident-filter halo kernel");
    TcOpPlan x = op;
    x.version = 7;
    x.wp = wp;
    x.row_bytes = row_bytes;
    x.cpb = cblocks;
    x.rows_box = BLOCK_M / wp;                         // rpt: output rows per CTA tile (kept in rows_box; the box is rpt + kh - 1 rows)
    x.tiles_per_img = (p.ho + 2 * x.rows_box - 1) / (2 * x.rows_box);  // pair tiles per image
    const int box_rows = x.rows_box + p.kh - 1;
    const int reach = BLOCK_M + (p.kh - 1) * wp + p.kw - 1;  // rows the shifted views of a 128-row MMA may touch
    x.a_cb_bytes = ((box_rows * wp > reach ? box_rows * wp : reach) * row_bytes + 1023) / 1024 * 1024;
    x.a_stage_bytes = cblocks * x.a_cb_bytes;
    int fixed = 0;
    x.bn = 0;
    for (int parts = 1; parts <= 4 && x.bn == 0; ++parts) {
      if (p.cout % parts || (p.cout / parts) % 16 || p.cout / parts > 256) continue;
      if (tune.bn > 0 && p.cout / parts != tune.bn) continue;
      const int bn = p.cout / parts;
      const int b_bytes = (taps * cblocks * (bn / 2) * row_bytes + 1023) / 1024 * 1024;
      fixed = 1024 + b_bytes + bn * 4 + 16 + 512;
      if ((SMEM_LIMIT - fixed) / x.a_stage_bytes >= 1) x.bn = bn;
    }
    if (x.bn == 0 || box_rows > 256) return tc_fail("pair resident-filter halo kernel does not fit in shared memory for this shape");
    x.acc_stride = 32;
    while (x.acc_stride < x.bn) x.acc_stride *= 2;
    x.tmem_cols = x.acc_stride <= 128 ? 4 * x.acc_stride : 2 * x.acc_stride;
    x.a_stages = (SMEM_LIMIT - fixed) / x.a_stage_bytes;
    if (x.a_stages > 4) x.a_stages = 4;
    if (tune.stages > 0 && tune.stages < x.a_stages) x.a_stages = tune.stages;
    x.smem = (size_t)fixed + (size_t)x.a_stages * x.a_stage_bytes;
    x.m_capacity = n_capacity;
    const CUtensorMapSwizzle sw = row_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : (row_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
    cuuint64_t gdim[4] = {(cuuint64_t)p.cin, (cuuint64_t)p.wi, (cuuint64_t)p.hi, (cuuint64_t)n_capacity};
    cuuint64_t gstr[3] = {(cuuint64_t)p.in_pitch * 2, (cuuint64_t)p.wi * p.in_pitch * 2, (cuuint64_t)p.hi * p.wi * p.in_pitch * 2};
    cuuint32_t box[4] = {(cuuint32_t)kb, (cuuint32_t)wp, (cuuint32_t)box_rows, 1};
    cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r7 = g_encode_tiled(&x.map_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(p.in), gdim, gstr, box, es,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r7 != CUDA_SUCCESS) return tc_fail("cuTensorMapEncodeTiled(halo A, v6x) failed (%d): wp=%d rows=%d", (int)r7, wp, box_rows);
    const cuuint64_t ktot = (cuuint64_t)taps * p.cin_pad;
    cuuint64_t bdim[2] = {ktot, (cuuint64_t)p.cout};
    cuuint64_t bstr[1] = {ktot * 2};
    cuuint32_t bbox[2] = {(cuuint32_t)kb, (cuuint32_t)(x.bn / 2)};
    cuuint32_t bes[2] = {1, 1};
    r7 = g_encode_tiled(&x.map_b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(p.w), bdim, bstr, bbox, bes,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r7 != CUDA_SUCCESS) return tc_fail("cuTensorMapEncodeTiled(weights, v6x) failed (%d)", (int)r7);
    x.valid = true;
    *out = x;
    return 0;
  }
  // ---- v6: halo kernel with the whole filter resident (few-channel stride-1 multi-tap layers) ----
  if (tune.version == 6 || tune.version == 0) {
    // in_row_cells: the input rows are physically padded to that many pixels and pre-swizzled (16-channel space-to-depth stems)
    if (p.in_row_cells && (p.cin != 16 || p.in_pitch != 16 || p.in_row_cells % 8 || p.in_row_cells < p.wo + p.kw - 1 || p.in_row_cells > 256 * 8))
      return tc_fail("row-padded input needs 16 channels and a row of a multiple of 8 pixels that covers the padded width");
    const int wp = p.in_row_cells ? p.in_row_cells : p.wo + p.kw - 1;
    // K blocks of 32 channels (64-byte rows) on request: Cin = 80 / 96 / 160 pad to 96 / 96 / 160 channels instead of 128 / 128 / 192
    const int row_bytes = p.cin <= 16 ? 32 : ((p.cin <= 32 || tune.row_bytes == 64) ? 64 : 128);
    const int kb = row_bytes / 2;
    const int cblocks = (p.cin + kb - 1) / kb;
    const int taps = p.kh * p.kw;
    const bool eligible = p.stride == 1 && taps > 1 && p.res == nullptr && wp <= 256 && p.out != nullptr && p.cout % 16 == 0 &&
                          p.cout <= 256 && p.cin % 8 == 0 && (p.in_pitch * 2) % 16 == 0 && (p.out_pitch * 2) % 16 == 0;
    if (tune.version == 6 && !eligible) return tc_fail("shape is not eligible for the resident-filter halo kernel");
    if (eligible) {
      TcOpPlan v6 = op;
      v6.version = 6;
      v6.wp = wp;
      v6.row_bytes = row_bytes;
      v6.cpb = cblocks;
      v6.q_valid = (p.ho - 1) * wp + p.wo;
      v6.tiles_per_img = (v6.q_valid + BLOCK_M - 1) / BLOCK_M;
      const int halo = (p.kh - 1) * wp + p.kw - 1;
      v6.rows_box = (wp - 1 + BLOCK_M - 1 + halo) / wp + 1;
      v6.a_cb_bytes = (v6.rows_box * wp * row_bytes + 1023) / 1024 * 1024;
      v6.a_stage_bytes = cblocks * v6.a_cb_bytes;
      // N tile: the whole Cout if its filter fits next to two halo stages, else the largest slice (Cout / 2, / 3, / 4) that does
      int b_bytes = 0, fixed = 0;
      v6.bn = 0;
      for (int parts = 1; parts <= 4 && v6.bn == 0; ++parts) {
        if (p.cout % parts || (p.cout / parts) % 16) continue;
        if (tune.bn > 0 && p.cout / parts != tune.bn) continue;
        const int bn = p.cout / parts;
        b_bytes = (taps * cblocks * bn * row_bytes + 1023) / 1024 * 1024;
        fixed = 1024 + b_bytes + bn * 4 + 16 + 512;
        if ((SMEM_LIMIT - fixed) / v6.a_stage_bytes >= 2) v6.bn = bn;
      }
      if (v6.bn == 0) {
        if (tune.version == 6) return tc_fail("resident-filter halo kernel does not fit in shared memory for this shape");
      } else {
      v6.acc_stride = 32;
      while (v6.acc_stride < v6.bn) v6.acc_stride *= 2;
      v6.tmem_cols = v6.acc_stride <= 128 ? 4 * v6.acc_stride : 2 * v6.acc_stride;  // four accumulators when they fit
      v6.cw = v6.bn % 64 == 0 ? 64 : (v6.bn % 32 == 0 ? 32 : 16);
      v6.a_stages = (SMEM_LIMIT - fixed) / v6.a_stage_bytes;
      if (v6.a_stages > 4) v6.a_stages = 4;
      if (tune.stages > 0 && tune.stages < v6.a_stages) v6.a_stages = tune.stages;
      const double eff = (double)p.ho * p.wo / ((double)v6.tiles_per_img * BLOCK_M);
      const bool fits = v6.rows_box <= 256 && v6.a_stages >= 2 && (long long)v6.tiles_per_img * BLOCK_M < 65536;
      if (tune.version == 6 && !fits) return tc_fail("resident-filter halo kernel does not fit in shared memory for this shape");
      if (fits && (tune.version == 6 || (eff >= 0.70 && !getenv("DML_NO_V6")))) {
        v6.smem = (size_t)fixed + (size_t)v6.a_stages * v6.a_stage_bytes;
        v6.m_capacity = n_capacity;
        const CUtensorMapSwizzle sw = row_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : (row_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
        cuuint64_t gdim[4] = {(cuuint64_t)p.cin, (cuuint64_t)p.wi, (cuuint64_t)p.hi, (cuuint64_t)n_capacity};
        cuuint64_t gstr[3] = {(cuuint64_t)p.in_pitch * 2, (cuuint64_t)p.wi * p.in_pitch * 2, (cuuint64_t)p.hi * p.wi * p.in_pitch * 2};
        cuuint32_t box[4] = {(cuuint32_t)kb, (cuuint32_t)wp, (cuuint32_t)v6.rows_box, 1};
        cuuint32_t es[4] = {1, 1, 1, 1};
        CUresult r6;
        if (p.in_row_cells) {
          // rows of wp pixels in memory: view (8 pixels x 16 channels = 256 bytes, wp / 8 groups, rows, images); one box brings the
          // whole halo in 256-byte lines, copied as is (the pixels are stored swizzled already)
          cuuint64_t pdim[4] = {128, (cuuint64_t)(wp / 8), (cuuint64_t)p.hi, (cuuint64_t)n_capacity};
          cuuint64_t pstr[3] = {256, (cuuint64_t)wp * 32, (cuuint64_t)p.hi * wp * 32};
          cuuint32_t pbox[4] = {128, (cuuint32_t)(wp / 8), (cuuint32_t)v6.rows_box, 1};
          r6 = g_encode_tiled(&v6.map_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(p.in), pdim, pstr, pbox, es,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        } else
        r6 = g_encode_tiled(&v6.map_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(p.in), gdim, gstr, box, es,
                                     CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r6 != CUDA_SUCCESS) return tc_fail("cuTensorMapEncodeTiled(halo A, v6) failed (%d): wp=%d rows=%d", (int)r6, wp, v6.rows_box);
        const cuuint64_t ktot = (cuuint64_t)taps * p.cin_pad;
        cuuint64_t bdim[2] = {ktot, (cuuint64_t)p.cout};
        cuuint64_t bstr[1] = {ktot * 2};
        cuuint32_t bbox[2] = {(cuuint32_t)kb, (cuuint32_t)v6.bn};
        cuuint32_t bes[2] = {1, 1};
        r6 = g_encode_tiled(&v6.map_b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(p.w), bdim, bstr, bbox, bes,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r6 != CUDA_SUCCESS) return tc_fail("cuTensorMapEncodeTiled(weights, v6) failed (%d)", (int)r6);
        v6.valid = true;
        *out = v6;
        return 0;
      }
      }
    }
  }
  // ---- v3: halo kernel for stride-1 multi-tap convs whose Cin is a multiple of 64 ----
  {
    const bool eligible = p.stride == 1 && p.kh * p.kw > 1 && p.cin % BLOCK_K == 0 && p.cin_pad == p.cin && p.res == nullptr &&
                          p.wi + 2 * p.pad_w <= 256 && p.out != nullptr;
    const int wp = p.wi + 2 * p.pad_w;
    const int q_valid = (p.ho - 1) * wp + p.wo;
    const int tiles_img = (q_valid + BLOCK_M - 1) / BLOCK_M;
    const double eff = (double)p.ho * p.wo / ((double)tiles_img * BLOCK_M);
    // measured on B200: the layers this kernel targets are bound by tcgen05.mma issue (~190 cycles per
    // instruction whatever N is), not by operand traffic, so it is opt-in (DML_USE_V3=1) for now
    const bool want = tune.version == 3 || (tune.version == 0 && eff >= 0.70 && getenv("DML_USE_V3"));
    if (tune.version == 3 && !eligible) return tc_fail("shape is not eligible for the halo kernel");
    if (eligible && want) {
      op.version = 3;
      op.wp = wp; op.q_valid = q_valid; op.tiles_per_img = tiles_img;
      const int halo = (p.kh - 1) * wp + p.kw - 1;
      op.rows_box = (wp - 1 + BLOCK_M - 1 + halo) / wp + 1;
      if (op.rows_box > 256) return tc_fail("halo box too tall");
      op.a_stage_bytes = (op.rows_box * wp * 128 + 1023) / 1024 * 1024;
      op.cpb = p.cin / BLOCK_K;
      const long long m_tiles = (long long)n_capacity * tiles_img;
      op.bn = pick_bn_v2(p.cout, m_tiles, tune.bn);
      if (op.bn == 0) return tc_fail("no N tile for cout = %d", p.cout);
      op.acc_stride = 32;
      while (op.acc_stride < op.bn) op.acc_stride *= 2;
      op.tmem_cols = 2 * op.acc_stride;
      op.cw = op.bn % 64 == 0 ? 64 : (op.bn % 32 == 0 ? 32 : 16);
      const int fixed = 1024 + 2 * 16384 + 1024 + 512;
      op.a_stages = op.cpb >= 2 ? 3 : 2;
      int b_bytes = op.bn * 128;
      while (op.a_stages > 1 && fixed + op.a_stages * op.a_stage_bytes + 2 * b_bytes > SMEM_LIMIT) --op.a_stages;
      int room = SMEM_LIMIT - fixed - op.a_stages * op.a_stage_bytes;
      op.b_stages = room / b_bytes;
      if (op.b_stages > 12) op.b_stages = 12;
      if (tune.stages > 0 && tune.stages < op.b_stages) op.b_stages = tune.stages;
      if (op.b_stages >= 2) {
        op.smem = (size_t)fixed + (size_t)op.a_stages * op.a_stage_bytes + (size_t)op.b_stages * b_bytes;
        op.m_capacity = n_capacity;
        cuuint64_t gdim[4] = {(cuuint64_t)p.cin, (cuuint64_t)p.wi, (cuuint64_t)p.hi, (cuuint64_t)n_capacity};
        cuuint64_t gstr[3] = {(cuuint64_t)p.in_pitch * 2, (cuuint64_t)p.wi * p.in_pitch * 2, (cuuint64_t)p.hi * p.wi * p.in_pitch * 2};
        cuuint32_t box[4] = {BLOCK_K, (cuuint32_t)wp, (cuuint32_t)op.rows_box, 1};
        cuuint32_t es[4] = {1, 1, 1, 1};
        CUresult r3 = g_encode_tiled(&op.map_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(p.in), gdim, gstr, box, es,
                                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r3 != CUDA_SUCCESS) return tc_fail("cuTensorMapEncodeTiled(halo A) failed (%d): wp=%d rows=%d", (int)r3, wp, op.rows_box);
        const cuuint64_t ktot = (cuuint64_t)p.kh * p.kw * p.cin_pad;
        cuuint64_t bdim[2] = {ktot, (cuuint64_t)p.cout};
        cuuint64_t bstr[1] = {ktot * 2};
        cuuint32_t bbox[2] = {BLOCK_K, (cuuint32_t)op.bn};
        cuuint32_t bes[2] = {1, 1};
        r3 = g_encode_tiled(&op.map_b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(p.w), bdim, bstr, bbox, bes,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r3 != CUDA_SUCCESS) return tc_fail("cuTensorMapEncodeTiled(weights) failed (%d)", (int)r3);
        op.valid = true;
        *out = op;
        return 0;
      }
      if (tune.version == 3) return tc_fail("halo kernel does not fit in shared memory for this shape");
    }
  }
  op.version = tune.version == 1 ? 1 : (tune.version == 4 ? 4 : 2);
  // default: CTA pairs (3-8 % faster on almost every layer) when there are enough 256-row tiles
  if (tune.version == 0 && !getenv("DML_NO_PAIR") && (long long)n_capacity * p.ho * p.wo >= 2 * BLOCK_M) op.version = 4;
  if (p.in2) {  // second GEMM source: pair kernel only
    if (tune.version != 0 && tune.version != 4) return tc_fail("only the CTA-pair kernel reads a second GEMM source");
    if (p.kh != 1 || p.kw != 1 || p.cin % BLOCK_K || p.cin2 % BLOCK_K || p.cin2_pad != p.cin2 || p.res)
      return tc_fail("second GEMM source needs a 1x1 convolution, channel counts that are multiples of 64 and no residual");
    op.version = 4;
  }
  const bool pair = op.version == 4;  // CTA pairs: 256-row tiles, each CTA stages half of the weight tile
  const long long M_cap = (long long)n_capacity * p.ho * p.wo;
  const long long m_tiles_cap = pair ? (M_cap + 2 * BLOCK_M - 1) / (2 * BLOCK_M) : (M_cap + BLOCK_M - 1) / BLOCK_M;
  op.has_res = op.version != 1 && p.res != nullptr;
  // residual in ONE tile buffer whose column chunks are handed back one by one (pair kernel) instead of two whole tile
  // buffers: frees 128 x bn x 2 bytes of shared memory, so the N tile can stay at 256 with a 3-stage operand ring (the A
  // tile is then re-read half as often; these layers are bound by the bytes an SM can ingest)
  op.res_direct = op.has_res && pair && (tune.res_direct == 2 || (tune.res_direct == 0 && !getenv("DML_NO_RES_CHUNKS")));
  const int res_bufs = op.has_res ? (op.res_direct ? 1 : 2) : 0;
  const bool res_smem = op.has_res && !op.res_direct;
  op.bn = op.version == 1 ? pick_bn(p.cout, tune.bn) : pick_bn_v2(p.cout, pair ? 2 * m_tiles_cap : m_tiles_cap, tune.bn);
  // two residual tiles of 128 x bn bf16 live in shared memory: keep them at 64 KiB
  if (res_smem && tune.bn == 0 && op.bn > 128 && !pair) op.bn = pick_bn(p.cout, 0);
  if (op.bn == 0) return tc_fail("no N tile for cout = %d (override %d)", p.cout, tune.bn);
  if (p.cin_pad % BLOCK_K) return tc_fail("cin_pad = %d is not a multiple of %d", p.cin_pad, BLOCK_K);
  op.cpb = p.cin_pad / BLOCK_K;
  const int num_kb = p.kh * p.kw * op.cpb + p.cin2_pad / BLOCK_K;
  const int stage_bytes = A_STAGE_BYTES + (pair ? op.bn * 64 : op.bn * 128);
  if (op.version == 1) {
    int target = op.bn <= 64 ? 3 : (op.bn <= 128 ? 3 : 4);
    if (tune.stages > 0) target = tune.stages;
    if (target > num_kb) target = num_kb;
    while (target > 1 && target * stage_bytes + 2048 > SMEM_LIMIT) --target;
    op.stages = target;
    op.smem = (size_t)op.stages * stage_bytes + 1024 /*alignment slack*/ + 16 * op.stages + 64;
    op.tmem_cols = 32;
    while (op.tmem_cols < op.bn) op.tmem_cols *= 2;
  } else {
    int fixed = 1024 /*alignment slack*/ + OUT_STAGING_BYTES + 256 /*barriers*/ + res_bufs * BLOCK_M * op.bn * 2;
    int sb = stage_bytes;
    // the two residual tiles may not leave room for two stages at the widest N tile: narrow it (heuristic choice only)
    while (tune.bn == 0 && (SMEM_LIMIT - fixed) / sb < 2 && op.bn > 64) {
      int nb = 0;
      for (int c = op.bn - 16; c >= 16; c -= 16)
        if (p.cout % c == 0) { nb = c; break; }
      if (!nb) break;
      op.bn = nb;
      fixed = 1024 + OUT_STAGING_BYTES + 256 + res_bufs * BLOCK_M * op.bn * 2;
      sb = A_STAGE_BYTES + (pair ? op.bn * 64 : op.bn * 128);
    }
    // k-blocks per stage: the MMA warp pays one barrier wait + one commit per STAGE, so narrow N tiles
    // (short MMAs) want several k-blocks behind each barrier; ~48 KiB per stage by default
    int G = tune.group > 0 ? tune.group : 49152 / sb;
    if (const char* e = getenv("DML_GROUP")) if (tune.group == 0) G = atoi(e);
    if (G > num_kb) G = num_kb;
    if (G < 1) G = 1;
    while (G > 1 && (SMEM_LIMIT - fixed) / (G * sb) < 2) --G;
    int target = (SMEM_LIMIT - fixed) / (G * sb);
    if (target > 8) target = 8;
    if (tune.stages > 0 && tune.stages < target) target = tune.stages;
    if (target < 2) return tc_fail("not enough shared memory for 2 stages at bn = %d", op.bn);
    op.stages = target;
    op.group = G;
    op.smem = (size_t)op.stages * G * sb + fixed;
    op.acc_stride = 32;
    while (op.acc_stride < op.bn) op.acc_stride *= 2;
    op.tmem_cols = 2 * op.acc_stride;
    op.cw = op.bn % 64 == 0 ? 64 : (op.bn % 32 == 0 ? 32 : 16);
    if (op.res_direct && op.bn / op.cw > 4) return tc_fail("chunked residual buffer needs at most 4 column chunks (bn %d, cw %d)", op.bn, op.cw);
    op.m_capacity = (int)M_cap;
    if (!p.out) return tc_fail("output pointer needed to plan the persistent kernel");
    cuuint64_t odim[2] = {(cuuint64_t)p.cout, (cuuint64_t)M_cap};
    cuuint64_t ostr[1] = {(cuuint64_t)p.out_pitch * 2};
    cuuint32_t obox[2] = {(cuuint32_t)op.cw, 32};
    cuuint32_t oes[2] = {1, 1};
    CUtensorMapSwizzle osw = op.cw == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : (op.cw == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
    CUresult ro = g_encode_tiled(&op.map_o, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, p.out, odim, ostr, obox, oes,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, osw, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                                 CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (ro != CUDA_SUCCESS)
      return tc_fail("cuTensorMapEncodeTiled(output) failed (%d): cout=%d M=%lld pitch=%d cw=%d", (int)ro, p.cout, M_cap, p.out_pitch, op.cw);
    op.map_r = op.map_o;
    if (op.has_res) {
      cuuint64_t rstr[1] = {(cuuint64_t)p.res_pitch * 2};
      cuuint32_t rbox[2] = {(cuuint32_t)op.cw, (cuuint32_t)BLOCK_M};
      ro = g_encode_tiled(&op.map_r, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(p.res), odim, rstr, rbox, oes,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, osw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (ro != CUDA_SUCCESS)
        return tc_fail("cuTensorMapEncodeTiled(residual) failed (%d): cout=%d M=%lld pitch=%d cw=%d", (int)ro, p.cout, M_cap, p.res_pitch, op.cw);
    }
  }

  // A: NHWC view as (C, W, H, N), im2col mode.  Corner convention as in CUTLASS'
  // make_im2col_tma_copy_desc: lower = -pad, upper = pad - (filter - 1); index 0 = W.
  cuuint64_t gdim[4] = {(cuuint64_t)p.cin, (cuuint64_t)p.wi, (cuuint64_t)p.hi, (cuuint64_t)n_capacity};
  cuuint64_t gstr[3] = {(cuuint64_t)p.in_pitch * 2, (cuuint64_t)p.wi * p.in_pitch * 2,
                        (cuuint64_t)p.hi * p.wi * p.in_pitch * 2};
  int lower[2] = {-p.pad_w, -p.pad_h};
  int upper[2] = {p.pad_w - (p.kw - 1), p.pad_h - (p.kh - 1)};
  cuuint32_t estr[4] = {1, (cuuint32_t)p.stride, (cuuint32_t)p.stride, 1};
  op.a_tiled = op.version != 1 && p.kh == 1 && p.kw == 1 && p.stride == 1 && p.pad_h == 0 && p.pad_w == 0 &&
               !getenv("DML_NO_TILED_A");
  CUresult r = CUDA_SUCCESS;
  if (op.a_tiled) {
    cuuint64_t adim[2] = {(cuuint64_t)p.cin, (cuuint64_t)n_capacity * p.hi * p.wi};
    cuuint64_t astr[1] = {(cuuint64_t)p.in_pitch * 2};
    cuuint32_t abox[2] = {BLOCK_K, BLOCK_M};
    cuuint32_t aes[2] = {1, 1};
    r = g_encode_tiled(&op.map_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(p.in), adim, astr, abox, aes,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return tc_fail("cuTensorMapEncodeTiled(A) failed (%d)", (int)r);
  } else
  r = g_encode_im2col(&op.map_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(p.in), gdim, gstr, lower,
                               upper, BLOCK_K, BLOCK_M, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return tc_fail("cuTensorMapEncodeIm2col failed (%d): C=%d W=%d H=%d N=%d pitch=%d k=%dx%d s=%d pad=%d,%d", (int)r, p.cin,
                   p.wi, p.hi, n_capacity, p.in_pitch, p.kh, p.kw, p.stride, p.pad_h, p.pad_w);
  // Same small-tensor fix-up CUTLASS applies for drivers <= 13.1 (tensors < 128 KiB).
  if (g_driver_version <= 13010 && (size_t)n_capacity * p.hi * p.wi * p.in_pitch * 2 < 131072)
    reinterpret_cast<uint64_t*>(&op.map_a)[1] &= ~(1ull << 21);

  // B: packed weights [cout][taps*cin_pad (+ cin2_pad)], K contiguous.
  const cuuint64_t ktot = (cuuint64_t)p.kh * p.kw * p.cin_pad + (cuuint64_t)p.cin2_pad;
  cuuint64_t bdim[2] = {ktot, (cuuint64_t)p.cout};
  cuuint64_t bstr[1] = {ktot * 2};
  cuuint32_t bbox[2] = {BLOCK_K, (cuuint32_t)(pair ? op.bn / 2 : op.bn)};
  cuuint32_t bes[2] = {1, 1};
  r = g_encode_tiled(&op.map_b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(p.w), bdim, bstr, bbox, bes,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return tc_fail("cuTensorMapEncodeTiled(weights) failed (%d): K=%llu cout=%d bn=%d", (int)r,
                                        (unsigned long long)ktot, p.cout, op.bn);
  op.map_a2 = op.map_a;
  op.a2_tiled = false;
  if (p.in2) {
    // second source: 1x1 convolution of stride stride2, no padding, over (C2, W2, H2, N)
    if ((p.hi2 - 1) / p.stride2 + 1 != p.ho || (p.wi2 - 1) / p.stride2 + 1 != p.wo)
      return tc_fail("second GEMM source %dx%d at stride %d does not produce %dx%d", p.hi2, p.wi2, p.stride2, p.ho, p.wo);
    op.a2_tiled = p.stride2 == 1;
    if (op.a2_tiled) {
      cuuint64_t adim[2] = {(cuuint64_t)p.cin2, (cuuint64_t)n_capacity * p.hi2 * p.wi2};
      cuuint64_t astr[1] = {(cuuint64_t)p.in2_pitch * 2};
      cuuint32_t abox[2] = {BLOCK_K, BLOCK_M};
      cuuint32_t aes[2] = {1, 1};
      r = g_encode_tiled(&op.map_a2, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(p.in2), adim, astr, abox, aes,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    } else {
      cuuint64_t g2dim[4] = {(cuuint64_t)p.cin2, (cuuint64_t)p.wi2, (cuuint64_t)p.hi2, (cuuint64_t)n_capacity};
      cuuint64_t g2str[3] = {(cuuint64_t)p.in2_pitch * 2, (cuuint64_t)p.wi2 * p.in2_pitch * 2,
                             (cuuint64_t)p.hi2 * p.wi2 * p.in2_pitch * 2};
      int lo2[2] = {0, 0}, up2[2] = {0, 0};
      cuuint32_t es2[4] = {1, (cuuint32_t)p.stride2, (cuuint32_t)p.stride2, 1};
      r = g_encode_im2col(&op.map_a2, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(p.in2), g2dim, g2str, lo2, up2, BLOCK_K,
                          BLOCK_M, es2, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r == CUDA_SUCCESS && g_driver_version <= 13010 && (size_t)n_capacity * p.hi2 * p.wi2 * p.in2_pitch * 2 < 131072)
        reinterpret_cast<uint64_t*>(&op.map_a2)[1] &= ~(1ull << 21);
    }
    if (r != CUDA_SUCCESS) return tc_fail("tensor map of the second GEMM source failed (%d)", (int)r);
  }
  op.valid = true;
  *out = op;
  return 0;
}

int launch_conv_tc_plan(const TcOpPlan& op, const ConvParams& p, cudaStream_t s) {
  if (!op.valid) return tc_fail("conv plan is not valid");
  if (p.norelu_cols > 0 && op.version != 2 && op.version != 4) return tc_fail("only the persistent GEMM kernels take ReLU-free output columns");
  if (op.version == 5) {
    TcV5Params k{};
    k.bias = p.bias;
    k.M = p.n * p.ho * p.wo;
    k.ho = p.ho; k.wo = p.wo;
    k.stride = p.stride; k.pad_h = p.pad_h; k.pad_w = p.pad_w;
    k.kw = p.kw; k.cpb = op.cpb; k.num_kb = p.kh * p.kw * op.cpb;
    k.last_kmma = (p.cin - (op.cpb - 1) * BLOCK_K + 15) / 16;
    k.cout = p.cout; k.box_cols = op.cw; k.relu = p.relu;
    k.stages = op.stages;
    k.p_tiles = (k.M + V5_PIX - 1) / V5_PIX;
    k.c_tiles = (p.cout + BLOCK_M - 1) / BLOCK_M;
    k.out_row0 = p.out_row0; k.in_img0 = p.in_img0;
    k.a_tiled = op.a_tiled ? 1 : 0;
    k.in_row0 = p.in_img0 * p.hi * p.wi;
    if (p.res != nullptr) return tc_fail("the weights-as-A kernel takes no residual");
    if (k.out_row0 + k.M > op.m_capacity) return tc_fail("batch of %d images exceeds the planned capacity", p.n);
    const int tiles = k.p_tiles * k.c_tiles;
    const int grid = tiles < g_num_sms ? tiles : g_num_sms;
    launch_pdl(conv_tc5_kernel, dim3(grid), dim3(TC_THREADS), op.smem, s, op.map_b, op.map_a, op.map_o, k);
    cudaError_t e5 = cudaGetLastError();
    if (e5 != cudaSuccess) return tc_fail("conv_tc5_kernel launch: %s", cudaGetErrorString(e5));
    return 0;
  }
  if (op.version == 7) {
    TcV6xParams k{};
    k.bias = p.bias;
    k.out = reinterpret_cast<__nv_bfloat16*>(p.out);
    if (p.n > op.m_capacity) return tc_fail("batch of %d images exceeds the planned capacity", p.n);
    k.n_img = p.n; k.ho = p.ho; k.wo = p.wo; k.wp = op.wp; k.out_pitch = p.out_pitch;
    k.kh = p.kh; k.kw = p.kw; k.pad_h = p.pad_h; k.pad_w = p.pad_w; k.cblocks = op.cpb;
    k.rpt = op.rows_box; k.ptiles_per_img = op.tiles_per_img;
    k.rows_box = op.rows_box + p.kh - 1;
    k.a_cb_bytes = op.a_cb_bytes; k.a_stage_bytes = op.a_stage_bytes; k.a_stages = op.a_stages;
    k.bn = op.bn; k.n_tiles = p.cout / op.bn; k.relu = p.relu; k.tmem_cols = op.tmem_cols; k.acc_stride = op.acc_stride;
    k.n_acc = op.tmem_cols / op.acc_stride;
    k.in_img0 = p.in_img0;
    k.out_img0 = p.out_row0 / (p.ho * p.wo);
    k.row_bytes = op.row_bytes;
    k.last_kmma = (p.cin - (op.cpb - 1) * (op.row_bytes / 2) + 15) / 16;
    k.cin_pad = p.cin_pad;
    k.inv_wp = 1.0f / (float)op.wp;
    k.wide_store = (p.out_pitch % 16 == 0 && reinterpret_cast<uintptr_t>(p.out) % 32 == 0) ? 1 : 0;
    k.rev = p.rev;
    const int tiles = p.n * op.tiles_per_img;
    const int pairs = g_num_sms / 2 / k.n_tiles;
    const int vn = tiles < pairs ? tiles : pairs;
    const int grid = 2 * vn * k.n_tiles;
    bool launched = false;
#define DML_V6X_TRY(KH_, KW_, CB_, RB_)                                                                                         \
    if (!launched && p.kh == KH_ && p.kw == KW_ && op.cpb == CB_ && op.row_bytes == RB_) {                                        \
      launch_pdl(conv_tc6x_kernel<KH_, KW_, CB_, RB_>, dim3(grid), dim3(TC6_THREADS), op.smem, s, op.map_a, op.map_b, k);        \
      launched = true;                                                                                                          \
    }
    DML_V6X_SHAPES(DML_V6X_TRY)
#undef DML_V6X_TRY
    if (!launched) launch_pdl(conv_tc6x_kernel<0, 0, 0, 0>, dim3(grid), dim3(TC6_THREADS), op.smem, s, op.map_a, op.map_b, k);
    cudaError_t e7 = cudaGetLastError();
    if (e7 != cudaSuccess) return tc_fail("conv_tc6x_kernel launch: %s", cudaGetErrorString(e7));
    return 0;
  }
  if (op.version == 6) {
    TcV6Params k{};
    k.bias = p.bias;
    k.out = reinterpret_cast<__nv_bfloat16*>(p.out);
    if (p.n > op.m_capacity) return tc_fail("batch of %d images exceeds the planned capacity", p.n);
    k.n_img = p.n; k.ho = p.ho; k.wo = p.wo; k.wp = op.wp; k.out_pitch = p.out_pitch;
    k.kh = p.kh; k.kw = p.kw; k.pad_h = p.pad_h; k.pad_w = p.in_row_cells ? 0 : p.pad_w; k.cblocks = op.cpb;  // (the left padding is in memory)
    k.tiles_per_img = op.tiles_per_img; k.rows_box = op.rows_box;
    k.a_cb_bytes = op.a_cb_bytes; k.a_stage_bytes = op.a_stage_bytes; k.a_stages = op.a_stages;
    k.bn = op.bn; k.relu = p.relu; k.tmem_cols = op.tmem_cols; k.acc_stride = op.acc_stride; k.cw = op.cw;
    k.in_img0 = p.in_img0;
    k.out_img0 = p.out_row0 / (p.ho * p.wo);
    k.row_bytes = op.row_bytes;
    const int kb = op.row_bytes / 2;
    k.last_kmma = (p.cin - (op.cpb - 1) * kb + 15) / 16;
    k.cin_pad = p.cin_pad;
    const int tiles = p.n * op.tiles_per_img;
    k.n_tiles = p.cout / op.bn;
    const int vgrid = tiles < g_num_sms / k.n_tiles ? tiles : g_num_sms / k.n_tiles;
    const int grid = vgrid * k.n_tiles;
    k.n_acc = op.tmem_cols / op.acc_stride;
    k.inv_wp = 1.0f / (float)op.wp;
    k.rev = p.rev;
    k.wide_store = (p.out_pitch % 16 == 0 && reinterpret_cast<uintptr_t>(p.out) % 32 == 0 && !getenv("DML_NO_WIDE_STORE")) ? 1 : 0;
    if (getenv("DML_V6_NOSTORE")) k.wide_store = 2;  // experiment: skip the epilogue's global stores
    static const bool prof6 = getenv("DML_TC_PROF") != nullptr;
    k.prof = prof6 ? 1 : 0;
    // compile-time shapes (fully unrolled MMA issue); anything else takes the generic loop
#define DML_V6_SHAPES(X) X(3, 3, 1, 128) X(3, 3, 1, 64) X(3, 3, 2, 128) X(4, 4, 1, 32) X(2, 2, 1, 32) X(5, 5, 1, 128) \
    X(1, 7, 2, 128) X(7, 1, 2, 128) X(1, 7, 3, 128) X(7, 1, 3, 128) X(1, 3, 6, 128) X(3, 1, 6, 128) X(3, 3, 3, 64) X(1, 7, 5, 64) X(7, 1, 5, 64)
    bool launched = false;
#define DML_V6_TRY(KH_, KW_, CB_, RB_)                                                                                         \
    if (!launched && p.kh == KH_ && p.kw == KW_ && op.cpb == CB_ && op.row_bytes == RB_) {                                       \
      launch_pdl(conv_tc6_kernel<KH_, KW_, CB_, RB_>, dim3(grid), dim3(TC6_THREADS), op.smem, s, op.map_a, op.map_b, k);        \
      launched = true;                                                                                                         \
    }
    DML_V6_SHAPES(DML_V6_TRY)
#undef DML_V6_TRY
    if (!launched) launch_pdl(conv_tc6_kernel<0, 0, 0, 0>, dim3(grid), dim3(TC6_THREADS), op.smem, s, op.map_a, op.map_b, k);
    if (prof6) {
      cudaStreamSynchronize(s);
      static long long h[160 * 8];
      cudaMemcpyFromSymbol(h, g_tc2_prof, sizeof h);
      double a[8] = {};
      for (int b2 = 0; b2 < grid; ++b2) for (int i = 0; i < 8; ++i) a[i] += (double)h[b2 * 8 + i];
      const double nt = tiles;
      fprintf(stderr, "[tc6 prof] bn %d a_stages %d a_stage_bytes %d | cycles per tile: total %.0f | producer wait empty %.0f | mma wait tmem-empty %.0f, wait A %.0f | epilogue wait accumulator %.0f\n",
              op.bn, op.a_stages, op.a_stage_bytes, a[4] / nt, a[0] / nt, a[1] / nt, a[2] / nt, a[3] / nt);
    }
    cudaError_t e6 = cudaGetLastError();
    if (e6 != cudaSuccess) return tc_fail("conv_tc6_kernel launch: %s", cudaGetErrorString(e6));
    return 0;
  }
#ifdef DML_LEGACY_VARIANTS
  if (op.version == 3) {
    TcV3Params k{};
    k.bias = p.bias;
    k.out = reinterpret_cast<__nv_bfloat16*>(p.out);
    k.n_img = p.n; k.ho = p.ho; k.wo = p.wo; k.wp = op.wp; k.out_pitch = p.out_pitch;
    k.kh = p.kh; k.kw = p.kw; k.pad_h = p.pad_h; k.pad_w = p.pad_w; k.cblocks = op.cpb;
    k.tiles_per_img = op.tiles_per_img; k.q_valid = op.q_valid; k.rows_box = op.rows_box;
    k.a_stage_bytes = op.a_stage_bytes; k.a_stages = op.a_stages; k.b_stages = op.b_stages;
    k.bn = op.bn; k.n_tiles = p.cout / op.bn; k.relu = p.relu;
    k.tmem_cols = op.tmem_cols; k.acc_stride = op.acc_stride; k.cw = op.cw;
    k.in_img0 = p.in_img0; k.out_img0 = 0;  // p.out already points at the launch's first image
    if (p.in_img0 + p.n > op.m_capacity) return tc_fail("batch of %d images exceeds the planned capacity", p.n);
    const long long tiles = (long long)p.n * op.tiles_per_img * k.n_tiles;
    const int grid = tiles < g_num_sms ? (int)tiles : g_num_sms;
    conv_tc3_kernel<<<grid, TC_THREADS, op.smem, s>>>(op.map_a, op.map_b, k);
    cudaError_t e3 = cudaGetLastError();
    if (e3 != cudaSuccess) return tc_fail("conv_tc3_kernel launch: %s", cudaGetErrorString(e3));
    return 0;
  }
#endif
  if (op.version == 2 || op.version == 4) {
    const bool pair = op.version == 4;
    TcV2Params k{};
    k.bias = p.bias;
    k.res = reinterpret_cast<const __nv_bfloat16*>(p.res);
    k.M = p.n * p.ho * p.wo;
    k.has_res = op.has_res ? (op.res_direct ? 2 : 1) : 0;
    k.M_cap = op.m_capacity;
    k.in_img0 = p.in_img0;
    k.a_tiled = op.a_tiled ? 1 : 0;
    k.last_kmma = (p.cin - (op.cpb - 1) * BLOCK_K + 15) / 16;
    k.in_row0 = p.in_img0 * p.hi * p.wi;
    k.out_row0 = p.out_row0;
    if ((p.res != nullptr) != op.has_res) return tc_fail("residual pointer does not match the plan");
    if (k.out_row0 + k.M > op.m_capacity) return tc_fail("batch of %d images exceeds the planned capacity", p.n);
    k.ho = p.ho; k.wo = p.wo;
    k.stride = p.stride; k.pad_h = p.pad_h; k.pad_w = p.pad_w;
    k.kw = p.kw; k.cpb = op.cpb; k.num_kb = p.kh * p.kw * op.cpb + p.cin2_pad / BLOCK_K;
    k.kb1 = p.kh * p.kw * op.cpb; k.stride2 = p.stride2; k.a2_tiled = op.a2_tiled ? 1 : 0;
    k.in2_row0 = p.in_img0 * p.hi2 * p.wi2;
    if (p.in2 && !pair) return tc_fail("only the CTA-pair kernel reads a second GEMM source");
    k.bn = op.bn; k.res_pitch = p.res_pitch; k.relu = p.relu;
    k.stages = op.stages; k.tmem_cols = op.tmem_cols; k.acc_stride = op.acc_stride;
    k.group = op.group;
    k.norelu_cols = p.norelu_cols;
    k.rev = p.rev;
    k.m_tiles = pair ? (k.M + 2 * BLOCK_M - 1) / (2 * BLOCK_M) : (k.M + BLOCK_M - 1) / BLOCK_M;
    k.n_tiles = p.cout / op.bn;
    k.cw = op.cw;
    const int tiles = k.m_tiles * k.n_tiles;
    static const int sm_limit = getenv("DML_SM_LIMIT") ? atoi(getenv("DML_SM_LIMIT")) : g_num_sms;  // experiments: use fewer SMs
    const int grid = tiles < sm_limit ? tiles : sm_limit;
    if (pair) {
      const int clusters = tiles < sm_limit / 2 ? tiles : sm_limit / 2;
      static const bool profx = getenv("DML_TC_PROF") != nullptr;
      k.prof = profx ? 1 : 0;
      launch_pdl(conv_tc2x_kernel, dim3(2 * clusters), dim3(TC2X_THREADS), op.smem, s, op.map_a, op.map_b, op.map_o, op.map_r, op.map_a2, k);
      if (profx) {
        cudaStreamSynchronize(s);
        static long long h[160 * 8];
        cudaMemcpyFromSymbol(h, g_tc2_prof, sizeof h);
        double a[8] = {};
        for (int b2 = 0; b2 < clusters; ++b2) for (int i = 0; i < 8; ++i) a[i] += (double)h[b2 * 8 + i];
        const double kb = a[7] * k.num_kb;
        fprintf(stderr, "[tc2x prof] bn %d group %d stages %d kb/tile %d tiles/cluster %.1f | cycles per k-block: producer total %.0f (wait empty %.0f) | mma total %.0f (wait full %.0f, wait tmem-empty %.0f)\n",
                op.bn, op.group, op.stages, k.num_kb, a[7] / clusters, a[1] / kb, a[0] / kb, a[4] / kb, a[3] / kb, a[2] / kb);
      }
    } else {
      static const bool prof = getenv("DML_TC_PROF") != nullptr;
      k.prof = prof ? 1 : 0;
      launch_pdl(conv_tc2_kernel, dim3(grid), dim3(TC2_THREADS), op.smem, s, op.map_a, op.map_b, op.map_o, op.map_r, k);
      if (prof) {
        cudaStreamSynchronize(s);
        static long long h[160 * 8];
        cudaMemcpyFromSymbol(h, g_tc2_prof, sizeof h);
        double a[8] = {};
        for (int b2 = 0; b2 < grid; ++b2) for (int i = 0; i < 8; ++i) a[i] += (double)h[b2 * 8 + i];
        const double kb = a[7] * k.num_kb;  // k-blocks over all CTAs
        fprintf(stderr, "[tc2 prof] bn %d group %d stages %d kb/tile %d tiles/cta %.1f | cycles per k-block: producer total %.0f (wait empty %.0f) | mma total %.0f (wait full %.0f, wait tmem-empty %.0f)\n",
                op.bn, op.group, op.stages, k.num_kb, a[7] / grid, a[1] / kb, a[0] / kb, a[4] / kb, a[3] / kb, a[2] / kb);
      }
    }
    cudaError_t e2 = cudaGetLastError();
    if (e2 != cudaSuccess) return tc_fail("conv_tc2_kernel launch: %s", cudaGetErrorString(e2));
    return 0;
  }
#ifndef DML_LEGACY_VARIANTS
  return tc_fail("kernel variant %d is not in this build (compile with -DDML_LEGACY_VARIANTS)", op.version);
#else
  TcKernelParams k{};
  k.bias = p.bias;
  k.res = reinterpret_cast<const __nv_bfloat16*>(p.res);
  k.out = reinterpret_cast<__nv_bfloat16*>(p.out);
  k.M = p.n * p.ho * p.wo;
  k.ho = p.ho; k.wo = p.wo;
  k.stride = p.stride; k.pad_h = p.pad_h; k.pad_w = p.pad_w;
  k.kw = p.kw; k.cpb = op.cpb; k.num_kb = p.kh * p.kw * op.cpb;
  k.bn = op.bn; k.out_pitch = p.out_pitch; k.res_pitch = p.res_pitch; k.relu = p.relu;
  k.stages = op.stages; k.tmem_cols = op.tmem_cols;
  dim3 grid((unsigned)((k.M + BLOCK_M - 1) / BLOCK_M), (unsigned)(p.cout / op.bn));
  conv_tc_kernel<<<grid, TC_THREADS, op.smem, s>>>(op.map_a, op.map_b, k);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return tc_fail("conv_tc_kernel launch: %s", cudaGetErrorString(e));
  return 0;
#endif
}


// ---- conv_chain_kernel: plan + launch over explicit pointers ----
bool chain_supported(const ChainArgs& a) {
  if (a.h <= 0 || a.w <= 0 || a.w + 2 > 64) return false;
  const int rpt = BLOCK_M / (a.w + 2);
  const int rows = rpt * a.w;
  if (rpt < 1 || a.h % (2 * rpt) != 0 || rows % 16 != 0 || rows > BLOCK_M) return false;
  if (!a.y1 || !a.w3 || !a.b3 || !a.we || !a.be || !a.out) return false;
  if ((a.res != nullptr) == (a.src2 != nullptr)) return false;   // exactly one of: residual tensor, second GEMM source
  if ((a.wr != nullptr) != (a.y1n != nullptr) || (a.wr && !a.br)) return false;
  if (a.src2 && !a.wr) return false;                              // variants built: res+next, src2+next, res only
  if (a.y1_pitch % 8 || a.out_pitch % 8 || (a.res && a.res_pitch % 8) || (a.src2 && a.src2_pitch % 8) || (a.y1n && a.y1n_pitch % 8)) return false;
  return true;
}

int chain_plan(ChainPlan* out, const ChainArgs& a, int n_capacity) {
  if (!g_encode_tiled) return tc_fail("tc_init was not called");
  if (!chain_supported(a)) return tc_fail("shape is not eligible for the bottleneck-tail chain kernel");
  ChainPlan pl;
  pl.wp = a.w + 2;
  pl.rpt = BLOCK_M / pl.wp;
  pl.rows = pl.rpt * a.w;
  pl.ptiles_per_img = a.h / (2 * pl.rpt);
  pl.box_bytes = (pl.rpt + 2) * pl.wp * 128;
  const int reach = BLOCK_M + 2 * pl.wp + 2;  // rows the shifted views of a 128-row MMA may touch
  const int box_rows = (pl.rpt + 2) * pl.wp;
  pl.halo_bytes = ((box_rows > reach ? box_rows : reach) * 128 + 1023) / 1024 * 1024;
  pl.kb2 = a.src2 ? 2 : 1;
  pl.variant = a.src2 ? 1 : (a.wr ? 0 : 2);
  const ChainSmem L = chain_smem(pl.rows, pl.halo_bytes, pl.kb2, a.res != nullptr, a.src2 != nullptr, a.wr != nullptr);
  if (L.total > (uint32_t)SMEM_LIMIT) return tc_fail("chain kernel needs %u bytes of shared memory", L.total);
  pl.smem = L.total;
  pl.n_capacity = n_capacity;
  const cuuint64_t M = (cuuint64_t)n_capacity * a.h * a.w;
  CUresult r;
  {  // halo of the 3x3: NHWC as (C, W, H, N), box {64, Wp, rpt + 2, 1}, zero fill outside the image
    cuuint64_t gdim[4] = {(cuuint64_t)CH_C, (cuuint64_t)a.w, (cuuint64_t)a.h, (cuuint64_t)n_capacity};
    cuuint64_t gstr[3] = {(cuuint64_t)a.y1_pitch * 2, (cuuint64_t)a.w * a.y1_pitch * 2, (cuuint64_t)a.h * a.w * a.y1_pitch * 2};
    cuuint32_t box[4] = {(cuuint32_t)CH_C, (cuuint32_t)pl.wp, (cuuint32_t)(pl.rpt + 2), 1};
    cuuint32_t es[4] = {1, 1, 1, 1};
    r = g_encode_tiled(&pl.map_halo, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(a.y1), gdim, gstr, box, es,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return tc_fail("cuTensorMapEncodeTiled(chain halo) failed (%d)", (int)r);
  }
  auto weight_map = [&](CUtensorMap* m, const void* w, int k, int rows_total, int box_rows_) -> CUresult {
    cuuint64_t d[2] = {(cuuint64_t)k, (cuuint64_t)rows_total};
    cuuint64_t st[1] = {(cuuint64_t)k * 2};
    cuuint32_t b[2] = {(cuuint32_t)CH_C, (cuuint32_t)box_rows_};
    cuuint32_t es[2] = {1, 1};
    return g_encode_tiled(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(w), d, st, b, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  };
  if ((r = weight_map(&pl.map_w3, a.w3, 9 * CH_C, CH_C, 32)) != CUDA_SUCCESS) return tc_fail("chain: 3x3 weight map failed (%d)", (int)r);
  if ((r = weight_map(&pl.map_we, a.we, pl.kb2 * CH_C, CH_N2, 128)) != CUDA_SUCCESS) return tc_fail("chain: expand weight map failed (%d)", (int)r);
  pl.map_wr = pl.map_we;
  if (a.wr && (r = weight_map(&pl.map_wr, a.wr, CH_N2, CH_C, 32)) != CUDA_SUCCESS) return tc_fail("chain: reduce weight map failed (%d)", (int)r);
  auto pixel_map = [&](CUtensorMap* m, const void* base, int channels, int pitch, int box_rows_) -> CUresult {
    cuuint64_t d[2] = {(cuuint64_t)channels, M};
    cuuint64_t st[1] = {(cuuint64_t)pitch * 2};
    cuuint32_t b[2] = {(cuuint32_t)CH_C, (cuuint32_t)box_rows_};
    cuuint32_t es[2] = {1, 1};
    return g_encode_tiled(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), d, st, b, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  };
  if ((r = pixel_map(&pl.map_out, a.out, CH_N2, a.out_pitch, 16)) != CUDA_SUCCESS) return tc_fail("chain: output map failed (%d)", (int)r);
  pl.map_res = pl.map_out; pl.map_s2 = pl.map_out;
  if (a.res && (r = pixel_map(&pl.map_res, a.res, CH_N2, a.res_pitch, pl.rows)) != CUDA_SUCCESS) return tc_fail("chain: residual map failed (%d)", (int)r);
  if (a.src2 && (r = pixel_map(&pl.map_s2, a.src2, CH_C, a.src2_pitch, pl.rows)) != CUDA_SUCCESS) return tc_fail("chain: second-source map failed (%d)", (int)r);
  if (a.y1n && (a.y1n_pitch % 16 || reinterpret_cast<uintptr_t>(a.y1n) % 32)) return tc_fail("chain: next-reduce output must be 32-byte aligned per pixel");
  pl.valid = true;
  *out = pl;
  return 0;
}

int launch_chain(const ChainPlan& pl, const ChainArgs& a, cudaStream_t s) {
  if (!pl.valid) return tc_fail("chain plan is not valid");
  if (a.n > pl.n_capacity) return tc_fail("batch of %d images exceeds the planned capacity", a.n);
  ChainParams k{};
  k.bias1 = a.b3; k.bias2 = a.be; k.bias3 = a.br;
  k.y1n = reinterpret_cast<__nv_bfloat16*>(a.y1n); k.y1n_pitch = a.y1n_pitch;
  k.out = reinterpret_cast<__nv_bfloat16*>(a.out); k.out_pitch = a.out_pitch;
  k.n_img = a.n; k.ho = a.h; k.wo = a.w; k.wp = pl.wp; k.rpt = pl.rpt; k.ptiles_per_img = pl.ptiles_per_img; k.rows = pl.rows;
  k.halo_bytes = pl.halo_bytes; k.box_bytes = pl.box_bytes; k.kb2 = pl.kb2; k.in_img0 = 0; k.pix_base = 0;
  k.inv_wp = 1.0f / (float)pl.wp;
  static const int ahead_env = getenv("DML_CHAIN_AHEAD") ? atoi(getenv("DML_CHAIN_AHEAD")) : 2;
  k.ahead = ahead_env == 1 ? 1 : 2;
  k.rev = a.rev;
  static const bool prof = getenv("DML_CHAIN_PROF") != nullptr;
  if (prof) {
    static long long zero[160 * 40] = {};
    cudaMemcpyToSymbol(g_chain_prof, zero, sizeof zero);
  }
  const int tiles = a.n * pl.ptiles_per_img;
  const int pairs = g_num_sms / 2;
  const int clusters = tiles < pairs ? tiles : pairs;
  const dim3 grid(2 * clusters), block(CHAIN_THREADS);
  static const bool dbg = getenv("DML_CHAIN_DEBUG") != nullptr;
  static unsigned long long* h_word = nullptr;
  if (dbg && !h_word) {
    unsigned long long* d_word = nullptr;
    if (cudaHostAlloc(&h_word, sizeof *h_word, cudaHostAllocMapped) == cudaSuccess && cudaHostGetDevicePointer(&d_word, h_word, 0) == cudaSuccess) {
      *h_word = 0;
      cudaMemcpyToSymbol(g_chain_dbg_ptr, &d_word, sizeof d_word);
    }
  }
#define DML_CHAIN_LAUNCH(R_, S_, N_)                                                                                                           \
  do {                                                                                                                                         \
    if (prof) launch_pdl(conv_chain_kernel<R_, S_, N_, true>, grid, block, pl.smem, s, pl.map_halo, pl.map_w3, pl.map_we, pl.map_wr, pl.map_res, pl.map_out, pl.map_s2, k); \
    else launch_pdl(conv_chain_kernel<R_, S_, N_, false>, grid, block, pl.smem, s, pl.map_halo, pl.map_w3, pl.map_we, pl.map_wr, pl.map_res, pl.map_out, pl.map_s2, k); \
  } while (0)
  if (pl.variant == 0) DML_CHAIN_LAUNCH(true, false, true);
  else if (pl.variant == 1) DML_CHAIN_LAUNCH(false, true, true);
  else DML_CHAIN_LAUNCH(true, false, false);
#undef DML_CHAIN_LAUNCH
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return tc_fail("conv_chain_kernel launch: %s", cudaGetErrorString(e));
  if (prof) {
    cudaStreamSynchronize(s);
    static long long h[160 * 40];
    cudaMemcpyFromSymbol(h, g_chain_prof, sizeof h);
    double a[40] = {};
    double nt = 0, loop = 0, ntl = 0;
    for (int b = 0; b < 2 * clusters && b < 160; ++b) {
      for (int i = 0; i < 38; ++i) a[i] += (double)h[b * 40 + i];
      nt += (double)h[b * 40 + 39];
      if (b % 2 == 0) { loop += (double)h[b * 40 + 38]; ntl += (double)h[b * 40 + 39]; }
    }
    if (nt < 1) nt = 1;
    fprintf(stderr, "[chain prof] variant %d, %.1f tiles per cluster, %.0f cycles per tile | wait cycles per tile and CTA (one warp per role):", pl.variant,
            ntl / clusters, loop / (ntl > 0 ? ntl : 1));
    struct { const char* name; int first, count; } fam[] = {
        {"producer:halo_empty", 3, 2}, {"producer|B:outc_empty", 19, 4}, {"producer|B:slot_free", 27, 4}, {"producer:s2_empty", 13, 2},
        {"mma:halo_full", 1, 2}, {"mma:acc1_empty", 7, 2}, {"mma:acc2_empty", 33, 1}, {"mma:y2_full", 9, 1}, {"mma:s2_full", 11, 2},
        {"mma:acc3_empty", 32, 1}, {"mma:outc_full", 15, 1}, {"A:acc1_full", 5, 2}, {"A:y2_free", 34, 1}, {"A:acc3_full", 31, 1},
        {"B:acc2_full", 10, 1}, {"B:res_full", 23, 4}, {"B:compute", 35, 1}, {"B:fence", 36, 1}, {"B:arrives", 16, 1}, {"B:store_issue", 17, 1}, {"B:store_read", 37, 1}};
    for (const auto& f : fam) {
      double v = 0;
      for (int i = 0; i < f.count; ++i) v += a[f.first + i];
      if (v > 0) fprintf(stderr, " %s %.0f", f.name, v / nt);
    }
    fprintf(stderr, "\n");
  }
  if (dbg) {  // debugging aid: run to completion and report a wait that timed out (id, tile, block, warp)
    cudaError_t se = cudaStreamSynchronize(s);
    fprintf(stderr, "[chain debug] variant %d launch: sync -> %s, word %llx\n", pl.variant, cudaGetErrorString(se), h_word ? *h_word : 0ull);
    if (h_word && *h_word)
      fprintf(stderr, "[chain debug] variant %d: wait id %llu timed out at local tile %llu, block %llu, warp %llu (%s)\n", pl.variant,
              *h_word & 0xff, (*h_word >> 8) & 0xffff, (*h_word >> 24) & 0xffff, (*h_word >> 40) & 0xff, cudaGetErrorString(se));
  }
  return 0;
}

// ---- conv_duo_kernel: plan + launch over explicit pointers ----
bool duo_supported(const DuoArgs& a) {
  if (a.M <= 0 || !a.y2 || !a.we || !a.be || !a.out || !a.wr || !a.br || !a.y1n) return false;
  if ((a.res != nullptr) == (a.src2 != nullptr)) return false;  // exactly one of: identity residual, second GEMM source
  if (a.kmid % 64 || a.kmid < 64 || a.kmid > 384) return false;
  if (a.s2_stride != 1 && (!a.src2 || a.s2_stride != 2 || a.n <= 0 || a.h <= 0 || a.w <= 0 || a.n * a.h * a.w != a.M)) return false;
  if (a.k1 % 64 || a.k1 < 64 || a.k1 > a.kmid || (a.src2 ? a.k1 == a.kmid : a.k1 != a.kmid)) return false;
  if (a.src2 && a.src2_pitch % 8) return false;
  if (a.nexp % 256 || a.nexp < 256 || a.nexp > 2048) return false;
  if (a.nred % 64 || a.nred < 64 || a.nred > 256) return false;
  if (a.y2_pitch % 8 || a.out_pitch % 8 || (a.res && a.res_pitch % 8) || a.y1n_pitch % 16) return false;
  if (reinterpret_cast<uintptr_t>(a.y1n) % 32) return false;  // 32-byte vector stores per pixel
  return true;
}

int duo_plan(DuoPlan* out, const DuoArgs& a, int m_capacity) {
  if (!g_encode_tiled) return tc_fail("tc_init was not called");
  if (!duo_supported(a)) return tc_fail("shape is not eligible for the expand + reduce kernel");
  DuoPlan pl;
  const int kmb = a.kmid / 64, nj = a.nexp / 64;
  // Buffering, best first: eight `out` slots (twice the residual / store traffic in flight; the ring of slots is what bounds the
  // kernel), a second y2 tile (the next tile's GEMM2 starts without waiting for its load), then weight stages.
  const int opts[9][4] = {{8, 2, 2, 2}, {8, 1, 3, 3}, {8, 1, 2, 2}, {4, 2, 4, 4}, {4, 2, 3, 2}, {4, 2, 2, 2}, {4, 1, 4, 3}, {4, 1, 3, 2}, {4, 1, 2, 2}};
  static const int slots_env = getenv("DML_DUO_SLOTS") ? atoi(getenv("DML_DUO_SLOTS")) : 8;  // experiments: 4 = never more than four slots
  bool fit = false;
  for (int pass = 0; pass < (g_smem_aligned == 1 ? 2 : 1) && !fit; ++pass)  // second pass: without the 1 KiB of slack for aligning the base (conv3_block1: exactly 227 KiB)
    for (const auto& o : opts) {
      if (o[0] > slots_env) continue;
      const DuoSmem L = duo_smem(kmb, nj, a.nred, o[1], o[2], o[3], o[0]);
      const uint32_t need = pass == 0 ? L.total : L.total - 1024u;
      if (need <= (uint32_t)SMEM_LIMIT) {
        pl.slots = o[0]; pl.y2_bufs = o[1]; pl.we_stages = o[2]; pl.wr_stages = o[3]; pl.smem = need; pl.no_slack = pass == 1;
        fit = true;
        break;
      }
    }
  if (!fit) return tc_fail("expand + reduce kernel does not fit in shared memory (kmid %d, nexp %d, nred %d)", a.kmid, a.nexp, a.nred);
  pl.m_capacity = m_capacity;
  auto map2d = [&](CUtensorMap* m, const void* base, int cols, int rows, int pitch, int box_rows, bool weights) -> CUresult {
    cuuint64_t d[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t st[1] = {(cuuint64_t)pitch * 2};
    cuuint32_t b[2] = {64u, (cuuint32_t)box_rows};
    cuuint32_t es[2] = {1, 1};
    return g_encode_tiled(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), d, st, b, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_128B, weights ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  };
  CUresult r;
  if ((r = map2d(&pl.map_y2, a.y2, a.k1, m_capacity, a.y2_pitch, BLOCK_M, false)) != CUDA_SUCCESS) return tc_fail("duo: y2 map failed (%d)", (int)r);
  pl.map_s2 = pl.map_y2;
  if (a.src2 && a.s2_stride == 1 && (r = map2d(&pl.map_s2, a.src2, a.kmid - a.k1, m_capacity, a.src2_pitch, BLOCK_M, false)) != CUDA_SUCCESS)
    return tc_fail("duo: second-source map failed (%d)", (int)r);
  if (a.src2 && a.s2_stride != 1) {
    // the shortcut's input (C2, W * stride, H * stride, N) in im2col mode: 128 output pixels per load, element strides (1, s, s, 1)
    if (!g_encode_im2col) return tc_fail("tc_init was not called");
    const int c2 = a.kmid - a.k1, hi2 = a.h * a.s2_stride, wi2 = a.w * a.s2_stride, n_cap = m_capacity / (a.h * a.w);
    cuuint64_t g2dim[4] = {(cuuint64_t)c2, (cuuint64_t)wi2, (cuuint64_t)hi2, (cuuint64_t)n_cap};
    cuuint64_t g2str[3] = {(cuuint64_t)a.src2_pitch * 2, (cuuint64_t)wi2 * a.src2_pitch * 2, (cuuint64_t)hi2 * wi2 * a.src2_pitch * 2};
    int lo2[2] = {0, 0}, up2[2] = {0, 0};
    cuuint32_t es2[4] = {1, (cuuint32_t)a.s2_stride, (cuuint32_t)a.s2_stride, 1};
    r = g_encode_im2col(&pl.map_s2, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(a.src2), g2dim, g2str, lo2, up2, BLOCK_K, BLOCK_M, es2,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r == CUDA_SUCCESS && g_driver_version <= 13010 && (size_t)n_cap * hi2 * wi2 * a.src2_pitch * 2 < 131072)
      reinterpret_cast<uint64_t*>(&pl.map_s2)[1] &= ~(1ull << 21);
    if (r != CUDA_SUCCESS) return tc_fail("duo: strided second-source map failed (%d)", (int)r);
  }
  if ((r = map2d(&pl.map_we, a.we, a.kmid, a.nexp, a.kmid, 32, true)) != CUDA_SUCCESS) return tc_fail("duo: expand weight map failed (%d)", (int)r);
  if ((r = map2d(&pl.map_wr, a.wr, a.nexp, a.nred, a.nexp, a.nred / 2, true)) != CUDA_SUCCESS) return tc_fail("duo: reduce weight map failed (%d)", (int)r);
  if ((r = map2d(&pl.map_out, a.out, a.nexp, m_capacity, a.out_pitch, 16, false)) != CUDA_SUCCESS) return tc_fail("duo: output map failed (%d)", (int)r);
  pl.map_res = pl.map_out;
  if (a.res && (r = map2d(&pl.map_res, a.res, a.nexp, m_capacity, a.res_pitch, BLOCK_M, false)) != CUDA_SUCCESS) return tc_fail("duo: residual map failed (%d)", (int)r);
  pl.valid = true;
  *out = pl;
  return 0;
}

int launch_duo(const DuoPlan& pl, const DuoArgs& a, cudaStream_t s) {
  if (!pl.valid) return tc_fail("duo plan is not valid");
  if (a.M > pl.m_capacity) return tc_fail("%d pixels exceed the planned capacity %d", a.M, pl.m_capacity);
  DuoParams k{};
  k.bias2 = a.be; k.bias3 = a.br;
  k.y1n = reinterpret_cast<__nv_bfloat16*>(a.y1n); k.y1n_pitch = a.y1n_pitch;
  k.M = a.M; k.m_tiles = (a.M + 2 * BLOCK_M - 1) / (2 * BLOCK_M);
  k.kmb = a.kmid / 64; k.kb1 = a.k1 / 64; k.nj = a.nexp / 64; k.nred = a.nred;
  k.y2_bufs = pl.y2_bufs; k.we_stages = pl.we_stages; k.wr_stages = pl.wr_stages; k.ns = pl.slots;
  k.s2_stride = a.src2 ? a.s2_stride : 1; k.ho = a.h; k.wo = a.w; k.no_slack = pl.no_slack ? 1 : 0;
  k.rev = a.rev;
  static const int pf_env = getenv("DML_DUO_PF") ? atoi(getenv("DML_DUO_PF")) : 8;
  k.pf = pf_env;
  static const bool prof = getenv("DML_CHAIN_PROF") != nullptr;
  if (prof) {
    static long long zero[160 * 40] = {};
    cudaMemcpyToSymbol(g_chain_prof, zero, sizeof zero);
  }
  const int pairs = g_num_sms / 2;
  const int clusters = k.m_tiles < pairs ? k.m_tiles : pairs;
  static const bool dbg = getenv("DML_CHAIN_DEBUG") != nullptr;  // debugging aid shared with the chain kernel (wait ids are 64 + barrier index here)
  static unsigned long long* h_word = nullptr;
  if (dbg && !h_word) {
    unsigned long long* d_word = nullptr;
    if (cudaHostAlloc(&h_word, sizeof *h_word, cudaHostAllocMapped) == cudaSuccess && cudaHostGetDevicePointer(&d_word, h_word, 0) == cudaSuccess) {
      *h_word = 0;
      cudaMemcpyToSymbol(g_chain_dbg_ptr, &d_word, sizeof d_word);
    }
  }
  const dim3 grid(2 * clusters), block(DUO_THREADS);
  if (a.res) {
    if (prof) launch_pdl(conv_duo_kernel<true, true>, grid, block, pl.smem, s, pl.map_y2, pl.map_we, pl.map_wr, pl.map_res, pl.map_out, pl.map_s2, k);
    else launch_pdl(conv_duo_kernel<false, true>, grid, block, pl.smem, s, pl.map_y2, pl.map_we, pl.map_wr, pl.map_res, pl.map_out, pl.map_s2, k);
  } else {
    if (prof) launch_pdl(conv_duo_kernel<true, false>, grid, block, pl.smem, s, pl.map_y2, pl.map_we, pl.map_wr, pl.map_res, pl.map_out, pl.map_s2, k);
    else launch_pdl(conv_duo_kernel<false, false>, grid, block, pl.smem, s, pl.map_y2, pl.map_we, pl.map_wr, pl.map_res, pl.map_out, pl.map_s2, k);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return tc_fail("conv_duo_kernel launch: %s", cudaGetErrorString(e));
  if (prof) {
    cudaStreamSynchronize(s);
    static long long h[160 * 40];
    cudaMemcpyFromSymbol(h, g_chain_prof, sizeof h);
    double v[40] = {};
    double nt = 0, loop = 0, ntl = 0;
    for (int b = 0; b < 2 * clusters && b < 160; ++b) {
      for (int i = 0; i < 38; ++i) v[i] += (double)h[b * 40 + i];
      nt += (double)h[b * 40 + 39];
      if (b % 2 == 0) { loop += (double)h[b * 40 + 38]; ntl += (double)h[b * 40 + 39]; }
    }
    if (nt < 1) nt = 1;
    const char* names[18] = {"g2:y2_full", "res:y2_empty", "g2:we_full", "g3:wr_full", "B|w:acc2_full", "g2:acc2_empty", "B:res_full", "g3:outc_full",
                             "res|w:outc_empty", "res:slot_free", "A:acc3_full", "g3:acc3_empty", "B:compute", "B:store", "g2:issue", "g2:commit",
                             "g3:issue", "g3:commit"};
    fprintf(stderr, "[duo prof] kmid %d nexp %d, %d slots, y2 x%d, stages %d/%d: %.1f tiles per cluster, %.0f cycles per tile | cycles per tile and CTA (one warp per role; the issuers' are per pair: double them):", a.kmid, a.nexp,
            pl.slots, pl.y2_bufs, pl.we_stages, pl.wr_stages, ntl / clusters, loop / (ntl > 0 ? ntl : 1));
    for (int i = 0; i < 18; ++i)
      if (v[i] > 0) fprintf(stderr, " %s %.0f", names[i], v[i] / nt);
    fprintf(stderr, "\n");
  }
  if (dbg) {
    cudaError_t se = cudaStreamSynchronize(s);
    fprintf(stderr, "[duo debug] launch (M %d, kmid %d, nexp %d, nred %d, y2 bufs %d, stages %d/%d): sync -> %s, word %llx\n", a.M, a.kmid, a.nexp, a.nred,
            pl.y2_bufs, pl.we_stages, pl.wr_stages, cudaGetErrorString(se), h_word ? *h_word : 0ull);
    if (h_word && *h_word)
      fprintf(stderr, "[duo debug] wait id %llu timed out at index %llu, block %llu, warp %llu (%s)\n", *h_word & 0xff, (*h_word >> 8) & 0xffff,
              (*h_word >> 24) & 0xffff, (*h_word >> 40) & 0xff, cudaGetErrorString(se));
  }
  return 0;
}

// ---- conv_tc6p_kernel: plan + launch ----
int conv_pool_plan(ConvPoolPlan* out, const ConvParams& p, int hp, int wq, int pool_pad, int n_capacity) {
  if (p.cout != V6P_BN || p.stride != 1 || p.res != nullptr || p.in2 != nullptr || p.norelu_cols != 0)
    return tc_fail("conv + pool kernel: needs a stride-1 convolution with 64 filters and no residual");
  if (pool_pad != 0 && pool_pad != 1) return tc_fail("conv + pool kernel: pool padding %d", pool_pad);
  if (pool_pad && !p.relu) return tc_fail("conv + pool kernel: the pool's zero padding needs the ReLU in front of it");
  if (hp != (p.ho + 2 * pool_pad - 3) / 2 + 1 || wq != (p.wo + 2 * pool_pad - 3) / 2 + 1) return tc_fail("conv + pool kernel: pooled size %d x %d does not follow from %d x %d", hp, wq, p.ho, p.wo);
  ConvPoolPlan pl;
  TcTune tune;
  tune.version = 6;
  if (tc_plan_one(&pl.conv, p, n_capacity, tune) != 0) return -1;
  const TcOpPlan& c = pl.conv;
  if (c.version != 6 || c.bn != V6P_BN) return tc_fail("conv + pool kernel: the filter does not fit in one CTA's shared memory");
  const bool shape_ok = (p.kh == 4 && p.kw == 4 && c.cpb == 1 && c.row_bytes == 32) || (p.kh == 3 && p.kw == 3 && c.cpb == 1 && c.row_bytes == 64);
  if (!shape_ok) return tc_fail("conv + pool kernel: no instance for %dx%d taps, %d channel blocks of %d-byte rows", p.kh, p.kw, c.cpb, c.row_bytes);
  const int taps = p.kh * p.kw;
  const size_t b_bytes = ((size_t)taps * c.cpb * V6P_BN * c.row_bytes + 1023) / 1024 * 1024;
  pl.hp = hp; pl.wq = wq; pl.pool_pad = pool_pad; pl.a_stages = 2;
  // pooled rows per band: the most useful rows per issued MMA row among the band buffers that fit beside two halo stages.  A band
  // that continues its predecessor (most do: a CTA walks consecutive bands) computes 2R rows in T = ceil(2R * Wp / 128) tiles.
  double best = 0;
  for (int R = 1; R <= 16 && R <= hp; ++R) {
    const size_t band = (size_t)(2 * R + 1) * p.wo * 128;
    const size_t smem = 1024 + b_bytes + (size_t)pl.a_stages * c.a_stage_bytes + band + V6P_BN * 4 + 16 + 256;
    if (smem > (size_t)SMEM_LIMIT || (2 * R + 2) * c.wp + BLOCK_M >= 65536) break;
    const int T = (2 * R * c.wp + BLOCK_M - 1) / BLOCK_M;
    // new convolution rows per issued MMA row; half a tile per band for its pooling pass, row move and barriers
    const double eff = (double)2 * R * p.wo / (((double)T + 0.5) * BLOCK_M);
    if (eff >= best) { best = eff; pl.R = R; pl.T = T; pl.bands = (hp + R - 1) / R; pl.smem = smem; }
  }
  if (pl.R == 0) return tc_fail("conv + pool kernel: no band of pooled rows fits in shared memory (row of %d pixels)", p.wo);
  pl.valid = true;
  *out = pl;
  return 0;
}

int launch_conv_pool(const ConvPoolPlan& pl, const ConvParams& p, void* pool_out, int pool_out_pitch, int pool_out_img0, cudaStream_t s) {
  if (!pl.valid) return tc_fail("conv + pool plan is not valid");
  const TcOpPlan& c = pl.conv;
  if (p.n > c.m_capacity) return tc_fail("batch of %d images exceeds the planned capacity", p.n);
  if (pool_out_pitch % 8 || reinterpret_cast<uintptr_t>(pool_out) % 16) return tc_fail("conv + pool kernel: the pooled rows must be 16-byte aligned");
  TcV6pParams k{};
  k.bias = p.bias;
  k.out = reinterpret_cast<__nv_bfloat16*>(pool_out);
  k.n_img = p.n; k.ho = p.ho; k.wo = p.wo; k.wp = c.wp;
  k.hp = pl.hp; k.wq = pl.wq; k.out_pitch = pool_out_pitch; k.pool_pad = pl.pool_pad;
  k.pad_h = p.pad_h; k.pad_w = p.in_row_cells ? 0 : p.pad_w;
  k.R = pl.R; k.bands = pl.bands;
  k.rows_box = c.rows_box; k.a_cb_bytes = c.a_cb_bytes; k.a_stage_bytes = c.a_stage_bytes; k.a_stages = pl.a_stages;
  k.relu = p.relu; k.tmem_cols = 4 * V6P_BN;
  k.in_img0 = p.in_img0; k.out_img0 = pool_out_img0;
  k.last_kmma = (p.cin - (c.cpb - 1) * (c.row_bytes / 2) + 15) / 16;
  k.cin_pad = p.cin_pad;
  k.inv_wp = 1.0f / (float)c.wp;
  const int units = p.n * pl.bands;
  const int grid = units < g_num_sms ? units : g_num_sms;
  if (p.kh == 4) launch_pdl(conv_tc6p_kernel<4, 4, 1, 32>, dim3(grid), dim3(TC6_THREADS), pl.smem, s, c.map_a, c.map_b, k);
  else launch_pdl(conv_tc6p_kernel<3, 3, 1, 64>, dim3(grid), dim3(TC6_THREADS), pl.smem, s, c.map_a, c.map_b, k);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return tc_fail("conv_tc6p_kernel launch: %s", cudaGetErrorString(e));
  return 0;
}

namespace {
// milliseconds of one launch, or a huge number if the candidate cannot run: one warm-up, then the best of
// three bursts of four launches (the first bursts after an idle period run at ramping clocks)
float time_plan(const TcOpPlan& op, const ConvParams& p, cudaEvent_t e0, cudaEvent_t e1) {
  if (launch_conv_tc_plan(op, p, nullptr) != 0) return 1e30f;
  float best = 1e30f;
  static const int reps = getenv("DML_TUNE_REPS") ? atoi(getenv("DML_TUNE_REPS")) : 3;      // experiments only
  static const int burst = getenv("DML_TUNE_BURST") ? atoi(getenv("DML_TUNE_BURST")) : 4;
  for (int rep = 0; rep < reps; ++rep) {
    cudaEventRecord(e0, nullptr);
    for (int i = 0; i < burst; ++i)
      if (launch_conv_tc_plan(op, p, nullptr) != 0) return 1e30f;
    cudaEventRecord(e1, nullptr);
    if (cudaEventSynchronize(e1) != cudaSuccess) return 1e30f;
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms / (float)burst < best) best = ms / (float)burst;
  }
  return best;
}
}  // namespace

// ---- autotune results: remembered per (library, model, micro-batch, SM count) ----
// The table chosen by timing is kept in the process (a second engine of the same shape plans without timing anything, and gets
// the same kernels) and in a small text file (DML_TUNE_CACHE_DIR, default .tune_cache/ next to the library; DML_TUNE_CACHE=0 turns the file
// off), so that later processes on the same machine -- the ranks of a job, the next run -- start with it.  Timing itself is
// bounded by DML_TUNE_BUDGET_MS of wall time (default 30000): ops past the budget keep the heuristic plan.
struct TuneChoice {
  int version = 0, bn = 0, group = 0, res_direct = 0, chain = -1;  // chain: -1 not a chain lead, 0 separate, 1 fused
  int row_bytes = 0;
  float ms = 0.f;
};
static std::map<std::string, std::vector<TuneChoice>> g_tune_cache;

static std::string tune_key(const Net& net, int micro_batch) {
  char buf[160];
  snprintf(buf, sizeof buf, "v5_m%d_mb%d_ops%zu_sm%d_w%lld", net.model, micro_batch, net.ops.size(), g_num_sms, (long long)net.weights_bytes);
  return buf;
}
static std::string tune_file(const std::string& key) {
  const char* off = getenv("DML_TUNE_CACHE");
  if (off && off[0] == '0') return "";
  std::string dir;
  if (const char* d = getenv("DML_TUNE_CACHE_DIR")) {
    dir = d;
  } else {  // next to the library itself
    Dl_info info;
    if (!dladdr(reinterpret_cast<const void*>(&tune_key), &info) || !info.dli_fname) return "";
    const std::string lib = info.dli_fname;
    const size_t slash = lib.find_last_of('/');
    dir = (slash == std::string::npos ? std::string(".") : lib.substr(0, slash)) + "/.tune_cache";
  }
  mkdir(dir.c_str(), 0755);
  return dir + "/tune_" + key + ".txt";
}
static bool tune_load(const std::string& key, size_t nops, std::vector<TuneChoice>* out) {
  auto it = g_tune_cache.find(key);
  if (it != g_tune_cache.end() && it->second.size() == nops) { *out = it->second; return true; }
  const std::string fn = tune_file(key);
  if (fn.empty()) return false;
  std::ifstream f(fn);
  if (!f) return false;
  std::vector<TuneChoice> v(nops);
  size_t n = 0, idx;
  TuneChoice c;
  while (f >> idx >> c.version >> c.bn >> c.group >> c.res_direct >> c.chain >> c.row_bytes >> c.ms) {
    if (idx >= nops) return false;
    v[idx] = c;
    ++n;
  }
  if (n != nops) return false;
  g_tune_cache[key] = v;
  *out = v;
  return true;
}
static void tune_store(const std::string& key, const std::vector<TuneChoice>& v) {
  g_tune_cache[key] = v;
  const std::string fn = tune_file(key);
  if (fn.empty()) return;
  const std::string tmp = fn + "." + std::to_string((long long)getpid()) + ".tmp";
  {
    std::ofstream f(tmp);
    if (!f) return;
    for (size_t i = 0; i < v.size(); ++i)
      f << i << ' ' << v[i].version << ' ' << v[i].bn << ' ' << v[i].group << ' ' << v[i].res_direct << ' ' << v[i].chain << ' ' << v[i].row_bytes << ' ' << v[i].ms << '\n';
  }
  rename(tmp.c_str(), fn.c_str());  // atomic: concurrent ranks write the same table
}

// ChainArgs of the chain led by op `lead` over the engine's tensors (ptr(tensor id) -> device pointer of the view)
template <typename PtrFn>
static void chain_args(const Net& net, uint8_t* wts, int lead, int n, PtrFn ptr, ChainArgs* a) {
  const Op& o3 = net.ops[lead];
  const Op& oe = net.ops[o3.chain_exp];
  const Tensor& ty1 = net.tensors[o3.in];
  const Tensor& tout = net.tensors[oe.out];
  a->y1 = ptr(o3.in); a->y1_pitch = ty1.c_pitch;
  a->w3 = wts + o3.wpack_off; a->b3 = reinterpret_cast<const float*>(wts + o3.bias_off);
  a->we = wts + oe.wpack_off; a->be = reinterpret_cast<const float*>(wts + oe.bias_off);
  if (oe.res >= 0) { a->res = ptr(oe.res); a->res_pitch = net.tensors[oe.res].c_pitch; }
  if (oe.in2 >= 0) { a->src2 = ptr(oe.in2); a->src2_pitch = net.tensors[oe.in2].c_pitch; }
  a->out = ptr(oe.out); a->out_pitch = tout.c_pitch;
  if (o3.chain_red >= 0) {
    const Op& orr = net.ops[o3.chain_red];
    a->wr = wts + orr.wpack_off; a->br = reinterpret_cast<const float*>(wts + orr.bias_off);
    a->y1n = ptr(orr.out); a->y1n_pitch = net.tensors[orr.out].c_pitch;
  }
  a->n = n; a->h = ty1.h; a->w = ty1.w;
}

template <class PtrFn>
static void duo_args(const Net& net, uint8_t* wts, int lead, int n, PtrFn ptr, DuoArgs* a) {
  const Op& oe = net.ops[lead];
  const Op& orr = net.ops[oe.duo_red];
  const Tensor& ty2 = net.tensors[oe.in];
  a->y2 = ptr(oe.in); a->y2_pitch = ty2.c_pitch;
  a->we = wts + oe.wpack_off; a->be = reinterpret_cast<const float*>(wts + oe.bias_off);
  if (oe.res >= 0) { a->res = ptr(oe.res); a->res_pitch = net.tensors[oe.res].c_pitch; }
  if (oe.in2 >= 0) { a->src2 = ptr(oe.in2); a->src2_pitch = net.tensors[oe.in2].c_pitch; a->s2_stride = oe.stride2; }
  a->n = n; a->h = ty2.h; a->w = ty2.w;
  a->out = ptr(oe.out); a->out_pitch = net.tensors[oe.out].c_pitch;
  a->wr = wts + orr.wpack_off; a->br = reinterpret_cast<const float*>(wts + orr.bias_off);
  a->y1n = ptr(orr.out); a->y1n_pitch = net.tensors[orr.out].c_pitch;
  a->M = n * ty2.h * ty2.w;
  a->k1 = oe.cin; a->kmid = oe.cin + (oe.in2 >= 0 ? oe.cin2 : 0); a->nexp = oe.cout; a->nred = orr.cout;
}

int tc_plan(TcPlan* plan, const Net& net, uint8_t* ws, size_t off_arena, size_t off_input, size_t off_tc, uint8_t* wts,
            int micro_batch, int max_batch) {
  (void)off_input; (void)off_tc; (void)max_batch;
  plan->ops.assign(net.ops.size(), TcOpPlan());
  plan->micro_batch = micro_batch;
  const bool autotune = getenv("DML_NO_AUTOTUNE") == nullptr;
  const bool tune_log = getenv("DML_TUNE_LOG") != nullptr;
  bool warmed = false;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  if (autotune && (cudaEventCreate(&ev0) != cudaSuccess || cudaEventCreate(&ev1) != cudaSuccess))
    return tc_fail("autotune: cudaEventCreate failed");
  const std::string key = tune_key(net, micro_batch);
  std::vector<TuneChoice> cached, chosen(net.ops.size());
  const bool have_cache = autotune && !getenv("DML_TUNE_FRESH") && tune_load(key, net.ops.size(), &cached);
  if (tune_log && have_cache) fprintf(stderr, "[tune] using the remembered table %s\n", key.c_str());
  const double budget_ms = getenv("DML_TUNE_BUDGET_MS") ? atof(getenv("DML_TUNE_BUDGET_MS")) : 30000.0;
  const auto tune_t0 = std::chrono::steady_clock::now();
  bool over_budget = false;
  for (size_t i = 0; i < net.ops.size(); ++i) {
    const Op& op = net.ops[i];
    if (!tc_supported(op) || op.skip) continue;  // a follower of a merged 1x1 pair is computed by its leader
    const Tensor& ti = net.tensors[op.in];
    const Tensor& to = net.tensors[op.merged_out >= 0 ? op.merged_out : op.out];
    const int cout_x = op.merged_cout > 0 ? op.merged_cout : op.cout;
    if (ti.buf < 0) continue;
    ConvParams p{};
    p.in = ws + off_arena + ((size_t)net.bufs[ti.buf].offset * micro_batch + ti.c_off) * 2;
    p.w = wts + op.merged_w_off;
    p.out = ws + off_arena + ((size_t)net.bufs[to.buf].offset * micro_batch + to.c_off) * 2;
    if (op.res >= 0) {
      const Tensor& tr = net.tensors[op.res];
      p.res = ws + off_arena + ((size_t)net.bufs[tr.buf].offset * micro_batch + tr.c_off) * 2;
      p.res_pitch = tr.c_pitch;
    }
    p.n = micro_batch; p.hi = ti.h; p.wi = ti.w; p.cin = op.cin; p.in_pitch = ti.c_pitch;
    p.ho = to.h; p.wo = to.w; p.cout = cout_x; p.out_pitch = to.c_pitch;
    p.kh = op.kh; p.kw = op.kw; p.stride = op.stride; p.pad_h = op.pad_h; p.pad_w = op.pad_w;
    p.cin_pad = op.cin_pad;
    p.bias = reinterpret_cast<const float*>(wts + op.merged_b_off);
    p.relu = op.relu;
    p.norelu_cols = op.norelu_cols;
    if (op.in2 >= 0) {
      const Tensor& t2 = net.tensors[op.in2];
      if (t2.buf < 0) return tc_fail("second GEMM source of %s is not an arena tensor", op.name.c_str());
      p.in2 = ws + off_arena + ((size_t)net.bufs[t2.buf].offset * micro_batch + t2.c_off) * 2;
      p.hi2 = t2.h; p.wi2 = t2.w; p.cin2 = op.cin2; p.in2_pitch = t2.c_pitch; p.stride2 = op.stride2; p.cin2_pad = op.cin2_pad;
    }
    if (have_cache && cached[i].version != 0) {
      TcTune t;
      t.version = cached[i].version; t.bn = cached[i].bn; t.group = cached[i].group; t.res_direct = cached[i].res_direct;
      t.row_bytes = cached[i].version >= 6 && cached[i].row_bytes == 64 ? 64 : 0;
      if (tc_plan_one(&plan->ops[i], p, micro_batch, t) == 0 && plan->ops[i].version == cached[i].version && plan->ops[i].bn == cached[i].bn) {
        plan->ops[i].best_ms = cached[i].ms;
        chosen[i] = cached[i];
        continue;
      }
    }
    if (tc_plan_one(&plan->ops[i], p, micro_batch, TcTune()) != 0) return -1;
    if (!autotune) continue;
    if (!over_budget && std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tune_t0).count() > budget_ms) {
      over_budget = true;
      if (tune_log) fprintf(stderr, "[tune] time budget of %.0f ms used up at %s: the remaining layers keep the heuristic plan\n", budget_ms, op.name.c_str());
    }
    if (over_budget) continue;
    // Time the candidate kernels / N tiles on the real buffers (contents do not matter) and keep the
    // fastest.  Every candidate produces bit-identical results (tests/test_gpu_kernels.py).
    if (!warmed) {  // bring the clocks up before the first comparison (~20 ms of launches)
      for (int w = 0; w < 200; ++w)
        if (launch_conv_tc_plan(plan->ops[i], p, nullptr) != 0) break;
      cudaDeviceSynchronize();
      warmed = true;
    }
    float best = time_plan(plan->ops[i], p, ev0, ev1);
    const float first = best;
    // candidates: kernel variant x N tile x k-blocks per stage (v2 / v4); duplicates of an already timed plan are skipped
    std::vector<std::array<int, 4>> seen;
    seen.push_back({plan->ops[i].version, plan->ops[i].bn, plan->ops[i].group, plan->ops[i].row_bytes});
    for (int version : {4, 2, 5, 6, 7}) {
      if (op.in2 >= 0 && version != 4) continue;  // only the pair kernel reads a second GEMM source
      if (op.norelu_cols > 0 && version != 4 && version != 2) continue;  // ReLU-free leading columns: persistent GEMM kernels only
      for (int bn : {0, 256, 128, 64}) {
        if (version == 5 && (bn != 0 || cout_x > 128 || op.res >= 0)) continue;
        if (version == 6 && bn != 0) continue;
        if (bn && (cout_x % bn != 0 || bn > cout_x)) continue;
        for (int g : {0, 1, 2, 3, 4, 64}) {  // 64: the halo kernels with 64-byte rows (32-channel K blocks) where that pads Cin less
          if (version >= 5 && g != 0 && g != 64) continue;
          if (g == 64 && (version < 6 || op.cin <= 32 || (op.cin + 31) / 32 * 32 >= (op.cin + 63) / 64 * 64)) continue;
          TcTune t;
          t.version = version; t.bn = bn; t.group = g == 64 ? 0 : g; t.row_bytes = g == 64 ? 64 : 0;
          TcOpPlan cand;
          if (tc_plan_one(&cand, p, micro_batch, t) != 0) continue;
          const std::array<int, 4> key = {cand.version, cand.bn, cand.group, cand.row_bytes};
          bool dup = false;
          for (const auto& k : seen) dup = dup || k == key;
          if (dup) continue;
          seen.push_back(key);
          float ms = time_plan(cand, p, ev0, ev1);
          if (tune_log) fprintf(stderr, "[tune]   %s v%d bn %d group %d stages %d row %d: %.4f ms\n", op.name.c_str(), cand.version, cand.bn, cand.group, cand.stages, cand.row_bytes, ms);
          if (ms < best * 0.98f) { best = ms; plan->ops[i] = cand; }
        }
      }
    }
    if (op.res >= 0) {  // residual layers: also with two whole residual tile buffers instead of the chunked single one
      for (int bn : {0, 256, 128}) {
        if (bn && (cout_x % bn != 0 || bn > cout_x)) continue;
        TcTune t;
        t.version = 4; t.bn = bn; t.res_direct = 1;
        TcOpPlan cand;
        if (tc_plan_one(&cand, p, micro_batch, t) != 0) continue;
        float ms = time_plan(cand, p, ev0, ev1);
        if (tune_log) fprintf(stderr, "[tune]   %s v%d bn %d group %d stages %d (two residual tile buffers): %.4f ms\n", op.name.c_str(), cand.version, cand.bn, cand.group, cand.stages, ms);
        if (ms < best * 0.98f) { best = ms; plan->ops[i] = cand; }
      }
    }
    plan->ops[i].best_ms = best;
    chosen[i].version = plan->ops[i].version; chosen[i].bn = plan->ops[i].bn; chosen[i].group = plan->ops[i].group;
    chosen[i].row_bytes = plan->ops[i].row_bytes;
    chosen[i].res_direct = plan->ops[i].has_res ? (plan->ops[i].res_direct ? 2 : 1) : 0; chosen[i].ms = best;
    if (tune_log)
      fprintf(stderr, "[tune] %s -> v%d bn %d group %d stages %d: %.4f ms (heuristic %.4f)\n", op.name.c_str(), plan->ops[i].version,
              plan->ops[i].bn, plan->ops[i].group, plan->ops[i].stages, best, first);
  }
  // ---- bottleneck-tail chains (conv_chain.inc): fused kernel against the three separate ones ----
  plan->chains.assign(net.ops.size(), ChainPlan());
  plan->chained.assign(net.ops.size(), 0);
  if (!getenv("DML_NO_CHAIN")) {
    auto tptr = [&](int t) -> uint8_t* {
      const Tensor& tt = net.tensors[t];
      return ws + off_arena + ((size_t)net.bufs[tt.buf].offset * micro_batch + tt.c_off) * 2;
    };
    for (size_t i = 0; i < net.ops.size(); ++i) {
      const Op& o3 = net.ops[i];
      if (o3.chain_exp < 0) continue;
      const Op& oe = net.ops[o3.chain_exp];
      const Tensor& ty1 = net.tensors[o3.in];
      if (ty1.buf < 0 || !plan->ops[i].valid || !plan->ops[o3.chain_exp].valid) continue;
      ChainArgs a;
      chain_args(net, wts, (int)i, micro_batch, tptr, &a);
      ChainPlan cp;
      if (chain_plan(&cp, a, micro_batch) != 0) {
        if (tune_log) fprintf(stderr, "[tune] chain at %s not planned: %s\n", o3.name.c_str(), tc_last_error());
        continue;
      }
      bool use = true;
      if (have_cache && cached[i].chain >= 0) {
        use = cached[i].chain == 1;
        chosen[i].chain = cached[i].chain;
      } else if (autotune && !over_budget) {
        float best = 1e30f;
        if (launch_chain(cp, a, nullptr) != 0 || cudaDeviceSynchronize() != cudaSuccess) return tc_fail("chain kernel failed at %s: %s", o3.name.c_str(), cudaGetErrorString(cudaGetLastError()));
        for (int rep = 0; rep < 3; ++rep) {
          cudaEventRecord(ev0, nullptr);
          for (int k = 0; k < 4; ++k) launch_chain(cp, a, nullptr);
          cudaEventRecord(ev1, nullptr);
          if (cudaEventSynchronize(ev1) != cudaSuccess) return tc_fail("chain kernel failed at %s: %s", o3.name.c_str(), cudaGetErrorString(cudaGetLastError()));
          float ms = 0.f;
          cudaEventElapsedTime(&ms, ev0, ev1);
          if (ms / 4.f < best) best = ms / 4.f;
        }
        cp.ms = best;
        cp.ms_separate = plan->ops[i].best_ms + plan->ops[o3.chain_exp].best_ms + (o3.chain_red >= 0 ? plan->ops[o3.chain_red].best_ms : 0.f);
        use = cp.ms < cp.ms_separate || getenv("DML_FORCE_CHAIN");
        if (tune_log)
          fprintf(stderr, "[tune] chain %s + %s%s%s: fused %.4f ms, separate %.4f ms -> %s\n", o3.name.c_str(), oe.name.c_str(),
                  o3.chain_red >= 0 ? " + " : "", o3.chain_red >= 0 ? net.ops[o3.chain_red].name.c_str() : "", cp.ms, cp.ms_separate,
                  use ? "fused" : "separate");
        chosen[i].chain = use ? 1 : 0;
      }
      if (!use) continue;
      plan->chains[i] = cp;
      plan->chained[o3.chain_exp] = 1;
      if (o3.chain_red >= 0) plan->chained[o3.chain_red] = 1;
    }
  }
  // ---- expand + next reduce duos (conv_duo.inc): fused kernel against the two separate ones ----
  plan->duos.assign(net.ops.size(), DuoPlan());
  if (!getenv("DML_NO_DUO")) {
    auto tptr = [&](int t) -> uint8_t* {
      const Tensor& tt = net.tensors[t];
      return ws + off_arena + ((size_t)net.bufs[tt.buf].offset * micro_batch + tt.c_off) * 2;
    };
    for (size_t i = 0; i < net.ops.size(); ++i) {
      const Op& oe = net.ops[i];
      if (oe.duo_red < 0 || plan->chained[i] || plan->chained[oe.duo_red]) continue;
      if (net.tensors[oe.in].buf < 0 || !plan->ops[i].valid || !plan->ops[oe.duo_red].valid || oe.cin_pad != oe.cin) continue;
      if (oe.in2 >= 0 && (oe.cin2_pad != oe.cin2 || net.tensors[oe.in2].buf < 0)) continue;
      DuoArgs a;
      duo_args(net, wts, (int)i, micro_batch, tptr, &a);
      DuoPlan dp;
      if (duo_plan(&dp, a, a.M) != 0) {
        if (tune_log) fprintf(stderr, "[tune] duo at %s not planned: %s\n", oe.name.c_str(), tc_last_error());
        continue;
      }
      bool use = true;
      if (have_cache && cached[i].chain >= 0) {
        use = cached[i].chain == 1;
        chosen[i].chain = cached[i].chain;
      } else if (autotune && !over_budget) {
        float best = 1e30f;
        if (launch_duo(dp, a, nullptr) != 0 || cudaDeviceSynchronize() != cudaSuccess) return tc_fail("duo kernel failed at %s: %s", oe.name.c_str(), cudaGetErrorString(cudaGetLastError()));
        for (int rep = 0; rep < 3; ++rep) {
          cudaEventRecord(ev0, nullptr);
          for (int k = 0; k < 4; ++k) launch_duo(dp, a, nullptr);
          cudaEventRecord(ev1, nullptr);
          if (cudaEventSynchronize(ev1) != cudaSuccess) return tc_fail("duo kernel failed at %s: %s", oe.name.c_str(), cudaGetErrorString(cudaGetLastError()));
          float ms = 0.f;
          cudaEventElapsedTime(&ms, ev0, ev1);
          if (ms / 4.f < best) best = ms / 4.f;
        }
        dp.ms = best;
        dp.ms_separate = plan->ops[i].best_ms + plan->ops[oe.duo_red].best_ms;
        use = dp.ms < dp.ms_separate || getenv("DML_FORCE_DUO");
        if (tune_log)
          fprintf(stderr, "[tune] duo %s + %s: fused %.4f ms, separate %.4f ms -> %s\n", oe.name.c_str(), net.ops[oe.duo_red].name.c_str(), dp.ms,
                  dp.ms_separate, use ? "fused" : "separate");
        chosen[i].chain = use ? 1 : 0;
      }
      if (!use) continue;
      plan->duos[i] = dp;
      plan->chained[oe.duo_red] = 1;
    }
  }
  if (autotune && !have_cache && !over_budget) tune_store(key, chosen);
  if (autotune) {
    cudaEventDestroy(ev0);
    cudaEventDestroy(ev1);
    if (cudaDeviceSynchronize() != cudaSuccess) return tc_fail("autotune: %s", cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}

int launch_chain_tc(const TcPlan& plan, const Net& net, uint8_t* wts, int lead, int n, uint8_t* ws, size_t off_arena, cudaStream_t s, int rev) {
  if (lead < 0 || lead >= (int)plan.chains.size() || !plan.chains[lead].valid) return tc_fail("no chain plan for op %d", lead);
  auto tptr = [&](int t) -> uint8_t* {
    const Tensor& tt = net.tensors[t];
    return ws + off_arena + ((size_t)net.bufs[tt.buf].offset * plan.micro_batch + tt.c_off) * 2;
  };
  ChainArgs a;
  chain_args(net, wts, lead, n, tptr, &a);
  a.rev = rev;
  return launch_chain(plan.chains[lead], a, s);
}

int launch_duo_tc(const TcPlan& plan, const Net& net, uint8_t* wts, int lead, int n, uint8_t* ws, size_t off_arena, cudaStream_t s, int rev) {
  if (lead < 0 || lead >= (int)plan.duos.size() || !plan.duos[lead].valid) return tc_fail("no duo plan for op %d", lead);
  auto tptr = [&](int t) -> uint8_t* {
    const Tensor& tt = net.tensors[t];
    return ws + off_arena + ((size_t)net.bufs[tt.buf].offset * plan.micro_batch + tt.c_off) * 2;
  };
  DuoArgs a;
  duo_args(net, wts, lead, n, tptr, &a);
  a.rev = rev;
  return launch_duo(plan.duos[lead], a, s);
}

int launch_conv_tc(const TcPlan& plan, int op_index, const ConvParams& p, int img0, cudaStream_t s, int* nlaunch) {
  (void)img0;
  if (op_index < 0 || op_index >= (int)plan.ops.size() || !plan.ops[op_index].valid)
    return tc_fail("no tcgen05 plan for op %d", op_index);
  if (nlaunch) *nlaunch = 1;
  return launch_conv_tc_plan(plan.ops[op_index], p, s);
}

}  // namespace dml
