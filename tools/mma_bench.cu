// mma_bench.cu -- how long does one tcgen05.mma take on B200, by N and by shared-memory operand layout?
// Build + run on the GPU box:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/mma_bench tools/mma_bench.cu && /tmp/mma_bench
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// layout: 2 = SWIZZLE_128B (rows 128 B, SBO 1024), 4 = SWIZZLE_64B (rows 64 B, SBO 512), 6 = SWIZZLE_32B (rows 32 B, SBO 256)
__device__ __forceinline__ uint64_t make_desc(uint32_t addr, int layout) {
  const uint32_t sbo = layout == 2 ? 1024 : (layout == 4 ? 512 : 256);
  return (uint64_t)((addr >> 4) & 0x3FFF) | ((uint64_t)1 << 16) | ((uint64_t)(sbo >> 4) << 32) | ((uint64_t)1 << 46) |
         ((uint64_t)layout << 61);
}

__global__ void __launch_bounds__(128, 1) bench(int n, int layout, int iters, int a_from_rows, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint32_t tmem_slot;
  __shared__ uint64_t bar;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem_raw + (base - smem_u32(smem_raw)))[i] = 0;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_slot;
  if (threadIdx.x < 32) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint32_t row_bytes = layout == 2 ? 128 : (layout == 4 ? 64 : 32);
    const uint32_t a_bytes = 128 * row_bytes, b_bytes = (uint32_t)n * row_bytes;
    const int ksteps = row_bytes / 32;  // K=16 slices inside one tile
    long long t0 = clock64();
    uint32_t pred;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
    if (pred) {
      for (int i = 0; i < iters; ++i) {
        const uint32_t slot = (uint32_t)(i / ksteps) & 1u;  // alternate between two operand buffers
        // mode 0: two fixed operand buffers, 1024-byte aligned; mode m > 0: the A start is shifted by (m - 1) rows
        // (the halo kernels' shifted views: row-aligned only), B stays aligned
        const uint32_t a_addr = base + slot * (a_bytes + b_bytes + 1024u) + (a_from_rows ? (uint32_t)(a_from_rows - 1) * row_bytes : 0u);
        const uint32_t b_addr = base + slot * (a_bytes + b_bytes + 1024u) + a_bytes + 1024u;
        const uint64_t da = make_desc(a_addr, layout) + (uint64_t)(2 * (i % ksteps));
        const uint64_t db = make_desc(b_addr, layout) + (uint64_t)(2 * (i % ksteps));
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem),
            "l"(da), "l"(db), "r"(idesc), "r"((uint32_t)(i != 0))
            : "memory");
      }
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    }
    __syncwarp();
    uint32_t done = 0;
    while (!done)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                   : "=r"(done) : "r"(smem_u32(&bar)), "r"(0u) : "memory");
    long long t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
}

// The MMA warp's per-k-block loop of the conv kernels, piece by piece: each iteration issues 4 MMAs (one 64-channel
// k-block) and optionally (bit 0) reads clock64, (bit 1) try_waits on an already completed mbarrier phase,
// (bit 2) tcgen05.fence::after_thread_sync, (bit 3) tcgen05.commit to a scratch mbarrier, (bit 4) elects the
// issuing lane per iteration + __syncwarp (otherwise one lane runs the whole loop).
template <int OPS>
__global__ void __launch_bounds__(128, 1) loop_bench(int n, int iters, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint32_t tmem_slot;
  __shared__ uint64_t bar, ready, scratch;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem_raw + (base - smem_u32(smem_raw)))[i] = 0;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)) : "memory");
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&ready)) : "memory");
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1000000;" ::"r"(smem_u32(&scratch)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&ready)) : "memory");  // phase 0 of `ready` is complete
  }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_slot;
  if (threadIdx.x < 32) {  // the whole warp walks the loop; one elected lane issues (as in the conv kernels)
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint32_t a_bytes = 128 * 128, b_bytes = (uint32_t)n * 128;
    const uint32_t ready_a = smem_u32(&ready), scratch_a = smem_u32(&scratch);
    long long t0 = clock64();
    long long sink = 0;
    for (int i = 0; i < iters; ++i) {
      const uint32_t slot = (uint32_t)i & 1u;
      if (OPS & 1) sink += clock64();
      if (OPS & 2) {
        uint32_t done = 0;
        while (!done)
          asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                       : "=r"(done) : "r"(ready_a), "r"(0u) : "memory");
      }
      if (OPS & 4) asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint64_t da = make_desc(base + slot * (a_bytes + b_bytes), 2);
      const uint64_t db = make_desc(base + slot * (a_bytes + b_bytes) + a_bytes, 2);
      uint32_t lane_ok;
      asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(lane_ok));
      if (lane_ok) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          asm volatile(
              "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
              "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem),
              "l"(da + (uint64_t)(2 * k)), "l"(db + (uint64_t)(2 * k)), "r"(idesc), "r"((uint32_t)((i | k) != 0))
              : "memory");
        if (OPS & 8)
          asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(scratch_a) : "memory");
      }
      __syncwarp();
    }
    uint32_t last;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(last));
    if (last) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    __syncwarp();
    uint32_t done = 0;
    while (!done)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                   : "=r"(done) : "r"(smem_u32(&bar)), "r"(0u) : "memory");
    long long t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0 + (sink == 12345 ? 1 : 0);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
}

template <int OPS>
void run_loop(int n, long long* d) {
  cudaFuncSetAttribute(loop_bench<OPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int rep = 0; rep < 2; ++rep) loop_bench<OPS><<<148, 128, 180 * 1024>>>(n, 1000, d);
  if (cudaDeviceSynchronize() != cudaSuccess) { printf("error %s\n", cudaGetErrorString(cudaGetLastError())); exit(1); }
  long long h[148];
  cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost);
  double mx = 0;
  for (int i = 0; i < 148; ++i) mx = h[i] > mx ? (double)h[i] : mx;
  printf("N %d ops %2d: %.1f\n", n, OPS, mx / 1000);
}

int main() {
  long long* d;
  cudaMalloc(&d, 148 * sizeof(long long));
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int iters = 4000;
  printf("layout N cycles_per_mma MAC_per_cycle_per_SM\n");
  for (int mode : {0, 2, 4, 5, 8})
  for (int layout : {2, 4})
    for (int n : {32, 64, 128, 256}) {
      printf("A shifted by %d rows: ", mode ? mode - 1 : 0);
      for (int rep = 0; rep < 2; ++rep) bench<<<148, 128, 180 * 1024>>>(n, layout, iters, mode, d);
      if (cudaDeviceSynchronize() != cudaSuccess) { printf("error %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
      long long h[148];
      cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost);
      double mx = 0;
      for (int i = 0; i < 148; ++i) mx = h[i] > mx ? (double)h[i] : mx;
      printf("%s %d %.1f %.0f\n", layout == 2 ? "SW128" : (layout == 4 ? "SW64" : "SW32"), n, mx / iters, 128.0 * n * 16 / (mx / iters));
    }
  printf("\nloop pieces: cycles per k-block iteration (4 MMAs); ops bits: 1 clock64, 2 try_wait(ready), 4 fence::after, 8 commit\n");
  for (int n : {64, 128, 256}) {
    run_loop<0>(n, d); run_loop<1>(n, d); run_loop<2>(n, d); run_loop<4>(n, d); run_loop<8>(n, d);
    run_loop<6>(n, d); run_loop<14>(n, d); run_loop<15>(n, d);
  }
  return 0;
}
