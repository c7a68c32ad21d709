// tma_mc_bench.cu -- is the per-SM TMA cost paid by the ISSUER only?  A cluster of CS CTAs streams 16 KiB boxes
// (128 rows x 128 B, SWIZZLE_128B) into shared memory, (a) every CTA loading all of its own boxes, (b) the boxes
// split between the CTAs and multicast to the whole cluster.  Prints cycles per box RECEIVED by an SM.
// Build + run on the GPU box:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/tma_mc tools/tma_mc_bench.cu -lcuda && /tmp/tma_mc
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {  // bounded: a protocol slip traps instead of hanging
  uint32_t done = 0;
  const long long t0 = clock64();
  while (!done) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (!done && clock64() - t0 > 2000000000LL) __trap();
  }
}
__device__ __forceinline__ void cluster_barrier() {
  asm volatile("barrier.cluster.arrive.release;\n\tbarrier.cluster.wait.acquire;" ::: "memory");
}

constexpr int SLOTS = 8, BOX_BYTES = 16384;

template <int CS>
__global__ void __launch_bounds__(128, 1) bench(const __grid_constant__ CUtensorMap map, int boxes, int multicast, int rows_total,
                                                long long* out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full[SLOTS];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  if (threadIdx.x == 0) {
    for (int i = 0; i < SLOTS; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&full[i])) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  {
    const long long t0 = clock64();
    // every CTA receives `boxes` boxes; slot s of round r is waited for before it is refilled (no consumer: the data is dropped).
    // All threads walk the loop (thread 0 issues) so that the cluster can re-synchronise every SLOTS / 2 boxes: a CTA never
    // runs a whole ring ahead of a peer whose slots it writes with multicast.
    for (int b = 0; b < boxes; ++b) {
      if (b && b % (SLOTS / 2) == 0) cluster_barrier();
      if (threadIdx.x != 0) continue;
      const int s = b % SLOTS;
      if (b >= SLOTS) mbar_wait(smem_u32(&full[s]), ((b / SLOTS) - 1) & 1);
      // NOTE: with multicast the other CTAs write into this slot too; their waits keep them within SLOTS boxes of us
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&full[s])), "r"(BOX_BYTES) : "memory");
      const int row = ((blockIdx.x * 131 + b) * 128) % rows_total;
      if (!multicast) {
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
                         base + s * BOX_BYTES),
                     "l"(&map), "r"(smem_u32(&full[s])), "r"(0), "r"(row)
                     : "memory");
      } else if (b % CS == (int)rank) {
        const int mrow = (((blockIdx.x / CS) * 131 + b) * 128) % rows_total;
        asm volatile(
            "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(
                base + s * BOX_BYTES),
            "l"(&map), "r"(smem_u32(&full[s])), "r"(0), "r"(mrow), "h"((uint16_t)((1u << CS) - 1u))
            : "memory");
      }
    }
    if (threadIdx.x == 0) {
      for (int b = boxes < SLOTS ? 0 : boxes - SLOTS; b < boxes; ++b) mbar_wait(smem_u32(&full[b % SLOTS]), (b / SLOTS) & 1);
      out[blockIdx.x] = clock64() - t0;
    }
  }
  __syncthreads();
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <int CS>
void run(const CUtensorMap& map, int rows_total, long long* d) {
  cudaFuncSetAttribute(bench<CS>, cudaFuncAttributeMaxDynamicSharedMemorySize, SLOTS * BOX_BYTES + 1024);
  cudaFuncSetAttribute(bench<CS>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  const int boxes = 2048;
  for (int mc = 0; mc < 2; ++mc) {
    const int grid = (148 / CS) * CS;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = SLOTS * BOX_BYTES + 1024;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    for (int rep = 0; rep < 2; ++rep) {
      cudaError_t e = cudaLaunchKernelEx(&cfg, bench<CS>, map, boxes, mc, rows_total, d);
      if (e != cudaSuccess) { printf("cluster %d launch: %s\n", CS, cudaGetErrorString(e)); return; }
    }
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("error %s\n", cudaGetErrorString(cudaGetLastError())); exit(1); }
    long long h[148];
    cudaMemcpy(h, d, grid * sizeof(long long), cudaMemcpyDeviceToHost);
    double mx = 0;
    for (int i = 0; i < grid; ++i) mx = h[i] > mx ? (double)h[i] : mx;
    printf("cluster %d %s: %.0f cycles per 16 KiB box received (%.1f B/clk/SM), %d CTAs\n", CS, mc ? "multicast" : "unicast  ", mx / boxes,
           BOX_BYTES / (mx / boxes), grid);
  }
}

int main() {
  const int rows_total = 1 << 18;  // 32 MiB of bf16 [rows][64]: L2 resident
  void* buf;
  cudaMalloc(&buf, (size_t)rows_total * 128);
  cudaMemset(buf, 0, (size_t)rows_total * 128);
  long long* d;
  cudaMalloc(&d, 148 * sizeof(long long));
  EncodeTiled enc = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&enc, cudaEnableDefault, &q);
  CUtensorMap map;
  cuuint64_t dim[2] = {64, (cuuint64_t)rows_total};
  cuuint64_t str[1] = {128};
  cuuint32_t box[2] = {64, 128};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, buf, dim, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
  run<1>(map, rows_total, d);
  run<2>(map, rows_total, d);
  run<4>(map, rows_total, d);
  return 0;
}
