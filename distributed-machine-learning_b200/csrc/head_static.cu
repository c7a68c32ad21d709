// head_static.cu -- dml_head_static: the dense(1000) + softmax + top-5 head for
// callers that must not allocate device memory.
//
// The regular head (dml_forward / dml_infer) keeps its n x 1000 fp32 logits in the
// workspace the caller allocated and bound to an engine.  Some callers have no way
// to provide that: code running between cudaStreamBeginCapture/EndCapture, a
// real-time serving thread that may not take cudaMalloc's implicit device-wide
// synchronisation, or a plugin loaded into a process whose memory is entirely owned
// by another framework's pool.  For them the logits live in a fixed array that is
// part of this library's device image.
#include "../../include/dml_b200.h"

#include <cuda_runtime.h>

#include <mutex>
#include <string>

#include "kernels.h"

namespace {

// DML_HEAD_STATIC_MAX_IMAGES x 1000 fp32 = 4,096,000 bytes, resident on every GPU that
// loads this library until the process exits.
static __device__ __align__(256) float g_head_logits[DML_HEAD_STATIC_MAX_IMAGES * DML_NUM_CLASSES];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
float* g_scratch_ptr[kMaxDevices] = {};  // device address of g_head_logits, per device

}  // namespace

extern "C" int dml_set_error(int code, const char* msg);

extern "C" int dml_head_static(const float* d_pooled, const float* d_dense_w, const float* d_dense_b, int n, int cin,
                               float* d_probs, int32_t* d_top5_idx, float* d_top5_prob, void* stream) {
  if (n < 0) return dml_set_error(DML_EINVAL, "dml_head_static: n is negative");
  if (n > DML_HEAD_STATIC_MAX_IMAGES)
    return dml_set_error(DML_ESIZE, ("dml_head_static: n = " + std::to_string(n) + " exceeds the static scratch limit of " +
                                     std::to_string(DML_HEAD_STATIC_MAX_IMAGES) + " images per call").c_str());
  if (cin < 8 || cin % 8 != 0 || cin > 3072)
    return dml_set_error(DML_ESIZE, "dml_head_static: cin must be a multiple of 8 in [8, 3072]");
  if (n == 0) return 0;
  if (!d_pooled || !d_dense_w || !d_dense_b || !d_top5_idx || !d_top5_prob)
    return dml_set_error(DML_EINVAL, "dml_head_static: NULL device pointer");
  int dev = -1;
  cudaError_t ce = cudaGetDevice(&dev);
  if (ce != cudaSuccess) return dml_set_error(DML_ECUDA, cudaGetErrorString(ce));
  if (dev < 0 || dev >= kMaxDevices) return dml_set_error(DML_EINVAL, "dml_head_static: device index out of range");
  float* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_ptr[dev]) {
      // One-time set-up for this device (it is the current one): the array's name is only a
      // host-side handle, the kernels need its device address.
      void* p = nullptr;
      ce = cudaGetSymbolAddress(&p, g_head_logits);
      if (ce != cudaSuccess)
        return dml_set_error(DML_ECUDA, (std::string("cudaGetSymbolAddress(g_head_logits): ") + cudaGetErrorString(ce)).c_str());
      dml::init_kernels();
      g_scratch_ptr[dev] = static_cast<float*>(p);
    }
    scratch = g_scratch_ptr[dev];
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // Rows [0, n) of the scratch are fully rewritten by the dense kernel before the softmax kernel
  // reads them; nothing in it has to start at zero.
  dml::launch_fc(d_pooled, d_dense_w, d_dense_b, scratch, n, cin, true, s);  // the default (bf16) engines' dense kernel
  dml::launch_softmax_top5(scratch, d_probs, d_top5_idx, d_top5_prob, n, s);
  ce = cudaGetLastError();
  if (ce != cudaSuccess) return dml_set_error(DML_ECUDA, cudaGetErrorString(ce));
  return 0;
}
