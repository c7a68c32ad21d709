/*
 * dml_b200.h -- C ABI of libdml_b200.so, the B200 (sm_100a) batch-inference
 * engine behind the reference's models.py predict path.
 *
 * Every entry point replaces a piece of what /root/reference/models.py does
 * through TensorFlow/Keras; the citation on each declaration names the
 * reference line(s) it stands in for.  Plain pointers and sizes only; no torch
 * types.  All functions return 0 on success and a negative DML_E* code on
 * failure; dml_last_error() gives the message (thread-local).
 *
 * Ownership: the caller (Python + torch in this repo) allocates every device
 * buffer and passes raw device pointers; the library allocates no device
 * memory in any of these entry points.  Kernels are enqueued on the stream the
 * caller passes (a cudaStream_t cast to void*; NULL = default stream) and the
 * calls return without waiting, except dml_sync and dml_load_weights.
 */
#ifndef DML_B200_H_
#define DML_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dml_engine dml_engine;

enum {
  DML_OK = 0,
  DML_EINVAL = -1,   /* bad argument */
  DML_ESTATE = -2,   /* call order: workspace/weights not bound or loaded */
  DML_ECUDA = -3,    /* a CUDA runtime/driver call failed */
  DML_ESIZE = -4,    /* size beyond a stated limit */
  DML_ECANCELLED = -5 /* dml_cancel was called: the call stopped between two micro-batches */
};

enum { DML_MODEL_RESNET50 = 0, DML_MODEL_INCEPTIONV3 = 1 };

/* dml_create flags */
enum {
  DML_FLAG_FP32_CHECK = 1,   /* fp32 activations+weights, CUDA-core convs (the "fp32-accumulate check mode") */
  DML_FLAG_NO_TENSOR_CORE = 2, /* bf16 storage but CUDA-core convs (debug twin of the tcgen05 path) */
  DML_FLAG_KEEP_ALL = 4        /* no activation-buffer reuse: every tensor stays readable (parity taps) */
};

#define DML_NUM_CLASSES 1000
#define DML_TOP 5

/* ---- model description (no GPU needed) ---------------------------------- */

/* Input side S (224 / 299) and preprocess mode (0 = caffe, 1 = tf) of a model:
 * the target_size / preprocess_input pair at models.py:34,38 and :59,63. */
int dml_model_info(int model, int* input_size, int* preprocess_mode);

/* Number of weight arrays and total fp32 element count of the Keras-layout
 * blob (ResNet50(...) / InceptionV3(...) construction, models.py:26, :51). */
int dml_model_num_weights(int model, int* num_arrays, int64_t* total_floats);

/* Name and shape of weight array i in blob order (layer creation order, each
 * layer in Keras get_weights() order).  shape[4], rank <= 4. */
int dml_model_weight_info(int model, int i, char* name, int name_cap, int64_t shape[4], int* rank);

/* The three 256-entry fp32 look-up tables preprocess_input reduces to for uint8
 * input: out[c*256 + v] is the value of OUTPUT channel c for source byte v
 * (models.py:38 tf mode, :63 caffe mode).  For caffe, output channel c reads
 * source channel 2-c. */
int dml_preprocess_lut(int model, float out[768]);

/* ---- engine lifetime ------------------------------------------------------ */

/* One engine per (process, device, model); replaces the per-task model build
 * at models.py:25-26 / :50-51.  max_batch = largest n a later call may pass;
 * micro_batch = images pushed through the layer stack together (0 = default). */
int dml_create(int device, int model, int max_batch, int micro_batch, int flags, dml_engine** out);
void dml_destroy(dml_engine* e);

/* Sizes of the two device buffers the caller must allocate (256-byte aligned)
 * and bind before any compute call. */
int dml_workspace_bytes(dml_engine* e, size_t* bytes);
int dml_weights_bytes(dml_engine* e, size_t* bytes);
int dml_bind_workspace(dml_engine* e, void* dev_ptr, size_t bytes);
int dml_bind_weights(dml_engine* e, void* dev_ptr, size_t bytes);

/* Keras-layout fp32 host blob -> fold BN (+bias) into each conv -> pack
 * (bf16 K-major, or fp32 in check mode) -> copy into the bound weights buffer.
 * Stands in for ResNet50(weights=...)/InceptionV3(weights=...) at models.py:26,:51.
 * Synchronous. */
int dml_load_weights(dml_engine* e, const float* host_blob, size_t num_floats);

/* Marks the bound weights buffer as already holding packed weights (after the
 * caller replicated another rank's buffer, e.g. with an NCCL broadcast). */
int dml_weights_ready(dml_engine* e);

/* ---- the hot path --------------------------------------------------------- */

/* load_img's NEAREST resize + img_to_array + preprocess_input for n images
 * (models.py:34-38, :59-63).  d_pixels: concatenated uint8 RGB HxWx3 images;
 * d_img_off[n]: byte offset of each image; d_hw[2n]: (H, W) per image;
 * d_xtab[n*S], d_ytab[n*S]: source column/row per output column/row (Pillow's
 * running-sum rule, computed on the host).  Writes the engine's input tensor. */
int dml_preprocess(dml_engine* e, const uint8_t* d_pixels, const int64_t* d_img_off, const int32_t* d_hw,
                   const int32_t* d_xtab, const int32_t* d_ytab, int n, void* stream);

/* modelObj.predict + decode_predictions(top=5) on the n preprocessed images
 * (models.py:40-42, :65-67).  d_probs: n x 1000 softmax (may be NULL);
 * d_top5_idx / d_top5_prob: n x 5, descending probability, ties -> larger
 * class index first. */
int dml_forward(dml_engine* e, int n, float* d_probs, int32_t* d_top5_idx, float* d_top5_prob, void* stream);

/* dml_preprocess + dml_forward fused per micro-batch (the production path:
 * the whole per-file loop body of models.py:30-44 / :55-69 for n files). */
int dml_infer(dml_engine* e, const uint8_t* d_pixels, const int64_t* d_img_off, const int32_t* d_hw,
              const int32_t* d_xtab, const int32_t* d_ytab, int n, float* d_probs, int32_t* d_top5_idx,
              float* d_top5_prob, void* stream);

/* Preemption: the reference's worker cancels the task it is running when a new
 * WORKER_TASK_REQUEST arrives (worker.py:944-953).  dml_cancel may be called from
 * any thread; the dml_infer / dml_forward call in progress on this engine (or the
 * next one, if none is) stops before its next micro-batch and returns
 * DML_ECANCELLED.  Kernels already enqueued still run; the outputs of a cancelled
 * call are undefined.  The flag clears when it has been honoured, or through
 * dml_cancel_clear. */
int dml_cancel(dml_engine* e);
int dml_cancel_clear(dml_engine* e);

/* ---- inspection (parity taps, counters) ----------------------------------- */

/* Number of tensors in the engine's plan; tensor i's name, logical shape
 * (h, w, c), and where it lives: device pointer of image 0 of the CURRENT
 * micro-batch, channel pitch (elements) and element size in bytes.  Tensors
 * are valid after dml_forward/dml_infer only if the engine was created with
 * micro_batch >= n (buffers are reused between micro-batches). */
int dml_num_tensors(dml_engine* e, int* n);
int dml_tensor_info(dml_engine* e, int i, char* name, int name_cap, int* h, int* w, int* c, void** dev_ptr,
                    int* c_pitch, int* elem_bytes, int* keep);

/* The engine's preprocessed input tensor: n x S x S x 3 (bf16, or fp32 in check mode). */
int dml_input_tensor(dml_engine* e, void** dev_ptr, int* elem_bytes);
/* How K1 lays the input tensor out.  layout 0: [n][S][S][3] (what preprocess_input
 * returns, models.py:37/:62).  layout 1 (bf16 tensor-core engines): space-to-depth
 * [n][S2][S2][16] bf16 with S2 = ceil(S/2); channel (py*2+px)*3 + c of cell
 * (y2, x2) is channel c of pixel (2*y2+py, 2*x2+px), channels 12..15 and pixels past
 * the edge are zero -- the form in which the stride-2 stem convolution is a
 * stride-1 convolution.  cells = S or S2, cell_channels = 3 or 16. */
int dml_input_layout(dml_engine* e, int* layout, int* cells, int* cell_channels);
/* Physical rows of the layout-1 input tensor: every row holds row_cells cells (>= cells), data cell x sits at column
 * first_cell + x and the other columns are zero (the stem's horizontal padding, kept in memory so that a whole halo tile
 * is one TMA box); in the cells whose physical column has bit swap_bit set, the two 16-byte halves (channels 0..7 / 8..15)
 * are stored swapped -- the shared-memory swizzle the tensor cores expect, applied when the cell is written.
 * Layout 0: row_cells = cells, first_cell = 0, swap_bit = -1 (no swap).  Still the preprocess_input result of
 * models.py:38 / :63, value for value. */
int dml_input_rows(dml_engine* e, int* row_cells, int* first_cell, int* swap_bit);

/* Logits (pre-softmax), n x 1000 fp32, valid after dml_forward / dml_infer. */
int dml_logits(dml_engine* e, float** dev_ptr);

/* Number of ops in the layer plan and, for op i: name, kind (0 conv, 1 max-pool,
 * 2 avg-pool, 3 global-avg-pool, 4 dense, 5 softmax+top5), which kernel runs it
 * (0 CUDA-core, 1 tcgen05), algorithmic MACs per image and algorithmic bytes
 * per image (activations in + out, residual) plus weight bytes per launch. */
int dml_num_ops(dml_engine* e, int* n);
int dml_op_info(dml_engine* e, int i, char* name, int name_cap, int* kind, int* impl, int64_t* macs_per_img,
                int64_t* act_bytes_per_img, int64_t* weight_bytes);

/* Branch schedule of op i (0 <= i < dml_num_ops minus the two head ops): independent branches of an InceptionV3 block
 * (the Keras functional graph behind /root/reference/models.py:26 runs them in whatever order TensorFlow's executor picks) are
 * launched on engine-owned CUDA streams that fork from and join the stream passed to dml_forward / dml_infer.  stream: 0 = the
 * caller's stream; num_waits: events of other streams the op waits for; num_streams: streams in use (1 = serial: ResNet50, the
 * fp32 check mode, DML_BRANCH_STREAMS=0).  Results do not depend on it. */
int dml_op_schedule(dml_engine* e, int i, int* stream, int* num_waits, int* num_streams);

/* dml_forward with a CUDA-event pair around every launch: op_ms[i] receives the
 * milliseconds op i took, summed over the micro-batches (cap >= dml_num_ops).
 * Synchronous; for bench.py's roofline, not for production. */
int dml_profile_forward(dml_engine* e, int n, int32_t* d_top5_idx, float* d_top5_prob, float* op_ms, int cap,
                        void* stream);

/* Kernels launched by this engine since creation (all of them ours). */
int dml_launch_count(dml_engine* e, int64_t* launches);

/* One Conv2D + folded bias (+ residual) (+ ReLU) of predict (models.py:40, :65)
 * over explicit device buffers: bf16 NHWC activations (channel pitches in
 * elements), packed bf16 weights [cout][kh*kw][cin_pad], fp32 bias.
 * impl 0 = CUDA-core kernel, 1 = tcgen05 one-tile-per-CTA kernel, 2 = persistent
 * tcgen05 kernel, 3 = C=3 stride-2 stem, 4 = halo kernel with streamed weights,
 * 5 = persistent kernel on CTA pairs, 6 = weights as the A operand, 7 = halo kernel
 * with the filter resident in shared memory, 8 = the same on CTA pairs (the engine
 * picks per layer among 2, 5, 6, 7, 8); tile_n / stages: 0 = default;
 * bits 8..15 of `stages` = 64-channel k-blocks per pipeline stage for impl 2 / 5 (0 = default),
 * bits 16..17 = residual staging of impl 5 (0 default, 1 two tile buffers, 2 one buffer with per-chunk hand-back),
 * bit 18 = impl 7 / 8 with 32-channel K blocks (64-byte shared-memory rows) whatever Cin.
 * impl 1, 3, 4 are superseded A/B twins: only in libraries compiled with -DDML_LEGACY_VARIANTS (dml_version() then
 * ends in "+legacy-variants"); otherwise they return DML_EINVAL.
 * The unit-test and tuning hook for the convolution kernels. */
int dml_conv2d_bf16(int impl, const void* d_in, const void* d_w, const float* d_bias, const void* d_res, void* d_out,
                    int n, int hi, int wi, int cin, int in_pitch, int cout, int out_pitch, int res_pitch, int kh,
                    int kw, int stride, int pad_h, int pad_w, int cin_pad, int relu, int tile_n, int stages,
                    void* stream);

/* The tail of a ResNet50 conv2 bottleneck block of predict (models.py:65; Keras block1: _2_conv 3x3 -> _3_conv 1x1
 * -> Add -> ReLU, then the next block's _1_conv) as ONE kernel over explicit device buffers, all bf16 NHWC with
 * contiguous channels: y1 [n][h][w][64] -> y2 = relu(conv3x3(y1) + b3) (never stored) -> out [n][h][w][256] =
 * relu(We y2 (+ Wp src2) + be (+ res)) -> y1n [n][h][w][64] = relu(Wr out + br).  Exactly one of d_res ([n][h][w][256])
 * and d_src2 ([n][h][w][64], the projection shortcut's input; We then holds [256][64 + 64]) is given; d_wr / d_br /
 * d_y1n may all be NULL (no next reduce; only with d_res).  Weights packed as for dml_conv2d_bf16.  Needs w <= 62,
 * h a multiple of 2 * floor(128 / (w + 2)) and floor(128 / (w + 2)) * w a multiple of 16 (56 x 56 qualifies).
 * Results are bit-identical to the three dml_conv2d_bf16 calls it replaces.  The unit-test hook of that kernel. */
int dml_chain_bf16(const void* d_y1, const void* d_w3, const float* d_b3, const void* d_we, const float* d_be,
                   const void* d_res, const void* d_src2, void* d_out, const void* d_wr, const float* d_br, void* d_y1n,
                   int n, int h, int w, void* stream);

/* A ResNet50 block's 1x1 expand with its shortcut and the NEXT block's 1x1 reduce of predict (models.py:65; Keras block1:
 * _3_conv -> Add -> ReLU, then _1_conv) as ONE kernel over explicit device buffers, all bf16 with contiguous channels:
 * y2 [n][h][w][kmid] -> out [n][h][w][nexp] = relu(We y2 (+ Wp src2) + be (+ res)) -> y1n [n][h][w][nred] = relu(Wr out + br)
 * (any n * h * w: partial tiles are clipped).  Exactly one of d_res ([n][h][w][nexp], the identity shortcut: conv3 / conv4) and
 * d_src2 (the projection shortcut's input folded in as a second GEMM source, [n][h * s2_stride][w * s2_stride][k2] read at stride
 * s2_stride = 1 (conv2_block1) or 2 (conv3_block1); We then holds [nexp][kmid + k2]) is given.  kmid, k2 multiples of 64 with
 * kmid + k2 <= 384, nexp a multiple of 256, nred in {64, 128, 192, 256}.  Weights packed as for dml_conv2d_bf16.  Results are
 * bit-identical to the two dml_conv2d_bf16 calls it replaces.  The unit-test hook of that kernel. */
int dml_duo_bf16(const void* d_y2, const void* d_we, const float* d_be, const void* d_res, const void* d_src2, void* d_out,
                 const void* d_wr, const float* d_br, void* d_y1n, int n, int h, int w, int kmid, int k2, int s2_stride, int nexp,
                 int nred, void* stream);

/* A stride-1 Conv2D with 64 filters (+ folded bias, + ReLU) and the MaxPooling2D(3, strides 2) behind it of predict
 * (models.py:40 InceptionV3 conv2d_2 + max_pooling2d; models.py:65 ResNet50 conv1 + pool1_pad + pool1_pool) as ONE kernel over
 * explicit device buffers: bf16 NHWC input [n][hi][wi][cin] (cin 32 with 3 x 3 taps; the engine also runs the 4 x 4 form of
 * ResNet50's stem over its space-to-depth input), packed bf16 weights [64][kh*kw][cin_pad], fp32 bias; output
 * [n][hp][wq][64] with hp = (ho + 2 * pool_pad - 3) / 2 + 1.  pool_pad 1 = ZeroPadding2D(1) in front of the pool (needs relu).
 * Results are bit-identical to dml_conv2d_bf16 followed by the max-pool.  The unit-test hook of that kernel. */
int dml_conv_pool_bf16(const void* d_in, const void* d_w, const float* d_bias, void* d_pool_out, int n, int hi, int wi, int cin,
                       int cin_pad, int kh, int kw, int pad, int relu, int pool_pad, void* stream);

/* The dense(1000) + softmax + top-5 head with NO device allocation anywhere: no
 * engine, no workspace.  (No counterpart in the reference: its Keras model owns
 * all of its memory.)  It is for callers that cannot allocate device memory --
 * inside a stream capture, on a serving thread that may not take cudaMalloc's
 * device-wide synchronisation, or inside a process whose GPU memory belongs to
 * another framework's pool.  The n x 1000 fp32 logits, which dml_forward keeps in
 * the caller's workspace, live in a fixed array inside the library instead:
 * DML_HEAD_STATIC_MAX_IMAGES x 1000 floats = 4,096,000 bytes held on every GPU
 * that loads the library, until the process exits.
 *   d_pooled [n][cin] fp32 (global-average-pooled features), d_dense_w [cin][1000]
 *   fp32, d_dense_b [1000]; outputs as in dml_forward (d_probs may be NULL).
 * Runs on the CURRENT device.  n > DML_HEAD_STATIC_MAX_IMAGES -> DML_ESIZE.
 * The first call on each device looks up the array's device address and sets a
 * kernel attribute: make that first call outside any stream capture.  All calls on
 * a device share the one array (it keeps the previous call's logits; every row a
 * call reads is rewritten by that call first, nothing needs clearing): two calls
 * whose kernels can overlap on the GPU (different streams) race -- serialise them. */
#define DML_HEAD_STATIC_MAX_IMAGES 1024
int dml_head_static(const float* d_pooled, const float* d_dense_w, const float* d_dense_b, int n, int cin,
                    float* d_probs, int32_t* d_top5_idx, float* d_top5_prob, void* stream);

int dml_sync(dml_engine* e);
const char* dml_last_error(void);
const char* dml_version(void);

#ifdef __cplusplus
}
#endif
#endif /* DML_B200_H_ */
