/*
 * convnets_b200.h — C ABI of libconvnets_b200.so
 *
 * B200-native (sm_100a) replacement for the arithmetic layer ("L0") underneath the ConvNets
 * architecture zoo.  The reference has no native interface of its own: every entry point below
 * replaces a PyTorch library call the reference reaches from its layer factories
 * (reference src/basemodel.py:966-1067) and model files; the call site each one replaces is cited
 * on the declaration.  The reference-side binding a maintainer adds is the ctypes stub shown in
 * INTEGRATION.md (the reference is pure Python, so ctypes is its FFI).
 *
 * Conventions
 *   - plain pointers and sizes only; no torch types.  All tensor pointers are DEVICE pointers
 *     owned by the caller; the library never frees or retains them past the call.
 *   - activations are NHWC ("channels_last"): x[n][h][w][c]; conv weights are KRSC:
 *     w[k][r][s][c/groups]; both in `dtype` (CNB_F32 or CNB_BF16).  Weight gradients, BatchNorm
 *     parameters / statistics and reduction outputs are always fp32.
 *   - every function enqueues work on `stream` (a cudaStream_t passed as void*) and returns
 *     immediately; no implicit device synchronisation.
 *   - return value: 0 = ok, negative = cnb_status error; cnb_last_error() gives a thread-local
 *     message.  There is NO silent fallback: an unsupported shape is an error.
 *   - scratch memory comes from the caller, sized by the *_workspace_bytes query (one exception:
 *     cnb_bn_fwd_train_static, whose scratch is a static device array inside the library).
 */
#ifndef CONVNETS_B200_H
#define CONVNETS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum { CNB_F32 = 0, CNB_BF16 = 1 } cnb_dtype;

typedef enum {
  CNB_OK = 0,
  CNB_ERR_INVALID = -1,     /* bad descriptor / null pointer / misaligned pointer */
  CNB_ERR_UNSUPPORTED = -2, /* valid but not implemented for this shape/dtype */
  CNB_ERR_CUDA = -3,        /* launch or driver failure (message has the CUDA error string) */
  CNB_ERR_WORKSPACE = -4    /* workspace too small */
} cnb_status;

/* cnb_conv_desc.flags (forward epilogue) */
#define CNB_CONV_BIAS 1
#define CNB_CONV_RELU 2
/* cnb_conv_desc.algo */
#define CNB_ALGO_AUTO 0 /* tcgen05 implicit GEMM when the shape qualifies, else CUDA-core */
#define CNB_ALGO_SIMT 1 /* force the CUDA-core implicit GEMM (exact fp32 FMA accumulate) */
#define CNB_ALGO_TC 2   /* force the tcgen05 path; error if the shape does not qualify */
/* direction argument of the queries */
#define CNB_FWD 0
#define CNB_DGRAD 1
#define CNB_WGRAD 2

/* 2-D convolution problem; replaces the arguments of nn.Conv2d as built by
 * reference src/basemodel.py:966-979 (conv2d) and :981-997 (conv2d_depthwise). */
typedef struct cnb_conv_desc {
  int32_t N, H, W, C; /* input batch, height, width, channels */
  int32_t K;          /* output channels */
  int32_t R, S;       /* filter height, width */
  int32_t stride_h, stride_w;
  int32_t pad_h, pad_w;
  int32_t dil_h, dil_w;
  int32_t groups;
  int32_t dtype; /* cnb_dtype of x / w / y */
  int32_t flags;
  int32_t algo;
} cnb_conv_desc;

/* Pooling problem; replaces nn.MaxPool2d / nn.AvgPool2d as built by
 * reference src/basemodel.py:1017-1033. */
typedef struct cnb_pool_desc {
  int32_t N, H, W, C;
  int32_t kh, kw, stride_h, stride_w, pad_h, pad_w;
  int32_t dtype;
} cnb_pool_desc;

/* BatchNorm2d problem over an NHWC tensor viewed as [M = N*H*W, C];
 * replaces nn.BatchNorm2d at reference src/basemodel.py:1008,1061, src/densenet.py:80,117,124,182,
 * src/mobilenet_v1.py:108,119. */
#define CNB_BN_RELU 1 /* y = relu(bn(x) [+ residual])  (fuses nn.ReLU, basemodel.py:1012) */
#define CNB_BN_ADD 2  /* a residual tensor is added before the activation */
typedef struct cnb_bn_desc {
  int64_t M; /* N*H*W */
  int32_t C;
  int32_t dtype;
  int32_t flags;
  float eps;
  float momentum;
  /* nn.BatchNorm2d.num_batches_tracked (device int64 scalar, may be NULL): the training forward adds 1
   * in its statistics epilogue, so the module's per-layer `add_(1)` launch disappears */
  int64_t* num_batches_tracked;
} cnb_bn_desc;

/* ---- library ------------------------------------------------------------------------- */
const char* cnb_version(void);
const char* cnb_last_error(void);
/* number of kernels this library has launched in the calling process (all threads) */
uint64_t cnb_launch_count(void);

/* ---- convolution: aten::convolution / convolution_backward (reference src/basemodel.py:974,992)
 * Dense, grouped (src/resnext.py:93,168; src/shufflenet_v1.py:119-141), dilated
 * (src/sknet.py:192-201) and depthwise (src/basemodel.py:992) are all served by these three. */
int cnb_conv2d_out_dims(const cnb_conv_desc* d, int32_t* P, int32_t* Q);
/* 1 if the tcgen05 tensor-core path would run for this direction, 0 for CUDA cores, <0 error */
int cnb_conv2d_uses_tc(const cnb_conv_desc* d, int direction);
size_t cnb_conv2d_workspace_bytes(const cnb_conv_desc* d, int direction);
/* rows of the BatchNorm partial-sum table the forward epilogue can emit for this problem (two rows
 * per 128-pixel output tile), or 0 when this conv runs on a path that cannot (CUDA cores). */
int cnb_conv2d_bn_partial_rows(const cnb_conv_desc* d);
/* y[N,P,Q,K] = conv(x[N,H,W,C], w[K,R,S,C/g]) (+ bias[K] fp32) (ReLU).
 * bn_partials (may be NULL): fp32 [rows][2][K] receives, per output tile, the per-channel sum and
 * sum of squares of the STORED (rounded) outputs — the statistics pass of the BatchNorm2d that
 * follows the conv in conv2d_block (reference src/basemodel.py:1004-1010), fused into the epilogue. */
int cnb_conv2d_fwd(const cnb_conv_desc* d, const void* x, const void* w, const float* bias, void* y,
                   float* bn_partials, void* workspace, size_t workspace_bytes, void* stream);
/* dx[N,H,W,C] = conv_transpose(dy[N,P,Q,K], w).  w_t (may be NULL): the filter with in/out
 * channels exchanged, [C][R*S][K], as written by cnb_pack_filter; without it the tensor-core path
 * transposes w into the workspace on every call. */
int cnb_conv2d_dgrad(const cnb_conv_desc* d, const void* dy, const void* w, const void* w_t, void* dx,
                     void* workspace, size_t workspace_bytes, void* stream);
/* dx = conv_transpose(dy, w) + addend[N,H,W,C]: the sum autograd forms where a residual block's
 * input feeds both the first conv and the identity shortcut (reference src/resnet.py:139-156,
 * 176-196), folded into the data-grad epilogue.  Stride-1 convolutions on the tensor-core path only
 * (anything else is CNB_ERR_UNSUPPORTED); dx may alias addend.
 * addend_mask (may be NULL; needs C % 32 == 0): 1 bit per addend element, bit j of byte (m*C + c)/8
 * for element (m, c + j) — the ReLU mask cnb_bn_fwd_train wrote for the block output.  With it the
 * addend is the block-output gradient ITSELF and the epilogue zeroes the elements whose ReLU was
 * inactive, so BatchNorm backward does not have to write the skip gradient as a tensor. */
int cnb_conv2d_dgrad_add(const cnb_conv_desc* d, const void* dy, const void* w, const void* w_t,
                         const void* addend, const void* addend_mask, void* dx, void* workspace,
                         size_t workspace_bytes, void* stream);
/* Data-grad whose output IS the gradient of a BatchNorm2d + ReLU output (conv -> BN -> ReLU -> conv,
 * reference src/basemodel.py:1004-1015 inside src/resnet.py:159-196): besides dx (optionally + addend,
 * as cnb_conv2d_dgrad_add) the epilogue forms that BatchNorm's backward sums from the tile it just
 * stored — per CTA row of bn_partials[rows][2][C] (fp32): sum dx', sum dx' * (bn_x - bn_mean), dx' =
 * dx where the ReLU bit is set — so cnb_bn_bwd_from_partials needs no reduction pass over x and dy.
 * bn_x [N,H,W,C]: the BatchNorm's input; bn_relu_mask: the 1-bit mask cnb_bn_fwd_train wrote; rows =
 * cnb_conv2d_dgrad_partial_rows(d) (0: this shape has no such path). */
int cnb_conv2d_dgrad_partial_rows(const cnb_conv_desc* d);
int cnb_conv2d_dgrad_bn(const cnb_conv_desc* d, const void* dy, const void* w, const void* w_t,
                        const void* addend, const void* addend_mask, const void* bn_x,
                        const void* bn_relu_mask, const float* bn_mean, float* bn_partials, void* dx,
                        void* workspace, size_t workspace_bytes, void* stream);
/* dw[K,R,S,C/g] (fp32) = x (*) dy ; dbias[K] (fp32, may be NULL) = sum over N,P,Q of dy */
int cnb_conv2d_wgrad(const cnb_conv_desc* d, const void* x, const void* dy, float* dw, float* dbias,
                     void* workspace, size_t workspace_bytes, void* stream);

/* ---- BatchNorm2d: aten::native_batch_norm / _backward ------------------------------------ */
size_t cnb_bn_workspace_bytes(const cnb_bn_desc* d);
/* training forward: batch statistics (biased var for normalisation, unbiased for running_var),
 * running-stat update, y = [relu]((x-mean)*rstd*gamma + beta [+ residual]).
 * save_mean / save_rstd: fp32 [C] outputs kept for backward.
 * conv_partials / n_partials (may be NULL / 0): the table written by cnb_conv2d_fwd for x; when
 *   given, x is not read for the statistics.
 * relu_mask (may be NULL; needs CNB_BN_RELU, bf16, C % 8 == 0): M*C/8 bytes, bit j of byte
 *   (m*C + c)/8 = (y[m][c + j] > 0); backward can take it instead of y. */
int cnb_bn_fwd_train(const cnb_bn_desc* d, const void* x, const void* residual, const float* gamma,
                     const float* beta, float* running_mean, float* running_var, void* y,
                     float* save_mean, float* save_rstd, const float* conv_partials,
                     int32_t n_partials, void* relu_mask, void* workspace, size_t workspace_bytes,
                     void* stream);
/* Allocation-free training forward.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA-graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, kernels, launch geometry and results as
 * cnb_bn_fwd_train with conv_partials = NULL, but the scratch (per-CTA partial sums, scale / shift,
 * arrival counters) is a 1 MiB + 4 KiB static __device__ array inside the library — held by every
 * GPU that loads it until exit — instead of a caller workspace; the call makes no cudaMalloc /
 * cudaFree and calls no allocating routine.  Limits: C <= 16384 and 8*C*(splits+1) bytes <=
 * cnb_bn_static_scratch_bytes() and at most 1024 channel blocks (on a 148-SM B200 the scratch need
 * peaks at ~434 KB: every C <= 16384 with C % 8 == 0 in bf16 / C % 4 == 0 in fp32 fits, unvectorised
 * rows up to C = 8192); beyond them it returns CNB_ERR_WORKSPACE with a message and launches nothing.  All calls share the array: calls that
 * overlap on the GPU (different streams without an ordering between them) race on it — serialise
 * them, or use cnb_bn_fwd_train with one workspace per stream. */
size_t cnb_bn_static_scratch_bytes(void);
int cnb_bn_fwd_train_static(const cnb_bn_desc* d, const void* x, const void* residual, const float* gamma,
                            const float* beta, float* running_mean, float* running_var, void* y,
                            float* save_mean, float* save_rstd, void* relu_mask, void* stream);
/* backward from the sums a data-grad epilogue already formed (cnb_conv2d_dgrad_bn): finish + dx pass
 * only.  bf16, CNB_BN_RELU with the 1-bit mask, C % 8 == 0. */
int cnb_bn_bwd_from_partials(const cnb_bn_desc* d, const void* x, const void* relu_mask, const void* dy,
                             const float* gamma, const float* save_mean, const float* save_rstd,
                             const float* partials, int32_t n_partials, void* dx, void* d_residual,
                             float* dgamma, float* dbeta, void* workspace, size_t workspace_bytes,
                             void* stream);
/* ---- concat-aware BatchNorm: DenseNet's BatchNorm2d over torch.cat([x, out], 1) (reference
 * src/densenet.py:117,124,131-140) when the block keeps its feature maps in one [M, C_total] buffer.
 * Layer l normalises the channel prefix [0, C_l) of the buffer: x has a row stride != C.
 *   cnb_bn_stats_from_partials: mean / rstd of the C channels a convolution just produced, from the
 *     partial-sum table its epilogue wrote (cnb_conv2d_fwd bn_partials); nothing else is touched.
 *   cnb_bn_fwd_train_prefix: batch statistics are GIVEN (a channel's statistics do not depend on the
 *     layer that reads it, so a block computes them once): running-statistics update + apply.
 *   cnb_bn_bwd_prefix: dgamma, dbeta and dx ACCUMULATED into dx_accum (row stride dx_row_stride) —
 *     the gradient of the shared buffer, which every later layer of the block adds to. */
int cnb_bn_stats_from_partials(const cnb_bn_desc* d, const float* conv_partials, int32_t n_partials,
                               float* mean, float* rstd, void* workspace, size_t workspace_bytes,
                               void* stream);
int cnb_bn_fwd_train_prefix(const cnb_bn_desc* d, const void* x, int64_t x_row_stride, const float* mean,
                            const float* rstd, const float* gamma, const float* beta,
                            float* running_mean, float* running_var, void* y, void* relu_mask,
                            void* workspace, size_t workspace_bytes, void* stream);
int cnb_bn_bwd_prefix(const cnb_bn_desc* d, const void* x, int64_t x_row_stride, const void* y,
                      const void* relu_mask, const void* dy, const float* mean, const float* rstd,
                      const float* gamma, void* dx_accum, int64_t dx_row_stride, float* dgamma,
                      float* dbeta, void* workspace, size_t workspace_bytes, void* stream);
/* inference forward with running statistics */
int cnb_bn_fwd_eval(const cnb_bn_desc* d, const void* x, const void* residual, const float* gamma,
                    const float* beta, const float* running_mean, const float* running_var, void* y,
                    void* workspace, size_t workspace_bytes, void* stream);
/* backward of the training forward.  With CNB_BN_RELU the mask y > 0 comes from relu_mask when it
 * is given (1 bit per element), else from the forward output y; d_residual (may be NULL) receives the masked upstream gradient. */
int cnb_bn_bwd(const cnb_bn_desc* d, const void* x, const void* y, const void* relu_mask,
               const void* dy, const float* gamma, const float* save_mean, const float* save_rstd, void* dx,
               void* d_residual, float* dgamma, float* dbeta, void* workspace,
               size_t workspace_bytes, void* stream);

/* ---- element-wise: aten::relu_/threshold_backward (basemodel.py:1012; resnet.py:154,194),
 * aten::add_ (resnet.py:151,191) ---------------------------------------------------------- */
int cnb_relu_fwd(const void* x, void* y, int64_t n, int dtype, void* stream);
int cnb_relu_bwd(const void* y, const void* dy, void* dx, int64_t n, int dtype, void* stream);
/* y = relu(a + b) */
int cnb_add_relu_fwd(const void* a, const void* b, void* y, int64_t n, int dtype, void* stream);
/* dx = dy * (y > 0) (serves both addends) */
int cnb_add_relu_bwd(const void* y, const void* dy, void* dx, int64_t n, int dtype, void* stream);
/* y = a + b (gradient accumulation of branching graphs) */
int cnb_add(const void* a, const void* b, void* y, int64_t n, int dtype, void* stream);

/* ---- pooling: aten::max_pool2d_with_indices / avg_pool2d / adaptive_avg_pool2d ------------ */
/* idx[N,P,Q,C] int32 = h*W + w of the first maximum in row-major window scan (PyTorch's rule;
 * padding is never selected; NaN propagates). */
/* code[N,P,Q,C] uint8 (may be NULL, as may idx): the same argmax as a window-relative position
 * (h - (p*stride_h - pad_h)) * kw + (w - (q*stride_w - pad_w)) — 1 byte instead of 4 for the
 * training path, where only backward consumes it.  Backward takes whichever is given (code wins). */
int cnb_maxpool_fwd(const cnb_pool_desc* d, const void* x, void* y, int32_t* idx, uint8_t* code,
                    void* stream);
int cnb_maxpool_bwd(const cnb_pool_desc* d, const void* dy, const int32_t* idx, const uint8_t* code,
                    void* dx, void* stream);
/* count_include_pad=True, ceil_mode=False (reference src/shufflenet_v1.py:148, densenet.py:188) */
int cnb_avgpool_fwd(const cnb_pool_desc* d, const void* x, void* y, void* stream);
int cnb_avgpool_bwd(const cnb_pool_desc* d, const void* dy, void* dx, void* stream);
/* global average pool: y[N,C] (fp32) = mean over HW of x[N,HW,C]  (basemodel.py:1035-1042) */
int cnb_gap_fwd(const void* x, float* y, int32_t N, int32_t HW, int32_t C, int dtype, void* stream);
int cnb_gap_bwd(const float* dy, void* dx, int32_t N, int32_t HW, int32_t C, int dtype,
                void* stream);

/* ---- data movement ------------------------------------------------------------------------ */
/* channel shuffle (reference src/shufflenet_v1.py:173-190): y[:, j*g+i] = x[:, i*(C/g)+j] */
int cnb_channel_shuffle(const void* x, void* y, int64_t M, int32_t C, int32_t groups, int dtype,
                        void* stream);
/* strided channel-slab copy: dst[m, dst_off + c] = src[m, src_off + c], c < Ccopy.  Serves
 * torch.cat(dim=1) (densenet.py:140, inceptionnet_v1.py:142, squeezenet.py:153,
 * shufflenet_v1.py:167) forward (dst is the concat buffer) and backward (src is its gradient). */
int cnb_copy_channels(const void* src, int32_t src_C, int32_t src_off, void* dst, int32_t dst_C,
                      int32_t dst_off, int64_t M, int32_t Ccopy, int dtype, void* stream);
/* layout / precision conversion of the network input: src fp32 NCHW [N,C,H,W] -> dst `dtype`
 * NHWC [N,H,W,C_dst], channels C..C_dst-1 zero-filled (C_dst = 8 lets an RGB stem take the
 * tensor-core path, whose TMA rows must be multiples of 16 bytes) */
int cnb_nchw_to_nhwc(const float* src, void* dst, int32_t N, int32_t C, int32_t H, int32_t W,
                     int32_t C_dst, int dtype, void* stream);
/* network input from uint8 pixels: src [N,C,H,W] (src_nhwc = 0) or [N,H,W,C] (src_nhwc = 1, what a
 * decoded image is) -> dst fp32 NCHW [N,C,H_out,W_out] = ((float)u8 / 255 - mean[c]) / std[c], the
 * fp32 arithmetic of transforms.ToTensor + transforms.Normalize (reference src/mngrdata.py:151-153,
 * :177-179) bit for bit, so the batch crosses PCIe as bytes and `DeviceMngr.move`
 * (src/mngrdevice.py:30-36) ships a quarter of the traffic.  `aug` (device, int32 [N][3] = top, left,
 * flip; may be NULL when H_out = H, W_out = W, pad = 0) applies RandomCrop(H_out, padding=pad)
 * (zero fill) and RandomHorizontalFlip (src/mngrdata.py:146-147) with offsets drawn by the caller.
 * mean / std are HOST arrays of C <= 4 floats. */
int cnb_normalize_u8(const uint8_t* src, float* dst, int32_t N, int32_t C, int32_t H, int32_t W,
                     int32_t src_nhwc, int32_t H_out, int32_t W_out, int32_t pad,
                     const float* mean_host, const float* std_host, const int32_t* aug,
                     void* stream);
/* explicit im2col of the network input for a first layer whose channel count cannot feed the
 * tensor-core path (RGB stems: reference src/resnet.py:66 7x7 s2; src/vggnet.py 3x3; ...):
 * src fp32 NCHW [N,C,H,W] -> dst `d->dtype` [N*P*Q, K_dst], column (r*S + s)*C + c, zero for padding
 * taps and for columns >= R*S*C.  The conv then runs as cnb_conv2d_{fwd,wgrad} with a 1x1
 * descriptor over C = K_dst "channels" (no data-grad: the network input needs none). */
int cnb_im2col_nchw(const cnb_conv_desc* d, const float* src, void* dst, int32_t K_dst,
                    void* stream);
/* Row im2col of the fp32 NCHW batch: dst[N][H][Q][C_dst], dst[n][h][q][s*C + c] =
 * src[n][c][h][q*stride_w - pad_w + s*dil_w] (zeros outside and in columns >= S*C).  The R x S stem
 * is then the R x 1 convolution with vertical stride stride_h over C_dst channels — a third of the
 * bytes of the full im2col matrix.  Only the W-side fields of d are used. */
int cnb_rowim2col_nchw(const cnb_conv_desc* d, const float* src, void* dst, int32_t C_dst,
                       void* stream);
/* fp32 master filter w[K][R*S][C] (the KRSC storage of a dense nn.Conv2d weight) -> bf16 operand
 * copies: w_bf16 in the same layout (forward / weight-grad) and, when wt_bf16 is not NULL,
 * [C][R*S][K] (data-grad).  One pass per optimiser step. */
int cnb_pack_filter(const float* w, void* w_bf16, void* wt_bf16, int32_t K, int32_t RS, int32_t C,
                    void* stream);
/* The same for every dense filter of a model in one launch (the per-layer launches are KB-sized
 * work).  `items_device` is a device array, sorted by first_block: item i owns the blocks
 * [first_block_i, first_block_i + cnb_pack_filter_blocks(K_i, RS_i, C_i)); total_blocks is their
 * sum.  wt_bf16 may be NULL per item.  The table is built once per model by the caller. */
typedef struct cnb_pack_item {
  const float* w;
  void* w_bf16;
  void* wt_bf16;
  int32_t K, RS, C;
  int32_t first_block;
} cnb_pack_item;
int64_t cnb_pack_filter_blocks(int32_t K, int32_t RS, int32_t C);
int cnb_pack_filters(const cnb_pack_item* items_device, int32_t n_items, int32_t total_blocks,
                     void* stream);
/* the optimiser step (reference src/basemodel.py:58-66 `optim.Adam(lr, weight_decay)`, stepped at
 * :284) for every parameter of a model in one call, writing the bf16 operand copies of the
 * tensor-core filters in the same pass (replaces cnb_pack_filters + torch's Adam kernels).
 * `items_host` is a HOST array (it travels to the device as kernel parameters, 56 items per launch,
 * so a CUDA-graph capture of the step holds it by value); first_block is filled in by the library.
 * Item: p / g / m / v fp32, same physical layout.  RS = 0: flat tensor of K*C elements (w_bf16, if
 * given, receives the cast copy; wt_bf16 ignored).  RS > 0: filter [K][RS][C] as in cnb_pack_item.
 * `state_device`: 16 doubles; the caller sets [0] step count so far, [1] lr, [2] beta1, [3] beta2,
 * [4] eps, [5] weight_decay; the call increments [0] and uses [8..15] as scratch.  Arithmetic is
 * fp32 in the order of torch's single-tensor Adam (the reference's CPU path); bias corrections in
 * double. */
typedef struct cnb_adam_item {
  float* p;
  const float* g;
  float* m;
  float* v;
  void* w_bf16;
  void* wt_bf16;
  int32_t K, RS, C;
  int32_t first_block;
} cnb_adam_item;
int64_t cnb_adam_item_blocks(int32_t K, int32_t RS, int32_t C);
int cnb_adam_step(const cnb_adam_item* items_host, int32_t n_items, double* state_device,
                  void* stream);
/* element-wise precision conversion (fp32 master weights -> bf16 operand copy, etc.) */
int cnb_cast(const void* src, int src_dtype, void* dst, int dst_dtype, int64_t n, void* stream);

/* ---- gating --------------------------------------------------------------------------------
 * SE scale (reference src/se_resnet.py:253): y[n,hw,c] = x[n,hw,c] * s[n,c]   (s fp32) */
int cnb_se_scale_fwd(const void* x, const float* s, void* y, int32_t N, int32_t HW, int32_t C,
                     int dtype, void* stream);
/* dx = dy * s ; ds[n,c] = sum_hw dy * x */
int cnb_se_scale_bwd(const void* x, const float* s, const void* dy, void* dx, float* ds, int32_t N,
                     int32_t HW, int32_t C, int dtype, void* stream);
/* SK fuse (reference src/sknet.py:226-231): g[n,c] (fp32) = mean_hw (u0 + u1) */
int cnb_sk_fuse_gap_fwd(const void* u0, const void* u1, float* g, int32_t N, int32_t HW, int32_t C,
                        int dtype, void* stream);
/* SK select (reference src/sknet.py:238-240): a = softmax over the 2 paths of logits[n,2,C]
 * (fp32); y = u0*a0 + u1*a1.  attn[n,2,C] (fp32) is kept for backward. */
int cnb_sk_select_fwd(const void* u0, const void* u1, const float* logits, float* attn, void* y,
                      int32_t N, int32_t HW, int32_t C, int dtype, void* stream);
/* du_i = dy*a_i + dg[n,c]/HW ; dlogits = softmax backward of (sum_hw dy*u_i) */
int cnb_sk_select_bwd(const void* u0, const void* u1, const float* attn, const void* dy,
                      const float* dg, void* du0, void* du1, float* dlogits, int32_t N, int32_t HW,
                      int32_t C, int dtype, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CONVNETS_B200_H */
