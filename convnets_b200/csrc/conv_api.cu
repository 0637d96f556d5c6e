// conv_api.cu — C-ABI entry points for convolution: descriptor validation and dispatch between
// the tcgen05 implicit-GEMM path (conv_tc.cu), the depthwise kernels (dwconv.cu) and the
// CUDA-core implicit GEMM (conv_simt.cu).  No silent fallbacks: CNB_ALGO_TC on a shape the
// tensor-core path cannot take is an error.
#include "common.cuh"

using namespace cnb;

static int validate_conv(const cnb_conv_desc* d, const char* who) {
  CNB_REQUIRE(d != nullptr, CNB_ERR_INVALID, "%s: null descriptor", who);
  CNB_REQUIRE(d->N > 0 && d->H > 0 && d->W > 0 && d->C > 0 && d->K > 0 && d->R > 0 && d->S > 0,
              CNB_ERR_INVALID, "%s: non-positive dimension (N=%d H=%d W=%d C=%d K=%d R=%d S=%d)", who,
              d->N, d->H, d->W, d->C, d->K, d->R, d->S);
  CNB_REQUIRE(d->stride_h > 0 && d->stride_w > 0 && d->dil_h > 0 && d->dil_w > 0 && d->pad_h >= 0 &&
                  d->pad_w >= 0,
              CNB_ERR_INVALID, "%s: bad stride/dilation/padding", who);
  CNB_REQUIRE(d->groups > 0 && d->C % d->groups == 0 && d->K % d->groups == 0, CNB_ERR_INVALID,
              "%s: in_channels (%d) and out_channels (%d) must be divisible by groups (%d)", who, d->C,
              d->K, d->groups);
  CNB_REQUIRE(valid_dtype(d->dtype), CNB_ERR_INVALID, "%s: bad dtype %d", who, d->dtype);
  CNB_REQUIRE(d->algo >= CNB_ALGO_AUTO && d->algo <= CNB_ALGO_TC, CNB_ERR_INVALID, "%s: bad algo %d",
              who, d->algo);
  int P, Q;
  conv_out_dims(*d, P, Q);
  CNB_REQUIRE(P > 0 && Q > 0, CNB_ERR_INVALID,
              "%s: kernel size can't be greater than actual input size (output %dx%d)", who, P, Q);
  return CNB_OK;
}

// 1 = tcgen05, 0 = CUDA cores (simt or depthwise), <0 = error
static int pick_algo(const cnb_conv_desc& d, int direction, const char* who) {
  if (is_depthwise(d)) {
    CNB_REQUIRE(d.algo != CNB_ALGO_TC, CNB_ERR_UNSUPPORTED,
                "%s: CNB_ALGO_TC requested for a depthwise convolution", who);
    return 0;
  }
  const char* why = "";
  bool ok = conv_tc_supported(d, direction, &why);
  if (d.algo == CNB_ALGO_TC) {
    CNB_REQUIRE(ok, CNB_ERR_UNSUPPORTED, "%s: CNB_ALGO_TC requested but: %s", who, why);
    return 1;
  }
  if (d.algo == CNB_ALGO_SIMT) return 0;
  return ok ? 1 : 0;
}

extern "C" {

int cnb_conv2d_out_dims(const cnb_conv_desc* d, int32_t* P, int32_t* Q) {
  CNB_TRY(validate_conv(d, "cnb_conv2d_out_dims"));
  CNB_REQUIRE(P && Q, CNB_ERR_INVALID, "cnb_conv2d_out_dims: null output");
  int p, q;
  conv_out_dims(*d, p, q);
  *P = p;
  *Q = q;
  return CNB_OK;
}

int cnb_conv2d_uses_tc(const cnb_conv_desc* d, int direction) {
  CNB_TRY(validate_conv(d, "cnb_conv2d_uses_tc"));
  CNB_REQUIRE(direction >= CNB_FWD && direction <= CNB_WGRAD, CNB_ERR_INVALID,
              "cnb_conv2d_uses_tc: bad direction %d", direction);
  return pick_algo(*d, direction, "cnb_conv2d_uses_tc");
}

size_t cnb_conv2d_workspace_bytes(const cnb_conv_desc* d, int direction) {
  if (validate_conv(d, "cnb_conv2d_workspace_bytes") != CNB_OK) return 0;
  if (is_depthwise(*d)) return direction == CNB_WGRAD ? dwconv_wgrad_ws(*d) : 0;
  int a = pick_algo(*d, direction, "cnb_conv2d_workspace_bytes");
  if (a < 0) return 0;
  if (a == 1) return conv_tc_workspace(*d, direction);
  return direction == CNB_WGRAD ? conv_simt_wgrad_ws(*d) : 0;
}

int cnb_conv2d_bn_partial_rows(const cnb_conv_desc* d) {
  CNB_TRY(validate_conv(d, "cnb_conv2d_bn_partial_rows"));
  if (is_depthwise(*d)) return dwconv_stats_rows(*d);
  int a = pick_algo(*d, CNB_FWD, "cnb_conv2d_bn_partial_rows");
  if (a < 0) return a;
  return a == 1 ? conv_tc_stats_rows(*d) : 0;
}

int cnb_conv2d_fwd(const cnb_conv_desc* d, const void* x, const void* w, const float* bias, void* y,
                   float* bn_partials, void* workspace, size_t workspace_bytes, void* stream) {
  CNB_TRY(validate_conv(d, "cnb_conv2d_fwd"));
  CNB_REQUIRE(x && w && y, CNB_ERR_INVALID, "cnb_conv2d_fwd: null tensor pointer");
  CNB_REQUIRE(!(d->flags & CNB_CONV_BIAS) || bias, CNB_ERR_INVALID,
              "cnb_conv2d_fwd: CNB_CONV_BIAS set but bias is NULL");
  cudaStream_t st = (cudaStream_t)stream;
  if (is_depthwise(*d)) {
    CNB_REQUIRE(d->algo != CNB_ALGO_TC, CNB_ERR_UNSUPPORTED, "cnb_conv2d_fwd: depthwise has no TC path");
    return dwconv_fwd(*d, x, w, bias, y, bn_partials, st);
  }
  int a = pick_algo(*d, CNB_FWD, "cnb_conv2d_fwd");
  if (a < 0) return a;
  if (a == 1) {
    CNB_REQUIRE(((uintptr_t)x | (uintptr_t)w | (uintptr_t)y) % 16 == 0, CNB_ERR_INVALID,
                "cnb_conv2d_fwd: tensor pointers must be 16-byte aligned for TMA");
    return conv_tc_fwd(*d, x, w, bias, y, bn_partials, workspace, workspace_bytes, st);
  }
  CNB_REQUIRE(!bn_partials, CNB_ERR_UNSUPPORTED,
              "cnb_conv2d_fwd: bn_partials requested on the CUDA-core path (cnb_conv2d_bn_partial_rows() == 0)");
  return conv_simt_fwd(*d, x, w, (d->flags & CNB_CONV_BIAS) ? bias : nullptr, y, st);
}

int cnb_conv2d_dgrad(const cnb_conv_desc* d, const void* dy, const void* w, const void* w_t, void* dx,
                     void* workspace, size_t workspace_bytes, void* stream) {
  CNB_TRY(validate_conv(d, "cnb_conv2d_dgrad"));
  CNB_REQUIRE(dy && w && dx, CNB_ERR_INVALID, "cnb_conv2d_dgrad: null tensor pointer");
  cudaStream_t st = (cudaStream_t)stream;
  if (is_depthwise(*d)) return dwconv_dgrad(*d, dy, w, dx, st);
  int a = pick_algo(*d, CNB_DGRAD, "cnb_conv2d_dgrad");
  if (a < 0) return a;
  if (a == 1) {
    CNB_REQUIRE(((uintptr_t)dy | (uintptr_t)w | (uintptr_t)dx) % 16 == 0, CNB_ERR_INVALID,
                "cnb_conv2d_dgrad: tensor pointers must be 16-byte aligned for TMA");
    CNB_REQUIRE(!w_t || (uintptr_t)w_t % 16 == 0, CNB_ERR_INVALID,
                "cnb_conv2d_dgrad: w_t must be 16-byte aligned");
    return conv_tc_dgrad(*d, dy, w, w_t, dx, workspace, workspace_bytes, st);
  }
  return conv_simt_dgrad(*d, dy, w, dx, st);
}

int cnb_conv2d_dgrad_add(const cnb_conv_desc* d, const void* dy, const void* w, const void* w_t,
                         const void* addend, const void* addend_mask, void* dx, void* workspace,
                         size_t workspace_bytes, void* stream) {
  CNB_TRY(validate_conv(d, "cnb_conv2d_dgrad_add"));
  CNB_REQUIRE(dy && w && dx && addend, CNB_ERR_INVALID, "cnb_conv2d_dgrad_add: null tensor pointer");
  CNB_REQUIRE(!is_depthwise(*d) && d->stride_h == 1 && d->stride_w == 1, CNB_ERR_UNSUPPORTED,
              "cnb_conv2d_dgrad_add: stride-1 tensor-core data-grad only (stride %dx%d, groups %d)",
              d->stride_h, d->stride_w, d->groups);
  const char* why = "";
  CNB_REQUIRE(d->algo != CNB_ALGO_SIMT && conv_tc_supported(*d, CNB_DGRAD, &why), CNB_ERR_UNSUPPORTED,
              "cnb_conv2d_dgrad_add: no tensor-core path: %s", why);
  CNB_REQUIRE(((uintptr_t)dy | (uintptr_t)w | (uintptr_t)dx | (uintptr_t)addend) % 16 == 0,
              CNB_ERR_INVALID, "cnb_conv2d_dgrad_add: tensor pointers must be 16-byte aligned");
  CNB_REQUIRE(!w_t || (uintptr_t)w_t % 16 == 0, CNB_ERR_INVALID,
              "cnb_conv2d_dgrad_add: w_t must be 16-byte aligned");
  CNB_REQUIRE(!addend_mask || d->C % 32 == 0, CNB_ERR_UNSUPPORTED,
              "cnb_conv2d_dgrad_add: a masked addend needs C %% 32 == 0 (C=%d)", d->C);
  return conv_tc_dgrad(*d, dy, w, w_t, dx, workspace, workspace_bytes, (cudaStream_t)stream, addend, addend_mask);
}

int cnb_conv2d_dgrad_partial_rows(const cnb_conv_desc* d) {
  CNB_TRY(validate_conv(d, "cnb_conv2d_dgrad_partial_rows"));
  const char* why = "";
  if (is_depthwise(*d) || d->stride_h != 1 || d->stride_w != 1 || d->algo == CNB_ALGO_SIMT ||
      !conv_tc_supported(*d, CNB_DGRAD, &why))
    return 0;
  return conv_tc_dgrad_stats_rows(*d);
}

int cnb_conv2d_dgrad_bn(const cnb_conv_desc* d, const void* dy, const void* w, const void* w_t,
                        const void* addend, const void* addend_mask, const void* bn_x,
                        const void* bn_relu_mask, const float* bn_mean, float* bn_partials, void* dx,
                        void* workspace, size_t workspace_bytes, void* stream) {
  CNB_TRY(validate_conv(d, "cnb_conv2d_dgrad_bn"));
  CNB_REQUIRE(dy && w && dx && bn_x && bn_relu_mask && bn_mean && bn_partials, CNB_ERR_INVALID,
              "cnb_conv2d_dgrad_bn: null tensor pointer");
  CNB_REQUIRE(!is_depthwise(*d) && d->stride_h == 1 && d->stride_w == 1, CNB_ERR_UNSUPPORTED,
              "cnb_conv2d_dgrad_bn: stride-1 tensor-core data-grad only (stride %dx%d, groups %d)",
              d->stride_h, d->stride_w, d->groups);
  const char* why = "";
  CNB_REQUIRE(d->algo != CNB_ALGO_SIMT && conv_tc_supported(*d, CNB_DGRAD, &why), CNB_ERR_UNSUPPORTED,
              "cnb_conv2d_dgrad_bn: no tensor-core path: %s", why);
  CNB_REQUIRE(((uintptr_t)dy | (uintptr_t)w | (uintptr_t)dx | (uintptr_t)addend | (uintptr_t)bn_x) % 16 == 0,
              CNB_ERR_INVALID, "cnb_conv2d_dgrad_bn: tensor pointers must be 16-byte aligned");
  CNB_REQUIRE(!w_t || (uintptr_t)w_t % 16 == 0, CNB_ERR_INVALID, "cnb_conv2d_dgrad_bn: w_t must be 16-byte aligned");
  CNB_REQUIRE(!addend_mask || (addend && d->C % 32 == 0), CNB_ERR_UNSUPPORTED,
              "cnb_conv2d_dgrad_bn: a masked addend needs the addend and C %% 32 == 0 (C=%d)", d->C);
  return conv_tc_dgrad(*d, dy, w, w_t, dx, workspace, workspace_bytes, (cudaStream_t)stream, addend, addend_mask,
                       bn_x, bn_relu_mask, bn_mean, bn_partials);
}

int cnb_conv2d_wgrad(const cnb_conv_desc* d, const void* x, const void* dy, float* dw, float* dbias,
                     void* workspace, size_t workspace_bytes, void* stream) {
  CNB_TRY(validate_conv(d, "cnb_conv2d_wgrad"));
  CNB_REQUIRE(x && dy && dw, CNB_ERR_INVALID, "cnb_conv2d_wgrad: null tensor pointer");
  cudaStream_t st = (cudaStream_t)stream;
  if (dbias) CNB_TRY(conv_bias_grad(*d, dy, dbias, st));
  if (is_depthwise(*d)) return dwconv_wgrad(*d, x, dy, dw, workspace, workspace_bytes, st);
  int a = pick_algo(*d, CNB_WGRAD, "cnb_conv2d_wgrad");
  if (a < 0) return a;
  if (a == 1) {
    CNB_REQUIRE(((uintptr_t)x | (uintptr_t)dy) % 16 == 0, CNB_ERR_INVALID,
                "cnb_conv2d_wgrad: tensor pointers must be 16-byte aligned for TMA");
    return conv_tc_wgrad(*d, x, dy, dw, workspace, workspace_bytes, st);
  }
  return conv_simt_wgrad(*d, x, dy, dw, workspace, workspace_bytes, st);
}

}  // extern "C"
