// common.cuh — shared host/device helpers for libconvnets_b200 (error state, dtype dispatch,
// vector packs, launch accounting).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>

#include "../../include/convnets_b200.h"

namespace cnb {

// ------------------------------------------------------------------ error state (thread-local)
extern thread_local char g_last_error[512];
extern std::atomic<uint64_t> g_launches;

inline int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
  va_end(ap);
  return code;
}

inline int check_launch(const char* what) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return set_error(CNB_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
  return CNB_OK;
}

#define CNB_REQUIRE(cond, code, ...) \
  do {                               \
    if (!(cond)) return ::cnb::set_error(code, __VA_ARGS__); \
  } while (0)

#define CNB_TRY(expr)        \
  do {                       \
    int _s = (expr);         \
    if (_s != CNB_OK) return _s; \
  } while (0)

inline size_t dtype_size(int dt) { return dt == CNB_BF16 ? 2 : 4; }
inline bool valid_dtype(int dt) { return dt == CNB_F32 || dt == CNB_BF16; }
inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

// number of SMs of the current device (cached per device)
int sm_count();

// ------------------------------------------------------------------ device scalar conversion
using bf16 = __nv_bfloat16;

template <typename T>
__device__ __forceinline__ float to_f(T v);
template <>
__device__ __forceinline__ float to_f<float>(float v) {
  return v;
}
template <>
__device__ __forceinline__ float to_f<bf16>(bf16 v) {
  return __bfloat162float(v);
}
template <typename T>
__device__ __forceinline__ T from_f(float v);
template <>
__device__ __forceinline__ float from_f<float>(float v) {
  return v;
}
template <>
__device__ __forceinline__ bf16 from_f<bf16>(float v) {
  return __float2bfloat16_rn(v);
}

// ------------------------------------------------------------------ aligned vector packs
template <typename T, int V>
struct alignas(sizeof(T) * V) Pack {
  T v[V];
};

template <typename T, int V>
__device__ __forceinline__ Pack<T, V> ld_pack(const T* p) {
  return *reinterpret_cast<const Pack<T, V>*>(p);
}
template <typename T, int V>
__device__ __forceinline__ void st_pack(T* p, const Pack<T, V>& v) {
  *reinterpret_cast<Pack<T, V>*>(p) = v;
}

// largest vector width (elements) usable for rows of `C` elements of size `esz` starting at
// 16B-aligned bases: V*esz <= 16 and C % V == 0
inline int pick_vec(int64_t C, size_t esz, const void* p0 = nullptr, const void* p1 = nullptr,
                    const void* p2 = nullptr, const void* p3 = nullptr) {
  int v = static_cast<int>(16 / esz);
  auto misaligned = [&](const void* p, int vv) {
    return p != nullptr && (reinterpret_cast<uintptr_t>(p) % (vv * esz)) != 0;
  };
  while (v > 1 && (C % v != 0 || misaligned(p0, v) || misaligned(p1, v) || misaligned(p2, v) ||
                   misaligned(p3, v)))
    v >>= 1;
  return v;
}

// Dispatch helpers: call f.template operator()<T, V>() for the runtime (dtype, vec) pair.
#define CNB_DISPATCH_TV(dtype, vec, ...)                                   \
  do {                                                                     \
    if ((dtype) == CNB_BF16) {                                             \
      using T = ::cnb::bf16;                                               \
      switch (vec) {                                                       \
        case 8: { constexpr int V = 8; __VA_ARGS__; } break;               \
        case 4: { constexpr int V = 4; __VA_ARGS__; } break;               \
        case 2: { constexpr int V = 2; __VA_ARGS__; } break;               \
        default: { constexpr int V = 1; __VA_ARGS__; } break;              \
      }                                                                    \
    } else {                                                               \
      using T = float;                                                     \
      switch (vec) {                                                       \
        case 4: { constexpr int V = 4; __VA_ARGS__; } break;               \
        case 2: { constexpr int V = 2; __VA_ARGS__; } break;               \
        default: { constexpr int V = 1; __VA_ARGS__; } break;              \
      }                                                                    \
    }                                                                      \
  } while (0)

#define CNB_DISPATCH_T(dtype, ...)        \
  do {                                    \
    if ((dtype) == CNB_BF16) {            \
      using T = ::cnb::bf16;              \
      __VA_ARGS__;                        \
    } else {                              \
      using T = float;                    \
      __VA_ARGS__;                        \
    }                                     \
  } while (0)

inline int grid_for(int64_t work_items, int threads, int max_waves = 8) {
  int64_t blocks = ceil_div(work_items, threads);
  int64_t cap = static_cast<int64_t>(sm_count()) * max_waves;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks);
}

// ------------------------------------------------------------------ internal entry points
// conv_simt.cu
int conv_simt_fwd(const cnb_conv_desc& d, const void* x, const void* w, const float* bias, void* y,
                  cudaStream_t st);
int conv_simt_dgrad(const cnb_conv_desc& d, const void* dy, const void* w, void* dx,
                    cudaStream_t st);
int conv_simt_wgrad(const cnb_conv_desc& d, const void* x, const void* dy, float* dw, void* ws,
                    size_t ws_bytes, cudaStream_t st);
size_t conv_simt_wgrad_ws(const cnb_conv_desc& d);
int conv_bias_grad(const cnb_conv_desc& d, const void* dy, float* dbias, cudaStream_t st);
// dwconv.cu
bool is_depthwise(const cnb_conv_desc& d);
int dwconv_fwd(const cnb_conv_desc& d, const void* x, const void* w, const float* bias, void* y,
               float* bn_partials, cudaStream_t st);
int dwconv_stats_rows(const cnb_conv_desc& d);
int dwconv_dgrad(const cnb_conv_desc& d, const void* dy, const void* w, void* dx, cudaStream_t st);
int dwconv_wgrad(const cnb_conv_desc& d, const void* x, const void* dy, float* dw, void* ws,
                 size_t ws_bytes, cudaStream_t st);
size_t dwconv_wgrad_ws(const cnb_conv_desc& d);
// conv_tc.cu
bool conv_tc_supported(const cnb_conv_desc& d, int direction, const char** why);
size_t conv_tc_workspace(const cnb_conv_desc& d, int direction);
int conv_tc_fwd(const cnb_conv_desc& d, const void* x, const void* w, const float* bias, void* y,
                float* stats, void* ws, size_t ws_bytes, cudaStream_t st);
int conv_tc_stats_rows(const cnb_conv_desc& d);
int conv_tc_dgrad(const cnb_conv_desc& d, const void* dy, const void* w, const void* w_t, void* dx,
                  void* ws, size_t ws_bytes, cudaStream_t st, const void* addend = nullptr,
                  const void* addend_mask = nullptr, const void* bn_x = nullptr, const void* bn_mask = nullptr,
                  const float* bn_mean = nullptr, float* bn_partials = nullptr);
int conv_tc_dgrad_stats_rows(const cnb_conv_desc& d);
int conv_tc_wgrad(const cnb_conv_desc& d, const void* x, const void* dy, float* dw, void* ws,
                  size_t ws_bytes, cudaStream_t st);

inline void conv_out_dims(const cnb_conv_desc& d, int& P, int& Q) {
  P = (d.H + 2 * d.pad_h - d.dil_h * (d.R - 1) - 1) / d.stride_h + 1;
  Q = (d.W + 2 * d.pad_w - d.dil_w * (d.S - 1) - 1) / d.stride_w + 1;
}

}  // namespace cnb
