// pool.cu — max / average pooling forward and backward over NHWC tensors (HBM-bound).
// Replaces aten::max_pool2d_with_indices(_backward) and aten::avg_pool2d(_backward) reached from
// reference src/basemodel.py:1017-1033.  One thread handles V contiguous channels of one output
// (forward) or input (backward) pixel; backward is a gather, so it needs neither atomics nor a
// zero-fill pass and is deterministic.
#include <math_constants.h>

#include "common.cuh"

namespace cnb {

constexpr int kPoolThreads = 256;

struct PoolP {
  int N, H, W, C, P, Q, kh, kw, sh, sw, ph, pw;
};

// PyTorch rule (aten/src/ATen/native/cuda/DilatedMaxPool2d.cu): scan the window row-major over
// in-bounds positions only, take `val > max || isnan(val)`; index = h*W + w.
// K > 0: the window is at most K x K — all K*K loads are issued before the first compare so
// that each thread keeps K*K independent 16-byte requests in flight; K == 0: runtime loops.
template <typename T, int V, int K>
__global__ void maxpool_fwd_k(const T* __restrict__ x, T* __restrict__ y, int32_t* __restrict__ idx,
                              uint8_t* __restrict__ code, PoolP p, int64_t total_vec) {
  // one output row (n, oh) per block, (q, channel vector) over the threads: the row decomposition
  // is block-uniform and each item costs a single integer division
  const int Cv = p.C / V;
  (void)total_vec;
  const int oh = blockIdx.x % p.P;
  const int n = blockIdx.x / p.P;
  for (int item = threadIdx.x; item < p.Q * Cv; item += blockDim.x) {
    const int q = item / Cv;
    const int cv = item - q * Cv;
    const int64_t i = ((int64_t)blockIdx.x * p.Q + q) * Cv + cv;
    int h0 = oh * p.sh - p.ph, w0 = q * p.sw - p.pw;
    const int hs = h0, ws = w0;  // un-clamped window origin (the 1-byte code is relative to it)
    int h1 = min(h0 + p.kh, p.H), w1 = min(w0 + p.kw, p.W);
    h0 = max(h0, 0);
    w0 = max(w0, 0);
    float best[V];
    int bi[V];
#pragma unroll
    for (int j = 0; j < V; ++j) {
      best[j] = -CUDART_INF_F;
      bi[j] = h0 * p.W + w0;
    }
    const T* base = x + (int64_t)n * p.H * p.W * p.C + cv * V;
    if (K > 0) {
      Pack<T, V> a[K > 0 ? K * K : 1];
#pragma unroll
      for (int dh = 0; dh < K; ++dh)
#pragma unroll
        for (int dw = 0; dw < K; ++dw)
          if (h0 + dh < h1 && w0 + dw < w1)
            a[dh * K + dw] = ld_pack<T, V>(base + ((int64_t)(h0 + dh) * p.W + w0 + dw) * p.C);
#pragma unroll
      for (int dh = 0; dh < K; ++dh)
#pragma unroll
        for (int dw = 0; dw < K; ++dw)
          if (h0 + dh < h1 && w0 + dw < w1) {
#pragma unroll
            for (int j = 0; j < V; ++j) {
              float f = to_f(a[dh * K + dw].v[j]);
              if (f > best[j] || f != f) {
                best[j] = f;
                bi[j] = (h0 + dh) * p.W + w0 + dw;
              }
            }
          }
    } else {
      for (int h = h0; h < h1; ++h)
        for (int w = w0; w < w1; ++w) {
          Pack<T, V> a = ld_pack<T, V>(base + ((int64_t)h * p.W + w) * p.C);
#pragma unroll
          for (int j = 0; j < V; ++j) {
            float f = to_f(a.v[j]);
            if (f > best[j] || f != f) {
              best[j] = f;
              bi[j] = h * p.W + w;
            }
          }
        }
    }
    Pack<T, V> o;
#pragma unroll
    for (int j = 0; j < V; ++j) o.v[j] = from_f<T>(best[j]);
    st_pack<T, V>(y + i * V, o);
    if (idx) {
#pragma unroll
      for (int j0 = 0; j0 < V; j0 += 4) {
        constexpr int W4 = V < 4 ? V : 4;
        Pack<int32_t, W4> oi;
#pragma unroll
        for (int j = 0; j < W4; ++j) oi.v[j] = bi[j0 + j];
        st_pack<int32_t, W4>(idx + i * V + j0, oi);
      }
    }
    if (code) {
      Pack<uint8_t, V> oc;
#pragma unroll
      for (int j = 0; j < V; ++j) {
        const int bh = bi[j] / p.W, bw = bi[j] - bh * p.W;
        oc.v[j] = (uint8_t)((bh - hs) * p.kw + (bw - ws));
      }
      st_pack<uint8_t, V>(code + i * V, oc);
    }
  }
}

// dx[n,h,w,c] = sum over output windows containing (h,w) of dy * (idx == h*W+w)
// NW > 0: at most NW x NW output windows contain an input pixel (NW = ceil(k / stride)); their
// index / gradient loads are all issued before the compares.  NW == 0: runtime loops.
template <typename T, int V, int NW>
__global__ void maxpool_bwd_k(const T* __restrict__ dy, const int32_t* __restrict__ idx,
                              const uint8_t* __restrict__ code, T* __restrict__ dx, PoolP p,
                              int64_t total_vec) {
  // one input row (n, h) per block: the candidate output rows are block-uniform
  const int Cv = p.C / V;
  (void)total_vec;
  const int h = blockIdx.x % p.H;
  const int n = blockIdx.x / p.H;
  // output rows oh with oh*sh - ph <= h < oh*sh - ph + kh
  int oh_lo = (h + p.ph - p.kh + p.sh) / p.sh;  // ceil((h+ph-kh+1)/sh)
  if (h + p.ph - p.kh + 1 <= 0) oh_lo = 0;
  const int oh_hi = min((h + p.ph) / p.sh, p.P - 1);
  for (int item = threadIdx.x; item < p.W * Cv; item += blockDim.x) {
    const int w = item / Cv;
    const int cv = item - w * Cv;
    const int64_t i = ((int64_t)blockIdx.x * p.W + w) * Cv + cv;
    int ow_lo = (w + p.pw - p.kw + p.sw) / p.sw;
    if (w + p.pw - p.kw + 1 <= 0) ow_lo = 0;
    int ow_hi = min((w + p.pw) / p.sw, p.Q - 1);
    const int me = h * p.W + w;
    float acc[V];
#pragma unroll
    for (int j = 0; j < V; ++j) acc[j] = 0.f;
    if (NW > 0) {
      Pack<uint8_t, V> cc[NW > 0 ? NW * NW : 1];
      Pack<T, V> g[NW > 0 ? NW * NW : 1];
      if (code) {
        // 1-byte window-relative argmax codes: this pixel is position
        // (h - (oh*sh - ph)) * kw + (w - (ow*sw - pw)) of window (oh, ow)
#pragma unroll
        for (int a = 0; a < NW; ++a)
#pragma unroll
          for (int b = 0; b < NW; ++b)
            if (oh_lo + a <= oh_hi && ow_lo + b <= ow_hi) {
              int64_t o = (((int64_t)n * p.P + oh_lo + a) * p.Q + ow_lo + b) * p.C + cv * V;
              cc[a * NW + b] = ld_pack<uint8_t, V>(code + o);
              g[a * NW + b] = ld_pack<T, V>(dy + o);
            }
#pragma unroll
        for (int a = 0; a < NW; ++a)
#pragma unroll
          for (int b = 0; b < NW; ++b)
            if (oh_lo + a <= oh_hi && ow_lo + b <= ow_hi) {
              const int mine = (h - ((oh_lo + a) * p.sh - p.ph)) * p.kw + (w - ((ow_lo + b) * p.sw - p.pw));
#pragma unroll
              for (int j = 0; j < V; ++j)
                if ((int)cc[a * NW + b].v[j] == mine) acc[j] += to_f(g[a * NW + b].v[j]);
            }
      } else {
#pragma unroll
        for (int a = 0; a < NW; ++a)
#pragma unroll
          for (int b = 0; b < NW; ++b)
            if (oh_lo + a <= oh_hi && ow_lo + b <= ow_hi) {
              int64_t o = (((int64_t)n * p.P + oh_lo + a) * p.Q + ow_lo + b) * p.C + cv * V;
#pragma unroll
              for (int j = 0; j < V; ++j)
                if (idx[o + j] == me) acc[j] += to_f(dy[o + j]);
            }
      }
    } else {
      for (int oh = oh_lo; oh <= oh_hi; ++oh)
        for (int ow = ow_lo; ow <= ow_hi; ++ow) {
          int64_t o = (((int64_t)n * p.P + oh) * p.Q + ow) * p.C + cv * V;
          const int mine = (h - (oh * p.sh - p.ph)) * p.kw + (w - (ow * p.sw - p.pw));
#pragma unroll
          for (int j = 0; j < V; ++j) {
            const bool hit = code ? ((int)code[o + j] == mine) : (idx[o + j] == me);
            if (hit) acc[j] += to_f(dy[o + j]);
          }
        }
    }
    Pack<T, V> out;
#pragma unroll
    for (int j = 0; j < V; ++j) out.v[j] = from_f<T>(acc[j]);
    st_pack<T, V>(dx + i * V, out);
  }
}

// k3 s2 p1 (the stem pool of every ResNet-family model) with 1-byte argmax codes: a thread owns a
// 2x2 input quad (rows 2a, 2a+1; columns 2b, 2b+1).  Only the windows (a..a+1) x (b..b+1) reach
// it, and each of them can hit the quad at fixed window positions only (window a: rows r = 1, 2;
// window a+1: row r = 0; same for columns), so 4 code + 4 gradient vectors produce 4 outputs with
// 9 compile-time (window, position) compares per channel instead of 4 x 4.
template <typename T, int V>
__global__ void maxpool_bwd_k3s2p1_k(const T* __restrict__ dy, const uint8_t* __restrict__ code,
                                     T* __restrict__ dx, PoolP p) {
  const int Cv = p.C / V;
  const int HQ = (p.H + 1) / 2, WQ = (p.W + 1) / 2;
  const int a = blockIdx.x % HQ;
  const int n = blockIdx.x / HQ;
  for (int item = threadIdx.x; item < WQ * Cv; item += blockDim.x) {
    const int b = item / Cv;
    const int cv = item - b * Cv;
    Pack<uint8_t, V> cc[2][2];
    Pack<T, V> g[2][2];
    bool ok[2][2];
#pragma unroll
    for (int wa = 0; wa < 2; ++wa)
#pragma unroll
      for (int wb = 0; wb < 2; ++wb) {
        ok[wa][wb] = a + wa < p.P && b + wb < p.Q;
        if (ok[wa][wb]) {
          const int64_t o = (((int64_t)n * p.P + a + wa) * p.Q + b + wb) * p.C + cv * V;
          cc[wa][wb] = ld_pack<uint8_t, V>(code + o);
          g[wa][wb] = ld_pack<T, V>(dy + o);
        }
      }
    float acc[2][2][V];
#pragma unroll
    for (int dr = 0; dr < 2; ++dr)
#pragma unroll
      for (int dc = 0; dc < 2; ++dc)
#pragma unroll
        for (int j = 0; j < V; ++j) acc[dr][dc][j] = 0.f;
    // quad row dr sits at window row r = dr + 1 of window a and (dr == 1 only) r = 0 of window a+1
#pragma unroll
    for (int wa = 0; wa < 2; ++wa)
#pragma unroll
      for (int wb = 0; wb < 2; ++wb) {
        if (!ok[wa][wb]) continue;
#pragma unroll
        for (int dr = wa; dr < 2; ++dr)
#pragma unroll
          for (int dc = wb; dc < 2; ++dc) {
            const int r = wa ? 0 : dr + 1, s_ = wb ? 0 : dc + 1;
            const int want = r * 3 + s_;
#pragma unroll
            for (int j = 0; j < V; ++j)
              if ((int)cc[wa][wb].v[j] == want) acc[dr][dc][j] += to_f(g[wa][wb].v[j]);
          }
      }
#pragma unroll
    for (int dr = 0; dr < 2; ++dr)
#pragma unroll
      for (int dc = 0; dc < 2; ++dc) {
        const int h = 2 * a + dr, w = 2 * b + dc;
        if (h < p.H && w < p.W) {
          Pack<T, V> out;
#pragma unroll
          for (int j = 0; j < V; ++j) out.v[j] = from_f<T>(acc[dr][dc][j]);
          st_pack<T, V>(dx + ((((int64_t)n * p.H + h) * p.W + w) * Cv + cv) * V, out);
        }
      }
  }
}

// count_include_pad=True: divisor is always kh*kw; padded positions contribute zero.
template <typename T, int V>
__global__ void avgpool_fwd_k(const T* __restrict__ x, T* __restrict__ y, PoolP p,
                              int64_t total_vec) {
  const int Cv = p.C / V;
  const float inv = 1.f / (float)(p.kh * p.kw);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total_vec;
       i += (int64_t)gridDim.x * blockDim.x) {
    int cv = (int)(i % Cv);
    int64_t pix = i / Cv;
    int q = (int)(pix % p.Q);
    int64_t t = pix / p.Q;
    int oh = (int)(t % p.P);
    int n = (int)(t / p.P);
    int h0 = oh * p.sh - p.ph, w0 = q * p.sw - p.pw;
    int h1 = min(h0 + p.kh, p.H), w1 = min(w0 + p.kw, p.W);
    h0 = max(h0, 0);
    w0 = max(w0, 0);
    float acc[V];
#pragma unroll
    for (int j = 0; j < V; ++j) acc[j] = 0.f;
    const T* base = x + (int64_t)n * p.H * p.W * p.C + cv * V;
    for (int h = h0; h < h1; ++h)
      for (int w = w0; w < w1; ++w) {
        Pack<T, V> a = ld_pack<T, V>(base + ((int64_t)h * p.W + w) * p.C);
#pragma unroll
        for (int j = 0; j < V; ++j) acc[j] += to_f(a.v[j]);
      }
    Pack<T, V> o;
#pragma unroll
    for (int j = 0; j < V; ++j) o.v[j] = from_f<T>(acc[j] * inv);
    st_pack<T, V>(y + i * V, o);
  }
}

template <typename T, int V>
__global__ void avgpool_bwd_k(const T* __restrict__ dy, T* __restrict__ dx, PoolP p,
                              int64_t total_vec) {
  const int Cv = p.C / V;
  const float inv = 1.f / (float)(p.kh * p.kw);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total_vec;
       i += (int64_t)gridDim.x * blockDim.x) {
    int cv = (int)(i % Cv);
    int64_t pix = i / Cv;
    int w = (int)(pix % p.W);
    int64_t t = pix / p.W;
    int h = (int)(t % p.H);
    int n = (int)(t / p.H);
    int oh_lo = (h + p.ph - p.kh + p.sh) / p.sh;
    if (h + p.ph - p.kh + 1 <= 0) oh_lo = 0;
    int oh_hi = min((h + p.ph) / p.sh, p.P - 1);
    int ow_lo = (w + p.pw - p.kw + p.sw) / p.sw;
    if (w + p.pw - p.kw + 1 <= 0) ow_lo = 0;
    int ow_hi = min((w + p.pw) / p.sw, p.Q - 1);
    float acc[V];
#pragma unroll
    for (int j = 0; j < V; ++j) acc[j] = 0.f;
    for (int oh = oh_lo; oh <= oh_hi; ++oh)
      for (int ow = ow_lo; ow <= ow_hi; ++ow) {
        int64_t o = (((int64_t)n * p.P + oh) * p.Q + ow) * p.C + cv * V;
        Pack<T, V> g = ld_pack<T, V>(dy + o);
#pragma unroll
        for (int j = 0; j < V; ++j) acc[j] += to_f(g.v[j]);
      }
    Pack<T, V> out;
#pragma unroll
    for (int j = 0; j < V; ++j) out.v[j] = from_f<T>(acc[j] * inv);
    st_pack<T, V>(dx + i * V, out);
  }
}

static int make_pool(const cnb_pool_desc* d, PoolP& p, const char* who) {
  CNB_REQUIRE(d != nullptr, CNB_ERR_INVALID, "%s: null descriptor", who);
  CNB_REQUIRE(d->N > 0 && d->H > 0 && d->W > 0 && d->C > 0 && d->kh > 0 && d->kw > 0 &&
                  d->stride_h > 0 && d->stride_w > 0 && d->pad_h >= 0 && d->pad_w >= 0 &&
                  valid_dtype(d->dtype),
              CNB_ERR_INVALID, "%s: bad descriptor", who);
  // PyTorch: pad should be at most half of kernel size
  CNB_REQUIRE(d->pad_h * 2 <= d->kh && d->pad_w * 2 <= d->kw, CNB_ERR_INVALID,
              "%s: pad should be at most half of kernel size", who);
  p = PoolP{d->N, d->H, d->W, d->C, 0, 0, d->kh, d->kw, d->stride_h, d->stride_w, d->pad_h, d->pad_w};
  p.P = (d->H + 2 * d->pad_h - d->kh) / d->stride_h + 1;
  p.Q = (d->W + 2 * d->pad_w - d->kw) / d->stride_w + 1;
  CNB_REQUIRE(p.P > 0 && p.Q > 0, CNB_ERR_INVALID, "%s: output size is too small (%dx%d)", who, p.P,
              p.Q);
  return CNB_OK;
}

}  // namespace cnb

using namespace cnb;

extern "C" {

int cnb_maxpool_fwd(const cnb_pool_desc* d, const void* x, void* y, int32_t* idx, uint8_t* code,
                    void* stream) {
  PoolP p;
  CNB_TRY(make_pool(d, p, "cnb_maxpool_fwd"));
  CNB_REQUIRE(x && y, CNB_ERR_INVALID, "cnb_maxpool_fwd: null pointer");
  CNB_REQUIRE(!code || p.kh * p.kw <= 255, CNB_ERR_UNSUPPORTED,
              "cnb_maxpool_fwd: 1-byte argmax codes need kh*kw <= 255");
  int v = pick_vec(p.C, dtype_size(d->dtype), x, y);
  while (v > 1 && idx && (uintptr_t)idx % 16 != 0) v >>= 1;
  while (v > 1 && code && (uintptr_t)code % v != 0) v >>= 1;
  int64_t total = (int64_t)p.N * p.P * p.Q * (p.C / v);
  const int k = p.kh > p.kw ? p.kh : p.kw;
  CNB_REQUIRE((int64_t)p.N * p.P < (1ll << 31), CNB_ERR_UNSUPPORTED, "cnb_maxpool_fwd: too many rows");
  const int grid = p.N * p.P;
  cudaStream_t st = (cudaStream_t)stream;
#define MPF(KK) CNB_DISPATCH_TV(d->dtype, v, (maxpool_fwd_k<T, V, KK><<<grid, kPoolThreads, 0, st>>>( \
                                   (const T*)x, (T*)y, idx, code, p, total)))
  if (k <= 2) MPF(2);
  else if (k == 3) MPF(3);
  else MPF(0);
#undef MPF
  return check_launch("maxpool_fwd");
}

int cnb_maxpool_bwd(const cnb_pool_desc* d, const void* dy, const int32_t* idx, const uint8_t* code,
                    void* dx, void* stream) {
  PoolP p;
  CNB_TRY(make_pool(d, p, "cnb_maxpool_bwd"));
  CNB_REQUIRE(dy && dx && (idx || code), CNB_ERR_INVALID,
              "cnb_maxpool_bwd: null pointer (needs idx or code)");
  int v = pick_vec(p.C, dtype_size(d->dtype), dy, dx);
  while (v > 1 && code && (uintptr_t)code % v != 0) v >>= 1;
  int64_t total = (int64_t)p.N * p.H * p.W * (p.C / v);
  const int nwh = (p.kh + p.sh - 1) / p.sh, nww = (p.kw + p.sw - 1) / p.sw;
  const int nw = nwh > nww ? nwh : nww;
  CNB_REQUIRE((int64_t)p.N * p.H < (1ll << 31), CNB_ERR_UNSUPPORTED, "cnb_maxpool_bwd: too many rows");
  const int grid = p.N * p.H;
  cudaStream_t st = (cudaStream_t)stream;
  if (code && p.kh == 3 && p.kw == 3 && p.sh == 2 && p.sw == 2 && p.ph == 1 && p.pw == 1) {
    const int items = ((p.W + 1) / 2) * (p.C / v);
    const int passes = (items + kPoolThreads - 1) / kPoolThreads;
    int threads = ((items + passes - 1) / passes + 31) / 32 * 32;
    threads = threads < 32 ? 32 : (threads > kPoolThreads ? kPoolThreads : threads);
    const int qgrid = p.N * ((p.H + 1) / 2);
    CNB_DISPATCH_TV(d->dtype, v, (maxpool_bwd_k3s2p1_k<T, V><<<qgrid, threads, 0, st>>>(
                                     (const T*)dy, code, (T*)dx, p)));
    return check_launch("maxpool_bwd(k3s2p1)");
  }
#define MPB(NN) CNB_DISPATCH_TV(d->dtype, v, (maxpool_bwd_k<T, V, NN><<<grid, kPoolThreads, 0, st>>>( \
                                   (const T*)dy, code ? nullptr : idx, code, (T*)dx, p, total)))
  if (nw == 1) MPB(1);
  else if (nw == 2) MPB(2);
  else if (nw == 3) MPB(3);
  else MPB(0);
#undef MPB
  return check_launch("maxpool_bwd");
}

int cnb_avgpool_fwd(const cnb_pool_desc* d, const void* x, void* y, void* stream) {
  PoolP p;
  CNB_TRY(make_pool(d, p, "cnb_avgpool_fwd"));
  CNB_REQUIRE(x && y, CNB_ERR_INVALID, "cnb_avgpool_fwd: null pointer");
  int v = pick_vec(p.C, dtype_size(d->dtype), x, y);
  int64_t total = (int64_t)p.N * p.P * p.Q * (p.C / v);
  CNB_DISPATCH_TV(d->dtype, v,
                  (avgpool_fwd_k<T, V><<<grid_for(total, kPoolThreads), kPoolThreads, 0,
                                         (cudaStream_t)stream>>>((const T*)x, (T*)y, p, total)));
  return check_launch("avgpool_fwd");
}

int cnb_avgpool_bwd(const cnb_pool_desc* d, const void* dy, void* dx, void* stream) {
  PoolP p;
  CNB_TRY(make_pool(d, p, "cnb_avgpool_bwd"));
  CNB_REQUIRE(dy && dx, CNB_ERR_INVALID, "cnb_avgpool_bwd: null pointer");
  int v = pick_vec(p.C, dtype_size(d->dtype), dy, dx);
  int64_t total = (int64_t)p.N * p.H * p.W * (p.C / v);
  CNB_DISPATCH_TV(d->dtype, v,
                  (avgpool_bwd_k<T, V><<<grid_for(total, kPoolThreads), kPoolThreads, 0,
                                         (cudaStream_t)stream>>>((const T*)dy, (T*)dx, p, total)));
  return check_launch("avgpool_bwd");
}

}  // extern "C"
