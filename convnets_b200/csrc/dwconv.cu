// dwconv.cu — depthwise convolution (groups == C == K) forward / data-grad / weight-grad on CUDA
// cores.  Replaces aten::convolution(groups=C) reached from reference src/basemodel.py:981-997
// (conv2d_depthwise), src/mobilenet_v1.py:102 and src/shufflenet_v1.py:126-133.
// Arithmetic intensity is ~R*S flops per 2 bytes -> HBM-bound, so the kernels are plain
// 128-bit-vectorised NHWC gathers: one thread owns V contiguous channels of one pixel; the
// R*S*V filter taps of a thread are read through L1 (the whole filter is <= 9*1024 elements).
#include <stdlib.h>

#include <type_traits>

#include "common.cuh"

namespace cnb {

constexpr int kDwThreads = 256;

struct DwP {
  int N, H, W, C, P, Q, R, S, sh, sw, ph, pw, dh, dw;
};

bool is_depthwise(const cnb_conv_desc& d) {
  return d.groups > 1 && d.groups == d.C && d.K == d.C;
}

static DwP make_dw(const cnb_conv_desc& d) {
  DwP p{d.N, d.H, d.W, d.C, 0, 0, d.R, d.S, d.stride_h, d.stride_w, d.pad_h, d.pad_w, d.dil_h, d.dil_w};
  conv_out_dims(d, p.P, p.Q);
  return p;
}

// w layout: [C][R][S] (KRSC with C/g == 1)
template <typename T, int V, bool BIAS, bool RELU>
__global__ void dw_fwd_k(const T* __restrict__ x, const T* __restrict__ w,
                         const float* __restrict__ bias, T* __restrict__ y, DwP p,
                         int64_t total_vec) {
  const int Cv = p.C / V;
  const int RS = p.R * p.S;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total_vec;
       i += (int64_t)gridDim.x * blockDim.x) {
    int cv = (int)(i % Cv);
    int64_t pix = i / Cv;
    int q = (int)(pix % p.Q);
    int64_t t = pix / p.Q;
    int oh = (int)(t % p.P);
    int n = (int)(t / p.P);
    float acc[V];
#pragma unroll
    for (int j = 0; j < V; ++j) acc[j] = BIAS ? bias[cv * V + j] : 0.f;
    const T* xb = x + (int64_t)n * p.H * p.W * p.C + cv * V;
    const T* wb = w + (int64_t)cv * V * RS;
    for (int r = 0; r < p.R; ++r) {
      int h = oh * p.sh - p.ph + r * p.dh;
      if (h < 0 || h >= p.H) continue;
      for (int s = 0; s < p.S; ++s) {
        int ww = q * p.sw - p.pw + s * p.dw;
        if (ww < 0 || ww >= p.W) continue;
        Pack<T, V> a = ld_pack<T, V>(xb + ((int64_t)h * p.W + ww) * p.C);
#pragma unroll
        for (int j = 0; j < V; ++j)
          acc[j] = fmaf(to_f(a.v[j]), to_f(wb[j * RS + r * p.S + s]), acc[j]);
      }
    }
    Pack<T, V> o;
#pragma unroll
    for (int j = 0; j < V; ++j) {
      float f = acc[j];
      if (RELU) f = (f > 0.f || f != f) ? f : 0.f;
      o.v[j] = from_f<T>(f);
    }
    st_pack<T, V>(y + i * V, o);
  }
}

// dx[n,h,w,c] = sum_{r,s} dy[n,(h+ph-r*dh)/sh,(w+pw-s*dw)/sw,c] * w[c,r,s]  (when divisible)
template <typename T, int V>
__global__ void dw_dgrad_k(const T* __restrict__ dy, const T* __restrict__ w, T* __restrict__ dx,
                           DwP p, int64_t total_vec) {
  const int Cv = p.C / V;
  const int RS = p.R * p.S;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total_vec;
       i += (int64_t)gridDim.x * blockDim.x) {
    int cv = (int)(i % Cv);
    int64_t pix = i / Cv;
    int ww = (int)(pix % p.W);
    int64_t t = pix / p.W;
    int h = (int)(t % p.H);
    int n = (int)(t / p.H);
    float acc[V];
#pragma unroll
    for (int j = 0; j < V; ++j) acc[j] = 0.f;
    const T* gb = dy + (int64_t)n * p.P * p.Q * p.C + cv * V;
    const T* wb = w + (int64_t)cv * V * RS;
    for (int r = 0; r < p.R; ++r) {
      int hh = h + p.ph - r * p.dh;
      if (hh < 0 || hh % p.sh != 0) continue;
      int oh = hh / p.sh;
      if (oh >= p.P) continue;
      for (int s = 0; s < p.S; ++s) {
        int wv = ww + p.pw - s * p.dw;
        if (wv < 0 || wv % p.sw != 0) continue;
        int ow = wv / p.sw;
        if (ow >= p.Q) continue;
        Pack<T, V> g = ld_pack<T, V>(gb + ((int64_t)oh * p.Q + ow) * p.C);
#pragma unroll
        for (int j = 0; j < V; ++j)
          acc[j] = fmaf(to_f(g.v[j]), to_f(wb[j * RS + r * p.S + s]), acc[j]);
      }
    }
    Pack<T, V> o;
#pragma unroll
    for (int j = 0; j < V; ++j) o.v[j] = from_f<T>(acc[j]);
    st_pack<T, V>(dx + i * V, o);
  }
}

// dw[c,r,s] = sum_{n,p,q} dy[n,p,q,c] * x[n, p*sh-ph+r*dh, q*sw-pw+s*dw, c]
// grid (channel blocks, pixel splits); block (cx, ry).  Each thread keeps R*S*V partial sums
// (V <= 4, R*S <= 9 on this path), block-reduced in shared memory, written as
// partial[split][c][rs]; a second kernel sums the splits (deterministic).
template <typename T, int V, int RS_MAX>
__global__ void dw_wgrad_k(const T* __restrict__ x, const T* __restrict__ dy,
                           float* __restrict__ part, DwP p, int64_t rows_per_split) {
  extern __shared__ float red[];  // [ry][cx*V*RS]
  const int RS = p.R * p.S;
  const int cv = blockIdx.x * blockDim.x + threadIdx.x;
  const int c0 = cv * V;
  const int64_t Mout = (int64_t)p.N * p.P * p.Q;
  const int64_t r0 = blockIdx.y * rows_per_split;
  int64_t r1 = r0 + rows_per_split;
  if (r1 > Mout) r1 = Mout;
  float acc[RS_MAX][V];
#pragma unroll
  for (int t = 0; t < RS_MAX; ++t)
#pragma unroll
    for (int j = 0; j < V; ++j) acc[t][j] = 0.f;
  if (c0 < p.C) {
    for (int64_t m = r0 + threadIdx.y; m < r1; m += blockDim.y) {
      int q = (int)(m % p.Q);
      int64_t t = m / p.Q;
      int oh = (int)(t % p.P);
      int n = (int)(t / p.P);
      Pack<T, V> g = ld_pack<T, V>(dy + m * p.C + c0);
      float gf[V];
#pragma unroll
      for (int j = 0; j < V; ++j) gf[j] = to_f(g.v[j]);
      const T* xb = x + (int64_t)n * p.H * p.W * p.C + c0;
#pragma unroll
      for (int tap = 0; tap < RS_MAX; ++tap) {
        if (tap < RS) {
          int r = tap / p.S, s = tap - r * p.S;
          int h = oh * p.sh - p.ph + r * p.dh;
          int ww = q * p.sw - p.pw + s * p.dw;
          if (h >= 0 && h < p.H && ww >= 0 && ww < p.W) {
            Pack<T, V> a = ld_pack<T, V>(xb + ((int64_t)h * p.W + ww) * p.C);
#pragma unroll
            for (int j = 0; j < V; ++j) acc[tap][j] = fmaf(gf[j], to_f(a.v[j]), acc[tap][j]);
          }
        }
      }
    }
  }
  const int Wd = blockDim.x * V * RS;
#pragma unroll
  for (int tap = 0; tap < RS_MAX; ++tap)
    if (tap < RS)
#pragma unroll
      for (int j = 0; j < V; ++j)
        red[threadIdx.y * Wd + (threadIdx.x * V + j) * RS + tap] = acc[tap][j];
  __syncthreads();
  // cooperative column sums over ry
  const int cbase = blockIdx.x * blockDim.x * V;
  for (int e = threadIdx.y * blockDim.x + threadIdx.x; e < Wd; e += blockDim.x * blockDim.y) {
    int c = cbase + e / RS;
    if (c < p.C) {
      float s = 0.f;
      for (int r = 0; r < (int)blockDim.y; ++r) s += red[r * Wd + e];
      part[((int64_t)blockIdx.y * p.C + c) * RS + (e % RS)] = s;
    }
  }
}

// ------------------------------------------------------------------------------------------
// 3x3 / pad 1 / dilation 1 strip kernels (every depthwise layer of MobileNet-v1 and ShuffleNet-v1).
// A thread owns one 16-byte channel vector and walks strips of TQ adjacent outputs of one row: the
// 9 x V filter taps stay in registers for the thread's whole lifetime (the channel vector of a
// thread never changes: blockDim.x spans channels, rows of the thread grid span strips), and each
// input vector is loaded once per strip and used by every (output, tap) pair it feeds — 4.5 loads
// per output at stride 1 instead of 9 vector + 72 scalar filter loads.
constexpr int kDwTQ = 4;
// 4 channels per thread (8-byte bf16 / 16-byte fp32 vectors): 36 filter + 16 accumulator + 36..72
// input registers leave room for two 256-thread CTAs per SM with ALL loads of a strip in flight
constexpr int kDwV = 4;

struct DwS {
  int N, IH, IW, OH, OW, C;   // input / output spatial extents of THIS launch (dgrad swaps them)
  int cv_total;               // channel vectors
  int strips_per_row;         // ceil(OW / kDwTQ)
  int64_t total_strips;       // N * OH * strips_per_row
  int flags;
};

// out(oh, q) = sum_{r,s} in(oh*STRIDE - 1 + r, q*STRIDE - 1 + s) * w[r][s]   (FLIP: w[2-r][2-s],
// which with STRIDE == 1 is exactly the data gradient of the pad-1 convolution)
// STATS: the kernel also emits, per CTA row (blockIdx.y), the per-channel sums of the STORED
// outputs and of their squares — the same [rows][2][C] table the tensor-core conv epilogue writes —
// so a BatchNorm2d that follows the depthwise conv (Conv2dBlockDW, reference
// src/mobilenet_v1.py:90-125; ShuffleUnit, src/shufflenet_v1.py:119-141) does not re-read the tensor
// for its statistics.  Threads keep their sums for their whole lifetime, the CTA combines its strip
// lanes through shared memory in fixed order (deterministic).
template <int V>
__device__ __forceinline__ void dw_emit_stats(const float (&ss)[V], const float (&sq)[V], float* red,
                                              float* __restrict__ stats, int C, int c0, bool active) {
  float* mine = red + ((size_t)threadIdx.y * blockDim.x + threadIdx.x) * 2 * V;
#pragma unroll
  for (int j = 0; j < V; ++j) { mine[j] = ss[j]; mine[V + j] = sq[j]; }
  __syncthreads();
  if (threadIdx.y == 0 && active) {
#pragma unroll
    for (int j = 0; j < V; ++j) {
      float S = 0.f, Q = 0.f;
      for (int y = 0; y < (int)blockDim.y; ++y) {
        const float* r = red + ((size_t)y * blockDim.x + threadIdx.x) * 2 * V;
        S += r[j];
        Q += r[V + j];
      }
      stats[((int64_t)blockIdx.y * 2 + 0) * C + c0 + j] = S;
      stats[((int64_t)blockIdx.y * 2 + 1) * C + c0 + j] = Q;
    }
  }
}

template <typename T, int STRIDE, bool FLIP, int TQ = kDwTQ, bool STATS = false>
__global__ void __launch_bounds__(256, sizeof(T) == 2 ? 2 : 1)
dw3x3_strip_k(const T* __restrict__ in, const T* __restrict__ w, const float* __restrict__ bias,
              T* __restrict__ out, DwS p, float* __restrict__ stats = nullptr) {
  constexpr int V = kDwV;
  constexpr int NCOL = TQ * STRIDE + 3 - STRIDE;
  extern __shared__ float dw_red[];
  const int cv = blockIdx.x * blockDim.x + threadIdx.x;
  const bool active = cv < p.cv_total;
  float ss[V], sq[V];
#pragma unroll
  for (int j = 0; j < V; ++j) ss[j] = sq[j] = 0.f;
  // Threads past the last channel vector leave at once — unless the CTA still has the statistics
  // barrier ahead of it: every thread of a warp must reach the SAME __syncthreads, so with STATS the
  // idle threads skip the loop (zero trips) and fall through to the one call site below.
  if (!STATS && !active) return;
  const int c0 = active ? cv * V : 0;
  float wr[9][V];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int j = 0; j < V; ++j) wr[t][j] = to_f(w[(int64_t)(c0 + j) * 9 + (FLIP ? 8 - t : t)]);
  float bz[V];
#pragma unroll
  for (int j = 0; j < V; ++j) bz[j] = (p.flags & CNB_CONV_BIAS) ? bias[c0 + j] : 0.f;
  const bool relu = p.flags & CNB_CONV_RELU;
  const int64_t step = (int64_t)gridDim.y * blockDim.y;
  const int64_t my_strips = active ? p.total_strips : 0;
  for (int64_t sidx = (int64_t)blockIdx.y * blockDim.y + threadIdx.y; sidx < my_strips; sidx += step) {
    const int qs = (int)(sidx % p.strips_per_row);
    const int64_t t2 = sidx / p.strips_per_row;
    const int oh = (int)(t2 % p.OH);
    const int n = (int)(t2 / p.OH);
    const int q0 = qs * TQ;
    float acc[TQ][V];
#pragma unroll
    for (int q = 0; q < TQ; ++q)
#pragma unroll
      for (int j = 0; j < V; ++j) acc[q][j] = bz[j];
    const T* ib = in + (int64_t)n * p.IH * p.IW * p.C + c0;
    const int iw0 = q0 * STRIDE - 1;
    // issue every load of the strip (3 rows x NCOL vectors) before the first FMA
    Pack<T, V> a[3][NCOL];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int ih = oh * STRIDE - 1 + r;
      const bool rok = ih >= 0 && ih < p.IH;
      const T* rowp = ib + (int64_t)ih * p.IW * p.C;
#pragma unroll
      for (int col = 0; col < NCOL; ++col) {
        const int iw = iw0 + col;
        if (rok && iw >= 0 && iw < p.IW) a[r][col] = ld_pack<T, V>(rowp + (int64_t)iw * p.C);
        else {
#pragma unroll
          for (int j = 0; j < V; ++j) a[r][col].v[j] = from_f<T>(0.f);
        }
      }
    }
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int q = 0; q < TQ; ++q)
#pragma unroll
        for (int s = 0; s < 3; ++s) {
#pragma unroll
          for (int j = 0; j < V; ++j)
            acc[q][j] = fmaf(to_f(a[r][q * STRIDE + s].v[j]), wr[r * 3 + s][j], acc[q][j]);
        }
    T* ob = out + (((int64_t)n * p.OH + oh) * p.OW + q0) * p.C + c0;
#pragma unroll
    for (int q = 0; q < TQ; ++q) {
      if (q0 + q < p.OW) {
        Pack<T, V> o;
#pragma unroll
        for (int j = 0; j < V; ++j) {
          float f = acc[q][j];
          if (relu) f = (f > 0.f || f != f) ? f : 0.f;
          o.v[j] = from_f<T>(f);
          if (STATS) {
            const float r = to_f(o.v[j]);
            ss[j] += r;
            sq[j] = fmaf(r, r, sq[j]);
          }
        }
        st_pack<T, V>(ob + (int64_t)q * p.C, o);
      }
    }
  }
  if (STATS) dw_emit_stats<V>(ss, sq, dw_red, stats, p.C, c0, active);
}

// Stride-1 column walk (bf16): a thread owns 4 channels x 4 output columns and walks kDwRH output
// rows DOWN the image with a 4-slot register ring of input rows — every input row is loaded once
// per walk instead of three times (once per filter row), and the row needed two steps ahead is
// requested before the current output row is computed.  Units = (image, row block, column strip).
constexpr int kDwRH = 8;

template <typename T, bool FLIP, bool STATS = false>
__global__ void __launch_bounds__(128, 3)
dw3x3_s1_walk_k(const T* __restrict__ in, const T* __restrict__ w, const float* __restrict__ bias,
                T* __restrict__ out, DwS p, int row_blocks, float* __restrict__ stats = nullptr) {
  constexpr int V = kDwV;
  constexpr int NCOL = kDwTQ + 2;
  extern __shared__ float dw_red[];
  const int cv = blockIdx.x * blockDim.x + threadIdx.x;
  const bool active = cv < p.cv_total;
  float ss[V], sq[V];
#pragma unroll
  for (int j = 0; j < V; ++j) ss[j] = sq[j] = 0.f;
  if (!STATS && !active) return;   // with STATS every thread reaches the one barrier below (see the strip kernel)
  const int c0 = active ? cv * V : 0;
  float wr[9][V];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int j = 0; j < V; ++j) wr[t][j] = to_f(w[(int64_t)(c0 + j) * 9 + (FLIP ? 8 - t : t)]);
  float bz[V];
#pragma unroll
  for (int j = 0; j < V; ++j) bz[j] = (p.flags & CNB_CONV_BIAS) ? bias[c0 + j] : 0.f;
  const bool relu = p.flags & CNB_CONV_RELU;
  const int64_t units = active ? (int64_t)p.N * row_blocks * p.strips_per_row : 0;
  const int64_t step = (int64_t)gridDim.y * blockDim.y;
  for (int64_t u = (int64_t)blockIdx.y * blockDim.y + threadIdx.y; u < units; u += step) {
    const int qs = (int)(u % p.strips_per_row);
    const int64_t t2 = u / p.strips_per_row;
    const int rb = (int)(t2 % row_blocks);
    const int n = (int)(t2 / row_blocks);
    const int q0 = qs * kDwTQ;
    const int oh0 = rb * kDwRH;
    const T* ib = in + (int64_t)n * p.IH * p.IW * p.C + c0;
    T* ob = out + ((int64_t)n * p.OH * p.OW + q0) * p.C + c0;
    const int iw0 = q0 - 1;
    Pack<T, V> S[4][NCOL];
    auto ld_row = [&](Pack<T, V>(&dst)[NCOL], int ih) {
      const bool rok = ih >= 0 && ih < p.IH;
      const T* rowp = ib + (int64_t)ih * p.IW * p.C;
#pragma unroll
      for (int col = 0; col < NCOL; ++col) {
        const int iw = iw0 + col;
        if (rok && iw >= 0 && iw < p.IW) dst[col] = ld_pack<T, V>(rowp + (int64_t)iw * p.C);
        else {
#pragma unroll
          for (int j = 0; j < V; ++j) dst[col].v[j] = from_f<T>(0.f);
        }
      }
    };
    ld_row(S[0], oh0 - 1);
    ld_row(S[1], oh0);
    ld_row(S[2], oh0 + 1);
#pragma unroll 1
    for (int kk = 0; kk < kDwRH; kk += 4)
#pragma unroll
    for (int k = 0; k < 4; ++k) {  // ring slots are compile-time constants inside this loop
      ld_row(S[(k + 3) & 3], oh0 + kk + k + 2);  // needed by output row kk + k + 1
      const int oh = oh0 + kk + k;
      if (oh < p.OH) {
        float acc[kDwTQ][V];
#pragma unroll
        for (int q = 0; q < kDwTQ; ++q)
#pragma unroll
          for (int j = 0; j < V; ++j) acc[q][j] = bz[j];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int q = 0; q < kDwTQ; ++q)
#pragma unroll
            for (int sx = 0; sx < 3; ++sx)
#pragma unroll
              for (int j = 0; j < V; ++j)
                acc[q][j] = fmaf(to_f(S[(k + r) & 3][q + sx].v[j]), wr[r * 3 + sx][j], acc[q][j]);
#pragma unroll
        for (int q = 0; q < kDwTQ; ++q) {
          if (q0 + q < p.OW) {
            Pack<T, V> o;
#pragma unroll
            for (int j = 0; j < V; ++j) {
              float f = acc[q][j];
              if (relu) f = (f > 0.f || f != f) ? f : 0.f;
              o.v[j] = from_f<T>(f);
              if (STATS) {
                const float r = to_f(o.v[j]);
                ss[j] += r;
                sq[j] = fmaf(r, r, sq[j]);
              }
            }
            st_pack<T, V>(ob + ((int64_t)oh * p.OW + q) * p.C, o);
          }
        }
      }
    }
  }
  if (STATS) dw_emit_stats<V>(ss, sq, dw_red, stats, p.C, c0, active);
}

// data gradient of the stride-2 / pad-1 3x3 depthwise conv: a strip of 4 dx columns (w0 % 4 == 0)
// of row h gathers from dy rows (h+1-r)/2, r of the parity of h+1, and dy columns w0/2 .. w0/2+2:
//   ow0 -> (s=1: dx0, s=2: dx1)   ow1 -> (s=0: dx1, s=1: dx2, s=2: dx3)   ow2 -> (s=0: dx3)
template <typename T>
__global__ void __launch_bounds__(256)
dw3x3_s2_dgrad_strip_k(const T* __restrict__ dy, const T* __restrict__ w, T* __restrict__ dx, DwS p) {
  constexpr int V = kDwV;
  const int cv = blockIdx.x * blockDim.x + threadIdx.x;
  if (cv >= p.cv_total) return;
  const int c0 = cv * V;
  float wr[9][V];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int j = 0; j < V; ++j) wr[t][j] = to_f(w[(int64_t)(c0 + j) * 9 + t]);
  const int64_t step = (int64_t)gridDim.y * blockDim.y;
  // here IH/IW are the dy extents (P, Q) and OH/OW the dx extents (H, W)
  for (int64_t sidx = (int64_t)blockIdx.y * blockDim.y + threadIdx.y; sidx < p.total_strips; sidx += step) {
    const int ws_ = (int)(sidx % p.strips_per_row);
    const int64_t t2 = sidx / p.strips_per_row;
    const int h = (int)(t2 % p.OH);
    const int n = (int)(t2 / p.OH);
    const int w0 = ws_ * 4;
    float acc[4][V];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
      for (int j = 0; j < V; ++j) acc[q][j] = 0.f;
    const T* gb = dy + (int64_t)n * p.IH * p.IW * p.C + c0;
    const int ow0 = w0 / 2;
    // filter row r is a compile-time constant in each call (registers, not local memory)
    auto do_row = [&](auto rc, int oh) {
      constexpr int r = decltype(rc)::value;
      if (oh < 0 || oh >= p.IH) return;
      const T* rowp = gb + (int64_t)oh * p.IW * p.C;
      Pack<T, V> g[3];
#pragma unroll
      for (int o = 0; o < 3; ++o) {
        if (ow0 + o < p.IW) g[o] = ld_pack<T, V>(rowp + (int64_t)(ow0 + o) * p.C);
        else {
#pragma unroll
          for (int j = 0; j < V; ++j) g[o].v[j] = from_f<T>(0.f);
        }
      }
#pragma unroll
      for (int j = 0; j < V; ++j) {
        const float g0 = to_f(g[0].v[j]), g1 = to_f(g[1].v[j]), g2 = to_f(g[2].v[j]);
        const float k0 = wr[r * 3 + 0][j], k1 = wr[r * 3 + 1][j], k2 = wr[r * 3 + 2][j];
        acc[0][j] = fmaf(g0, k1, acc[0][j]);
        acc[1][j] = fmaf(g0, k2, fmaf(g1, k0, acc[1][j]));
        acc[2][j] = fmaf(g1, k1, acc[2][j]);
        acc[3][j] = fmaf(g1, k2, fmaf(g2, k0, acc[3][j]));
      }
    };
    if (h & 1) {  // h odd: filter rows 0 and 2
      do_row(std::integral_constant<int, 0>{}, (h + 1) >> 1);
      do_row(std::integral_constant<int, 2>{}, (h - 1) >> 1);
    } else {      // h even: filter row 1
      do_row(std::integral_constant<int, 1>{}, h >> 1);
    }
    T* ob = dx + (((int64_t)n * p.OH + h) * p.OW + w0) * p.C + c0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      if (w0 + q < p.OW) {
        Pack<T, V> o;
#pragma unroll
        for (int j = 0; j < V; ++j) o.v[j] = from_f<T>(acc[q][j]);
        st_pack<T, V>(ob + (int64_t)q * p.C, o);
      }
    }
  }
}

// weight gradient, same strip walk: 4 dy vectors and 3 x (4*STRIDE + 3 - STRIDE) x vectors per strip
// feed 9 x V register accumulators; rows of the thread block are then summed through shared memory
// three taps at a time and written as partial[split][c][9] (fixed order -> deterministic).
template <typename T, int STRIDE>
__global__ void __launch_bounds__(256)
dw3x3_wgrad_strip_k(const T* __restrict__ x, const T* __restrict__ dy, float* __restrict__ part,
                    DwS p, int64_t strips_per_split) {
  constexpr int V = kDwV;
  constexpr int NCOL = kDwTQ * STRIDE + 3 - STRIDE;
  extern __shared__ float red[];  // [ry][cx * V * 3]
  const int cv = blockIdx.x * blockDim.x + threadIdx.x;
  const int c0 = cv * V;
  const bool live = cv < p.cv_total;
  float acc[9][V];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int j = 0; j < V; ++j) acc[t][j] = 0.f;
  // here IH/IW are the x extents and OH/OW the dy extents
  const int64_t s0 = (int64_t)blockIdx.y * strips_per_split;
  int64_t s1 = s0 + strips_per_split;
  if (s1 > p.total_strips) s1 = p.total_strips;
  if (live) {
    for (int64_t sidx = s0 + threadIdx.y; sidx < s1; sidx += blockDim.y) {
      const int qs = (int)(sidx % p.strips_per_row);
      const int64_t t2 = sidx / p.strips_per_row;
      const int oh = (int)(t2 % p.OH);
      const int n = (int)(t2 / p.OH);
      const int q0 = qs * kDwTQ;
      float gf[kDwTQ][V];
      const T* gb = dy + (((int64_t)n * p.OH + oh) * p.OW + q0) * p.C + c0;
#pragma unroll
      for (int q = 0; q < kDwTQ; ++q) {
        if (q0 + q < p.OW) {
          Pack<T, V> g = ld_pack<T, V>(gb + (int64_t)q * p.C);
#pragma unroll
          for (int j = 0; j < V; ++j) gf[q][j] = to_f(g.v[j]);
        } else {
#pragma unroll
          for (int j = 0; j < V; ++j) gf[q][j] = 0.f;
        }
      }
      const T* ib = x + (int64_t)n * p.IH * p.IW * p.C + c0;
      const int iw0 = q0 * STRIDE - 1;
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        const int ih = oh * STRIDE - 1 + r;
        if (ih < 0 || ih >= p.IH) continue;
        const T* rowp = ib + (int64_t)ih * p.IW * p.C;
        Pack<T, V> a[NCOL];
#pragma unroll
        for (int col = 0; col < NCOL; ++col) {
          const int iw = iw0 + col;
          if (iw >= 0 && iw < p.IW) a[col] = ld_pack<T, V>(rowp + (int64_t)iw * p.C);
          else {
#pragma unroll
            for (int j = 0; j < V; ++j) a[col].v[j] = from_f<T>(0.f);
          }
        }
#pragma unroll
        for (int q = 0; q < kDwTQ; ++q)
#pragma unroll
          for (int s = 0; s < 3; ++s)
#pragma unroll
            for (int j = 0; j < V; ++j)
              acc[r * 3 + s][j] = fmaf(gf[q][j], to_f(a[q * STRIDE + s].v[j]), acc[r * 3 + s][j]);
      }
    }
  }
  const int Wd = blockDim.x * V * 3;
  const int cbase = blockIdx.x * blockDim.x * V;
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    __syncthreads();
#pragma unroll
    for (int s = 0; s < 3; ++s)
#pragma unroll
      for (int j = 0; j < V; ++j) red[threadIdx.y * Wd + (threadIdx.x * V + j) * 3 + s] = acc[r * 3 + s][j];
    __syncthreads();
    for (int e = threadIdx.y * blockDim.x + threadIdx.x; e < Wd; e += blockDim.x * blockDim.y) {
      const int c = cbase + e / 3;
      if (c < p.C) {
        float sum = 0.f;
        for (int y = 0; y < (int)blockDim.y; ++y) sum += red[y * Wd + e];
        part[((int64_t)blockIdx.y * p.C + c) * 9 + r * 3 + (e % 3)] = sum;
      }
    }
  }
}

// thread-grid geometry of the strip kernels: blockDim = (cx channel vectors, 256 / cx strips)
struct DwStripGeom { dim3 grid, block; };
static DwStripGeom dw_strip_geom(int cv_total, int64_t strips, int waves, int threads = kDwThreads) {
  int cx = 1;
  while (cx * 2 <= cv_total && cx * 2 <= 32) cx *= 2;
  const int ry = threads / cx;
  const int cblocks = (int)ceil_div(cv_total, cx);
  int64_t rows = ceil_div((int64_t)sm_count() * waves, cblocks);
  const int64_t max_rows = ceil_div(strips, ry);
  if (rows > max_rows) rows = max_rows;
  if (rows < 1) rows = 1;
  if (rows > 65535) rows = 65535;
  DwStripGeom g;
  g.grid = dim3(cblocks, (unsigned)rows);
  g.block = dim3(cx, ry);
  return g;
}

// CNB_DW_WALK=0 keeps the per-strip kernel for stride-1 layers (A/B measurement)
static bool dw_walk_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("CNB_DW_WALK");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1;
}

static bool dw_strip_ok(const cnb_conv_desc& d, const void* a, const void* b) {
  return d.R == 3 && d.S == 3 && d.pad_h == 1 && d.pad_w == 1 && d.dil_h == 1 && d.dil_w == 1 &&
         d.stride_h == d.stride_w && (d.stride_h == 1 || d.stride_h == 2) && d.C % kDwV == 0 &&
         (uintptr_t)a % 16 == 0 && (uintptr_t)b % 16 == 0;
}

__global__ void sum_splits_k(const float* __restrict__ part, float* __restrict__ out, int splits,
                             int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    float s = 0.f;
    for (int k = 0; k < splits; ++k) s += part[(int64_t)k * n + i];
    out[i] = s;
  }
}

struct DwWgGeom {
  int V, cx, ry, cblocks, splits;
  int64_t rows_per_split;
};
static DwWgGeom dw_wg_geom(const cnb_conv_desc& d, int v) {
  DwWgGeom g;
  int P, Q;
  conv_out_dims(d, P, Q);
  int64_t M = (int64_t)d.N * P * Q;
  g.V = v;
  int cvs = d.C / v;
  int px = 1;
  while (px * 2 <= cvs && px * 2 <= 16) px *= 2;
  g.cx = px;
  g.ry = kDwThreads / g.cx;
  g.cblocks = (int)ceil_div(cvs, g.cx);
  int64_t splits = ceil_div((int64_t)sm_count() * 4, g.cblocks);
  int64_t max_splits = ceil_div(M, (int64_t)g.ry * 4);
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  if (splits > 1024) splits = 1024;
  g.rows_per_split = ceil_div(M, splits);
  g.splits = (int)ceil_div(M, g.rows_per_split);
  return g;
}

size_t dwconv_wgrad_ws(const cnb_conv_desc& d) {
  int64_t splits = (int64_t)sm_count() * 4;
  if (splits > 1024) splits = 1024;
  return (size_t)splits * d.C * d.R * d.S * sizeof(float) + 256;
}

// launch shape of the forward strip / walk kernels (also the number of BatchNorm partial rows)
struct DwFwdPlan { DwS q; DwStripGeom g; bool walk; int rbs; int tq; };
static DwFwdPlan dw_fwd_plan(const cnb_conv_desc& d, const float* bias) {
  DwP p = make_dw(d);
  DwFwdPlan f;
  f.q = DwS{p.N, p.H, p.W, p.P, p.Q, p.C, p.C / kDwV, (int)ceil_div(p.Q, kDwTQ), 0,
            d.flags & (((d.flags & CNB_CONV_BIAS) && bias ? CNB_CONV_BIAS : 0) | CNB_CONV_RELU)};
  f.q.total_strips = (int64_t)p.N * p.P * f.q.strips_per_row;
  f.walk = d.stride_h == 1 && d.dtype == CNB_BF16 && dw_walk_enabled();
  f.rbs = 0;
  f.tq = kDwTQ;
  if (f.walk) {
    f.rbs = (int)ceil_div(p.P, kDwRH);
    f.g = dw_strip_geom(f.q.cv_total, (int64_t)p.N * f.rbs * f.q.strips_per_row, 12, 128);
  } else if (d.stride_h == 1) {
    f.g = dw_strip_geom(f.q.cv_total, f.q.total_strips, 8);
  } else {
    // stride 2: strips of 2 outputs (3 x 5 input vectors) keep two CTAs per SM resident
    f.tq = 2;
    f.q.strips_per_row = (int)ceil_div(p.Q, 2);
    f.q.total_strips = (int64_t)p.N * p.P * f.q.strips_per_row;
    f.g = dw_strip_geom(f.q.cv_total, f.q.total_strips, 8);
  }
  return f;
}

// rows of the BatchNorm partial table the forward kernel would emit (0: this shape takes the
// generic kernel, which emits none); pointer alignment is assumed (checked at launch)
int dwconv_stats_rows(const cnb_conv_desc& d) {
  // Opt-in (CNB_DW_STATS=1): measured SLOWER on MobileNet-v1 (r02f: depthwise forward 1.44 -> 2.12 ms,
  // BatchNorm forward 1.76 -> 1.49 ms per step) — the eight extra accumulators push the column-walk
  // kernel from 52 to 308 bytes of spills.  Kept, tested, and off until the walk is re-registered.
  const char* e = getenv("CNB_DW_STATS");   // read per call: the tests switch it in-process
  if (!(e && e[0] == '1')) return 0;
  {
    // only shapes whose channel vectors fill every CTA column (no idle threads next to working ones)
    const int cvt = d.C / kDwV;
    int cx = 1;
    while (cx * 2 <= cvt && cx * 2 <= 32) cx *= 2;
    if (cvt % cx != 0) return 0;
  }
  if (!(d.R == 3 && d.S == 3 && d.pad_h == 1 && d.pad_w == 1 && d.dil_h == 1 && d.dil_w == 1 &&
        d.stride_h == d.stride_w && (d.stride_h == 1 || d.stride_h == 2) && d.C % kDwV == 0))
    return 0;
  return (int)dw_fwd_plan(d, nullptr).g.grid.y;
}

int dwconv_fwd(const cnb_conv_desc& d, const void* x, const void* w, const float* bias, void* y,
               float* stats, cudaStream_t st) {
  DwP p = make_dw(d);
  if (dw_strip_ok(d, x, y)) {
    DwFwdPlan f = dw_fwd_plan(d, bias);
    const DwS& q = f.q;
    const DwStripGeom& g = f.g;
    const size_t red = stats ? (size_t)g.block.x * g.block.y * 2 * kDwV * sizeof(float) : 0;
    if (f.walk) {
      if (stats)
        dw3x3_s1_walk_k<bf16, false, true><<<g.grid, g.block, red, st>>>((const bf16*)x, (const bf16*)w, bias,
                                                                         (bf16*)y, q, f.rbs, stats);
      else
        dw3x3_s1_walk_k<bf16, false><<<g.grid, g.block, 0, st>>>((const bf16*)x, (const bf16*)w, bias,
                                                                 (bf16*)y, q, f.rbs);
      return check_launch("dwconv_fwd(walk)");
    }
    if (d.stride_h == 1) {
      if (stats)
        CNB_DISPATCH_T(d.dtype, (dw3x3_strip_k<T, 1, false, kDwTQ, true><<<g.grid, g.block, red, st>>>(
                                    (const T*)x, (const T*)w, bias, (T*)y, q, stats)));
      else
        CNB_DISPATCH_T(d.dtype, (dw3x3_strip_k<T, 1, false><<<g.grid, g.block, 0, st>>>(
                                    (const T*)x, (const T*)w, bias, (T*)y, q)));
    } else {
      if (stats)
        CNB_DISPATCH_T(d.dtype, (dw3x3_strip_k<T, 2, false, 2, true><<<g.grid, g.block, red, st>>>(
                                    (const T*)x, (const T*)w, bias, (T*)y, q, stats)));
      else
        CNB_DISPATCH_T(d.dtype, (dw3x3_strip_k<T, 2, false, 2><<<g.grid, g.block, 0, st>>>(
                                    (const T*)x, (const T*)w, bias, (T*)y, q)));
    }
    return check_launch("dwconv_fwd(strip)");
  }
  CNB_REQUIRE(!stats, CNB_ERR_UNSUPPORTED,
              "dwconv_fwd: BatchNorm partials are emitted by the 3x3 strip kernels only (16-byte aligned tensors)");
  int v = pick_vec(p.C, dtype_size(d.dtype), x, y);
  int64_t total = (int64_t)p.N * p.P * p.Q * (p.C / v);
  int gr = grid_for(total, kDwThreads);
  const bool hb = (d.flags & CNB_CONV_BIAS) && bias, relu = d.flags & CNB_CONV_RELU;
#define FWD(B, R)                                                                                  \
  CNB_DISPATCH_TV(d.dtype, v, (dw_fwd_k<T, V, B, R><<<gr, kDwThreads, 0, st>>>(                    \
                                  (const T*)x, (const T*)w, bias, (T*)y, p, total)))
  if (hb && relu) FWD(true, true);
  else if (hb) FWD(true, false);
  else if (relu) FWD(false, true);
  else FWD(false, false);
#undef FWD
  return check_launch("dwconv_fwd");
}

int dwconv_dgrad(const cnb_conv_desc& d, const void* dy, const void* w, void* dx, cudaStream_t st) {
  DwP p = make_dw(d);
  if (dw_strip_ok(d, dy, dx)) {
    const int v = kDwV;
    // input of this launch = dy (P x Q), output = dx (H x W)
    DwS q{p.N, p.P, p.Q, p.H, p.W, p.C, p.C / v, (int)ceil_div(p.W, kDwTQ), 0, 0};
    q.total_strips = (int64_t)p.N * p.H * q.strips_per_row;
    DwStripGeom g = dw_strip_geom(q.cv_total, q.total_strips, 8);
    if (d.stride_h == 1 && d.dtype == CNB_BF16 && dw_walk_enabled()) {
      const int rbs = (int)ceil_div(p.H, kDwRH);
      DwStripGeom gw = dw_strip_geom(q.cv_total, (int64_t)p.N * rbs * q.strips_per_row, 12, 128);
      dw3x3_s1_walk_k<bf16, true><<<gw.grid, gw.block, 0, st>>>((const bf16*)dy, (const bf16*)w, nullptr,
                                                                (bf16*)dx, q, rbs);
      return check_launch("dwconv_dgrad(walk)");
    }
    if (d.stride_h == 1)  // H == P, W == Q: correlation with the flipped filter, pad 2 - 1 = 1
      CNB_DISPATCH_T(d.dtype, (dw3x3_strip_k<T, 1, true><<<g.grid, g.block, 0, st>>>(
                                  (const T*)dy, (const T*)w, nullptr, (T*)dx, q)));
    else
      CNB_DISPATCH_T(d.dtype, (dw3x3_s2_dgrad_strip_k<T><<<g.grid, g.block, 0, st>>>(
                                  (const T*)dy, (const T*)w, (T*)dx, q)));
    return check_launch("dwconv_dgrad(strip)");
  }
  int v = pick_vec(p.C, dtype_size(d.dtype), dy, dx);
  int64_t total = (int64_t)p.N * p.H * p.W * (p.C / v);
  CNB_DISPATCH_TV(d.dtype, v, (dw_dgrad_k<T, V><<<grid_for(total, kDwThreads), kDwThreads, 0, st>>>(
                                  (const T*)dy, (const T*)w, (T*)dx, p, total)));
  return check_launch("dwconv_dgrad");
}

int dwconv_wgrad(const cnb_conv_desc& d, const void* x, const void* dy, float* dw, void* ws,
                 size_t ws_bytes, cudaStream_t st) {
  DwP p = make_dw(d);
  const int RS = p.R * p.S;
  CNB_REQUIRE(RS <= 25, CNB_ERR_UNSUPPORTED, "dwconv_wgrad: filter %dx%d larger than 5x5", p.R, p.S);
  CNB_REQUIRE(ws && ws_bytes >= dwconv_wgrad_ws(d), CNB_ERR_WORKSPACE,
              "dwconv_wgrad: workspace %zu < %zu", ws_bytes, dwconv_wgrad_ws(d));
  if (dw_strip_ok(d, x, dy)) {
    const int v = kDwV;
    DwS q{p.N, p.H, p.W, p.P, p.Q, p.C, p.C / v, (int)ceil_div(p.Q, kDwTQ), 0, 0};
    q.total_strips = (int64_t)p.N * p.P * q.strips_per_row;
    DwStripGeom g = dw_strip_geom(q.cv_total, q.total_strips, 2);
    int64_t max_splits = (int64_t)sm_count() * 4;
    if (max_splits > 1024) max_splits = 1024;
    int64_t splits = g.grid.y;
    if (splits > max_splits) splits = max_splits;
    const int64_t per = ceil_div(q.total_strips, splits);
    splits = ceil_div(q.total_strips, per);
    g.grid.y = (unsigned)splits;
    const size_t smem = (size_t)g.block.y * g.block.x * v * 3 * sizeof(float);
    float* part = (float*)ws;
    if (d.stride_h == 1)
      CNB_DISPATCH_T(d.dtype, (dw3x3_wgrad_strip_k<T, 1><<<g.grid, g.block, smem, st>>>(
                                  (const T*)x, (const T*)dy, part, q, per)));
    else
      CNB_DISPATCH_T(d.dtype, (dw3x3_wgrad_strip_k<T, 2><<<g.grid, g.block, smem, st>>>(
                                  (const T*)x, (const T*)dy, part, q, per)));
    CNB_TRY(check_launch("dwconv_wgrad(strip)"));
    int64_t n = (int64_t)p.C * 9;
    sum_splits_k<<<grid_for(n, 256), 256, 0, st>>>(part, dw, (int)splits, n);
    return check_launch("dwconv_wgrad_sum");
  }
  int v = pick_vec(p.C, dtype_size(d.dtype), x, dy);
  if (v > 4) v = 4;
  if (RS > 9 && v > 2) v = 2;
  DwWgGeom g = dw_wg_geom(d, v);
  dim3 grid(g.cblocks, g.splits), block(g.cx, g.ry);
  size_t smem = (size_t)g.ry * g.cx * v * RS * sizeof(float);
  CNB_REQUIRE(smem <= 48 * 1024, CNB_ERR_UNSUPPORTED, "dwconv_wgrad: smem %zu too large", smem);
  float* part = (float*)ws;
  if (RS <= 9)
    CNB_DISPATCH_TV(d.dtype, v, (dw_wgrad_k<T, V, 9><<<grid, block, smem, st>>>(
                                    (const T*)x, (const T*)dy, part, p, g.rows_per_split)));
  else
    CNB_DISPATCH_TV(d.dtype, v, (dw_wgrad_k<T, V, 25><<<grid, block, smem, st>>>(
                                    (const T*)x, (const T*)dy, part, p, g.rows_per_split)));
  CNB_TRY(check_launch("dwconv_wgrad"));
  int64_t n = (int64_t)p.C * RS;
  sum_splits_k<<<grid_for(n, 256), 256, 0, st>>>(part, dw, g.splits, n);
  return check_launch("dwconv_wgrad_sum");
}

}  // namespace cnb
