// dwconv.cu — depthwise convolution (groups == C == K) forward / data-grad / weight-grad on CUDA
// cores.  Replaces aten::convolution(groups=C) reached from reference src/basemodel.py:981-997
// (conv2d_depthwise), src/mobilenet_v1.py:102 and src/shufflenet_v1.py:126-133.
// Arithmetic intensity is ~R*S flops per 2 bytes -> HBM-bound, so the kernels are plain
// 128-bit-vectorised NHWC gathers: one thread owns V contiguous channels of one pixel; the
// R*S*V filter taps of a thread are read through L1 (the whole filter is <= 9*1024 elements).
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

int dwconv_fwd(const cnb_conv_desc& d, const void* x, const void* w, const float* bias, void* y,
               cudaStream_t st) {
  DwP p = make_dw(d);
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
