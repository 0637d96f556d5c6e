// eltwise.cu — bandwidth-bound element-wise, data-movement, pooling-to-vector and gating kernels.
// All are 128-bit vectorised over the innermost (channel) dimension of NHWC tensors and sized as
// grid-stride loops over a multiple of the SM count.  Roofline: HBM.
#include "common.cuh"

namespace cnb {

thread_local char g_last_error[512] = "";
std::atomic<uint64_t> g_launches{0};

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

constexpr int kThreads = 256;

// ------------------------------------------------------------------ relu / add
template <typename T, int V>
__global__ void relu_fwd_k(const T* __restrict__ x, T* __restrict__ y, int64_t nvec) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * blockDim.x) {
    Pack<T, V> a = ld_pack<T, V>(x + i * V), o;
#pragma unroll
    for (int j = 0; j < V; ++j) {
      float f = to_f(a.v[j]);
      // torch relu: NaN propagates (clamp_min semantics)
      o.v[j] = (f > 0.f || f != f) ? a.v[j] : from_f<T>(0.f);
    }
    st_pack<T, V>(y + i * V, o);
  }
}

// dx = dy * (y > 0)
template <typename T, int V>
__global__ void relu_bwd_k(const T* __restrict__ y, const T* __restrict__ dy, T* __restrict__ dx,
                           int64_t nvec) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * blockDim.x) {
    Pack<T, V> a = ld_pack<T, V>(y + i * V), g = ld_pack<T, V>(dy + i * V), o;
#pragma unroll
    for (int j = 0; j < V; ++j) o.v[j] = to_f(a.v[j]) > 0.f ? g.v[j] : from_f<T>(0.f);
    st_pack<T, V>(dx + i * V, o);
  }
}

template <typename T, int V, bool RELU>
__global__ void add_k(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ y,
                      int64_t nvec) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * blockDim.x) {
    Pack<T, V> p = ld_pack<T, V>(a + i * V), q = ld_pack<T, V>(b + i * V), o;
#pragma unroll
    for (int j = 0; j < V; ++j) {
      // the reference adds in the storage dtype (output += identity), then F.relu
      T s = from_f<T>(to_f(p.v[j]) + to_f(q.v[j]));
      float f = to_f(s);
      o.v[j] = (!RELU || f > 0.f || f != f) ? s : from_f<T>(0.f);
    }
    st_pack<T, V>(y + i * V, o);
  }
}

// ------------------------------------------------------------------ cast
template <typename TS, typename TD>
__global__ void cast_k(const TS* __restrict__ s, TD* __restrict__ d, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    d[i] = from_f<TD>(to_f(s[i]));
}
// fp32 -> bf16, 4 elements per thread (16B load, 8B store)
__global__ void cast_f32_bf16_v4(const float* __restrict__ s, bf16* __restrict__ d, int64_t nvec) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * blockDim.x) {
    Pack<float, 4> a = ld_pack<float, 4>(s + i * 4);
    Pack<bf16, 4> o;
#pragma unroll
    for (int j = 0; j < 4; ++j) o.v[j] = __float2bfloat16_rn(a.v[j]);
    st_pack<bf16, 4>(d + i * 4, o);
  }
}

// ------------------------------------------------------------------ uint8 image batch -> fp32 NCHW
// The reference normalises on the host (transforms.ToTensor + Normalize inside the DataLoader
// workers, src/mngrdata.py:139-165) and ships fp32 over PCIe; here the uint8 pixels cross the bus
// (a quarter of the bytes) and this kernel restates the two transforms bit for bit:
// ToTensor = float(u8) / 255, Normalize = (t - mean[c]) / std[c], both IEEE fp32, plus the index
// arithmetic of RandomCrop(size, padding) (zero fill) and RandomHorizontalFlip with the offsets /
// flips drawn by the caller.  A thread owns 4 adjacent output pixels of every channel.
struct U8NormP {
  int N, C, H, W, Ho, Wo, pad, nhwc;
  float mean[4], std[4];
};
__global__ void u8_normalize_k(const uint8_t* __restrict__ src, float* __restrict__ dst,
                               const int32_t* __restrict__ aug, U8NormP p, int64_t total) {
  const int wq = (p.Wo + 3) >> 2;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int jq = (int)(i % wq);
    const int64_t t = i / wq;
    const int oi = (int)(t % p.Ho);
    const int n = (int)(t / p.Ho);
    int top = 0, left = 0, flip = 0;
    if (aug) {
      top = aug[n * 3 + 0];
      left = aug[n * 3 + 1];
      flip = aug[n * 3 + 2];
    }
    const int h = oi + top - p.pad;
    const bool row_in = h >= 0 && h < p.H;
    const uint8_t* img = src + (int64_t)n * p.C * p.H * p.W;
    for (int c = 0; c < p.C; ++c) {
      float v[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int oj = jq * 4 + k;
        const int cj = flip ? p.Wo - 1 - oj : oj;   // column of the crop this output shows
        const int w = cj + left - p.pad;
        uint8_t u = 0;
        if (row_in && oj < p.Wo && w >= 0 && w < p.W)
          u = p.nhwc ? img[((int64_t)h * p.W + w) * p.C + c] : img[((int64_t)c * p.H + h) * p.W + w];
        v[k] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)u, 255.f), p.mean[c]), p.std[c]);
      }
      float* o = dst + (((int64_t)n * p.C + c) * p.Ho + oi) * p.Wo + jq * 4;
      if ((p.Wo & 3) == 0) {
        *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (jq * 4 + k < p.Wo) o[k] = v[k];
      }
    }
  }
}

// ------------------------------------------------------------------ NCHW fp32 -> NHWC T
// One block transposes a [C x 32] (channels x pixels) slab through shared memory so both the
// NCHW reads (pixel-contiguous) and the NHWC writes (channel-contiguous) are coalesced.
template <typename T>
__global__ void nchw_to_nhwc_k(const float* __restrict__ src, T* __restrict__ dst, int C, int Cd,
                               int HW) {
  extern __shared__ float tile[];  // [C][33]
  const int n = blockIdx.y;
  const int p0 = blockIdx.x * 32;
  const float* s = src + (int64_t)n * C * HW;
  T* o = dst + (int64_t)n * HW * Cd;
  for (int i = threadIdx.x; i < C * 32; i += blockDim.x) {
    int c = i / 32, p = i % 32;
    tile[c * 33 + p] = (p0 + p < HW) ? s[(int64_t)c * HW + p0 + p] : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < Cd * 32; i += blockDim.x) {
    int p = i / Cd, c = i % Cd;
    if (p0 + p < HW) o[(int64_t)(p0 + p) * Cd + c] = from_f<T>(c < C ? tile[c * 33 + p] : 0.f);
  }
}

// Explicit im2col of the fp32 NCHW network input: dst[m][(r*S + s)*C + c] = x[n][c][h][w] with
// m = (n*P + i)*Q + j, h = i*sh - ph + r*dh, w = j*sw - pw + s*dw; zero for padding taps and for
// the alignment columns >= R*S*C.  One thread writes one 16-byte column vector.
struct Im2colP {
  int N, C, H, W, R, S, P, Q, sh, sw, ph, pw, dh, dw, Kd;
};
template <typename T, int V>
__global__ void im2col_nchw_k(const float* __restrict__ src, T* __restrict__ dst, Im2colP p,
                              int64_t total_vec) {
  const int kv = p.Kd / V;
  const int RSC = p.R * p.S * p.C;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total_vec;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int cv = (int)(i % kv);
    const int64_t m = i / kv;
    const int j = (int)(m % p.Q);
    const int64_t t = m / p.Q;
    const int ii = (int)(t % p.P);
    const int n = (int)(t / p.P);
    Pack<T, V> o;
#pragma unroll
    for (int e = 0; e < V; ++e) {
      const int col = cv * V + e;
      float v = 0.f;
      if (col < RSC) {
        const int c = col % p.C;
        const int rs = col / p.C;
        const int s_ = rs % p.S, r = rs / p.S;
        const int h = ii * p.sh - p.ph + r * p.dh;
        const int w = j * p.sw - p.pw + s_ * p.dw;
        if (h >= 0 && h < p.H && w >= 0 && w < p.W)
          v = __ldg(src + (((int64_t)n * p.C + c) * p.H + h) * p.W + w);
      }
      o.v[e] = from_f<T>(v);
    }
    st_pack<T, V>(dst + i * V, o);
  }
}

// Row-staged variant: one CTA per (image, output row).  The R (dilated) input rows that row needs
// are loaded once, coalesced, into a zero-padded smem patch [C][R][W + 2*pad_w] already converted
// to T, together with a column table (offset of every im2col column inside the patch, -1 for the
// zero columns); each thread then assembles 16-byte output vectors from smem and the CTA writes
// one contiguous Q x Kd row block.  No integer division or bounds check in the inner loop.
template <typename T, int V>
__global__ void __launch_bounds__(256)
im2col_nchw_rows_k(const float* __restrict__ src, T* __restrict__ dst, Im2colP p) {
  extern __shared__ __align__(16) uint8_t sm_raw[];
  const int Wp = p.W + 2 * p.pw;
  int* coltab = reinterpret_cast<int*>(sm_raw);                       // [Kd]
  T* patch = reinterpret_cast<T*>(sm_raw + (size_t)((p.Kd * 4 + 15) / 16) * 16);  // [C][R][Wp]
  const int ii = blockIdx.x % p.P;
  const int n = blockIdx.x / p.P;
  const int RSC = p.R * p.S * p.C;
  for (int col = threadIdx.x; col < p.Kd; col += blockDim.x) {
    int off = -1;
    if (col < RSC) {
      const int c = col % p.C, rs = col / p.C;
      const int s_ = rs % p.S, r = rs / p.S;
      off = (c * p.R + r) * Wp + s_ * p.dw;
    }
    coltab[col] = off;
  }
  const int h0 = ii * p.sh - p.ph;
  const int rows = p.C * p.R;
  for (int i = threadIdx.x; i < rows * Wp; i += blockDim.x) {
    const int wp = i % Wp, cr = i / Wp;
    const int r = cr % p.R, c = cr / p.R;
    const int h = h0 + r * p.dh, w = wp - p.pw;
    float v = 0.f;
    if (h >= 0 && h < p.H && w >= 0 && w < p.W)
      v = __ldg(src + (((int64_t)n * p.C + c) * p.H + h) * p.W + w);
    patch[i] = from_f<T>(v);
  }
  __syncthreads();
  const int kv = p.Kd / V;
  T* out = dst + ((int64_t)n * p.P + ii) * p.Q * p.Kd;
  const T zero = from_f<T>(0.f);
  for (int i = threadIdx.x; i < p.Q * kv; i += blockDim.x) {
    const int cv = i % kv, j = i / kv;
    const int base = j * p.sw;
    Pack<T, V> o;
#pragma unroll
    for (int e = 0; e < V; ++e) {
      const int off = coltab[cv * V + e];
      o.v[e] = off >= 0 ? patch[off + base] : zero;
    }
    st_pack<T, V>(out + (int64_t)i * V, o);
  }
}

// Output-staged variant (the one the RGB stems use).  The patch is kept pixel-major, [R][Wp][C], so
// the columns of one filter row are one contiguous run starting at j*stride*C.  Phase A: a warp
// takes one 16-byte column block `cv` (its V patch offsets live in registers) and its lanes take
// consecutive output pixels j — patch reads are sw*C elements apart across lanes (odd word stride
// for RGB, conflict-free) and the assembled vectors go to a smem copy of the output row block.
// Phase B streams that block to global memory with fully coalesced 16-byte stores.
template <typename T, int V>
__global__ void __launch_bounds__(256)
im2col_nchw_staged_k(const float* __restrict__ src, T* __restrict__ dst, Im2colP p, int row_stride) {
  extern __shared__ __align__(16) uint8_t sm_raw[];
  const int Wp = p.W + 2 * p.pw;
  const int patch_elems = p.R * Wp * p.C;
  int* coltab = reinterpret_cast<int*>(sm_raw);                                          // [Kd]
  T* patch = reinterpret_cast<T*>(sm_raw + (size_t)((p.Kd * 4 + 15) / 16) * 16);        // [R][Wp][C]
  T* outs = patch + (size_t)((patch_elems + V - 1) / V) * V;                             // [Q][row_stride]
  const int ii = blockIdx.x % p.P;
  const int n = blockIdx.x / p.P;
  const int RSC = p.R * p.S * p.C;
  for (int col = threadIdx.x; col < p.Kd; col += blockDim.x) {
    int off = -1;
    if (col < RSC) {
      const int c = col % p.C, rs = col / p.C;
      const int s_ = rs % p.S, r = rs / p.S;
      off = (r * Wp + s_ * p.dw) * p.C + c;
    }
    coltab[col] = off;
  }
  const int h0 = ii * p.sh - p.ph;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  // one warp per (channel, filter row): global reads run along w (coalesced), one division per
  // row instead of three per element; the smem store interleaves the channels
  for (int cr = warp; cr < p.C * p.R; cr += nwarps) {
    const int c = cr / p.R, r = cr - c * p.R;
    const int h = h0 + r * p.dh;
    const bool row_ok = h >= 0 && h < p.H;
    const float* srow = src + (((int64_t)n * p.C + c) * p.H + (row_ok ? h : 0)) * p.W;
    T* prow = patch + (size_t)r * Wp * p.C + c;
    for (int wp = lane; wp < Wp; wp += 32) {
      const int w = wp - p.pw;
      float v = 0.f;
      if (row_ok && w >= 0 && w < p.W) v = __ldg(srow + w);
      prow[wp * p.C] = from_f<T>(v);
    }
  }
  __syncthreads();
  const int kv = p.Kd / V;
  const T zero = from_f<T>(0.f);
  const int jstep = p.sw * p.C;
  for (int cv = warp; cv < kv; cv += nwarps) {
    int off[V];
#pragma unroll
    for (int e = 0; e < V; ++e) off[e] = coltab[cv * V + e];
    for (int j = lane; j < p.Q; j += 32) {
      const int base = j * jstep;
      Pack<T, V> o;
#pragma unroll
      for (int e = 0; e < V; ++e) o.v[e] = off[e] >= 0 ? patch[off[e] + base] : zero;
      st_pack<T, V>(outs + (size_t)j * row_stride + cv * V, o);
    }
  }
  __syncthreads();
  T* out = dst + ((int64_t)n * p.P + ii) * p.Q * p.Kd;
  int j = threadIdx.x / kv, cv = threadIdx.x - j * kv;
  const int dj = blockDim.x / kv, dcv = blockDim.x - dj * kv;
  for (int i = threadIdx.x; i < p.Q * kv; i += blockDim.x) {
    st_pack<T, V>(out + (int64_t)i * V, ld_pack<T, V>(outs + (size_t)j * row_stride + cv * V));
    j += dj;
    cv += dcv;
    if (cv >= kv) { cv -= kv; ++j; }
  }
}

// Row im2col for the RGB stems (cnb_rowim2col_nchw): only the S horizontal taps are unrolled into
// channels, dst[n][h][q][s*C + c] = src[n][c][h][q*sw - pw + s*dw] (zero outside / in the alignment
// columns).  The R x S stem then runs on the im2col-mode TMA path as an R x 1 convolution over
// S*C (21 -> 24) channels with vertical stride sh: a [N, H, Q, 24] tensor (316 MB at B = 256)
// replaces the [N*P*Q, 152] matrix (976 MB).  One CTA per input row (n, h).
template <typename T, int V>
__global__ void __launch_bounds__(256)
rowim2col_nchw_k(const float* __restrict__ src, T* __restrict__ dst, Im2colP p) {
  extern __shared__ __align__(16) uint8_t sm_raw[];
  const int Wp = p.W + 2 * p.pw;
  int* coltab = reinterpret_cast<int*>(sm_raw);                                   // [Kd]
  T* patch = reinterpret_cast<T*>(sm_raw + (size_t)((p.Kd * 4 + 15) / 16) * 16);  // [Wp][C]
  const int h = blockIdx.x % p.H;
  const int n = blockIdx.x / p.H;
  const int SC = p.S * p.C;
  for (int col = threadIdx.x; col < p.Kd; col += blockDim.x)
    coltab[col] = col < SC ? (col / p.C) * p.dw * p.C + (col % p.C) : -1;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  for (int c = warp; c < p.C; c += nwarps) {
    const float* srow = src + (((int64_t)n * p.C + c) * p.H + h) * p.W;
    for (int wp = lane; wp < Wp; wp += 32) {
      const int w = wp - p.pw;
      patch[wp * p.C + c] = from_f<T>((w >= 0 && w < p.W) ? __ldg(srow + w) : 0.f);
    }
  }
  __syncthreads();
  const int kv = p.Kd / V;
  T* out = dst + ((int64_t)n * p.H + h) * p.Q * p.Kd;
  const T zero = from_f<T>(0.f);
  int q = threadIdx.x / kv, cv = threadIdx.x - q * kv;
  const int dq = blockDim.x / kv, dcv = blockDim.x - dq * kv;
  for (int i = threadIdx.x; i < p.Q * kv; i += blockDim.x) {
    const int base = q * p.sw * p.C;
    Pack<T, V> o;
#pragma unroll
    for (int e = 0; e < V; ++e) {
      const int off = coltab[cv * V + e];
      o.v[e] = off >= 0 ? patch[off + base] : zero;
    }
    st_pack<T, V>(out + (int64_t)i * V, o);
    q += dq;
    cv += dcv;
    if (cv >= kv) { cv -= kv; ++q; }
  }
}

// ------------------------------------------------------------------ channel shuffle / slab copy
// y[m, j*g + i] = x[m, i*(C/g) + j]; one thread per output element pair is fine: rows are short
// and both sides stay within one row (L1/L2 absorb the permutation), traffic = 2 * numel.
template <typename T>
__global__ void shuffle_k(const T* __restrict__ x, T* __restrict__ y, int64_t M, int C, int g) {
  const int cpg = C / g;
  const int64_t total = M * C;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    int64_t m = i / C;
    int co = (int)(i - m * C);
    int j = co / g, gi = co - j * g;
    y[i] = x[m * C + (int64_t)gi * cpg + j];
  }
}

// Vector form for the two shapes the zoo produces: a small factor G (the group count in the forward
// shuffle, the channels per group in its inverse) and a large factor L, L % V == 0.  A thread moves
// a [G x V] block: G vector loads, a register permutation with compile-time indices, G vector
// stores — no per-element division, every access V*sizeof(T) wide.
//   INV == false: groups = G, C/groups = L:  y[j*G + i] = x[i*L + j]
//   INV == true : groups = L, C/groups = G:  y[j*L + i] = x[i*G + j]
template <typename T, int V, int G, bool INV>
__global__ void shuffle_vec_k(const T* __restrict__ x, T* __restrict__ y, int64_t M, int L) {
  const int nb = L / V;
  const int C = G * L;
  const int64_t total = M * nb;
  for (int64_t it = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; it < total;
       it += (int64_t)gridDim.x * blockDim.x) {
    const int64_t m = it / nb;
    const int b = (int)(it - m * nb);
    const T* xr = x + m * C;
    T* yr = y + m * C;
    Pack<T, V> in[G], out[G];
    if (!INV) {
#pragma unroll
      for (int i = 0; i < G; ++i) in[i] = ld_pack<T, V>(xr + (int64_t)i * L + b * V);
#pragma unroll
      for (int jj = 0; jj < V; ++jj)
#pragma unroll
        for (int i = 0; i < G; ++i) out[(jj * G + i) / V].v[(jj * G + i) % V] = in[i].v[jj];
#pragma unroll
      for (int k = 0; k < G; ++k) st_pack<T, V>(yr + (int64_t)b * V * G + k * V, out[k]);
    } else {
#pragma unroll
      for (int k = 0; k < G; ++k) in[k] = ld_pack<T, V>(xr + (int64_t)b * V * G + k * V);
#pragma unroll
      for (int ii = 0; ii < V; ++ii)
#pragma unroll
        for (int j = 0; j < G; ++j) out[j].v[ii] = in[(ii * G + j) / V].v[(ii * G + j) % V];
#pragma unroll
      for (int j = 0; j < G; ++j) st_pack<T, V>(yr + (int64_t)j * L + b * V, out[j]);
    }
  }
}

template <typename T, int V, bool INV>
static bool launch_shuffle_vec(const void* x, void* y, int64_t M, int G, int L, cudaStream_t st) {
  const int grid = grid_for(M * (L / V), kThreads);
#define SHV(GG)                                                                                   \
  case GG:                                                                                        \
    shuffle_vec_k<T, V, GG, INV><<<grid, kThreads, 0, st>>>((const T*)x, (T*)y, M, L);            \
    return true
  switch (G) {
    SHV(1);
    SHV(2);
    SHV(3);
    SHV(4);
    SHV(8);
    default: return false;
  }
#undef SHV
}

// true when a vector kernel took the shuffle
static bool shuffle_vec(const void* x, void* y, int64_t M, int C, int groups, int dtype,
                        cudaStream_t st) {
  const int cpg = C / groups;
  const bool fwd_shape = groups <= 8;            // small group count, wide groups
  const bool inv_shape = !fwd_shape && cpg <= 8; // the inverse permutation (backward pass)
  if (!fwd_shape && !inv_shape) return false;
  const int G = fwd_shape ? groups : cpg, L = fwd_shape ? cpg : groups;
  const int v = pick_vec(L, dtype_size(dtype), x, y);
  if (v < 2) return false;
  bool done = false;
  CNB_DISPATCH_TV(dtype, v, {
    if constexpr (V >= 2)
      done = fwd_shape ? launch_shuffle_vec<T, V, false>(x, y, M, G, L, st)
                       : launch_shuffle_vec<T, V, true>(x, y, M, G, L, st);
  });
  return done;
}

template <typename T, int V>
__global__ void copy_channels_k(const T* __restrict__ src, int srcC, int srcOff,
                                T* __restrict__ dst, int dstC, int dstOff, int64_t M, int Cv) {
  const int64_t total = M * Cv;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    int64_t m = i / Cv;
    int c = (int)(i - m * Cv) * V;
    st_pack<T, V>(dst + m * dstC + dstOff + c, ld_pack<T, V>(src + m * srcC + srcOff + c));
  }
}

// ------------------------------------------------------------------ global average pool
// grid (N, ceil(C / (V*CV))), block (CV, RY): thread (cx, ry) sums rows ry, ry+RY, ... of its V
// channels; cross-ry reduction in shared memory.  One pass over x.
template <typename T, int V>
__global__ void gap_fwd_k(const T* __restrict__ x, const T* __restrict__ x2, float* __restrict__ y,
                          int HW, int C) {
  extern __shared__ float red[];  // [RY][CV*V]
  const int n = blockIdx.x;
  const int cv = blockIdx.y * blockDim.x + threadIdx.x;
  const int c0 = cv * V;
  float acc[V];
#pragma unroll
  for (int j = 0; j < V; ++j) acc[j] = 0.f;
  if (c0 < C) {
    const T* base = x + (int64_t)n * HW * C + c0;
    const T* base2 = x2 ? x2 + (int64_t)n * HW * C + c0 : nullptr;
    for (int r = threadIdx.y; r < HW; r += blockDim.y) {
      Pack<T, V> a = ld_pack<T, V>(base + (int64_t)r * C);
#pragma unroll
      for (int j = 0; j < V; ++j) acc[j] += to_f(a.v[j]);
      if (base2) {
        Pack<T, V> b = ld_pack<T, V>(base2 + (int64_t)r * C);
#pragma unroll
        for (int j = 0; j < V; ++j) acc[j] += to_f(b.v[j]);
      }
    }
  }
  const int W = blockDim.x * V;
#pragma unroll
  for (int j = 0; j < V; ++j) red[threadIdx.y * W + threadIdx.x * V + j] = acc[j];
  __syncthreads();
  if (threadIdx.y == 0 && c0 < C) {
    const float inv = 1.f / (float)HW;
#pragma unroll
    for (int j = 0; j < V; ++j) {
      float s = 0.f;
      for (int r = 0; r < (int)blockDim.y; ++r) s += red[r * W + threadIdx.x * V + j];
      y[(int64_t)n * C + c0 + j] = s * inv;
    }
  }
}

template <typename T, int V>
__global__ void gap_bwd_k(const float* __restrict__ dy, T* __restrict__ dx, int64_t total_vec,
                          int HW, int Cv) {
  const float inv = 1.f / (float)HW;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total_vec;
       i += (int64_t)gridDim.x * blockDim.x) {
    int64_t row = i / Cv;
    int cv = (int)(i - row * Cv);
    int64_t n = row / HW;
    Pack<T, V> o;
#pragma unroll
    for (int j = 0; j < V; ++j) o.v[j] = from_f<T>(dy[n * (int64_t)Cv * V + cv * V + j] * inv);
    st_pack<T, V>(dx + i * V, o);
  }
}

// ------------------------------------------------------------------ SE scale
template <typename T, int V>
__global__ void se_scale_fwd_k(const T* __restrict__ x, const float* __restrict__ s,
                               T* __restrict__ y, int64_t total_vec, int HW, int Cv) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total_vec;
       i += (int64_t)gridDim.x * blockDim.x) {
    int64_t row = i / Cv;
    int cv = (int)(i - row * Cv);
    int64_t n = row / HW;
    const float* sp = s + n * (int64_t)Cv * V + cv * V;
    Pack<T, V> a = ld_pack<T, V>(x + i * V), o;
#pragma unroll
    for (int j = 0; j < V; ++j) o.v[j] = from_f<T>(to_f(a.v[j]) * to_f(from_f<T>(sp[j])));
    st_pack<T, V>(y + i * V, o);
  }
}

// dx = dy * s ; ds[n,c] = sum_hw dy*x.  Same thread layout as gap_fwd_k (one pass over x, dy).
template <typename T, int V>
__global__ void se_scale_bwd_k(const T* __restrict__ x, const float* __restrict__ s,
                               const T* __restrict__ dy, T* __restrict__ dx,
                               float* __restrict__ ds, int HW, int C) {
  extern __shared__ float red[];
  const int n = blockIdx.x;
  const int cv = blockIdx.y * blockDim.x + threadIdx.x;
  const int c0 = cv * V;
  float acc[V], sv[V];
#pragma unroll
  for (int j = 0; j < V; ++j) acc[j] = 0.f;
  if (c0 < C) {
#pragma unroll
    for (int j = 0; j < V; ++j) sv[j] = to_f(from_f<T>(s[(int64_t)n * C + c0 + j]));
    const int64_t base = (int64_t)n * HW * C + c0;
    for (int r = threadIdx.y; r < HW; r += blockDim.y) {
      Pack<T, V> a = ld_pack<T, V>(x + base + (int64_t)r * C);
      Pack<T, V> g = ld_pack<T, V>(dy + base + (int64_t)r * C), o;
#pragma unroll
      for (int j = 0; j < V; ++j) {
        float gf = to_f(g.v[j]);
        acc[j] += gf * to_f(a.v[j]);
        o.v[j] = from_f<T>(gf * sv[j]);
      }
      st_pack<T, V>(dx + base + (int64_t)r * C, o);
    }
  }
  const int W = blockDim.x * V;
#pragma unroll
  for (int j = 0; j < V; ++j) red[threadIdx.y * W + threadIdx.x * V + j] = acc[j];
  __syncthreads();
  if (threadIdx.y == 0 && c0 < C) {
#pragma unroll
    for (int j = 0; j < V; ++j) {
      float t = 0.f;
      for (int r = 0; r < (int)blockDim.y; ++r) t += red[r * W + threadIdx.x * V + j];
      ds[(int64_t)n * C + c0 + j] = t;
    }
  }
}

// ------------------------------------------------------------------ SK select
// attn = softmax over 2 paths; written once per (n,c) by a tiny kernel, then applied.
__global__ void sk_softmax2_k(const float* __restrict__ logits, float* __restrict__ attn, int N,
                              int C) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N * C) return;
  int n = i / C, c = i % C;
  float l0 = logits[((int64_t)n * 2 + 0) * C + c], l1 = logits[((int64_t)n * 2 + 1) * C + c];
  float m = fmaxf(l0, l1);
  float e0 = expf(l0 - m), e1 = expf(l1 - m);
  float inv = 1.f / (e0 + e1);
  attn[((int64_t)n * 2 + 0) * C + c] = e0 * inv;
  attn[((int64_t)n * 2 + 1) * C + c] = e1 * inv;
}

template <typename T, int V>
__global__ void sk_select_fwd_k(const T* __restrict__ u0, const T* __restrict__ u1,
                                const float* __restrict__ attn, T* __restrict__ y,
                                int64_t total_vec, int HW, int Cv) {
  const int C = Cv * V;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total_vec;
       i += (int64_t)gridDim.x * blockDim.x) {
    int64_t row = i / Cv;
    int cv = (int)(i - row * Cv);
    int64_t n = row / HW;
    const float* a0 = attn + (n * 2 + 0) * C + cv * V;
    const float* a1 = attn + (n * 2 + 1) * C + cv * V;
    Pack<T, V> p = ld_pack<T, V>(u0 + i * V), q = ld_pack<T, V>(u1 + i * V), o;
#pragma unroll
    for (int j = 0; j < V; ++j) o.v[j] = from_f<T>(to_f(p.v[j]) * a0[j] + to_f(q.v[j]) * a1[j]);
    st_pack<T, V>(y + i * V, o);
  }
}

// du_i = dy*a_i + dg/HW ; t_i[n,c] = sum_hw dy*u_i ; dlogit_i = a_i*(t_i - (a0*t0 + a1*t1))
template <typename T, int V>
__global__ void sk_select_bwd_k(const T* __restrict__ u0, const T* __restrict__ u1,
                                const float* __restrict__ attn, const T* __restrict__ dy,
                                const float* __restrict__ dg, T* __restrict__ du0,
                                T* __restrict__ du1, float* __restrict__ dlogits, int HW, int C) {
  extern __shared__ float red[];  // [RY][2][W]
  const int n = blockIdx.x;
  const int cv = blockIdx.y * blockDim.x + threadIdx.x;
  const int c0 = cv * V;
  float t0[V], t1[V], a0[V], a1[V], g[V];
#pragma unroll
  for (int j = 0; j < V; ++j) t0[j] = t1[j] = 0.f;
  if (c0 < C) {
    const float inv = 1.f / (float)HW;
#pragma unroll
    for (int j = 0; j < V; ++j) {
      a0[j] = attn[((int64_t)n * 2 + 0) * C + c0 + j];
      a1[j] = attn[((int64_t)n * 2 + 1) * C + c0 + j];
      g[j] = dg ? dg[(int64_t)n * C + c0 + j] * inv : 0.f;
    }
    const int64_t base = (int64_t)n * HW * C + c0;
    for (int r = threadIdx.y; r < HW; r += blockDim.y) {
      const int64_t o = base + (int64_t)r * C;
      Pack<T, V> p = ld_pack<T, V>(u0 + o), q = ld_pack<T, V>(u1 + o), d = ld_pack<T, V>(dy + o);
      Pack<T, V> o0, o1;
#pragma unroll
      for (int j = 0; j < V; ++j) {
        float df = to_f(d.v[j]);
        t0[j] += df * to_f(p.v[j]);
        t1[j] += df * to_f(q.v[j]);
        o0.v[j] = from_f<T>(df * a0[j] + g[j]);
        o1.v[j] = from_f<T>(df * a1[j] + g[j]);
      }
      st_pack<T, V>(du0 + o, o0);
      st_pack<T, V>(du1 + o, o1);
    }
  }
  const int W = blockDim.x * V;
#pragma unroll
  for (int j = 0; j < V; ++j) {
    red[(threadIdx.y * 2 + 0) * W + threadIdx.x * V + j] = t0[j];
    red[(threadIdx.y * 2 + 1) * W + threadIdx.x * V + j] = t1[j];
  }
  __syncthreads();
  if (threadIdx.y == 0 && c0 < C) {
#pragma unroll
    for (int j = 0; j < V; ++j) {
      float s0 = 0.f, s1 = 0.f;
      for (int r = 0; r < (int)blockDim.y; ++r) {
        s0 += red[(r * 2 + 0) * W + threadIdx.x * V + j];
        s1 += red[(r * 2 + 1) * W + threadIdx.x * V + j];
      }
      float dot = a0[j] * s0 + a1[j] * s1;
      dlogits[((int64_t)n * 2 + 0) * C + c0 + j] = a0[j] * (s0 - dot);
      dlogits[((int64_t)n * 2 + 1) * C + c0 + j] = a1[j] * (s1 - dot);
    }
  }
}

// block geometry for the (n, channel-chunk) x rows reductions
struct RedGeom {
  dim3 grid, block;
  size_t smem;
};
static RedGeom red_geom(int N, int C, int V, int planes) {
  int cvs = C / V;
  int cx = cvs < 32 ? cvs : 32;
  // keep cx a power of two <= cvs for simple coverage; gridDim.y covers the remainder
  int p = 1;
  while (p * 2 <= cx) p *= 2;
  cx = p;
  int ry = 256 / cx;
  RedGeom g;
  g.block = dim3(cx, ry);
  g.grid = dim3(N, (unsigned)ceil_div(cvs, cx));
  g.smem = (size_t)ry * cx * V * sizeof(float) * planes;
  return g;
}

// fp32 master filter [K][RS][C] -> bf16 operand copy in the same layout and, optionally, the copy
// with in/out channels exchanged [C][RS][K] that data-grad consumes: one pass over the weights
// per optimiser step instead of a cast kernel plus a transpose inside every data-grad call.
__global__ void pack_filter_k(const float* __restrict__ w, bf16* __restrict__ wb, bf16* __restrict__ wt,
                              int K, int RS, int C) {
  const int64_t total = (int64_t)K * RS * C;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const bf16 v = __float2bfloat16_rn(w[i]);
    wb[i] = v;
    if (wt) {
      const int c = (int)(i % C);
      const int64_t t = i / C;
      const int rs = (int)(t % RS);
      const int k = (int)(t / RS);
      wt[((int64_t)c * RS + rs) * K + k] = v;
    }
  }
}

// Every filter of a model in ONE launch (cnb_pack_filters): block -> item by binary search over the
// items' first_block, then one [32 out-ch x 32 in-ch] tile of one tap: coalesced fp32 reads, the
// bf16 copy in place, the channel-exchanged copy through a padded shared-memory transpose (64-byte
// runs along K instead of 2-byte scattered stores).
__global__ void __launch_bounds__(256)
pack_filters_batched_k(const cnb_pack_item* __restrict__ items, int n_items) {
  __shared__ bf16 tile[32][34];
  int lo = 0, hi = n_items - 1;
  const int b = blockIdx.x;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (items[mid].first_block <= b) lo = mid;
    else hi = mid - 1;
  }
  const cnb_pack_item it = items[lo];
  const int lb = b - it.first_block;
  const int ctiles = (it.C + 31) / 32, ktiles = (it.K + 31) / 32;
  const int ct = lb % ctiles;
  const int t2 = lb / ctiles;
  const int kt = t2 % ktiles;
  const int rs = t2 / ktiles;
  if (rs >= it.RS) return;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int k0 = kt * 32, c0 = ct * 32;
  const float* w = it.w;
  bf16* wb = (bf16*)it.w_bf16;
  bf16* wt = (bf16*)it.wt_bf16;
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    const int k = k0 + ty + 8 * p, c = c0 + tx;
    bf16 v = __float2bfloat16_rn(0.f);
    if (k < it.K && c < it.C) {
      const int64_t i = ((int64_t)k * it.RS + rs) * it.C + c;
      v = __float2bfloat16_rn(w[i]);
      wb[i] = v;
    }
    tile[ty + 8 * p][tx] = v;
  }
  if (!wt) return;
  __syncthreads();
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    const int c = c0 + ty + 8 * p, k = k0 + tx;
    if (c < it.C && k < it.K) wt[((int64_t)c * it.RS + rs) * it.K + k] = tile[tx][ty + 8 * p];
  }
}

// ------------------------------------------------------------------ multi-tensor Adam (cnb_adam_step)
// The whole optimiser step of a model in two launches: `adam_tick_k` advances the step counter and
// derives the step's scalars in double (as the reference's Python does, torch/optim/adam.py
// _single_tensor_adam reached from src/basemodel.py:58-66,284), `adam_step_k` updates every
// parameter and writes the bf16 operand copies of the tensor-core filters in the same pass — the
// separate pack pass over the weights disappears.  Per element, fp32, in the reference's order:
//   g' = fma(wd, p, g);  m = fma(1 - b1, g' - m, m);  v = fma((1 - b2) * g', g', v * b2);
//   p += (-step_size * m) / (sqrt(v) / sqrt(1 - b2^t) + eps),  step_size = lr / (1 - b1^t)
// state (doubles, device): [0] t  [1] lr  [2] b1  [3] b2  [4] eps  [5] wd  | derived, written by the tick:
// [8] 1-b1  [9] b2  [10] 1-b2  [11] step_size  [12] sqrt(1-b2^t)  [13] eps  [14] wd
__global__ void adam_tick_k(double* __restrict__ st) {
  const double t = st[0] + 1.0;
  st[0] = t;
  const double b1 = st[2], b2 = st[3];
  st[8] = (double)(float)(1.0 - b1);
  st[9] = (double)(float)b2;
  st[10] = (double)(float)(1.0 - b2);
  st[11] = (double)(float)(st[1] / (1.0 - pow(b1, t)));
  st[12] = (double)(float)sqrt(1.0 - pow(b2, t));
  st[13] = (double)(float)st[4];
  st[14] = (double)(float)st[5];
}

struct AdamK {
  float w1, b2, w2, step_size, bc2_sqrt, eps, wd;
};
__device__ __forceinline__ float adam_update(float& p, float g, float& m, float& v, const AdamK& k) {
  // operation order and fusion of torch's CPU vector kernels (pinned in tests/test_optim.py)
  if (k.wd != 0.f) g = __fmaf_rn(k.wd, p, g);
  m = __fmaf_rn(k.w1, __fsub_rn(g, m), m);                                   // lerp_, weight < 0.5
  v = __fmaf_rn(__fmul_rn(k.w2, g), g, __fmul_rn(v, k.b2));                  // mul_ then addcmul_
  const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), k.bc2_sqrt), k.eps);
  p = __fadd_rn(p, __fdiv_rn(__fmul_rn(-k.step_size, m), denom));            // addcdiv_
  return p;
}

constexpr int kAdamFlat = 4096;   // elements per block of a flat item
constexpr int kAdamChunk = 56;    // items per launch: the table travels as a kernel parameter (< 4 KB),
struct AdamChunk {                // so a captured step holds it by value — no device table to keep alive
  cnb_adam_item items[kAdamChunk];
  int n;
};
__global__ void __launch_bounds__(256)
adam_step_k(const __grid_constant__ AdamChunk ch, const double* __restrict__ st) {
  __shared__ bf16 tile[32][34];
  const cnb_adam_item* items = ch.items;
  int lo = 0, hi = ch.n - 1;
  const int b = blockIdx.x;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (items[mid].first_block <= b) lo = mid;
    else hi = mid - 1;
  }
  const cnb_adam_item it = items[lo];
  const int lb = b - it.first_block;
  AdamK k;
  k.w1 = (float)st[8]; k.b2 = (float)st[9]; k.w2 = (float)st[10]; k.step_size = (float)st[11];
  k.bc2_sqrt = (float)st[12]; k.eps = (float)st[13]; k.wd = (float)st[14];
  bf16* wb = (bf16*)it.w_bf16;
  if (it.RS == 0) {   // flat item: K * C contiguous elements (K = 1 for plain tensors)
    const int64_t n = (int64_t)it.K * it.C;
    const int64_t e0 = (int64_t)lb * kAdamFlat;
    const bool vec = ((((uintptr_t)it.p | (uintptr_t)it.g | (uintptr_t)it.m | (uintptr_t)it.v) & 15) == 0);
#pragma unroll
    for (int u = 0; u < kAdamFlat / 1024; ++u) {
      const int64_t e = e0 + (int64_t)u * 1024 + threadIdx.x * 4;
      if (e >= n) break;
      if (vec && e + 4 <= n) {
        float4 p4 = *reinterpret_cast<const float4*>(it.p + e);
        const float4 g4 = *reinterpret_cast<const float4*>(it.g + e);
        float4 m4 = *reinterpret_cast<const float4*>(it.m + e);
        float4 v4 = *reinterpret_cast<const float4*>(it.v + e);
        adam_update(p4.x, g4.x, m4.x, v4.x, k);
        adam_update(p4.y, g4.y, m4.y, v4.y, k);
        adam_update(p4.z, g4.z, m4.z, v4.z, k);
        adam_update(p4.w, g4.w, m4.w, v4.w, k);
        *reinterpret_cast<float4*>(it.p + e) = p4;
        *reinterpret_cast<float4*>(it.m + e) = m4;
        *reinterpret_cast<float4*>(it.v + e) = v4;
        if (wb) {
          wb[e] = __float2bfloat16_rn(p4.x); wb[e + 1] = __float2bfloat16_rn(p4.y);
          wb[e + 2] = __float2bfloat16_rn(p4.z); wb[e + 3] = __float2bfloat16_rn(p4.w);
        }
      } else {
        for (int j = 0; j < 4 && e + j < n; ++j) {
          float p = it.p[e + j], m = it.m[e + j], v = it.v[e + j];
          adam_update(p, it.g[e + j], m, v, k);
          it.p[e + j] = p; it.m[e + j] = m; it.v[e + j] = v;
          if (wb) wb[e + j] = __float2bfloat16_rn(p);
        }
      }
    }
    return;
  }
  // filter item [K][RS][C]: one [32 out-ch x 32 in-ch] tile of one tap (layout of pack_filters_batched_k)
  const int ctiles = (it.C + 31) / 32, ktiles = (it.K + 31) / 32;
  const int ct = lb % ctiles;
  const int t2 = lb / ctiles;
  const int kt = t2 % ktiles;
  const int rs = t2 / ktiles;
  if (rs >= it.RS) return;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int k0 = kt * 32, c0 = ct * 32;
  bf16* wt = (bf16*)it.wt_bf16;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int kk = k0 + ty + 8 * q, c = c0 + tx;
    bf16 o = __float2bfloat16_rn(0.f);
    if (kk < it.K && c < it.C) {
      const int64_t i = ((int64_t)kk * it.RS + rs) * it.C + c;
      float p = it.p[i], m = it.m[i], v = it.v[i];
      adam_update(p, it.g[i], m, v, k);
      it.p[i] = p; it.m[i] = m; it.v[i] = v;
      o = __float2bfloat16_rn(p);
      if (wb) wb[i] = o;
    }
    tile[ty + 8 * q][tx] = o;
  }
  if (!wt) return;
  __syncthreads();
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int c = c0 + ty + 8 * q, kk = k0 + tx;
    if (c < it.C && kk < it.K) wt[((int64_t)c * it.RS + rs) * it.K + kk] = tile[tx][ty + 8 * q];
  }
}

}  // namespace cnb

using namespace cnb;

// ======================================================================= C ABI
extern "C" {

const char* cnb_version(void) { return "convnets_b200 0.1.0 (sm_100a)"; }
const char* cnb_last_error(void) { return cnb::g_last_error; }
uint64_t cnb_launch_count(void) { return cnb::g_launches.load(); }

int cnb_relu_fwd(const void* x, void* y, int64_t n, int dtype, void* stream) {
  CNB_REQUIRE(x && y && n >= 0 && valid_dtype(dtype), CNB_ERR_INVALID, "cnb_relu_fwd: bad args");
  if (n == 0) return CNB_OK;
  int v = pick_vec(n, dtype_size(dtype), x, y);
  CNB_DISPATCH_TV(dtype, v, (relu_fwd_k<T, V><<<grid_for(n / V, kThreads), kThreads, 0,
                                                 (cudaStream_t)stream>>>((const T*)x, (T*)y, n / V)));
  return check_launch("relu_fwd");
}

int cnb_relu_bwd(const void* y, const void* dy, void* dx, int64_t n, int dtype, void* stream) {
  CNB_REQUIRE(y && dy && dx && n >= 0 && valid_dtype(dtype), CNB_ERR_INVALID,
              "cnb_relu_bwd: bad args");
  if (n == 0) return CNB_OK;
  int v = pick_vec(n, dtype_size(dtype), y, dy, dx);
  CNB_DISPATCH_TV(dtype, v,
                  (relu_bwd_k<T, V><<<grid_for(n / V, kThreads), kThreads, 0, (cudaStream_t)stream>>>(
                      (const T*)y, (const T*)dy, (T*)dx, n / V)));
  return check_launch("relu_bwd");
}

int cnb_add_relu_fwd(const void* a, const void* b, void* y, int64_t n, int dtype, void* stream) {
  CNB_REQUIRE(a && b && y && n >= 0 && valid_dtype(dtype), CNB_ERR_INVALID,
              "cnb_add_relu_fwd: bad args");
  if (n == 0) return CNB_OK;
  int v = pick_vec(n, dtype_size(dtype), a, b, y);
  CNB_DISPATCH_TV(dtype, v,
                  (add_k<T, V, true><<<grid_for(n / V, kThreads), kThreads, 0, (cudaStream_t)stream>>>(
                      (const T*)a, (const T*)b, (T*)y, n / V)));
  return check_launch("add_relu_fwd");
}

int cnb_add_relu_bwd(const void* y, const void* dy, void* dx, int64_t n, int dtype, void* stream) {
  return cnb_relu_bwd(y, dy, dx, n, dtype, stream);
}

int cnb_add(const void* a, const void* b, void* y, int64_t n, int dtype, void* stream) {
  CNB_REQUIRE(a && b && y && n >= 0 && valid_dtype(dtype), CNB_ERR_INVALID, "cnb_add: bad args");
  if (n == 0) return CNB_OK;
  int v = pick_vec(n, dtype_size(dtype), a, b, y);
  CNB_DISPATCH_TV(dtype, v,
                  (add_k<T, V, false><<<grid_for(n / V, kThreads), kThreads, 0, (cudaStream_t)stream>>>(
                      (const T*)a, (const T*)b, (T*)y, n / V)));
  return check_launch("add");
}

int cnb_pack_filter(const float* w, void* w_bf16, void* wt_bf16, int32_t K, int32_t RS, int32_t C,
                    void* stream) {
  CNB_REQUIRE(w && w_bf16 && K > 0 && RS > 0 && C > 0, CNB_ERR_INVALID, "cnb_pack_filter: bad args");
  const int64_t n = (int64_t)K * RS * C;
  pack_filter_k<<<grid_for(n, kThreads), kThreads, 0, (cudaStream_t)stream>>>(
      w, (bf16*)w_bf16, (bf16*)wt_bf16, K, RS, C);
  return check_launch("pack_filter");
}

int64_t cnb_pack_filter_blocks(int32_t K, int32_t RS, int32_t C) {
  if (K <= 0 || RS <= 0 || C <= 0) return 0;
  return (int64_t)RS * ((K + 31) / 32) * ((C + 31) / 32);
}

int cnb_pack_filters(const cnb_pack_item* items_device, int32_t n_items, int32_t total_blocks,
                     void* stream) {
  CNB_REQUIRE(items_device && n_items > 0 && total_blocks > 0, CNB_ERR_INVALID,
              "cnb_pack_filters: bad args (items=%p n=%d blocks=%d)", (const void*)items_device, n_items,
              total_blocks);
  pack_filters_batched_k<<<(unsigned)total_blocks, 256, 0, (cudaStream_t)stream>>>(items_device, n_items);
  return check_launch("pack_filters");
}

int64_t cnb_adam_item_blocks(int32_t K, int32_t RS, int32_t C) {
  if (K <= 0 || RS < 0 || C <= 0) return 0;
  if (RS == 0) return ceil_div((int64_t)K * C, (int64_t)kAdamFlat);
  return (int64_t)RS * ceil_div(K, 32) * ceil_div(C, 32);
}

int cnb_adam_step(const cnb_adam_item* items_host, int32_t n_items, double* state_device,
                  void* stream) {
  CNB_REQUIRE(items_host && state_device && n_items > 0, CNB_ERR_INVALID,
              "cnb_adam_step: bad args (items=%p state=%p n=%d)", (const void*)items_host,
              (const void*)state_device, n_items);
  for (int i = 0; i < n_items; ++i) {
    const cnb_adam_item& it = items_host[i];
    CNB_REQUIRE(it.p && it.g && it.m && it.v && it.K > 0 && it.C > 0 && it.RS >= 0, CNB_ERR_INVALID,
                "cnb_adam_step: item %d is incomplete", i);
  }
  adam_tick_k<<<1, 1, 0, (cudaStream_t)stream>>>(state_device);
  int rc = check_launch("adam_tick");
  for (int i0 = 0; i0 < n_items && rc == CNB_OK; i0 += kAdamChunk) {
    AdamChunk ch;
    ch.n = n_items - i0 < kAdamChunk ? n_items - i0 : kAdamChunk;
    int64_t blocks = 0;
    for (int i = 0; i < ch.n; ++i) {
      ch.items[i] = items_host[i0 + i];
      ch.items[i].first_block = (int32_t)blocks;
      blocks += cnb_adam_item_blocks(ch.items[i].K, ch.items[i].RS, ch.items[i].C);
    }
    CNB_REQUIRE(blocks > 0 && blocks < (int64_t)INT32_MAX, CNB_ERR_INVALID, "cnb_adam_step: %lld blocks",
                (long long)blocks);
    adam_step_k<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(ch, state_device);
    rc = check_launch("adam_step");
  }
  return rc;
}

int cnb_cast(const void* src, int sdt, void* dst, int ddt, int64_t n, void* stream) {
  CNB_REQUIRE(src && dst && n >= 0 && valid_dtype(sdt) && valid_dtype(ddt), CNB_ERR_INVALID,
              "cnb_cast: bad args");
  if (n == 0) return CNB_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (sdt == CNB_F32 && ddt == CNB_BF16) {
    if (n % 4 == 0 && (uintptr_t)src % 16 == 0 && (uintptr_t)dst % 8 == 0)
      cast_f32_bf16_v4<<<grid_for(n / 4, kThreads), kThreads, 0, st>>>((const float*)src,
                                                                       (bf16*)dst, n / 4);
    else
      cast_k<float, bf16><<<grid_for(n, kThreads), kThreads, 0, st>>>((const float*)src, (bf16*)dst, n);
  } else if (sdt == CNB_BF16 && ddt == CNB_F32) {
    cast_k<bf16, float><<<grid_for(n, kThreads), kThreads, 0, st>>>((const bf16*)src, (float*)dst, n);
  } else if (sdt == CNB_F32) {
    cast_k<float, float><<<grid_for(n, kThreads), kThreads, 0, st>>>((const float*)src, (float*)dst, n);
  } else {
    cast_k<bf16, bf16><<<grid_for(n, kThreads), kThreads, 0, st>>>((const bf16*)src, (bf16*)dst, n);
  }
  return check_launch("cast");
}

int cnb_nchw_to_nhwc(const float* src, void* dst, int32_t N, int32_t C, int32_t H, int32_t W,
                     int32_t C_dst, int dtype, void* stream) {
  CNB_REQUIRE(src && dst && N > 0 && C > 0 && H > 0 && W > 0 && C_dst >= C && valid_dtype(dtype),
              CNB_ERR_INVALID, "cnb_nchw_to_nhwc: bad args");
  CNB_REQUIRE((size_t)C * 33 * 4 <= 48 * 1024, CNB_ERR_UNSUPPORTED,
              "cnb_nchw_to_nhwc: C=%d too large (input-image conversion only)", C);
  int HW = H * W;
  dim3 grid((unsigned)ceil_div(HW, 32), N);
  size_t smem = (size_t)C * 33 * sizeof(float);
  CNB_DISPATCH_T(dtype, (nchw_to_nhwc_k<T><<<grid, 256, smem, (cudaStream_t)stream>>>(src, (T*)dst, C, C_dst, HW)));
  return check_launch("nchw_to_nhwc");
}

int cnb_normalize_u8(const uint8_t* src, float* dst, int32_t N, int32_t C, int32_t H, int32_t W,
                     int32_t src_nhwc, int32_t H_out, int32_t W_out, int32_t pad,
                     const float* mean_host, const float* std_host, const int32_t* aug,
                     void* stream) {
  CNB_REQUIRE(src && dst && mean_host && std_host, CNB_ERR_INVALID, "cnb_normalize_u8: null pointer");
  CNB_REQUIRE(N > 0 && C > 0 && C <= 4 && H > 0 && W > 0 && H_out > 0 && W_out > 0 && pad >= 0,
              CNB_ERR_INVALID, "cnb_normalize_u8: bad args (C must be 1..4)");
  CNB_REQUIRE(aug || (pad == 0 && H_out == H && W_out == W), CNB_ERR_INVALID,
              "cnb_normalize_u8: a crop (output %dx%d of %dx%d, pad %d) needs the per-image offsets",
              H_out, W_out, H, W, pad);
  CNB_REQUIRE(H_out <= H + 2 * pad && W_out <= W + 2 * pad, CNB_ERR_INVALID,
              "cnb_normalize_u8: crop larger than the padded image");
  CNB_REQUIRE(((uintptr_t)dst & 15) == 0, CNB_ERR_INVALID, "cnb_normalize_u8: dst must be 16-byte aligned");
  U8NormP p;
  p.N = N; p.C = C; p.H = H; p.W = W; p.Ho = H_out; p.Wo = W_out; p.pad = pad; p.nhwc = src_nhwc ? 1 : 0;
  for (int c = 0; c < 4; ++c) {
    p.mean[c] = c < C ? mean_host[c] : 0.f;
    p.std[c] = c < C ? std_host[c] : 1.f;
    CNB_REQUIRE(p.std[c] != 0.f, CNB_ERR_INVALID, "cnb_normalize_u8: std[%d] is zero", c);
  }
  const int64_t total = (int64_t)N * H_out * ((W_out + 3) / 4);
  u8_normalize_k<<<grid_for(total, kThreads, 16), kThreads, 0, (cudaStream_t)stream>>>(src, dst, aug, p, total);
  return check_launch("u8_normalize");
}

int cnb_im2col_nchw(const cnb_conv_desc* d, const float* src, void* dst, int32_t K_dst,
                    void* stream) {
  CNB_REQUIRE(d && src && dst, CNB_ERR_INVALID, "cnb_im2col_nchw: null pointer");
  CNB_REQUIRE(d->N > 0 && d->H > 0 && d->W > 0 && d->C > 0 && d->R > 0 && d->S > 0 &&
                  d->stride_h > 0 && d->stride_w > 0 && d->dil_h > 0 && d->dil_w > 0 &&
                  d->pad_h >= 0 && d->pad_w >= 0 && valid_dtype(d->dtype),
              CNB_ERR_INVALID, "cnb_im2col_nchw: bad descriptor");
  int P, Q;
  conv_out_dims(*d, P, Q);
  CNB_REQUIRE(P > 0 && Q > 0, CNB_ERR_INVALID, "cnb_im2col_nchw: empty output");
  const int v = (int)(16 / dtype_size(d->dtype));
  CNB_REQUIRE(K_dst >= d->R * d->S * d->C && K_dst % v == 0 && (uintptr_t)dst % 16 == 0,
              CNB_ERR_INVALID, "cnb_im2col_nchw: K_dst=%d must be >= R*S*C=%d, a multiple of %d, dst 16B-aligned",
              K_dst, d->R * d->S * d->C, v);
  Im2colP p{d->N, d->C, d->H, d->W, d->R, d->S, P, Q, d->stride_h, d->stride_w, d->pad_h, d->pad_w,
            d->dil_h, d->dil_w, K_dst};
  const int64_t total = (int64_t)d->N * P * Q * (K_dst / v);
  {
    // output-staged kernel when patch + one output row block fit 96 KB of shared memory
    const size_t esz = dtype_size(d->dtype);
    const int kv = K_dst / v;
    const int row_stride = K_dst + (kv % 2 == 0 ? v : 0);  // odd 16-byte stride: conflict-free phase A
    const size_t patch_elems = (size_t)d->R * (d->W + 2 * d->pad_w) * d->C;
    const size_t smem = (size_t)((K_dst * 4 + 15) / 16) * 16 + ((patch_elems + v - 1) / v) * v * esz +
                        (size_t)Q * row_stride * esz;
    if (smem <= 96 * 1024 && kv <= 256 && (int64_t)d->N * P < (1ll << 31)) {
      const unsigned grid = (unsigned)((int64_t)d->N * P);
      static bool attr_bf16 = false, attr_f32 = false;  // benign race: same value from every thread
      if (d->dtype == CNB_BF16) {
        if (!attr_bf16) {
          cudaError_t e = cudaFuncSetAttribute(im2col_nchw_staged_k<bf16, 8>,
                                               cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
          CNB_REQUIRE(e == cudaSuccess, CNB_ERR_CUDA, "cnb_im2col_nchw: cudaFuncSetAttribute: %s",
                      cudaGetErrorString(e));
          attr_bf16 = true;
        }
        im2col_nchw_staged_k<bf16, 8><<<grid, 256, smem, (cudaStream_t)stream>>>(src, (bf16*)dst, p, row_stride);
      } else {
        if (!attr_f32) {
          cudaError_t e = cudaFuncSetAttribute(im2col_nchw_staged_k<float, 4>,
                                               cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
          CNB_REQUIRE(e == cudaSuccess, CNB_ERR_CUDA, "cnb_im2col_nchw: cudaFuncSetAttribute: %s",
                      cudaGetErrorString(e));
          attr_f32 = true;
        }
        im2col_nchw_staged_k<float, 4><<<grid, 256, smem, (cudaStream_t)stream>>>(src, (float*)dst, p, row_stride);
      }
      return check_launch("im2col_nchw_staged");
    }
  }
  {
    // row-staged kernel when its smem patch fits the default 48 KB
    const size_t smem = (size_t)((K_dst * 4 + 15) / 16) * 16 +
                        (size_t)d->C * d->R * (d->W + 2 * d->pad_w) * dtype_size(d->dtype);
    if (smem <= 48 * 1024 && (int64_t)d->N * P < (1ll << 31)) {
      const unsigned grid = (unsigned)((int64_t)d->N * P);
      if (d->dtype == CNB_BF16)
        im2col_nchw_rows_k<bf16, 8><<<grid, 256, smem, (cudaStream_t)stream>>>(src, (bf16*)dst, p);
      else
        im2col_nchw_rows_k<float, 4><<<grid, 256, smem, (cudaStream_t)stream>>>(src, (float*)dst, p);
      return check_launch("im2col_nchw_rows");
    }
  }
  if (d->dtype == CNB_BF16)
    im2col_nchw_k<bf16, 8><<<grid_for(total, kThreads, 16), kThreads, 0, (cudaStream_t)stream>>>(
        src, (bf16*)dst, p, total);
  else
    im2col_nchw_k<float, 4><<<grid_for(total, kThreads, 16), kThreads, 0, (cudaStream_t)stream>>>(
        src, (float*)dst, p, total);
  return check_launch("im2col_nchw");
}

int cnb_rowim2col_nchw(const cnb_conv_desc* d, const float* src, void* dst, int32_t C_dst,
                       void* stream) {
  CNB_REQUIRE(d && src && dst, CNB_ERR_INVALID, "cnb_rowim2col_nchw: null pointer");
  CNB_REQUIRE(d->N > 0 && d->H > 0 && d->W > 0 && d->C > 0 && d->S > 0 && d->stride_w > 0 &&
                  d->dil_w > 0 && d->pad_w >= 0 && valid_dtype(d->dtype),
              CNB_ERR_INVALID, "cnb_rowim2col_nchw: bad descriptor");
  const int Q = (d->W + 2 * d->pad_w - d->dil_w * (d->S - 1) - 1) / d->stride_w + 1;
  CNB_REQUIRE(Q > 0, CNB_ERR_INVALID, "cnb_rowim2col_nchw: empty output");
  const int v = (int)(16 / dtype_size(d->dtype));
  CNB_REQUIRE(C_dst >= d->S * d->C && C_dst % v == 0 && (uintptr_t)dst % 16 == 0, CNB_ERR_INVALID,
              "cnb_rowim2col_nchw: C_dst=%d must be >= S*C=%d, a multiple of %d, dst 16B-aligned", C_dst,
              d->S * d->C, v);
  CNB_REQUIRE(C_dst / v <= 256 && (int64_t)d->N * d->H < (1ll << 31), CNB_ERR_UNSUPPORTED,
              "cnb_rowim2col_nchw: C_dst=%d too wide (network-input stems only)", C_dst);
  Im2colP p{d->N, d->C, d->H, d->W, 1, d->S, d->H, Q, 1, d->stride_w, 0, d->pad_w, 1, d->dil_w, C_dst};
  const size_t smem = (size_t)((C_dst * 4 + 15) / 16) * 16 +
                      (size_t)(d->W + 2 * d->pad_w) * d->C * dtype_size(d->dtype);
  CNB_REQUIRE(smem <= 48 * 1024, CNB_ERR_UNSUPPORTED, "cnb_rowim2col_nchw: row of %d x %d too large", d->W,
              d->C);
  const unsigned grid = (unsigned)((int64_t)d->N * d->H);
  if (d->dtype == CNB_BF16)
    rowim2col_nchw_k<bf16, 8><<<grid, 256, smem, (cudaStream_t)stream>>>(src, (bf16*)dst, p);
  else
    rowim2col_nchw_k<float, 4><<<grid, 256, smem, (cudaStream_t)stream>>>(src, (float*)dst, p);
  return check_launch("rowim2col_nchw");
}

int cnb_channel_shuffle(const void* x, void* y, int64_t M, int32_t C, int32_t groups, int dtype,
                        void* stream) {
  CNB_REQUIRE(x && y && M >= 0 && C > 0 && groups > 0 && valid_dtype(dtype), CNB_ERR_INVALID,
              "cnb_channel_shuffle: bad args");
  CNB_REQUIRE(C % groups == 0, CNB_ERR_INVALID, "cnb_channel_shuffle: C=%d not divisible by groups=%d",
              C, groups);
  if (M == 0) return CNB_OK;
  if (shuffle_vec(x, y, M, C, groups, dtype, (cudaStream_t)stream)) return check_launch("channel_shuffle(vec)");
  CNB_DISPATCH_T(dtype, (shuffle_k<T><<<grid_for(M * C, kThreads), kThreads, 0, (cudaStream_t)stream>>>(
                            (const T*)x, (T*)y, M, C, groups)));
  return check_launch("channel_shuffle");
}

int cnb_copy_channels(const void* src, int32_t src_C, int32_t src_off, void* dst, int32_t dst_C,
                      int32_t dst_off, int64_t M, int32_t Ccopy, int dtype, void* stream) {
  CNB_REQUIRE(src && dst && M >= 0 && Ccopy > 0 && valid_dtype(dtype), CNB_ERR_INVALID,
              "cnb_copy_channels: bad args");
  CNB_REQUIRE(src_off >= 0 && dst_off >= 0 && src_off + Ccopy <= src_C && dst_off + Ccopy <= dst_C,
              CNB_ERR_INVALID, "cnb_copy_channels: slab [%d,+%d) / [%d,+%d) outside rows of %d / %d",
              src_off, Ccopy, dst_off, Ccopy, src_C, dst_C);
  if (M == 0) return CNB_OK;
  size_t esz = dtype_size(dtype);
  int v = (int)(16 / esz);
  auto ok = [&](int vv) {
    return Ccopy % vv == 0 && src_C % vv == 0 && dst_C % vv == 0 && src_off % vv == 0 &&
           dst_off % vv == 0 && (uintptr_t)src % (vv * esz) == 0 && (uintptr_t)dst % (vv * esz) == 0;
  };
  while (v > 1 && !ok(v)) v >>= 1;
  CNB_DISPATCH_TV(dtype, v,
                  (copy_channels_k<T, V><<<grid_for(M * (Ccopy / V), kThreads), kThreads, 0,
                                           (cudaStream_t)stream>>>((const T*)src, src_C, src_off,
                                                                   (T*)dst, dst_C, dst_off, M,
                                                                   Ccopy / V)));
  return check_launch("copy_channels");
}

int cnb_gap_fwd(const void* x, float* y, int32_t N, int32_t HW, int32_t C, int dtype, void* stream) {
  CNB_REQUIRE(x && y && N > 0 && HW > 0 && C > 0 && valid_dtype(dtype), CNB_ERR_INVALID,
              "cnb_gap_fwd: bad args");
  int v = pick_vec(C, dtype_size(dtype), x);
  RedGeom g = red_geom(N, C, v, 1);
  CNB_DISPATCH_TV(dtype, v, (gap_fwd_k<T, V><<<g.grid, g.block, g.smem, (cudaStream_t)stream>>>(
                                (const T*)x, (const T*)nullptr, y, HW, C)));
  return check_launch("gap_fwd");
}

int cnb_sk_fuse_gap_fwd(const void* u0, const void* u1, float* gout, int32_t N, int32_t HW,
                        int32_t C, int dtype, void* stream) {
  CNB_REQUIRE(u0 && u1 && gout && N > 0 && HW > 0 && C > 0 && valid_dtype(dtype), CNB_ERR_INVALID,
              "cnb_sk_fuse_gap_fwd: bad args");
  int v = pick_vec(C, dtype_size(dtype), u0, u1);
  RedGeom g = red_geom(N, C, v, 1);
  CNB_DISPATCH_TV(dtype, v, (gap_fwd_k<T, V><<<g.grid, g.block, g.smem, (cudaStream_t)stream>>>(
                                (const T*)u0, (const T*)u1, gout, HW, C)));
  return check_launch("sk_fuse_gap_fwd");
}

int cnb_gap_bwd(const float* dy, void* dx, int32_t N, int32_t HW, int32_t C, int dtype,
                void* stream) {
  CNB_REQUIRE(dy && dx && N > 0 && HW > 0 && C > 0 && valid_dtype(dtype), CNB_ERR_INVALID,
              "cnb_gap_bwd: bad args");
  int v = pick_vec(C, dtype_size(dtype), dx);
  int64_t total = (int64_t)N * HW * (C / v);
  CNB_DISPATCH_TV(dtype, v, (gap_bwd_k<T, V><<<grid_for(total, kThreads), kThreads, 0,
                                                (cudaStream_t)stream>>>(dy, (T*)dx, total, HW, C / V)));
  return check_launch("gap_bwd");
}

int cnb_se_scale_fwd(const void* x, const float* s, void* y, int32_t N, int32_t HW, int32_t C,
                     int dtype, void* stream) {
  CNB_REQUIRE(x && s && y && N > 0 && HW > 0 && C > 0 && valid_dtype(dtype), CNB_ERR_INVALID,
              "cnb_se_scale_fwd: bad args");
  int v = pick_vec(C, dtype_size(dtype), x, y);
  int64_t total = (int64_t)N * HW * (C / v);
  CNB_DISPATCH_TV(dtype, v,
                  (se_scale_fwd_k<T, V><<<grid_for(total, kThreads), kThreads, 0, (cudaStream_t)stream>>>(
                      (const T*)x, s, (T*)y, total, HW, C / V)));
  return check_launch("se_scale_fwd");
}

int cnb_se_scale_bwd(const void* x, const float* s, const void* dy, void* dx, float* ds, int32_t N,
                     int32_t HW, int32_t C, int dtype, void* stream) {
  CNB_REQUIRE(x && s && dy && dx && ds && N > 0 && HW > 0 && C > 0 && valid_dtype(dtype),
              CNB_ERR_INVALID, "cnb_se_scale_bwd: bad args");
  int v = pick_vec(C, dtype_size(dtype), x, dy, dx);
  RedGeom g = red_geom(N, C, v, 1);
  CNB_DISPATCH_TV(dtype, v, (se_scale_bwd_k<T, V><<<g.grid, g.block, g.smem, (cudaStream_t)stream>>>(
                                (const T*)x, s, (const T*)dy, (T*)dx, ds, HW, C)));
  return check_launch("se_scale_bwd");
}

int cnb_sk_select_fwd(const void* u0, const void* u1, const float* logits, float* attn, void* y,
                      int32_t N, int32_t HW, int32_t C, int dtype, void* stream) {
  CNB_REQUIRE(u0 && u1 && logits && attn && y && N > 0 && HW > 0 && C > 0 && valid_dtype(dtype),
              CNB_ERR_INVALID, "cnb_sk_select_fwd: bad args");
  cudaStream_t st = (cudaStream_t)stream;
  sk_softmax2_k<<<(unsigned)ceil_div((int64_t)N * C, 256), 256, 0, st>>>(logits, attn, N, C);
  CNB_TRY(check_launch("sk_softmax2"));
  int v = pick_vec(C, dtype_size(dtype), u0, u1, y);
  if (v > 4) v = 4;  // attn is read as scalars; keep register pressure low
  int64_t total = (int64_t)N * HW * (C / v);
  CNB_DISPATCH_TV(dtype, v, (sk_select_fwd_k<T, V><<<grid_for(total, kThreads), kThreads, 0, st>>>(
                                (const T*)u0, (const T*)u1, attn, (T*)y, total, HW, C / V)));
  return check_launch("sk_select_fwd");
}

int cnb_sk_select_bwd(const void* u0, const void* u1, const float* attn, const void* dy,
                      const float* dg, void* du0, void* du1, float* dlogits, int32_t N, int32_t HW,
                      int32_t C, int dtype, void* stream) {
  CNB_REQUIRE(u0 && u1 && attn && dy && du0 && du1 && dlogits && N > 0 && HW > 0 && C > 0 &&
                  valid_dtype(dtype),
              CNB_ERR_INVALID, "cnb_sk_select_bwd: bad args");
  int v = pick_vec(C, dtype_size(dtype), u0, u1, dy, du0);
  if (v > 4) v = 4;
  while (v > 1 && (uintptr_t)du1 % (v * dtype_size(dtype)) != 0) v >>= 1;
  RedGeom g = red_geom(N, C, v, 2);
  CNB_DISPATCH_TV(dtype, v, (sk_select_bwd_k<T, V><<<g.grid, g.block, g.smem, (cudaStream_t)stream>>>(
                                (const T*)u0, (const T*)u1, attn, (const T*)dy, dg, (T*)du0,
                                (T*)du1, dlogits, HW, C)));
  return check_launch("sk_select_bwd");
}

}  // extern "C"
