// conv_simt.cu — CUDA-core implicit-GEMM convolution (forward, data-grad, weight-grad) for every
// shape the tcgen05 path does not take: fp32-IO mode (exact fp32 FMA accumulation, the
// "<= 1e-4" parity mode of BASELINE.json), grouped convolutions whose per-group width cannot feed
// a 128-row MMA (ResNeXt/SKNet 4..32 channels per group, ShuffleNet 20..320), and C=3 stems.
// Replaces aten::convolution / convolution_backward (reference src/basemodel.py:974).
//
// One generic 64x64x16 shared-memory-tiled kernel; the three directions differ only in how the
// GEMM operands are gathered from the NHWC / KRSC tensors:
//   fwd  : M = N*P*Q          N = K/g        Kd = R*S*C/g   A = im2col(x)        B = w
//   dgrad: M = N*H*W          N = C/g        Kd = R*S*K/g   A = im2col^T(dy)     B = w (transposed)
//   wgrad: M = K/g            N = R*S*C/g    Kd = N*P*Q     A = dy^T             B = im2col(x)
// wgrad splits Kd over CTAs into fp32 partials which a second kernel sums (deterministic).
#include "common.cuh"

namespace cnb {

struct ConvP {
  int N, H, W, C, K, R, S, P, Q;
  int sh, sw, ph, pw, dh, dw;
  int G, Cg, Kg;
};

static ConvP make_convp(const cnb_conv_desc& d) {
  ConvP p;
  p.N = d.N; p.H = d.H; p.W = d.W; p.C = d.C; p.K = d.K; p.R = d.R; p.S = d.S;
  conv_out_dims(d, p.P, p.Q);
  p.sh = d.stride_h; p.sw = d.stride_w; p.ph = d.pad_h; p.pw = d.pad_w;
  p.dh = d.dil_h; p.dw = d.dil_w;
  p.G = d.groups; p.Cg = d.C / d.groups; p.Kg = d.K / d.groups;
  return p;
}

constexpr int BM = 64, BN = 64, BK = 16;
constexpr int MODE_FWD = 0, MODE_DGRAD = 1, MODE_WGRAD = 2;

template <typename T, int MODE>
__global__ void __launch_bounds__(256)
conv_simt_k(ConvP p, const T* __restrict__ src_a, const T* __restrict__ src_b,
            void* __restrict__ out, const float* __restrict__ bias, int flags, int splits,
            int64_t kd_per_split) {
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];

  const int t = threadIdx.x;
  const int g = blockIdx.z / splits;
  const int split = blockIdx.z - g * splits;
  const int64_t m0 = (int64_t)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;

  int64_t Mg, Kd;
  int Ng;
  if (MODE == MODE_FWD) { Mg = (int64_t)p.N * p.P * p.Q; Ng = p.Kg; Kd = (int64_t)p.R * p.S * p.Cg; }
  else if (MODE == MODE_DGRAD) { Mg = (int64_t)p.N * p.H * p.W; Ng = p.Cg; Kd = (int64_t)p.R * p.S * p.Kg; }
  else { Mg = p.Kg; Ng = p.R * p.S * p.Cg; Kd = (int64_t)p.N * p.P * p.Q; }

  int64_t k_begin = 0, k_end = Kd;
  if (MODE == MODE_WGRAD) {
    k_begin = (int64_t)split * kd_per_split;
    k_end = k_begin + kd_per_split;
    if (k_end > Kd) k_end = Kd;
  }

  // ---- per-thread operand coordinates that do not change over the k loop
  // A: fwd/dgrad -> k-contiguous mapping (kk = t%16, rows t/16 + 16*i)
  //    wgrad     -> m-contiguous mapping (mm = t%64, ks  t/64 + 4*i)
  int a_n[4], a_h[4], a_w[4];
  bool a_ok[4];
  if (MODE != MODE_WGRAD) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int64_t m = m0 + t / 16 + 16 * i;
      a_ok[i] = m < Mg;
      int64_t mm = a_ok[i] ? m : 0;
      int WW = (MODE == MODE_FWD) ? p.Q : p.W;
      int HH = (MODE == MODE_FWD) ? p.P : p.H;
      a_w[i] = (int)(mm % WW);
      int64_t r = mm / WW;
      a_h[i] = (int)(r % HH);
      a_n[i] = (int)(r / HH);
    }
  }
  // B: fwd -> k-contiguous mapping (kk = t%16, cols t/16 + 16*i); dgrad/wgrad -> n-contiguous
  int b_r = 0, b_s = 0, b_c = 0;
  bool b_ok = false;
  if (MODE == MODE_WGRAD) {
    int nn = n0 + t % 64;
    b_ok = nn < Ng;
    int nv = b_ok ? nn : 0;
    b_c = nv % p.Cg;
    int rs = nv / p.Cg;
    b_s = rs % p.S;
    b_r = rs / p.S;
  }

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  const int ty = t / 16, tx = t % 16;

  for (int64_t k0 = k_begin; k0 < k_end; k0 += BK) {
    // ------------------------------------------------ load A tile
    if (MODE == MODE_FWD) {
      int64_t k = k0 + t % 16;
      bool kok = k < k_end;
      int kv = kok ? (int)k : 0;
      int c = kv % p.Cg;
      int rs = kv / p.Cg;
      int s = rs % p.S, r = rs / p.S;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float v = 0.f;
        int h = a_h[i] * p.sh - p.ph + r * p.dh;
        int w = a_w[i] * p.sw - p.pw + s * p.dw;
        if (kok && a_ok[i] && h >= 0 && h < p.H && w >= 0 && w < p.W)
          v = to_f(src_a[(((int64_t)a_n[i] * p.H + h) * p.W + w) * p.C + g * p.Cg + c]);
        As[t % 16][t / 16 + 16 * i] = v;
      }
    } else if (MODE == MODE_DGRAD) {
      int64_t k = k0 + t % 16;
      bool kok = k < k_end;
      int kv = kok ? (int)k : 0;
      int kk = kv % p.Kg;
      int rs = kv / p.Kg;
      int s = rs % p.S, r = rs / p.S;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float v = 0.f;
        int hh = a_h[i] + p.ph - r * p.dh;
        int ww = a_w[i] + p.pw - s * p.dw;
        if (kok && a_ok[i] && hh >= 0 && ww >= 0 && hh % p.sh == 0 && ww % p.sw == 0) {
          int oh = hh / p.sh, ow = ww / p.sw;
          if (oh < p.P && ow < p.Q)
            v = to_f(src_a[(((int64_t)a_n[i] * p.P + oh) * p.Q + ow) * p.K + g * p.Kg + kk]);
        }
        As[t % 16][t / 16 + 16 * i] = v;
      }
    } else {
      int mm = (int)m0 + t % 64;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        int64_t k = k0 + t / 64 + 4 * i;
        float v = 0.f;
        if (mm < Mg && k < k_end) v = to_f(src_a[k * p.K + g * p.Kg + mm]);
        As[t / 64 + 4 * i][t % 64] = v;
      }
    }
    // ------------------------------------------------ load B tile
    if (MODE == MODE_FWD) {
      int64_t k = k0 + t % 16;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        int nn = n0 + t / 16 + 16 * i;
        float v = 0.f;
        if (k < k_end && nn < Ng) v = to_f(src_b[((int64_t)g * p.Kg + nn) * Kd + k]);
        Bs[t % 16][t / 16 + 16 * i] = v;
      }
    } else if (MODE == MODE_DGRAD) {
      int c = n0 + t % 64;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        int64_t k = k0 + t / 64 + 4 * i;
        float v = 0.f;
        if (k < k_end && c < Ng) {
          int kk = (int)(k % p.Kg);
          int rs = (int)(k / p.Kg);
          v = to_f(src_b[(((int64_t)g * p.Kg + kk) * p.R * p.S + rs) * p.Cg + c]);
        }
        Bs[t / 64 + 4 * i][t % 64] = v;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        int64_t k = k0 + t / 64 + 4 * i;
        float v = 0.f;
        if (b_ok && k < k_end) {
          int ow = (int)(k % p.Q);
          int64_t r2 = k / p.Q;
          int oh = (int)(r2 % p.P);
          int n = (int)(r2 / p.P);
          int h = oh * p.sh - p.ph + b_r * p.dh;
          int w = ow * p.sw - p.pw + b_s * p.dw;
          if (h >= 0 && h < p.H && w >= 0 && w < p.W)
            v = to_f(src_b[(((int64_t)n * p.H + h) * p.W + w) * p.C + g * p.Cg + b_c]);
        }
        Bs[t / 64 + 4 * i][t % 64] = v;
      }
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

  // ------------------------------------------------ store
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int64_t m = m0 + ty * 4 + i;
    if (m >= Mg) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int nn = n0 + tx * 4 + j;
      if (nn >= Ng) continue;
      float v = acc[i][j];
      if (MODE == MODE_FWD) {
        int ko = g * p.Kg + nn;
        if (flags & CNB_CONV_BIAS) v += bias[ko];
        if (flags & CNB_CONV_RELU) v = (v > 0.f || v != v) ? v : 0.f;
        reinterpret_cast<T*>(out)[m * p.K + ko] = from_f<T>(v);
      } else if (MODE == MODE_DGRAD) {
        reinterpret_cast<T*>(out)[m * p.C + g * p.Cg + nn] = from_f<T>(v);
      } else {
        int64_t wsz = (int64_t)p.K * p.R * p.S * p.Cg;
        reinterpret_cast<float*>(out)[(int64_t)split * wsz +
                                      ((int64_t)g * p.Kg + m) * ((int64_t)p.R * p.S * p.Cg) + nn] = v;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// 1x1 convolutions on [B, C, 1, 1] vectors (SK descriptor / attention convs, src/sknet.py:205-219):
// a skinny GEMM out[i][j] = sum_k A[i*sAi + k*sAk] * B[k*sBk + j*sBj] with a handful of rows.  The
// 64x64 tile kernel above would run it on one or two CTAs; here every output element gets KS = 8
// lanes that stride the reduction (fixed-order shuffle tree -> deterministic).
template <typename T, typename TO>
__global__ void __launch_bounds__(256)
vec_gemm_k(const T* __restrict__ A, int64_t sAi, int64_t sAk, const T* __restrict__ B, int64_t sBk,
           int64_t sBj, TO* __restrict__ out, int I, int J, int Kd, const float* __restrict__ bias,
           int flags) {
  constexpr int KS = 8;
  const int64_t gid = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int kl = (int)(gid % KS);
  const int64_t o = gid / KS;
  const bool live = o < (int64_t)I * J;
  const int i = live ? (int)(o / J) : 0, j = live ? (int)(o % J) : 0;
  const T* a = A + i * sAi;
  const T* b = B + j * sBj;
  float acc0 = 0.f, acc1 = 0.f;
  int k = kl;
  for (; k + KS < Kd; k += 2 * KS) {
    acc0 = fmaf(to_f(a[k * sAk]), to_f(b[k * sBk]), acc0);
    acc1 = fmaf(to_f(a[(k + KS) * sAk]), to_f(b[(k + KS) * sBk]), acc1);
  }
  if (k < Kd) acc0 = fmaf(to_f(a[k * sAk]), to_f(b[k * sBk]), acc0);
  float v = acc0 + acc1;
#pragma unroll
  for (int off = KS / 2; off > 0; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off, KS);
  if (live && kl == 0) {
    if (flags & CNB_CONV_BIAS) v += bias[j];
    if (flags & CNB_CONV_RELU) v = (v > 0.f || v != v) ? v : 0.f;
    out[o] = from_f<TO>(v);
  }
}

static bool is_vector_gemm(const ConvP& p) {
  return p.G == 1 && p.R == 1 && p.S == 1 && p.H == 1 && p.W == 1 && p.P == 1 && p.Q == 1 &&
         p.ph == 0 && p.pw == 0;
}
static unsigned vec_gemm_blocks(int I, int J) { return (unsigned)ceil_div((int64_t)I * J * 8, 256); }

__global__ void sum_splits2_k(const float* __restrict__ part, float* __restrict__ out, int splits,
                              int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    float s = 0.f;
    for (int k = 0; k < splits; ++k) s += part[(int64_t)k * n + i];
    out[i] = s;
  }
}

// dbias[k] = sum over pixels of dy[m, k]; one block per 32 channels, 8 row-lanes.
template <typename T>
__global__ void bias_grad_k(const T* __restrict__ dy, float* __restrict__ db, int64_t M, int K) {
  __shared__ float red[8][33];
  int k = blockIdx.x * 32 + threadIdx.x;
  float s = 0.f;
  if (k < K)
    for (int64_t m = threadIdx.y; m < M; m += 8) s += to_f(dy[m * K + k]);
  red[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && k < K) {
    float tsum = 0.f;
    for (int r = 0; r < 8; ++r) tsum += red[r][threadIdx.x];
    db[k] = tsum;
  }
}

static int wgrad_splits(const ConvP& p) {
  int64_t tiles = ceil_div(p.Kg, BM) * ceil_div((int64_t)p.R * p.S * p.Cg, BN) * p.G;
  int64_t Kd = (int64_t)p.N * p.P * p.Q;
  int64_t want = ceil_div((int64_t)sm_count() * 4, tiles);
  int64_t maxs = ceil_div(Kd, 4 * BK);
  if (want > maxs) want = maxs;
  if (want < 1) want = 1;
  if (want > 256) want = 256;
  return (int)want;
}

size_t conv_simt_wgrad_ws(const cnb_conv_desc& d) {
  ConvP p = make_convp(d);
  int splits = wgrad_splits(p);
  return (size_t)splits * p.K * p.R * p.S * p.Cg * sizeof(float) + 256;
}

int conv_simt_fwd(const cnb_conv_desc& d, const void* x, const void* w, const float* bias, void* y,
                  cudaStream_t st) {
  ConvP p = make_convp(d);
  int64_t Mg = (int64_t)p.N * p.P * p.Q;
  dim3 grid((unsigned)ceil_div(Mg, BM), (unsigned)ceil_div(p.Kg, BN), p.G);
  CNB_REQUIRE(grid.y <= 65535 && grid.z <= 65535, CNB_ERR_UNSUPPORTED, "conv_simt_fwd: grid too large");
  int flags = d.flags & ((bias ? CNB_CONV_BIAS : 0) | CNB_CONV_RELU);
  if (is_vector_gemm(p)) {  // y[n][k] = sum_c x[n][c] * w[k][c]
    CNB_DISPATCH_T(d.dtype, (vec_gemm_k<T, T><<<vec_gemm_blocks(p.N, p.K), 256, 0, st>>>(
                                (const T*)x, p.C, 1, (const T*)w, 1, p.C, (T*)y, p.N, p.K, p.C, bias, flags)));
    return check_launch("conv_simt_fwd(vector)");
  }
  CNB_DISPATCH_T(d.dtype, (conv_simt_k<T, MODE_FWD><<<grid, 256, 0, st>>>(p, (const T*)x, (const T*)w,
                                                                          y, bias, flags, 1, 0)));
  return check_launch("conv_simt_fwd");
}

int conv_simt_dgrad(const cnb_conv_desc& d, const void* dy, const void* w, void* dx,
                    cudaStream_t st) {
  ConvP p = make_convp(d);
  int64_t Mg = (int64_t)p.N * p.H * p.W;
  dim3 grid((unsigned)ceil_div(Mg, BM), (unsigned)ceil_div(p.Cg, BN), p.G);
  CNB_REQUIRE(grid.y <= 65535 && grid.z <= 65535, CNB_ERR_UNSUPPORTED, "conv_simt_dgrad: grid too large");
  if (is_vector_gemm(p)) {  // dx[n][c] = sum_k dy[n][k] * w[k][c]
    CNB_DISPATCH_T(d.dtype, (vec_gemm_k<T, T><<<vec_gemm_blocks(p.N, p.C), 256, 0, st>>>(
                                (const T*)dy, p.K, 1, (const T*)w, p.C, 1, (T*)dx, p.N, p.C, p.K, nullptr, 0)));
    return check_launch("conv_simt_dgrad(vector)");
  }
  CNB_DISPATCH_T(d.dtype, (conv_simt_k<T, MODE_DGRAD><<<grid, 256, 0, st>>>(
                              p, (const T*)dy, (const T*)w, dx, nullptr, 0, 1, 0)));
  return check_launch("conv_simt_dgrad");
}

int conv_simt_wgrad(const cnb_conv_desc& d, const void* x, const void* dy, float* dw, void* ws,
                    size_t ws_bytes, cudaStream_t st) {
  ConvP p = make_convp(d);
  if (is_vector_gemm(p)) {  // dw[k][c] = sum_n dy[n][k] * x[n][c]
    CNB_DISPATCH_T(d.dtype, (vec_gemm_k<T, float><<<vec_gemm_blocks(p.K, p.C), 256, 0, st>>>(
                                (const T*)dy, 1, p.K, (const T*)x, p.C, 1, dw, p.K, p.C, p.N, nullptr, 0)));
    return check_launch("conv_simt_wgrad(vector)");
  }
  int splits = wgrad_splits(p);
  int64_t wsz = (int64_t)p.K * p.R * p.S * p.Cg;
  int64_t Kd = (int64_t)p.N * p.P * p.Q;
  int64_t per = ceil_div(ceil_div(Kd, splits), BK) * BK;
  splits = (int)ceil_div(Kd, per);
  float* part = dw;
  if (splits > 1) {
    CNB_REQUIRE(ws && ws_bytes >= (size_t)splits * wsz * sizeof(float), CNB_ERR_WORKSPACE,
                "conv_simt_wgrad: workspace %zu < %zu", ws_bytes, (size_t)splits * wsz * sizeof(float));
    part = (float*)ws;
  }
  dim3 grid((unsigned)ceil_div(p.Kg, BM), (unsigned)ceil_div((int64_t)p.R * p.S * p.Cg, BN),
            (unsigned)(p.G * splits));
  CNB_REQUIRE(grid.y <= 65535 && grid.z <= 65535, CNB_ERR_UNSUPPORTED, "conv_simt_wgrad: grid too large");
  CNB_DISPATCH_T(d.dtype, (conv_simt_k<T, MODE_WGRAD><<<grid, 256, 0, st>>>(
                              p, (const T*)dy, (const T*)x, part, nullptr, 0, splits, per)));
  CNB_TRY(check_launch("conv_simt_wgrad"));
  if (splits > 1) {
    sum_splits2_k<<<grid_for(wsz, 256), 256, 0, st>>>(part, dw, splits, wsz);
    CNB_TRY(check_launch("conv_simt_wgrad_sum"));
  }
  return CNB_OK;
}

int conv_bias_grad(const cnb_conv_desc& d, const void* dy, float* dbias, cudaStream_t st) {
  int P, Q;
  conv_out_dims(d, P, Q);
  int64_t M = (int64_t)d.N * P * Q;
  dim3 grid((unsigned)ceil_div(d.K, 32)), block(32, 8);
  CNB_DISPATCH_T(d.dtype, (bias_grad_k<T><<<grid, block, 0, st>>>((const T*)dy, dbias, M, d.K)));
  return check_launch("conv_bias_grad");
}

}  // namespace cnb
