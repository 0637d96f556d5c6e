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

// ------------------------------------------------------------------------------------------
// Pointwise (1x1, stride 1) convolutions with narrow groups that the TMA path cannot address
// (C or K not a multiple of 8: ShuffleNet-g3 stage 2 has 60-channel tensors = 120-byte rows).
// They are skinny GEMMs — hundreds of thousands of pixels against <= 96 x 96 channels per group —
// and HBM-bound, but the generic 64x64 tile above fills 20 of its 64 columns and loads scalars.
//   pw_gemm_k  : out[m][n0 + j] = sum_k A[m][a0 + k] * B[k][j]   (forward and data-grad)
//                CTA = (128 pixels, one group); the group's whole filter and the A tile (stored
//                k-major) sit in shared memory as fp32; a thread owns 4 pixels x 4 channels and
//                reads two 16-byte vectors per 16 FMAs.
//   pw_wgrad_k : dW[k][c] = sum_m dy[m][k] * x[m][c] per group, split over pixel ranges; 64-pixel
//                chunks of dy and x are staged in shared memory, a thread owns a 4 x 4 tile of dW
//                (several thread lanes share a tile when the group is small and are summed in a
//                fixed order at the end) -> fp32 partials -> sum_splits2_k.
constexpr int PW_BM = 128, PW_MAXD = 96, PW_RC = 64;

template <typename T>
__global__ void __launch_bounds__(256)
pw_gemm_k(const T* __restrict__ A, int lda, const T* __restrict__ W, int64_t sWk, int64_t sWj,
          int64_t sWg, T* __restrict__ out, int ldo, int64_t M, int Kd, int Ng,
          const float* __restrict__ bias, int flags) {
  extern __shared__ __align__(16) float pw_smem[];
  const int Np = (Ng + 3) & ~3;
  constexpr int LDA = PW_BM + 4;
  float* Bs = pw_smem;                 // [Kd][Np]
  float* As = pw_smem + Kd * Np;       // [Kd][LDA]  (k-major: 4 pixels of one k are one float4)
  const int t = threadIdx.x;
  const int g = blockIdx.y;
  const int64_t m0 = (int64_t)blockIdx.x * PW_BM;
  const int a0 = g * Kd, n0 = g * Ng;
  const T* Wg = W + g * sWg;
  for (int i = t; i < Kd * Np; i += 256) {
    const int k = i / Np, j = i - k * Np;
    Bs[i] = j < Ng ? to_f(Wg[k * sWk + j * sWj]) : 0.f;
  }
  // 4-element vectors along the reduction when rows, group offsets and pointers allow it: a thread
  // keeps ONE pixel (conflict-free k-major stores) and strides the vectors — no divisions
  const bool avec = (Kd % 4 == 0) && (lda % 4 == 0) && ((uintptr_t)A % (4 * sizeof(T)) == 0);
  if (avec) {
    const int m = t & (PW_BM - 1);
    const int64_t row = m0 + m;
    const T* arow = A + row * lda + a0;
    for (int v = t / PW_BM; v < Kd / 4; v += 256 / PW_BM) {
      float f[4] = {0.f, 0.f, 0.f, 0.f};
      if (row < M) {
        const Pack<T, 4> q = ld_pack<T, 4>(arow + v * 4);
#pragma unroll
        for (int j = 0; j < 4; ++j) f[j] = to_f(q.v[j]);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) As[(v * 4 + j) * LDA + m] = f[j];
    }
  } else {
    for (int i = t; i < PW_BM * Kd; i += 256) {
      const int m = i / Kd, k = i - m * Kd;
      const int64_t row = m0 + m;
      As[k * LDA + m] = row < M ? to_f(A[row * lda + a0 + k]) : 0.f;
    }
  }
  __syncthreads();
  const bool ovec = (Ng % 4 == 0) && (ldo % 4 == 0) && ((uintptr_t)out % (4 * sizeof(T)) == 0);
  const int rgp = t & 31;
  for (int cg = t >> 5; cg < Np / 4; cg += 8) {
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    const float* ap = As + rgp * 4;
    const float* bp = Bs + cg * 4;
#pragma unroll 4
    for (int k = 0; k < Kd; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(ap + k * LDA);
      const float4 b = *reinterpret_cast<const float4*>(bp + k * Np);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int64_t row = m0 + rgp * 4 + i;
      if (row >= M) continue;
      Pack<T, 4> o;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int col = cg * 4 + j;
        float v = acc[i][j];
        if ((flags & CNB_CONV_BIAS) && col < Ng) v += bias[n0 + col];
        if (flags & CNB_CONV_RELU) v = (v > 0.f || v != v) ? v : 0.f;
        o.v[j] = from_f<T>(v);
        if (!ovec && col < Ng) out[row * ldo + n0 + col] = o.v[j];
      }
      if (ovec) st_pack<T, 4>(out + row * ldo + n0 + cg * 4, o);  // Ng % 4 == 0: the tile is inside
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
pw_wgrad_k(const T* __restrict__ dy, int ldy, const T* __restrict__ x, int ldx,
           float* __restrict__ part, int64_t M, int Kg, int Cg, int K, int64_t rows_per_split) {
  extern __shared__ __align__(16) float pw_smem[];
  const int Kp = (Kg + 3) & ~3, Cp = (Cg + 3) & ~3;
  float* dys = pw_smem;                  // [PW_RC][Kp]
  float* xs = pw_smem + PW_RC * Kp;      // [PW_RC][Cp]
  const int t = threadIdx.x;
  const int g = blockIdx.y;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_split;
  int64_t r1 = r0 + rows_per_split;
  if (r1 > M) r1 = M;
  const int ctiles = Cp / 4;
  const int tiles = (Kp / 4) * ctiles;
  // few tiles: several row lanes per tile; many tiles: up to 3 tiles per thread, one lane
  const int lanes = tiles < 256 ? 256 / tiles : 1;
  const int lane = tiles < 256 ? t / tiles : 0;
  const int tile0 = tiles < 256 ? t % tiles : t;
  const bool live = lane < lanes;
  float acc[3][4][4];
#pragma unroll
  for (int s = 0; s < 3; ++s)
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[s][i][j] = 0.f;
  for (int i = t; i < PW_RC * (Kp + Cp); i += 256) pw_smem[i] = 0.f;  // padding columns stay zero
  const bool vec = (Kg % 4 == 0) && (Cg % 4 == 0) && (ldy % 4 == 0) && (ldx % 4 == 0) &&
                   ((uintptr_t)dy % (4 * sizeof(T)) == 0) && ((uintptr_t)x % (4 * sizeof(T)) == 0);
  for (int64_t mb = r0; mb < r1; mb += PW_RC) {
    __syncthreads();
    if (vec) {
      const int m = t & (PW_RC - 1);
      const bool rok = mb + m < r1;
      const T* dyr = dy + (mb + m) * ldy + g * Kg;
      const T* xr = x + (mb + m) * ldx + g * Cg;
      for (int v = t / PW_RC; v < Kg / 4; v += 256 / PW_RC) {
        float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
        if (rok) {
          const Pack<T, 4> q = ld_pack<T, 4>(dyr + v * 4);
          f = make_float4(to_f(q.v[0]), to_f(q.v[1]), to_f(q.v[2]), to_f(q.v[3]));
        }
        *reinterpret_cast<float4*>(dys + m * Kp + v * 4) = f;
      }
      for (int v = t / PW_RC; v < Cg / 4; v += 256 / PW_RC) {
        float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
        if (rok) {
          const Pack<T, 4> q = ld_pack<T, 4>(xr + v * 4);
          f = make_float4(to_f(q.v[0]), to_f(q.v[1]), to_f(q.v[2]), to_f(q.v[3]));
        }
        *reinterpret_cast<float4*>(xs + m * Cp + v * 4) = f;
      }
    } else {
      for (int i = t; i < PW_RC * Kg; i += 256) {
        const int m = i / Kg, k = i - m * Kg;
        dys[m * Kp + k] = mb + m < r1 ? to_f(dy[(mb + m) * ldy + g * Kg + k]) : 0.f;
      }
      for (int i = t; i < PW_RC * Cg; i += 256) {
        const int m = i / Cg, c = i - m * Cg;
        xs[m * Cp + c] = mb + m < r1 ? to_f(x[(mb + m) * ldx + g * Cg + c]) : 0.f;
      }
    }
    __syncthreads();
    if (live) {
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int tile = tile0 + s * 256;
        if (tile < tiles && (s == 0 || tiles >= 256)) {
          const int kt = tile / ctiles, ct = tile - kt * ctiles;
          for (int m = lane; m < PW_RC; m += lanes) {
            const float4 a = *reinterpret_cast<const float4*>(dys + m * Kp + kt * 4);
            const float4 b = *reinterpret_cast<const float4*>(xs + m * Cp + ct * 4);
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
              for (int j = 0; j < 4; ++j) acc[s][i][j] = fmaf(av[i], bv[j], acc[s][i][j]);
          }
        }
      }
    }
  }
  __syncthreads();
  float* outp = part + ((int64_t)blockIdx.x * K + (int64_t)g * Kg) * Cg;
  if (tiles >= 256) {
#pragma unroll
    for (int s = 0; s < 3; ++s) {
      const int tile = tile0 + s * 256;
      if (tile < tiles) {
        const int kt = tile / ctiles, ct = tile - kt * ctiles;
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (kt * 4 + i < Kg && ct * 4 + j < Cg) outp[(kt * 4 + i) * Cg + ct * 4 + j] = acc[s][i][j];
      }
    }
    return;
  }
  // sum the row lanes of each tile in lane order through shared memory ([lanes][tiles][16] floats)
  float* red = pw_smem;
  if (live) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) red[(lane * tiles + tile0) * 16 + i * 4 + j] = acc[0][i][j];
  }
  __syncthreads();
  for (int e = t; e < tiles * 16; e += 256) {
    const int tile = e / 16, ij = e - tile * 16;
    float sum = 0.f;
    for (int l = 0; l < lanes; ++l) sum += red[(l * tiles + tile) * 16 + ij];
    const int kt = tile / ctiles, ct = tile - kt * ctiles;
    const int k = kt * 4 + ij / 4, c = ct * 4 + (ij & 3);
    if (k < Kg && c < Cg) outp[k * Cg + c] = sum;
  }
}

static bool is_pointwise_narrow(const ConvP& p) {
  return p.R == 1 && p.S == 1 && p.sh == 1 && p.sw == 1 && p.ph == 0 && p.pw == 0 &&
         p.Cg <= PW_MAXD && p.Kg <= PW_MAXD;
}
static size_t pw_gemm_smem(int Kd, int Ng) {
  return ((size_t)Kd * ((Ng + 3) & ~3) + (size_t)Kd * (PW_BM + 4)) * sizeof(float);
}
static size_t pw_wgrad_smem(int Kg, int Cg) {
  const size_t kp = (Kg + 3) & ~3, cp = (Cg + 3) & ~3;
  size_t stage = PW_RC * (kp + cp) * sizeof(float);
  size_t red = 256 * 16 * sizeof(float);
  return stage > red ? stage : red;
}
static int pw_wgrad_splits(const ConvP& p) {
  const int64_t M = (int64_t)p.N * p.P * p.Q;
  int64_t want = ceil_div((int64_t)sm_count() * 4, p.G);
  const int64_t maxs = ceil_div(M, 4 * PW_RC);
  if (want > maxs) want = maxs;
  if (want < 1) want = 1;
  return (int)want;
}
// opt in to 100 KB of dynamic shared memory, once per kernel instantiation (T = its element type)
template <typename T, int WHICH, typename K>
static int pw_set_smem(K kern, size_t bytes, const char* what) {
  static bool done = false;
  CNB_REQUIRE(bytes <= 100 * 1024, CNB_ERR_UNSUPPORTED, "%s: shared memory %zu > 100 KB", what, bytes);
  if (!done) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    CNB_REQUIRE(e == cudaSuccess, CNB_ERR_CUDA, "%s: cudaFuncSetAttribute: %s", what, cudaGetErrorString(e));
    done = true;
  }
  return CNB_OK;
}

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
  if (is_pointwise_narrow(p) && !is_vector_gemm(p)) splits = pw_wgrad_splits(p);
  return (size_t)splits * p.K * p.R * p.S * p.Cg * sizeof(float) + 256;
}

int conv_simt_fwd(const cnb_conv_desc& d, const void* x, const void* w, const float* bias, void* y,
                  cudaStream_t st) {
  ConvP p = make_convp(d);
  int64_t Mg = (int64_t)p.N * p.P * p.Q;
  dim3 grid((unsigned)ceil_div(Mg, BM), (unsigned)ceil_div(p.Kg, BN), p.G);
  CNB_REQUIRE(grid.y <= 65535 && grid.z <= 65535, CNB_ERR_UNSUPPORTED, "conv_simt_fwd: grid too large");
  int flags = d.flags & ((bias ? CNB_CONV_BIAS : 0) | CNB_CONV_RELU);
  if (!is_vector_gemm(p) && is_pointwise_narrow(p)) {
    // y[m][g*Kg + j] = sum_c x[m][g*Cg + c] * w[g*Kg + j][c]:  B[k=c][j] = w[(g*Kg + j)*Cg + c]
    const size_t smem = pw_gemm_smem(p.Cg, p.Kg);
    dim3 pgrid((unsigned)ceil_div(Mg, PW_BM), p.G);
    CNB_DISPATCH_T(d.dtype, {
      CNB_TRY((pw_set_smem<T, 0>(pw_gemm_k<T>, smem, "conv_simt_fwd(pointwise)")));
      pw_gemm_k<T><<<pgrid, 256, smem, st>>>((const T*)x, p.C, (const T*)w, 1, p.Cg, (int64_t)p.Kg * p.Cg,
                                             (T*)y, p.K, Mg, p.Cg, p.Kg, bias, flags);
    });
    return check_launch("conv_simt_fwd(pointwise)");
  }
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
  if (!is_vector_gemm(p) && is_pointwise_narrow(p)) {
    // dx[m][g*Cg + c] = sum_k dy[m][g*Kg + k] * w[g*Kg + k][c]:  B[k][j=c] = w[(g*Kg + k)*Cg + c]
    const size_t smem = pw_gemm_smem(p.Kg, p.Cg);
    dim3 pgrid((unsigned)ceil_div(Mg, PW_BM), p.G);
    CNB_DISPATCH_T(d.dtype, {
      CNB_TRY((pw_set_smem<T, 0>(pw_gemm_k<T>, smem, "conv_simt_dgrad(pointwise)")));
      pw_gemm_k<T><<<pgrid, 256, smem, st>>>((const T*)dy, p.K, (const T*)w, p.Cg, 1, (int64_t)p.Kg * p.Cg,
                                             (T*)dx, p.C, Mg, p.Kg, p.Cg, nullptr, 0);
    });
    return check_launch("conv_simt_dgrad(pointwise)");
  }
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
  if (is_pointwise_narrow(p)) {
    const int64_t M = (int64_t)p.N * p.P * p.Q;
    const int64_t wsz1 = (int64_t)p.K * p.Cg;
    int psplits = pw_wgrad_splits(p);
    const int64_t rows = ceil_div(M, psplits);
    psplits = (int)ceil_div(M, rows);
    CNB_REQUIRE(ws && ws_bytes >= (size_t)psplits * wsz1 * sizeof(float), CNB_ERR_WORKSPACE,
                "conv_simt_wgrad(pointwise): workspace %zu < %zu", ws_bytes,
                (size_t)psplits * wsz1 * sizeof(float));
    const size_t smem = pw_wgrad_smem(p.Kg, p.Cg);
    dim3 pgrid((unsigned)psplits, p.G);
    CNB_DISPATCH_T(d.dtype, {
      CNB_TRY((pw_set_smem<T, 1>(pw_wgrad_k<T>, smem, "conv_simt_wgrad(pointwise)")));
      pw_wgrad_k<T><<<pgrid, 256, smem, st>>>((const T*)dy, p.K, (const T*)x, p.C, (float*)ws, M, p.Kg, p.Cg,
                                              p.K, rows);
    });
    CNB_TRY(check_launch("conv_simt_wgrad(pointwise)"));
    sum_splits2_k<<<grid_for(wsz1, 256), 256, 0, st>>>((const float*)ws, dw, psplits, wsz1);
    return check_launch("conv_simt_wgrad_sum");
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
