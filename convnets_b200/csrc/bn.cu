// bn.cu — BatchNorm2d forward (training / inference) and backward over NHWC tensors viewed as
// [M = N*H*W, C].  Replaces aten::native_batch_norm(_backward) reached from reference
// src/basemodel.py:1008,1061; src/densenet.py:80,117,124,182; src/mobilenet_v1.py:108,119.
//
// Structure (all HBM-bound, 128-bit accesses along C):
//   stats   : per-(row-chunk, channel) shifted sums -> fp32 partials in the workspace; the last
//             CTA of each channel block to arrive merges the partials in split order (fp64),
//             updates the running statistics and emits the scale/shift vectors
//   apply   : y = [relu](x*scale + shift [+ residual])
//   bwd     : reduce (sum dy', sum dy'*(x-mean)) + last-CTA merge (dgamma, dbeta, coefficients) -> dx
// The merges sum in a fixed order (no floating-point atomics): results are bit-reproducible.
#include <string.h>

#include <mutex>

#include "common.cuh"

namespace cnb {

constexpr int kBnThreads = 256;
constexpr int kBnUnroll = 4;

struct BnGeom {
  int V;          // channels per thread
  int cx;         // threads along channels
  int ry;         // threads along rows
  int cblocks;    // blocks along channels
  int splits;     // blocks along rows
  int64_t rows_per_split;
};

// `waves`: target resident CTAs per SM.  `max_cx`: cap on the threads along channels.
// Streaming kernels (apply, dx) use full 512-byte row segments (max_cx = 32) and many CTAs.
// Reduction kernels use 128-byte row segments (max_cx = 8 -> at most 64 channels per CTA): their
// partials are merged by the last CTA of each channel block, and that serial tail is proportional
// to (CTAs per channel block) x (channels per CTA) / 256 threads — narrow channel blocks keep it
// to a few microseconds where a 256-channel-wide CTA column spent ~70 us in it.
static BnGeom bn_geom(int64_t M, int C, int V, int waves, int max_cx = 32) {
  BnGeom g;
  g.V = V;
  int cvs = C / V;
  int p = 1;
  while (p * 2 <= cvs && p * 2 <= max_cx) p *= 2;
  g.cx = p;
  g.ry = kBnThreads / g.cx;
  g.cblocks = (int)ceil_div(cvs, g.cx);
  int64_t target = (int64_t)sm_count() * waves;
  int64_t splits = target / g.cblocks;  // floor: never more CTAs than `waves` whole waves
  int64_t max_splits = ceil_div(M, (int64_t)g.ry * kBnUnroll);  // >= one unrolled trip per thread
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  if (splits > 2048) splits = 2048;
  g.rows_per_split = ceil_div(M, splits);
  g.splits = (int)ceil_div(M, g.rows_per_split);
  return g;
}

// Arrival counters for the "last CTA of a channel block merges the partials" step.  Library-owned,
// zero-initialised once per device and left at zero by every kernel that uses them; each call
// takes the next of kCounterSlots slots so that calls in flight on different streams do not share.
constexpr int kCounterSlots = 64, kCountersPerSlot = 1024;
static int* bn_counters(int cblocks) {
  static int* base[64] = {nullptr};
  static std::atomic<unsigned> next{0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64 || cblocks > kCountersPerSlot) return nullptr;
  if (!base[dev]) {
    int* p = nullptr;
    if (cudaMalloc(&p, kCounterSlots * kCountersPerSlot * sizeof(int)) != cudaSuccess) return nullptr;
    cudaMemset(p, 0, kCounterSlots * kCountersPerSlot * sizeof(int));
    base[dev] = p;
  }
  return base[dev] + (next.fetch_add(1) % kCounterSlots) * kCountersPerSlot;
}

// workspace layout (floats): [0, 2*splits*C) partials | then 5*C floats of per-channel
// coefficients (scale, shift / c1, c2, c3)

struct BnFwdFin {
  const float* gamma; const float* beta;
  float* rmean; float* rvar; float* save_mean; float* save_rstd; float* scale; float* shift;
  float eps, momentum;
  long long* nbt;   // num_batches_tracked (may be null): +1 per training forward, by channel 0's thread
};
struct BnBwdFin {
  const float* gamma; const float* rstd;
  float* dgamma; float* dbeta; float* c1; float* c2; float* c3;
};

// Block-level tail shared by the reduction kernels, part 1: tree-reduce the per-thread (s, q) over
// the row lanes and store the CTA's partial row ([split][2][C] floats).
template <int V>
__device__ __forceinline__ void bn_block_partial(float (&s)[V], float (&q)[V], float* red,
                                                 float* __restrict__ part, int C) {
  const int cx = blockDim.x, ry = blockDim.y;
  const int W = cx * V;
  const int t = threadIdx.y * cx + threadIdx.x;
#pragma unroll
  for (int j = 0; j < V; ++j) {
    red[(threadIdx.y * 2 + 0) * W + threadIdx.x * V + j] = s[j];
    red[(threadIdx.y * 2 + 1) * W + threadIdx.x * V + j] = q[j];
  }
  __syncthreads();
  for (int st = ry >> 1; st > 0; st >>= 1) {
    if ((int)threadIdx.y < st) {
#pragma unroll
      for (int j = 0; j < 2 * V; ++j) {
        const int col = threadIdx.x * V + (j % V);
        const int which = j / V;
        red[(threadIdx.y * 2 + which) * W + col] += red[((threadIdx.y + st) * 2 + which) * W + col];
      }
    }
    __syncthreads();
  }
  const int cbase = blockIdx.x * W;
  for (int i = t; i < 2 * W; i += kBnThreads) {
    const int which = i / W, col = i % W;
    if (cbase + col < C)
      part[((int64_t)blockIdx.y * 2 + which) * C + cbase + col] = red[which * W + col];
  }
}

// Part 2: sum the partial rows of this CTA's channel block in split order (fixed order => bit-
// reproducible, and identical in every CTA that runs it) in fp64.  Returns true in the threads that
// own a finished channel (S, Q valid).  `red` is re-used as 256 x 2 doubles.
template <int V>
__device__ __forceinline__ bool bn_merge_partials(float* red, const float* __restrict__ part, int C,
                                                  int splits, int& c_out, double& S, double& Q) {
  const int cx = blockDim.x;
  const int W = cx * V;
  const int t = threadIdx.y * cx + threadIdx.x;
  const int cbase = blockIdx.x * W;
  // 256 threads = (256 / W) split-lanes x W channels
  double* dred = reinterpret_cast<double*>(red);
  const int lanes = kBnThreads / W;
  const int col = t % W, lane = t / W;
  const int c = cbase + col;
  double a = 0.0, b = 0.0;
  if (c < C) {
    constexpr int MU = 8;  // independent L2 loads in flight per thread; summation order is fixed
    for (int sp = lane; sp < splits; sp += lanes * MU) {
      float va[MU], vb[MU];
#pragma unroll
      for (int u = 0; u < MU; ++u) {
        const int si = sp + u * lanes;
        va[u] = si < splits ? __ldcg(part + ((int64_t)si * 2 + 0) * C + c) : 0.f;
        vb[u] = si < splits ? __ldcg(part + ((int64_t)si * 2 + 1) * C + c) : 0.f;
      }
#pragma unroll
      for (int u = 0; u < MU; ++u) {
        a += (double)va[u];
        b += (double)vb[u];
      }
    }
  }
  __syncthreads();
  dred[(lane * 2 + 0) * W + col] = a;
  dred[(lane * 2 + 1) * W + col] = b;
  __syncthreads();
  if (lane != 0 || c >= C) return false;
  S = 0.0; Q = 0.0;
  for (int l = 0; l < lanes; ++l) {
    S += dred[(l * 2 + 0) * W + col];
    Q += dred[(l * 2 + 1) * W + col];
  }
  c_out = c;
  return true;
}

// Two-kernel form: the last-arriving CTA of a channel block merges (everyone else leaves).
template <int V>
__device__ __forceinline__ bool bn_block_merge(float (&s)[V], float (&q)[V], float* red,
                                               float* __restrict__ part, int C, int splits,
                                               int* __restrict__ counter, int& c_out, double& S,
                                               double& Q) {
  bn_block_partial<V>(s, q, red, part, C);
  const int t = threadIdx.y * blockDim.x + threadIdx.x;
  __threadfence();
  __syncthreads();
  __shared__ int s_last;
  if (t == 0) {
    int prev = atomicAdd(counter + blockIdx.x, 1);
    s_last = (prev == splits - 1);
    if (s_last) counter[blockIdx.x] = 0;  // leave the slot clean for the next call
  }
  __syncthreads();
  if (!s_last) return false;
  __threadfence();
  return bn_merge_partials<V>(red, part, C, splits, c_out, S, Q);
}

// Per-channel epilogue of the forward statistics: S = sum(x - K), Q = sum((x - K)^2).
__device__ __forceinline__ void bn_finish_fwd(int c, double S, double Q, double K, int64_t M,
                                              const BnFwdFin& f) {
  const double ms = S / (double)M;
  const double mean = K + ms;
  double var = Q / (double)M - ms * ms;  // biased: used for normalisation
  if (var < 0.0) var = 0.0;
  const float rstd = (float)(1.0 / sqrt(var + (double)f.eps));
  const float mu = (float)mean;
  f.save_mean[c] = mu;
  f.save_rstd[c] = rstd;
  if (f.nbt && c == 0) *f.nbt += 1;
  if (f.rmean) f.rmean[c] = (1.f - f.momentum) * f.rmean[c] + f.momentum * mu;
  if (f.rvar) {
    double unbiased = M > 1 ? var * (double)M / (double)(M - 1) : var;
    f.rvar[c] = (1.f - f.momentum) * f.rvar[c] + f.momentum * (float)unbiased;
  }
  const float g = f.gamma ? f.gamma[c] : 1.f, b = f.beta ? f.beta[c] : 0.f;
  f.scale[c] = g * rstd;
  f.shift[c] = b;  // bn_apply_k computes (x - mean) * scale + shift: no mean*scale cancellation
}

// ---------------------------------------------------------------- forward statistics
// Shifted sums about the pivot K_c = x[row 0][c] (kills the mean^2/var cancellation of a raw
// sum of squares): s = sum(x - K), q = sum((x - K)^2); mean = K + s/M, var = q/M - (s/M)^2.
template <typename T, int V>
__global__ void __launch_bounds__(kBnThreads)
bn_stats_k(const T* __restrict__ x, float* __restrict__ part, int64_t M, int C,
           int64_t rows_per_split, int splits, int* __restrict__ counter, BnFwdFin f) {
  extern __shared__ __align__(16) float red[];  // [ry][2][cx*V] floats (>= lanes*2*W doubles)
  const int cv = blockIdx.x * blockDim.x + threadIdx.x;
  const int c0 = cv * V;
  const int64_t r0 = blockIdx.y * rows_per_split;
  int64_t r1 = r0 + rows_per_split;
  if (r1 > M) r1 = M;
  float s[V], q[V], piv[V];
#pragma unroll
  for (int j = 0; j < V; ++j) s[j] = q[j] = piv[j] = 0.f;
  if (c0 < C) {
    Pack<T, V> p0 = ld_pack<T, V>(x + c0);
#pragma unroll
    for (int j = 0; j < V; ++j) piv[j] = to_f(p0.v[j]);
    const int64_t step = (int64_t)blockDim.y * kBnUnroll;
    for (int64_t r = r0 + threadIdx.y; r < r1; r += step) {
      Pack<T, V> a[kBnUnroll];
#pragma unroll
      for (int u = 0; u < kBnUnroll; ++u) {
        int64_t rr = r + (int64_t)u * blockDim.y;
        if (rr < r1) a[u] = ld_pack<T, V>(x + rr * C + c0);
      }
#pragma unroll
      for (int u = 0; u < kBnUnroll; ++u) {
        int64_t rr = r + (int64_t)u * blockDim.y;
        if (rr < r1) {
#pragma unroll
          for (int j = 0; j < V; ++j) {
            float d = to_f(a[u].v[j]) - piv[j];
            s[j] += d;
            q[j] = fmaf(d, d, q[j]);
          }
        }
      }
    }
  }
  int c;
  double S, Q;
  if (!bn_block_merge<V>(s, q, red, part, C, splits, counter, c, S, Q)) return;
  bn_finish_fwd(c, S, Q, (double)to_f(x[c]), M, f);
}

// Pointer-walk form of bn_stats_k (same sums, same merge; see the backward kernels for the loop
// shape): U rows in flight per thread, MINB resident CTAs per SM.  Opt-in through
// CNB_BN_STATS_VARIANT until it has been swept on the GPU (tools/gpu_bn_variants.sh <tag> STATS ...).
template <typename T, int V, int U, int MINB>
__global__ void __launch_bounds__(kBnThreads, MINB)
bn_stats_walk_k(const T* __restrict__ x, float* __restrict__ part, int64_t M, int C,
                int64_t rows_per_split, int splits, int* __restrict__ counter, BnFwdFin f) {
  extern __shared__ __align__(16) float red[];
  const int cv = blockIdx.x * blockDim.x + threadIdx.x;
  const int c0 = cv * V;
  const int64_t r0 = blockIdx.y * rows_per_split;
  int64_t r1 = r0 + rows_per_split;
  if (r1 > M) r1 = M;
  float s[V], q[V], piv[V];
#pragma unroll
  for (int j = 0; j < V; ++j) s[j] = q[j] = piv[j] = 0.f;
  if (c0 < C) {
    const Pack<T, V> p0 = ld_pack<T, V>(x + c0);
#pragma unroll
    for (int j = 0; j < V; ++j) piv[j] = to_f(p0.v[j]);
    const int ry = blockDim.y;
    const int64_t stride = (int64_t)ry * C;
    int64_t r = r0 + threadIdx.y;
    const T* xp = x + r * C + c0;
    auto accum = [&](const Pack<T, V>& a) {
#pragma unroll
      for (int j = 0; j < V; ++j) {
        const float dlt = to_f(a.v[j]) - piv[j];
        s[j] += dlt;
        q[j] = fmaf(dlt, dlt, q[j]);
      }
    };
    for (; r + (int64_t)(U - 1) * ry < r1; r += (int64_t)U * ry) {
      Pack<T, V> a[U];
#pragma unroll
      for (int u = 0; u < U; ++u) a[u] = ld_pack<T, V>(xp + u * stride);
#pragma unroll
      for (int u = 0; u < U; ++u) accum(a[u]);
      xp += U * stride;
    }
    for (; r < r1; r += ry) {
      accum(ld_pack<T, V>(xp));
      xp += stride;
    }
  }
  int c;
  double S, Q;
  if (!bn_block_merge<V>(s, q, red, part, C, splits, counter, c, S, Q)) return;
  bn_finish_fwd(c, S, Q, (double)to_f(x[c]), M, f);
}

// Statistics from per-tile partials written by the convolution epilogue (conv_tc.cu):
// tiles[n][2][C] fp32 = column sums / sums of squares of 128-row tiles of the stored activations.
// Same two-level, fixed-order merge as bn_stats_k (rows here are tiles, no pivot).
__global__ void __launch_bounds__(kBnThreads)
bn_partials_k(const float* __restrict__ tiles, float* __restrict__ part, int64_t ntiles, int64_t M,
              int C, int64_t rows_per_split, int splits, int* __restrict__ counter, BnFwdFin f) {
  extern __shared__ __align__(16) float red[];
  const int c0 = blockIdx.x * blockDim.x + threadIdx.x;  // V = 1: one channel per thread column
  const int64_t r0 = blockIdx.y * rows_per_split;
  int64_t r1 = r0 + rows_per_split;
  if (r1 > ntiles) r1 = ntiles;
  float s[1] = {0.f}, q[1] = {0.f};
  if (c0 < C) {
    constexpr int U = 2;
    const int64_t step = (int64_t)blockDim.y * U;
    for (int64_t r = r0 + threadIdx.y; r < r1; r += step) {
      float a[U], b[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t rr = r + (int64_t)u * blockDim.y;
        a[u] = rr < r1 ? __ldg(tiles + (rr * 2 + 0) * C + c0) : 0.f;
        b[u] = rr < r1 ? __ldg(tiles + (rr * 2 + 1) * C + c0) : 0.f;
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        s[0] += a[u];
        q[0] += b[u];
      }
    }
  }
  int c;
  double S, Q;
  if (!bn_block_merge<1>(s, q, red, part, C, splits, counter, c, S, Q)) return;
  bn_finish_fwd(c, S, Q, 0.0, M, f);
}

__global__ void bn_eval_coeff_k(const float* __restrict__ gamma, const float* __restrict__ beta,
                                const float* __restrict__ rmean, const float* __restrict__ rvar,
                                float* __restrict__ scale, float* __restrict__ shift, int C,
                                float eps) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float rstd = 1.f / sqrtf(rvar[c] + eps);
  float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
  scale[c] = g * rstd;
  shift[c] = b;
}

// y = [relu](x*scale + shift [+ res]); 2-D thread map (channel vector fixed per thread, so the
// coefficients live in registers), kBnUnroll independent 16-byte loads in flight per thread.
template <typename T, int V, bool RELU, bool ADD, int U, int MINB>
__global__ void __launch_bounds__(kBnThreads, MINB)
bn_apply_k(const T* __restrict__ x, const T* __restrict__ res, const float* __restrict__ scale,
           const float* __restrict__ shift, const float* __restrict__ mean, T* __restrict__ y,
           uint8_t* __restrict__ mask, int64_t M, int C, int64_t rows_per_split) {
  const int cv = blockIdx.x * blockDim.x + threadIdx.x;
  const int c0 = cv * V;
  if (c0 >= C) return;
  // Row blocks are walked from the END of the tensor: x was just written front-to-back by the
  // producing kernel, so its tail is what the 126 MB L2 still holds when this kernel starts.
  const int64_t r0 = (int64_t)(gridDim.y - 1 - blockIdx.y) * rows_per_split;
  int64_t r1 = r0 + rows_per_split;
  if (r1 > M) r1 = M;
  float sc[V], sh[V], mu[V];
#pragma unroll
  for (int j = 0; j < V; ++j) {
    sc[j] = scale[c0 + j];
    sh[j] = shift[c0 + j];
    mu[j] = mean[c0 + j];
  }
  const bool bits_out = RELU && V == 8 && mask != nullptr;
  const int ry = blockDim.y;
  const int64_t stride = (int64_t)ry * C, mstride = stride >> 3;
  int64_t r = r0 + threadIdx.y;
  const int64_t e0 = r * C + c0;
  const T* xp = x + e0;
  const T* rp = ADD ? res + e0 : nullptr;
  T* yp = y + e0;
  uint8_t* mp = bits_out ? mask + (e0 >> 3) : nullptr;
  auto emit = [&](const Pack<T, V>& a, const Pack<T, V>& rs, T* o_ptr, uint8_t* m_ptr) {
    Pack<T, V> o;
    uint32_t bits = 0;
#pragma unroll
    for (int j = 0; j < V; ++j) {
      float v = fmaf(to_f(a.v[j]) - mu[j], sc[j], sh[j]);
      if (ADD) {
        // the reference materialises bn(x) in the storage dtype before `output += identity`
        v = to_f(from_f<T>(v)) + to_f(rs.v[j]);
      }
      if (RELU) v = (v > 0.f || v != v) ? v : 0.f;
      o.v[j] = from_f<T>(v);
      if (RELU && V == 8) bits |= (to_f(o.v[j]) > 0.f ? 1u : 0u) << j;
    }
    st_pack<T, V>(o_ptr, o);
    // 1 bit per element (y > 0): backward reads this instead of y (1/16 of the bytes)
    if (bits_out) *m_ptr = (uint8_t)bits;
  };
  for (; r + (int64_t)(U - 1) * ry < r1; r += (int64_t)U * ry) {
    Pack<T, V> a[U], rs[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      a[u] = ld_pack<T, V>(xp + u * stride);
      if (ADD) rs[u] = ld_pack<T, V>(rp + u * stride);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) emit(a[u], rs[u], yp + u * stride, bits_out ? mp + u * mstride : nullptr);
    xp += U * stride;
    yp += U * stride;
    if (ADD) rp += U * stride;
    if (bits_out) mp += U * mstride;
  }
  for (; r < r1; r += ry) {
    Pack<T, V> a = ld_pack<T, V>(xp), rs;
    if (ADD) rs = ld_pack<T, V>(rp);
    emit(a, rs, yp, mp);
    xp += stride;
    yp += stride;
    if (ADD) rp += stride;
    if (bits_out) mp += mstride;
  }
}

// ---------------------------------------------------------------- backward reduce
// s = sum dy', q = sum dy' * (x - mean)   (dy' = dy masked by y > 0 when ReLU is fused), then
// dbeta = s ; dgamma = rstd * q ; dx = c1 * (dy' - c2 - (x - mean) * c3) with
// c1 = gamma*rstd, c2 = s/M, c3 = rstd^2 * q / M
// U rows of every operand are in flight per thread; MINB resident CTAs per SM bound the registers.
// The row loop walks pointers (one 64-bit add per operand and trip) and runs unguarded while U
// whole rows remain; a one-row tail loop finishes the split.
// RELU: 0 = none, 1 = mask from the forward output y > 0, 2 = 1-bit mask written by the forward pass
template <typename T, int V, int RELU, int U, int MINB>
__global__ void __launch_bounds__(kBnThreads, MINB)
bn_bwd_reduce_k(const T* __restrict__ x, const T* __restrict__ y, const uint8_t* __restrict__ mask,
                const T* __restrict__ dy, const float* __restrict__ mean, float* __restrict__ part,
                int64_t M, int C,
                int64_t rows_per_split, int splits, int* __restrict__ counter, BnBwdFin f) {
  extern __shared__ __align__(16) float red[];
  const int cv = blockIdx.x * blockDim.x + threadIdx.x;
  const int c0 = cv * V;
  // reversed walk: dy was just written front-to-back (its tail is in L2), and the rows this pass
  // touches last — the front — are the ones bn_bwd_dx_k, walking forward, asks for first
  const int64_t r0 = (int64_t)(gridDim.y - 1 - blockIdx.y) * rows_per_split;
  int64_t r1 = r0 + rows_per_split;
  if (r1 > M) r1 = M;
  float s[V], q[V], mu[V];
#pragma unroll
  for (int j = 0; j < V; ++j) s[j] = q[j] = mu[j] = 0.f;
  if (c0 < C) {
#pragma unroll
    for (int j = 0; j < V; ++j) mu[j] = mean[c0 + j];
    constexpr bool bits = RELU == 2 && V == 8;
    const int ry = blockDim.y;
    const int64_t stride = (int64_t)ry * C;
    int64_t r = r0 + threadIdx.y;
    const int64_t e0 = r * C + c0;
    const T* xp = x + e0;
    const T* gp = dy + e0;
    const T* yp = (RELU && !bits) ? y + e0 : nullptr;
    const uint8_t* mp = bits ? mask + (e0 >> 3) : nullptr;
    const int64_t mstride = stride >> 3;
    auto accum = [&](const Pack<T, V>& a, const Pack<T, V>& g, uint32_t mb, const Pack<T, V>& o) {
#pragma unroll
      for (int j = 0; j < V; ++j) {
        float gf = to_f(g.v[j]);
        if (RELU) {
          const bool pos = bits ? ((mb >> j) & 1u) : (to_f(o.v[j]) > 0.f);
          if (!pos) gf = 0.f;
        }
        s[j] += gf;
        q[j] = fmaf(gf, to_f(a.v[j]) - mu[j], q[j]);
      }
    };
    for (; r + (int64_t)(U - 1) * ry < r1; r += (int64_t)U * ry) {
      Pack<T, V> a[U], g[U], o[U];
      uint32_t mb[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        a[u] = ld_pack<T, V>(xp + u * stride);
        g[u] = ld_pack<T, V>(gp + u * stride);
        mb[u] = 0;
        if (RELU) {
          if (bits) mb[u] = mp[u * mstride];
          else o[u] = ld_pack<T, V>(yp + u * stride);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) accum(a[u], g[u], mb[u], o[u]);
      xp += U * stride;
      gp += U * stride;
      if (RELU) {
        if (bits) mp += U * mstride;
        else yp += U * stride;
      }
    }
    for (; r < r1; r += ry) {
      Pack<T, V> a = ld_pack<T, V>(xp), g = ld_pack<T, V>(gp), o;
      uint32_t mb = 0;
      if (RELU) {
        if (bits) mb = *mp;
        else o = ld_pack<T, V>(yp);
      }
      accum(a, g, mb, o);
      xp += stride;
      gp += stride;
      if (RELU) {
        if (bits) mp += mstride;
        else yp += stride;
      }
    }
  }
  int c;
  double S, Q;
  if (!bn_block_merge<V>(s, q, red, part, C, splits, counter, c, S, Q)) return;
  const float rs = f.rstd[c];
  const float g = f.gamma ? f.gamma[c] : 1.f;
  if (f.dbeta) f.dbeta[c] = (float)S;
  if (f.dgamma) f.dgamma[c] = (float)(Q * (double)rs);
  f.c1[c] = g * rs;
  f.c2[c] = (float)(S / (double)M);
  f.c3[c] = (float)(Q * (double)rs * (double)rs / (double)M);
}

template <typename T, int V, int RELU, bool DRES, int U, int MINB>
__global__ void __launch_bounds__(kBnThreads, MINB)
bn_bwd_dx_k(const T* __restrict__ x, const T* __restrict__ y, const uint8_t* __restrict__ mask,
            const T* __restrict__ dy, const float* __restrict__ mean, const float* __restrict__ c1,
            const float* __restrict__ c2, const float* __restrict__ c3, T* __restrict__ dx,
            T* __restrict__ dres, int64_t M, int C, int64_t rows_per_split) {
  const int cv = blockIdx.x * blockDim.x + threadIdx.x;
  const int c0 = cv * V;
  if (c0 >= C) return;
  const int64_t r0 = blockIdx.y * rows_per_split;
  int64_t r1 = r0 + rows_per_split;
  if (r1 > M) r1 = M;
  // dx = c1 * (dy' - c2 - (x - mean) * c3), with x centred first: folding the mean into a constant
  // term (k3*x + k0) cancels catastrophically in fp32 when |mean| is large against the spread
  float k1[V], k2[V], k3[V], mu[V];
#pragma unroll
  for (int j = 0; j < V; ++j) {
    k1[j] = c1[c0 + j];
    k2[j] = c2[c0 + j];
    k3[j] = c3[c0 + j];
    mu[j] = mean[c0 + j];
  }
  constexpr bool bits = RELU == 2 && V == 8;
  const int ry = blockDim.y;
  const int64_t stride = (int64_t)ry * C;
  int64_t r = r0 + threadIdx.y;
  const int64_t e0 = r * C + c0;
  const T* xp = x + e0;
  const T* gp = dy + e0;
  const T* yp = (RELU && !bits) ? y + e0 : nullptr;
  const uint8_t* mp = bits ? mask + (e0 >> 3) : nullptr;
  T* op = dx + e0;
  T* dp = DRES ? dres + e0 : nullptr;
  const int64_t mstride = stride >> 3;
  auto emit = [&](const Pack<T, V>& a, const Pack<T, V>& g, uint32_t mb, const Pack<T, V>& m, T* o_ptr,
                  T* d_ptr) {
    Pack<T, V> o, dr;
#pragma unroll
    for (int j = 0; j < V; ++j) {
      float gf = to_f(g.v[j]);
      if (RELU) {
        const bool pos = bits ? ((mb >> j) & 1u) : (to_f(m.v[j]) > 0.f);
        if (!pos) gf = 0.f;
      }
      if (DRES) dr.v[j] = from_f<T>(gf);
      o.v[j] = from_f<T>(k1[j] * (gf - k2[j] - (to_f(a.v[j]) - mu[j]) * k3[j]));
    }
    st_pack<T, V>(o_ptr, o);
    if (DRES) st_pack<T, V>(d_ptr, dr);
  };
  for (; r + (int64_t)(U - 1) * ry < r1; r += (int64_t)U * ry) {
    Pack<T, V> a[U], g[U], m[U];
    uint32_t mb[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      a[u] = ld_pack<T, V>(xp + u * stride);
      g[u] = ld_pack<T, V>(gp + u * stride);
      mb[u] = 0;
      if (RELU) {
        if (bits) mb[u] = mp[u * mstride];
        else m[u] = ld_pack<T, V>(yp + u * stride);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) emit(a[u], g[u], mb[u], m[u], op + u * stride, DRES ? dp + u * stride : nullptr);
    xp += U * stride;
    gp += U * stride;
    op += U * stride;
    if (DRES) dp += U * stride;
    if (RELU) {
      if (bits) mp += U * mstride;
      else yp += U * stride;
    }
  }
  for (; r < r1; r += ry) {
    Pack<T, V> a = ld_pack<T, V>(xp), g = ld_pack<T, V>(gp), m;
    uint32_t mb = 0;
    if (RELU) {
      if (bits) mb = *mp;
      else m = ld_pack<T, V>(yp);
    }
    emit(a, g, mb, m, op, dp);
    xp += stride;
    gp += stride;
    op += stride;
    if (DRES) dp += stride;
    if (RELU) {
      if (bits) mp += mstride;
      else yp += stride;
    }
  }
}

constexpr int kReduceWaves = 4, kStreamWaves = 8, kReduceCx = 8;

// ---------------------------------------------------------------- backward, single DRAM read
// One kernel for tensors whose x + dy fit the 126 MB L2 (ResNet-50 at batch 256: the eleven
// C = 256 @ 14x14 and five C = 512 @ 7x7 layers, which the two-kernel form runs at 1.1 - 1.9 TB/s
// because launch + last-CTA-merge latency is a third of their 35 - 40 us).  Every CTA reduces its
// rows (phase 1, exactly bn_bwd_reduce_k), meets the other CTAs OF ITS CHANNEL BLOCK at a counter
// (no grid-wide barrier: CTAs of different channel blocks never wait for each other), sums the
// partial rows of the channel block itself — every CTA in the same fixed order, so all of them hold
// bit-identical coefficients — and walks its rows again for dx (phase 2): the second read of x, dy
// and the mask is served by L2.  Launched cooperatively (all CTAs co-resident by construction).
// counter layout: [0, kCountersPerSlot/2) arrivals, [kCountersPerSlot/2, ..) departures.
template <typename T, int V, int RELU, bool DRES, int U, int MINB>
__global__ void __launch_bounds__(kBnThreads, MINB)
bn_bwd_fused_k(const T* __restrict__ x, const T* __restrict__ y, const uint8_t* __restrict__ mask,
               const T* __restrict__ dy, const float* __restrict__ mean, float* __restrict__ part,
               T* __restrict__ dx, T* __restrict__ dres, int64_t M, int C, int64_t rows_per_split,
               int splits, int* __restrict__ counter, BnBwdFin f) {
  extern __shared__ __align__(16) float red[];
  __shared__ float coef[3][kBnThreads];   // c1, c2, c3 of this CTA's channels (W <= 256)
  const int cv = blockIdx.x * blockDim.x + threadIdx.x;
  const int c0 = cv * V;
  const int t = threadIdx.y * blockDim.x + threadIdx.x;
  const int W = blockDim.x * V;
  const int64_t r0 = blockIdx.y * rows_per_split;
  int64_t r1 = r0 + rows_per_split;
  if (r1 > M) r1 = M;
  constexpr bool bits = RELU == 2 && V == 8;
  const int ry = blockDim.y;
  const int64_t stride = (int64_t)ry * C, mstride = stride >> 3;
  const int64_t e0 = (r0 + threadIdx.y) * C + c0;
  float s[V], q[V], mu[V];
#pragma unroll
  for (int j = 0; j < V; ++j) s[j] = q[j] = mu[j] = 0.f;
  // ---- phase 1: partial sums of dy' and dy' * (x - mean) over this CTA's rows
  if (c0 < C) {
#pragma unroll
    for (int j = 0; j < V; ++j) mu[j] = mean[c0 + j];
    int64_t r = r0 + threadIdx.y;
    const T* xp = x + e0;
    const T* gp = dy + e0;
    const T* yp = (RELU && !bits) ? y + e0 : nullptr;
    const uint8_t* mp = bits ? mask + (e0 >> 3) : nullptr;
    auto accum = [&](const Pack<T, V>& a, const Pack<T, V>& g, uint32_t mb, const Pack<T, V>& o) {
#pragma unroll
      for (int j = 0; j < V; ++j) {
        float gf = to_f(g.v[j]);
        if (RELU) {
          const bool pos = bits ? ((mb >> j) & 1u) : (to_f(o.v[j]) > 0.f);
          if (!pos) gf = 0.f;
        }
        s[j] += gf;
        q[j] = fmaf(gf, to_f(a.v[j]) - mu[j], q[j]);
      }
    };
    for (; r + (int64_t)(U - 1) * ry < r1; r += (int64_t)U * ry) {
      Pack<T, V> a[U], g[U], o[U];
      uint32_t mb[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        a[u] = ld_pack<T, V>(xp + u * stride);
        g[u] = ld_pack<T, V>(gp + u * stride);
        mb[u] = 0;
        if (RELU) {
          if (bits) mb[u] = mp[u * mstride];
          else o[u] = ld_pack<T, V>(yp + u * stride);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) accum(a[u], g[u], mb[u], o[u]);
      xp += U * stride;
      gp += U * stride;
      if (RELU) {
        if (bits) mp += U * mstride;
        else yp += U * stride;
      }
    }
    for (; r < r1; r += ry) {
      Pack<T, V> a = ld_pack<T, V>(xp), g = ld_pack<T, V>(gp), o;
      uint32_t mb = 0;
      if (RELU) {
        if (bits) mb = *mp;
        else o = ld_pack<T, V>(yp);
      }
      accum(a, g, mb, o);
      xp += stride;
      gp += stride;
      if (RELU) {
        if (bits) mp += mstride;
        else yp += stride;
      }
    }
  }
  bn_block_partial<V>(s, q, red, part, C);
  // ---- meet the other CTAs of this channel block
  __threadfence();
  __syncthreads();
  if (t == 0) {
    atomicAdd(counter + blockIdx.x, 1);
    int seen;
    do {
      asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(counter + blockIdx.x) : "memory");
      if (seen < splits) __nanosleep(32);
    } while (seen < splits);
  }
  __syncthreads();
  __threadfence();
  int c;
  double S, Q;
  if (bn_merge_partials<V>(red, part, C, splits, c, S, Q)) {
    const float rs = f.rstd[c];
    const float g = f.gamma ? f.gamma[c] : 1.f;
    const int col = c - blockIdx.x * W;
    coef[0][col] = g * rs;
    coef[1][col] = (float)(S / (double)M);
    coef[2][col] = (float)(Q * (double)rs * (double)rs / (double)M);
    if (blockIdx.y == 0) {
      if (f.dbeta) f.dbeta[c] = (float)S;
      if (f.dgamma) f.dgamma[c] = (float)(Q * (double)rs);
    }
  }
  __syncthreads();
  if (t == 0) {   // the last CTA to leave the barrier hands the counters back clean
    const int half = kCountersPerSlot / 2;
    const int prev = atomicAdd(counter + half + blockIdx.x, 1);
    if (prev == splits - 1) {
      counter[blockIdx.x] = 0;
      counter[half + blockIdx.x] = 0;
    }
  }
  if (c0 >= C) return;
  // ---- phase 2: dx = c1 * (dy' - c2 - (x - mean) * c3) over the same rows (L2-resident by now)
  float k1[V], k2[V], k3[V];
#pragma unroll
  for (int j = 0; j < V; ++j) {
    k1[j] = coef[0][threadIdx.x * V + j];
    k2[j] = coef[1][threadIdx.x * V + j];
    k3[j] = coef[2][threadIdx.x * V + j];
  }
  {
    int64_t r = r0 + threadIdx.y;
    const T* xp = x + e0;
    const T* gp = dy + e0;
    const T* yp = (RELU && !bits) ? y + e0 : nullptr;
    const uint8_t* mp = bits ? mask + (e0 >> 3) : nullptr;
    T* op = dx + e0;
    T* dp = DRES ? dres + e0 : nullptr;
    auto emit = [&](const Pack<T, V>& a, const Pack<T, V>& g, uint32_t mb, const Pack<T, V>& m, T* o_ptr,
                    T* d_ptr) {
      Pack<T, V> o, dr;
#pragma unroll
      for (int j = 0; j < V; ++j) {
        float gf = to_f(g.v[j]);
        if (RELU) {
          const bool pos = bits ? ((mb >> j) & 1u) : (to_f(m.v[j]) > 0.f);
          if (!pos) gf = 0.f;
        }
        if (DRES) dr.v[j] = from_f<T>(gf);
        o.v[j] = from_f<T>(k1[j] * (gf - k2[j] - (to_f(a.v[j]) - mu[j]) * k3[j]));
      }
      st_pack<T, V>(o_ptr, o);
      if (DRES) st_pack<T, V>(d_ptr, dr);
    };
    for (; r + (int64_t)(U - 1) * ry < r1; r += (int64_t)U * ry) {
      Pack<T, V> a[U], g[U], m[U];
      uint32_t mb[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        a[u] = ld_pack<T, V>(xp + u * stride);
        g[u] = ld_pack<T, V>(gp + u * stride);
        mb[u] = 0;
        if (RELU) {
          if (bits) mb[u] = mp[u * mstride];
          else m[u] = ld_pack<T, V>(yp + u * stride);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) emit(a[u], g[u], mb[u], m[u], op + u * stride, DRES ? dp + u * stride : nullptr);
      xp += U * stride;
      gp += U * stride;
      op += U * stride;
      if (DRES) dp += U * stride;
      if (RELU) {
        if (bits) mp += U * mstride;
        else yp += U * stride;
      }
    }
    for (; r < r1; r += ry) {
      Pack<T, V> a = ld_pack<T, V>(xp), g = ld_pack<T, V>(gp), m;
      uint32_t mb = 0;
      if (RELU) {
        if (bits) mb = *mp;
        else m = ld_pack<T, V>(yp);
      }
      emit(a, g, mb, m, op, dp);
      xp += stride;
      gp += stride;
      op += stride;
      if (DRES) dp += stride;
      if (RELU) {
        if (bits) mp += mstride;
        else yp += stride;
      }
    }
  }
}

// ================================================================ concat-aware BatchNorm (DenseNet)
// A dense block keeps its feature maps in ONE [M, C_total] buffer; layer l normalises the channel
// prefix [0, C_l) of it (reference src/densenet.py:117,131-140: BatchNorm2d over torch.cat([x, out])).
// Two facts make that cheap: (1) the batch statistics of a channel do not depend on which layer
// looks at it, so they are computed once — from the producing convolution's epilogue — and every
// later layer only applies its own gamma / beta; (2) the input gradient of layer l's BatchNorm is a
// contribution to the gradient of the shared buffer, so the backward pass accumulates in place.
// These kernels are the row-strided forms of bn_apply_k / bn_bwd_reduce_k / bn_bwd_dx_k (x and dx
// have a row stride != C); the dense layers keep their tuned kernels.
__global__ void bn_prefix_coeff_k(const float* __restrict__ gamma, const float* __restrict__ beta,
                                  const float* __restrict__ mean, const float* __restrict__ rstd,
                                  float* __restrict__ rmean, float* __restrict__ rvar,
                                  long long* __restrict__ nbt, float* __restrict__ scale,
                                  float* __restrict__ shift, int C, int64_t M, float eps, float momentum) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float rs = rstd[c];
  scale[c] = (gamma ? gamma[c] : 1.f) * rs;
  shift[c] = beta ? beta[c] : 0.f;
  if (nbt && c == 0) *nbt += 1;
  if (rmean) rmean[c] = (1.f - momentum) * rmean[c] + momentum * mean[c];
  if (rvar) {
    double var = 1.0 / ((double)rs * (double)rs) - (double)eps;   // biased batch variance
    if (var < 0.0) var = 0.0;
    const double unbiased = M > 1 ? var * (double)M / (double)(M - 1) : var;
    rvar[c] = (1.f - momentum) * rvar[c] + momentum * (float)unbiased;
  }
}

template <typename T, int V, bool RELU>
__global__ void __launch_bounds__(kBnThreads)
bn_apply_ld_k(const T* __restrict__ x, int64_t ldx, const float* __restrict__ scale,
              const float* __restrict__ shift, const float* __restrict__ mean, T* __restrict__ y,
              uint8_t* __restrict__ mask, int64_t M, int C, int64_t rows_per_split) {
  const int cv = blockIdx.x * blockDim.x + threadIdx.x;
  const int c0 = cv * V;
  if (c0 >= C) return;
  const int64_t r0 = blockIdx.y * rows_per_split;
  int64_t r1 = r0 + rows_per_split;
  if (r1 > M) r1 = M;
  float sc[V], sh[V], mu[V];
#pragma unroll
  for (int j = 0; j < V; ++j) { sc[j] = scale[c0 + j]; sh[j] = shift[c0 + j]; mu[j] = mean[c0 + j]; }
  const bool bits_out = RELU && V == 8 && mask != nullptr;
  const int ry = blockDim.y;
  auto one = [&](int64_t r, const Pack<T, V>& a) {
    Pack<T, V> o;
    uint32_t bits = 0;
#pragma unroll
    for (int j = 0; j < V; ++j) {
      float v = fmaf(to_f(a.v[j]) - mu[j], sc[j], sh[j]);
      if (RELU) v = (v > 0.f || v != v) ? v : 0.f;
      o.v[j] = from_f<T>(v);
      if (RELU && V == 8) bits |= (to_f(o.v[j]) > 0.f ? 1u : 0u) << j;
    }
    st_pack<T, V>(y + r * C + c0, o);
    if (bits_out) mask[(r * C + c0) >> 3] = (uint8_t)bits;
  };
  int64_t r = r0 + threadIdx.y;
  for (; r + 3 * (int64_t)ry < r1; r += 4 * (int64_t)ry) {
    Pack<T, V> a[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) a[u] = ld_pack<T, V>(x + (r + (int64_t)u * ry) * ldx + c0);
#pragma unroll
    for (int u = 0; u < 4; ++u) one(r + (int64_t)u * ry, a[u]);
  }
  for (; r < r1; r += ry) one(r, ld_pack<T, V>(x + r * ldx + c0));
}

// RELU: 0 = none, 1 = mask from y > 0, 2 = 1-bit mask
template <typename T, int V, int RELU>
__global__ void __launch_bounds__(kBnThreads)
bn_bwd_reduce_ld_k(const T* __restrict__ x, int64_t ldx, const T* __restrict__ y,
                   const uint8_t* __restrict__ mask, const T* __restrict__ dy,
                   const float* __restrict__ mean, float* __restrict__ part, int64_t M, int C,
                   int64_t rows_per_split, int splits, int* __restrict__ counter, BnBwdFin f) {
  extern __shared__ __align__(16) float red[];
  const int cv = blockIdx.x * blockDim.x + threadIdx.x;
  const int c0 = cv * V;
  const int64_t r0 = blockIdx.y * rows_per_split;
  int64_t r1 = r0 + rows_per_split;
  if (r1 > M) r1 = M;
  float s[V], q[V], mu[V];
#pragma unroll
  for (int j = 0; j < V; ++j) s[j] = q[j] = mu[j] = 0.f;
  if (c0 < C) {
#pragma unroll
    for (int j = 0; j < V; ++j) mu[j] = mean[c0 + j];
    constexpr bool bits = RELU == 2 && V == 8;
    const int ry = blockDim.y;
    auto one = [&](int64_t r, const Pack<T, V>& a, const Pack<T, V>& g) {
      uint32_t mb = 0;
      Pack<T, V> o;
      if (RELU) {
        if (bits) mb = mask[(r * C + c0) >> 3];
        else o = ld_pack<T, V>(y + r * C + c0);
      }
#pragma unroll
      for (int j = 0; j < V; ++j) {
        float gf = to_f(g.v[j]);
        if (RELU) {
          const bool pos = bits ? ((mb >> j) & 1u) : (to_f(o.v[j]) > 0.f);
          if (!pos) gf = 0.f;
        }
        s[j] += gf;
        q[j] = fmaf(gf, to_f(a.v[j]) - mu[j], q[j]);
      }
    };
    int64_t r = r0 + threadIdx.y;
    for (; r + (int64_t)ry < r1; r += 2 * (int64_t)ry) {
      const Pack<T, V> a0 = ld_pack<T, V>(x + r * ldx + c0), a1 = ld_pack<T, V>(x + (r + ry) * ldx + c0);
      const Pack<T, V> g0 = ld_pack<T, V>(dy + r * C + c0), g1 = ld_pack<T, V>(dy + (r + ry) * C + c0);
      one(r, a0, g0);
      one(r + ry, a1, g1);
    }
    for (; r < r1; r += ry) one(r, ld_pack<T, V>(x + r * ldx + c0), ld_pack<T, V>(dy + r * C + c0));
  }
  int c;
  double S, Q;
  if (!bn_block_merge<V>(s, q, red, part, C, splits, counter, c, S, Q)) return;
  const float rs = f.rstd[c];
  const float g = f.gamma ? f.gamma[c] : 1.f;
  if (f.dbeta) f.dbeta[c] = (float)S;
  if (f.dgamma) f.dgamma[c] = (float)(Q * (double)rs);
  f.c1[c] = g * rs;
  f.c2[c] = (float)(S / (double)M);
  f.c3[c] = (float)(Q * (double)rs * (double)rs / (double)M);
}

// dx_acc[r * lddx + c] += c1 * (dy' - c2 - (x - mean) * c3): the sum is formed in fp32 from the stored
// accumulator and rounded once
template <typename T, int V, int RELU>
__global__ void __launch_bounds__(kBnThreads)
bn_bwd_dx_ld_k(const T* __restrict__ x, int64_t ldx, const T* __restrict__ y,
               const uint8_t* __restrict__ mask, const T* __restrict__ dy,
               const float* __restrict__ mean, const float* __restrict__ c1,
               const float* __restrict__ c2, const float* __restrict__ c3, T* __restrict__ dx_acc,
               int64_t lddx, int64_t M, int C, int64_t rows_per_split) {
  const int cv = blockIdx.x * blockDim.x + threadIdx.x;
  const int c0 = cv * V;
  if (c0 >= C) return;
  const int64_t r0 = blockIdx.y * rows_per_split;
  int64_t r1 = r0 + rows_per_split;
  if (r1 > M) r1 = M;
  float k1[V], k2[V], k3[V], mu[V];
#pragma unroll
  for (int j = 0; j < V; ++j) { k1[j] = c1[c0 + j]; k2[j] = c2[c0 + j]; k3[j] = c3[c0 + j]; mu[j] = mean[c0 + j]; }
  constexpr bool bits = RELU == 2 && V == 8;
  const int ry = blockDim.y;
  auto one = [&](int64_t r, const Pack<T, V>& a, const Pack<T, V>& g, const Pack<T, V>& acc) {
    uint32_t mb = 0;
    Pack<T, V> o, out;
    if (RELU) {
      if (bits) mb = mask[(r * C + c0) >> 3];
      else o = ld_pack<T, V>(y + r * C + c0);
    }
#pragma unroll
    for (int j = 0; j < V; ++j) {
      float gf = to_f(g.v[j]);
      if (RELU) {
        const bool pos = bits ? ((mb >> j) & 1u) : (to_f(o.v[j]) > 0.f);
        if (!pos) gf = 0.f;
      }
      out.v[j] = from_f<T>(to_f(acc.v[j]) + k1[j] * (gf - k2[j] - (to_f(a.v[j]) - mu[j]) * k3[j]));
    }
    st_pack<T, V>(dx_acc + r * lddx + c0, out);
  };
  int64_t r = r0 + threadIdx.y;
  for (; r + (int64_t)ry < r1; r += 2 * (int64_t)ry) {
    const Pack<T, V> a0 = ld_pack<T, V>(x + r * ldx + c0), a1 = ld_pack<T, V>(x + (r + ry) * ldx + c0);
    const Pack<T, V> g0 = ld_pack<T, V>(dy + r * C + c0), g1 = ld_pack<T, V>(dy + (r + ry) * C + c0);
    const Pack<T, V> d0 = ld_pack<T, V>(dx_acc + r * lddx + c0), d1 = ld_pack<T, V>(dx_acc + (r + ry) * lddx + c0);
    one(r, a0, g0, d0);
    one(r + ry, a1, g1, d1);
  }
  for (; r < r1; r += ry)
    one(r, ld_pack<T, V>(x + r * ldx + c0), ld_pack<T, V>(dy + r * C + c0), ld_pack<T, V>(dx_acc + r * lddx + c0));
}

// Largest x + dy footprint (bytes) the single-read backward takes; 0 disables it.
// CNB_BN_BWD_FUSED_MB overrides (sweep: tools/gpu_bn_fused.sh).
static int64_t bn_bwd_fused_limit() {
  const char* e = getenv("CNB_BN_BWD_FUSED_MB");   // read per call: the tests switch paths in-process
  return (int64_t)(e ? atoll(e) : 64) << 20;
}

template <int RELU, bool DRES>
static int launch_bn_bwd_fused(const bf16* x, const bf16* y, const uint8_t* mask, const bf16* dy,
                               const float* mean, float* part, bf16* dx, bf16* dres, int64_t M, int C,
                               const BnBwdFin& fin, bool* taken, cudaStream_t st) {
  constexpr int U = 2, MINB = 3;
  auto kern = bn_bwd_fused_k<bf16, 8, RELU, DRES, U, MINB>;
  static int per_sm = -1;
  const int cvs = C / 8;
  int cx = 1;
  while (cx * 2 <= cvs && cx * 2 <= kReduceCx) cx *= 2;
  const int ry = kBnThreads / cx;
  size_t dyn = (size_t)ry * 2 * cx * 8 * sizeof(float);   // [row lanes][2][channels of the CTA] floats
  if (dyn < 4096) dyn = 4096;                              // the merge re-uses it for 256 x 2 doubles
  if (per_sm < 0) {
    int n = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, kBnThreads, 16384);
    CNB_REQUIRE(e == cudaSuccess, CNB_ERR_CUDA, "cnb_bn_bwd: occupancy query failed: %s", cudaGetErrorString(e));
    per_sm = n < MINB ? n : MINB;
  }
  *taken = false;
  if (per_sm < 1) return CNB_OK;
  const int cblocks = (int)ceil_div(cvs, cx);
  if (cblocks > kCountersPerSlot / 2) return CNB_OK;
  int64_t splits = (int64_t)sm_count() * per_sm / cblocks;
  const int64_t max_splits = ceil_div(M, (int64_t)ry * U);
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) return CNB_OK;   // more channel blocks than co-resident CTAs: two-kernel form
  const int64_t rows_per_split = ceil_div(M, splits);
  const int nsplit = (int)ceil_div(M, rows_per_split);
  int* counter = bn_counters(cblocks);
  CNB_REQUIRE(counter, CNB_ERR_CUDA, "cnb_bn_bwd: cannot allocate the arrival counters");
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(cblocks, nsplit);
  cfg.blockDim = dim3(cx, ry);
  cfg.dynamicSmemBytes = dyn;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, x, y, mask, dy, mean, part, dx, dres, M, C,
                                     rows_per_split, nsplit, counter, fin);
  CNB_REQUIRE(e == cudaSuccess, CNB_ERR_CUDA, "cnb_bn_bwd(fused): %s", cudaGetErrorString(e));
  *taken = true;
  return check_launch("bn_bwd_fused");
}


// Loop shape of the bf16 backward kernels: (rows in flight per thread, resident CTAs per SM).
// 0: (2, 2)  1: (2, 3)  2: (4, 2)  3: (3, 2)  4: (4, 3).  CNB_BN_BWD_VARIANT overrides the default.
// Loop shape of the bf16 apply kernel: 0: (4, 2)  1: (2, 3)  2: (4, 3)  3: (2, 4).
static int bn_fwd_variant() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("CNB_BN_FWD_VARIANT");
    v = e ? atoi(e) : 2;   // measured on ResNet-50 (profiles/r01v): 3.73 / 3.59 / 3.40 / 3.48 ms per step
    if (v < 0 || v > 3) v = 0;
  }
  return v;
}
// bf16 statistics kernel: 0 = bn_stats_k (validated), 1 = walk (4 rows, 4 CTAs/SM), 2 = walk (8, 3)
static int bn_stats_variant() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("CNB_BN_STATS_VARIANT");
    v = e ? atoi(e) : 0;
    if (v < 0 || v > 2) v = 0;
  }
  return v;
}
static int apply_waves(bool hot, int variant) {
  if (hot && (variant == 1 || variant == 2)) return 6;   // 3 resident CTAs per SM
  return kStreamWaves;                                    // 2 or 4 resident CTAs per SM
}
static int bn_bwd_variant() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("CNB_BN_BWD_VARIANT");
    v = e ? atoi(e) : 1;   // measured on ResNet-50 (profiles/r01u): 7.07 / 6.10 / 6.49 / - / 6.15 ms per step
    if (v < 0 || v > 4) v = 0;
  }
  return v;
}
static size_t bn_ws_bytes(int C) {
  // partials: splits <= min(2048, kReduceWaves*SMs) rows of 2*C floats, then 5*C coefficients
  int64_t splits = (int64_t)sm_count() * kReduceWaves;
  if (splits > 2048) splits = 2048;
  return (size_t)(2 * (int64_t)C * splits + 5 * (int64_t)C) * sizeof(float) + 256;
}

static int validate_bn(const cnb_bn_desc* d, const char* who) {
  CNB_REQUIRE(d != nullptr, CNB_ERR_INVALID, "%s: null descriptor", who);
  CNB_REQUIRE(d->M > 0 && d->C > 0 && valid_dtype(d->dtype), CNB_ERR_INVALID,
              "%s: bad descriptor (M=%lld C=%d dtype=%d)", who, (long long)d->M, d->C, d->dtype);
  return CNB_OK;
}

}  // namespace cnb

using namespace cnb;

// second half of BatchNorm backward: dx (+ d_residual) from the per-channel coefficients c1..c3
static int launch_bn_bwd_dx(const cnb_bn_desc* d, const void* x, const void* y, const uint8_t* mask,
                            const void* dy, const float* save_mean, const float* c1, const float* c2,
                            const float* c3, void* dx, void* d_residual, int v, bool hot, int variant,
                            int mode, bool three, cudaStream_t st) {
  const int C = d->C;
  const int64_t M = d->M;
  BnGeom ga = bn_geom(M, C, v, three ? 6 : kStreamWaves);
  dim3 agrid(ga.cblocks, ga.splits), ablock(ga.cx, ga.ry);
  const int64_t arows = ga.rows_per_split;
#define DX_ARGS (const T*)x, (const T*)y, mask, (const T*)dy, save_mean, c1, c2, c3, (T*)dx, \
                (T*)d_residual, M, C, arows
#define DX(R, D) \
  CNB_DISPATCH_TV(d->dtype, v, (bn_bwd_dx_k<T, V, R, D, 2, 2><<<agrid, ablock, 0, st>>>(DX_ARGS)))
#define DX_HOT(UU, MB)                                                                            \
  do {                                                                                            \
    using T = bf16;                                                                               \
    if (mode == 2 && d_residual) bn_bwd_dx_k<T, 8, 2, true, UU, MB><<<agrid, ablock, 0, st>>>(DX_ARGS); \
    else if (mode == 2) bn_bwd_dx_k<T, 8, 2, false, UU, MB><<<agrid, ablock, 0, st>>>(DX_ARGS);   \
    else if (mode == 1 && d_residual) bn_bwd_dx_k<T, 8, 1, true, UU, MB><<<agrid, ablock, 0, st>>>(DX_ARGS); \
    else if (mode == 1) bn_bwd_dx_k<T, 8, 1, false, UU, MB><<<agrid, ablock, 0, st>>>(DX_ARGS);   \
    else if (d_residual) bn_bwd_dx_k<T, 8, 0, true, UU, MB><<<agrid, ablock, 0, st>>>(DX_ARGS);   \
    else bn_bwd_dx_k<T, 8, 0, false, UU, MB><<<agrid, ablock, 0, st>>>(DX_ARGS);                  \
  } while (0)
  if (hot && variant == 1) DX_HOT(2, 3);
  else if (hot && variant == 2) DX_HOT(4, 2);
  else if (hot && variant == 3) DX_HOT(3, 2);
  else if (hot && variant == 4) DX_HOT(4, 3);
  else if (mode == 2 && d_residual) DX(2, true);
  else if (mode == 2) DX(2, false);
  else if (mode == 1 && d_residual) DX(1, true);
  else if (mode == 1) DX(1, false);
  else if (d_residual) DX(0, true);
  else DX(0, false);
#undef DX_HOT
#undef DX
#undef DX_ARGS
  return check_launch("bn_bwd_dx");
}

// Backward reduction from per-CTA partial rows written by the epilogue of the data-grad that
// produced dy (conv_tc.cu, bn_x != nullptr): rows[n][2][C] = (sum dy', sum dy' * (x - mean)).  Same
// fixed-order merge as bn_partials_k, finish of bn_bwd_reduce_k.
__global__ void __launch_bounds__(kBnThreads)
bn_bwd_partials_k(const float* __restrict__ rows, float* __restrict__ part, int64_t nrows, int64_t M,
                  int C, int64_t rows_per_split, int splits, int* __restrict__ counter, BnBwdFin f) {
  extern __shared__ __align__(16) float red[];
  const int c0 = blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t r0 = blockIdx.y * rows_per_split;
  int64_t r1 = r0 + rows_per_split;
  if (r1 > nrows) r1 = nrows;
  float s[1] = {0.f}, q[1] = {0.f};
  if (c0 < C) {
    for (int64_t r = r0 + threadIdx.y; r < r1; r += blockDim.y) {
      s[0] += __ldg(rows + (r * 2 + 0) * C + c0);
      q[0] += __ldg(rows + (r * 2 + 1) * C + c0);
    }
  }
  int c;
  double S, Q;
  if (!bn_block_merge<1>(s, q, red, part, C, splits, counter, c, S, Q)) return;
  const float rs = f.rstd[c];
  const float g = f.gamma ? f.gamma[c] : 1.f;
  if (f.dbeta) f.dbeta[c] = (float)S;
  if (f.dgamma) f.dgamma[c] = (float)(Q * (double)rs);
  f.c1[c] = g * rs;
  f.c2[c] = (float)(S / (double)M);
  f.c3[c] = (float)(Q * (double)rs * (double)rs / (double)M);
}

extern "C" {

size_t cnb_bn_workspace_bytes(const cnb_bn_desc* d) {
  if (!d || d->C <= 0) return 0;
  return bn_ws_bytes(d->C);
}

int cnb_bn_fwd_train(const cnb_bn_desc* d, const void* x, const void* residual, const float* gamma,
                     const float* beta, float* running_mean, float* running_var, void* y,
                     float* save_mean, float* save_rstd, const float* conv_partials,
                     int32_t n_partials, void* relu_mask, void* workspace, size_t workspace_bytes,
                     void* stream) {
  CNB_TRY(validate_bn(d, "cnb_bn_fwd_train"));
  CNB_REQUIRE(!conv_partials || n_partials > 0, CNB_ERR_INVALID,
              "cnb_bn_fwd_train: conv_partials given with n_partials=%d", n_partials);
  CNB_REQUIRE(!relu_mask || ((d->flags & CNB_BN_RELU) && d->dtype == CNB_BF16 && d->C % 8 == 0),
              CNB_ERR_INVALID,
              "cnb_bn_fwd_train: relu_mask needs CNB_BN_RELU, bf16 and C %% 8 == 0 (C=%d)", d->C);
  CNB_REQUIRE(x && y && save_mean && save_rstd && workspace, CNB_ERR_INVALID,
              "cnb_bn_fwd_train: null pointer");
  const bool add = d->flags & CNB_BN_ADD, relu = d->flags & CNB_BN_RELU;
  CNB_REQUIRE(!add || residual, CNB_ERR_INVALID, "cnb_bn_fwd_train: CNB_BN_ADD without residual");
  CNB_REQUIRE(d->M > 1, CNB_ERR_INVALID,
              "cnb_bn_fwd_train: expected more than 1 value per channel when training (M=%lld)",
              (long long)d->M);
  CNB_REQUIRE(workspace_bytes >= bn_ws_bytes(d->C), CNB_ERR_WORKSPACE,
              "cnb_bn_fwd_train: workspace %zu < %zu", workspace_bytes, bn_ws_bytes(d->C));
  cudaStream_t st = (cudaStream_t)stream;
  const int C = d->C;
  const int64_t M = d->M;
  int v = pick_vec(C, dtype_size(d->dtype), x, y, add ? residual : nullptr);
  CNB_REQUIRE(!relu_mask || v == 8, CNB_ERR_INVALID,
              "cnb_bn_fwd_train: relu_mask needs 16-byte aligned tensors");
  BnFwdFin fin{gamma, beta, running_mean, running_var, save_mean, save_rstd, nullptr, nullptr,
               d->eps, d->momentum, (long long*)d->num_batches_tracked};
  float* part = (float*)workspace;
  float* scale;
  if (conv_partials) {
    // statistics were accumulated by the producing convolution's epilogue: merge its per-tile rows
    BnGeom g;
    // <= 512 rows (one per persistent conv CTA): 16 channels x 16 row lanes per CTA keeps the
    // serial chain of dependent L2 loads per thread to ~5 trips and spreads C over C/16 CTAs
    g.V = 1; g.cx = n_partials <= 512 ? 16 : 64; g.ry = kBnThreads / g.cx;
    g.cblocks = (int)ceil_div(C, g.cx);
    int64_t splits = ceil_div((int64_t)sm_count() * 2, g.cblocks);
    int64_t max_splits = ceil_div((int64_t)n_partials, (int64_t)g.ry * 4);
    if (splits > max_splits) splits = max_splits;
    // one row per persistent conv CTA (<= SM count): a single CTA per 64 channels sums them
    if (n_partials <= 512) splits = 1;
    if (splits < 1) splits = 1;
    g.rows_per_split = ceil_div((int64_t)n_partials, splits);
    g.splits = (int)ceil_div((int64_t)n_partials, g.rows_per_split);
    scale = part + (size_t)2 * C * g.splits;
    fin.scale = scale; fin.shift = scale + C;
    int* counter = bn_counters(g.cblocks);
    CNB_REQUIRE(counter, CNB_ERR_CUDA, "cnb_bn_fwd_train: cannot allocate the arrival counters");
    dim3 grid(g.cblocks, g.splits), block(g.cx, g.ry);
    bn_partials_k<<<grid, block, 4096, st>>>(conv_partials, part, n_partials, M, C, g.rows_per_split,
                                             g.splits, counter, fin);
    CNB_TRY(check_launch("bn_partials"));
  } else {
    const int svariant = (d->dtype == CNB_BF16 && v == 8) ? bn_stats_variant() : 0;
    BnGeom g = bn_geom(M, C, v, svariant == 2 ? 3 : kReduceWaves, kReduceCx);
    scale = part + (size_t)2 * C * g.splits;
    fin.scale = scale; fin.shift = scale + C;
    int* counter = bn_counters(g.cblocks);
    CNB_REQUIRE(counter, CNB_ERR_CUDA, "cnb_bn_fwd_train: cannot allocate the arrival counters");
    dim3 grid(g.cblocks, g.splits), block(g.cx, g.ry);
    size_t smem = (size_t)g.ry * 2 * g.cx * v * sizeof(float);
    if (smem < 4096) smem = 4096;  // the merge step re-uses it for 256 x 2 doubles
    if (svariant == 1)
      bn_stats_walk_k<bf16, 8, 4, 4><<<grid, block, smem, st>>>((const bf16*)x, part, M, C, g.rows_per_split,
                                                                g.splits, counter, fin);
    else if (svariant == 2)
      bn_stats_walk_k<bf16, 8, 8, 3><<<grid, block, smem, st>>>((const bf16*)x, part, M, C, g.rows_per_split,
                                                                g.splits, counter, fin);
    else
      CNB_DISPATCH_TV(d->dtype, v, (bn_stats_k<T, V><<<grid, block, smem, st>>>(
                                       (const T*)x, part, M, C, g.rows_per_split, g.splits, counter,
                                       fin)));
    CNB_TRY(check_launch("bn_stats"));
  }
  float* shift = scale + C;
  const float* mean_ptr = save_mean;
  uint8_t* mask = (uint8_t*)relu_mask;
  const bool ahot = d->dtype == CNB_BF16 && v == 8;
  const int avariant = bn_fwd_variant();
  BnGeom ga = bn_geom(M, C, v, apply_waves(ahot, avariant));
  dim3 agrid(ga.cblocks, ga.splits), ablock(ga.cx, ga.ry);
  const int64_t arows = ga.rows_per_split;
#define APPLY_ARGS (const T*)x, (const T*)residual, scale, shift, mean_ptr, (T*)y, mask, d->M, C, arows
#define APPLY(R, A) \
  CNB_DISPATCH_TV(d->dtype, v, (bn_apply_k<T, V, R, A, 4, 2><<<agrid, ablock, 0, st>>>(APPLY_ARGS)))
#define APPLY_HOT(UU, MB)                                                                          \
  do {                                                                                             \
    using T = bf16;                                                                                \
    if (relu && add) bn_apply_k<T, 8, true, true, UU, MB><<<agrid, ablock, 0, st>>>(APPLY_ARGS);   \
    else if (relu) bn_apply_k<T, 8, true, false, UU, MB><<<agrid, ablock, 0, st>>>(APPLY_ARGS);    \
    else if (add) bn_apply_k<T, 8, false, true, UU, MB><<<agrid, ablock, 0, st>>>(APPLY_ARGS);     \
    else bn_apply_k<T, 8, false, false, UU, MB><<<agrid, ablock, 0, st>>>(APPLY_ARGS);             \
  } while (0)
#define APPLY_ANY()                                   \
  do {                                                \
    if (ahot && avariant == 1) APPLY_HOT(2, 3);       \
    else if (ahot && avariant == 2) APPLY_HOT(4, 3);  \
    else if (ahot && avariant == 3) APPLY_HOT(2, 4);  \
    else if (relu && add) APPLY(true, true);          \
    else if (relu) APPLY(true, false);                \
    else if (add) APPLY(false, true);                 \
    else APPLY(false, false);                         \
  } while (0)
  APPLY_ANY();
  return check_launch("bn_apply");
}

int cnb_bn_stats_from_partials(const cnb_bn_desc* d, const float* conv_partials, int32_t n_partials,
                               float* mean, float* rstd, void* workspace, size_t workspace_bytes,
                               void* stream) {
  CNB_TRY(validate_bn(d, "cnb_bn_stats_from_partials"));
  CNB_REQUIRE(conv_partials && n_partials > 0 && mean && rstd && workspace, CNB_ERR_INVALID,
              "cnb_bn_stats_from_partials: null pointer / no rows");
  CNB_REQUIRE(workspace_bytes >= bn_ws_bytes(d->C), CNB_ERR_WORKSPACE,
              "cnb_bn_stats_from_partials: workspace %zu < %zu", workspace_bytes, bn_ws_bytes(d->C));
  cudaStream_t st = (cudaStream_t)stream;
  const int C = d->C;
  BnFwdFin fin{nullptr, nullptr, nullptr, nullptr, mean, rstd, nullptr, nullptr, d->eps, d->momentum, nullptr};
  float* part = (float*)workspace;
  BnGeom g;
  g.V = 1; g.cx = n_partials <= 512 ? 16 : 64; g.ry = kBnThreads / g.cx;
  g.cblocks = (int)ceil_div(C, g.cx);
  int64_t splits = ceil_div((int64_t)sm_count() * 2, g.cblocks);
  int64_t max_splits = ceil_div((int64_t)n_partials, (int64_t)g.ry * 4);
  if (splits > max_splits) splits = max_splits;
  if (n_partials <= 512) splits = 1;
  if (splits < 1) splits = 1;
  g.rows_per_split = ceil_div((int64_t)n_partials, splits);
  g.splits = (int)ceil_div((int64_t)n_partials, g.rows_per_split);
  float* scale = part + (size_t)2 * C * g.splits;
  fin.scale = scale; fin.shift = scale + C;
  int* counter = bn_counters(g.cblocks);
  CNB_REQUIRE(counter, CNB_ERR_CUDA, "cnb_bn_stats_from_partials: cannot allocate the arrival counters");
  dim3 grid(g.cblocks, g.splits), block(g.cx, g.ry);
  bn_partials_k<<<grid, block, 4096, st>>>(conv_partials, part, n_partials, d->M, C, g.rows_per_split,
                                           g.splits, counter, fin);
  return check_launch("bn_partials(stats only)");
}

int cnb_bn_fwd_train_prefix(const cnb_bn_desc* d, const void* x, int64_t x_row_stride, const float* mean,
                            const float* rstd, const float* gamma, const float* beta,
                            float* running_mean, float* running_var, void* y, void* relu_mask,
                            void* workspace, size_t workspace_bytes, void* stream) {
  CNB_TRY(validate_bn(d, "cnb_bn_fwd_train_prefix"));
  CNB_REQUIRE(x && y && mean && rstd && workspace, CNB_ERR_INVALID, "cnb_bn_fwd_train_prefix: null pointer");
  CNB_REQUIRE(!(d->flags & CNB_BN_ADD), CNB_ERR_UNSUPPORTED, "cnb_bn_fwd_train_prefix: no residual input");
  CNB_REQUIRE(x_row_stride >= d->C, CNB_ERR_INVALID, "cnb_bn_fwd_train_prefix: row stride %lld < C=%d",
              (long long)x_row_stride, d->C);
  CNB_REQUIRE(workspace_bytes >= bn_ws_bytes(d->C), CNB_ERR_WORKSPACE,
              "cnb_bn_fwd_train_prefix: workspace %zu < %zu", workspace_bytes, bn_ws_bytes(d->C));
  const bool relu = d->flags & CNB_BN_RELU;
  CNB_REQUIRE(!relu_mask || (relu && d->dtype == CNB_BF16 && d->C % 8 == 0), CNB_ERR_INVALID,
              "cnb_bn_fwd_train_prefix: relu_mask needs CNB_BN_RELU, bf16 and C %% 8 == 0");
  cudaStream_t st = (cudaStream_t)stream;
  const int C = d->C;
  const int64_t M = d->M;
  const size_t es = dtype_size(d->dtype);
  int v = pick_vec(C, es, x, y);
  while (v > 1 && (x_row_stride % v) != 0) v >>= 1;
  CNB_REQUIRE(!relu_mask || v == 8, CNB_ERR_INVALID, "cnb_bn_fwd_train_prefix: relu_mask needs 16-byte aligned rows");
  float* scale = (float*)workspace;
  float* shift = scale + C;
  bn_prefix_coeff_k<<<(unsigned)ceil_div(C, 128), 128, 0, st>>>(gamma, beta, mean, rstd, running_mean, running_var,
                                                                (long long*)d->num_batches_tracked, scale, shift,
                                                                C, M, d->eps, d->momentum);
  CNB_TRY(check_launch("bn_prefix_coeff"));
  BnGeom ga = bn_geom(M, C, v, kStreamWaves);
  dim3 agrid(ga.cblocks, ga.splits), ablock(ga.cx, ga.ry);
  if (relu)
    CNB_DISPATCH_TV(d->dtype, v, (bn_apply_ld_k<T, V, true><<<agrid, ablock, 0, st>>>(
                                     (const T*)x, x_row_stride, scale, shift, mean, (T*)y, (uint8_t*)relu_mask, M, C,
                                     ga.rows_per_split)));
  else
    CNB_DISPATCH_TV(d->dtype, v, (bn_apply_ld_k<T, V, false><<<agrid, ablock, 0, st>>>(
                                     (const T*)x, x_row_stride, scale, shift, mean, (T*)y, nullptr, M, C,
                                     ga.rows_per_split)));
  return check_launch("bn_apply(prefix)");
}

int cnb_bn_bwd_prefix(const cnb_bn_desc* d, const void* x, int64_t x_row_stride, const void* y,
                      const void* relu_mask, const void* dy, const float* mean, const float* rstd,
                      const float* gamma, void* dx_accum, int64_t dx_row_stride, float* dgamma,
                      float* dbeta, void* workspace, size_t workspace_bytes, void* stream) {
  CNB_TRY(validate_bn(d, "cnb_bn_bwd_prefix"));
  CNB_REQUIRE(x && dy && mean && rstd && dx_accum && workspace, CNB_ERR_INVALID, "cnb_bn_bwd_prefix: null pointer");
  CNB_REQUIRE(!(d->flags & CNB_BN_ADD), CNB_ERR_UNSUPPORTED, "cnb_bn_bwd_prefix: no residual input");
  CNB_REQUIRE(x_row_stride >= d->C && dx_row_stride >= d->C, CNB_ERR_INVALID, "cnb_bn_bwd_prefix: row stride < C");
  CNB_REQUIRE(workspace_bytes >= bn_ws_bytes(d->C), CNB_ERR_WORKSPACE,
              "cnb_bn_bwd_prefix: workspace %zu < %zu", workspace_bytes, bn_ws_bytes(d->C));
  const bool relu = d->flags & CNB_BN_RELU;
  CNB_REQUIRE(!relu || y || relu_mask, CNB_ERR_INVALID, "cnb_bn_bwd_prefix: CNB_BN_RELU needs y or the bit mask");
  cudaStream_t st = (cudaStream_t)stream;
  const int C = d->C;
  const int64_t M = d->M;
  const size_t es = dtype_size(d->dtype);
  int v = pick_vec(C, es, x, dy, dx_accum, relu && !relu_mask ? y : nullptr);
  while (v > 1 && ((x_row_stride % v) != 0 || (dx_row_stride % v) != 0)) v >>= 1;
  const uint8_t* mask = (const uint8_t*)relu_mask;
  CNB_REQUIRE(!mask || (v == 8 && d->dtype == CNB_BF16), CNB_ERR_INVALID,
              "cnb_bn_bwd_prefix: the bit mask needs bf16 and 16-byte aligned rows");
  const int mode = !relu ? 0 : (mask ? 2 : 1);
  float* part = (float*)workspace;
  BnGeom g = bn_geom(M, C, v, kReduceWaves, kReduceCx);
  float* c1 = part + (size_t)2 * C * g.splits;
  float* c2 = c1 + C;
  float* c3 = c2 + C;
  int* counter = bn_counters(g.cblocks);
  CNB_REQUIRE(counter, CNB_ERR_CUDA, "cnb_bn_bwd_prefix: cannot allocate the arrival counters");
  {
    dim3 grid(g.cblocks, g.splits), block(g.cx, g.ry);
    size_t smem = (size_t)g.ry * 2 * g.cx * v * sizeof(float);
    if (smem < 4096) smem = 4096;
    BnBwdFin fin{gamma, rstd, dgamma, dbeta, c1, c2, c3};
#define RED(R)                                                                                           \
  CNB_DISPATCH_TV(d->dtype, v, (bn_bwd_reduce_ld_k<T, V, R><<<grid, block, smem, st>>>(                  \
                                   (const T*)x, x_row_stride, (const T*)y, mask, (const T*)dy, mean, part, M, C, \
                                   g.rows_per_split, g.splits, counter, fin)))
    if (mode == 2) RED(2);
    else if (mode == 1) RED(1);
    else RED(0);
#undef RED
    CNB_TRY(check_launch("bn_bwd_reduce(prefix)"));
  }
  BnGeom ga = bn_geom(M, C, v, kStreamWaves);
  dim3 agrid(ga.cblocks, ga.splits), ablock(ga.cx, ga.ry);
#define DXL(R)                                                                                            \
  CNB_DISPATCH_TV(d->dtype, v, (bn_bwd_dx_ld_k<T, V, R><<<agrid, ablock, 0, st>>>(                        \
                                   (const T*)x, x_row_stride, (const T*)y, mask, (const T*)dy, mean, c1, c2, c3, \
                                   (T*)dx_accum, dx_row_stride, M, C, ga.rows_per_split)))
  if (mode == 2) DXL(2);
  else if (mode == 1) DXL(1);
  else DXL(0);
#undef DXL
  return check_launch("bn_bwd_dx(prefix)");
}

int cnb_bn_fwd_eval(const cnb_bn_desc* d, const void* x, const void* residual, const float* gamma,
                    const float* beta, const float* running_mean, const float* running_var, void* y,
                    void* workspace, size_t workspace_bytes, void* stream) {
  CNB_TRY(validate_bn(d, "cnb_bn_fwd_eval"));
  CNB_REQUIRE(x && y && running_mean && running_var && workspace, CNB_ERR_INVALID,
              "cnb_bn_fwd_eval: null pointer");
  CNB_REQUIRE(workspace_bytes >= (size_t)2 * d->C * sizeof(float), CNB_ERR_WORKSPACE,
              "cnb_bn_fwd_eval: workspace %zu < %zu", workspace_bytes,
              (size_t)2 * d->C * sizeof(float));
  const bool add = d->flags & CNB_BN_ADD, relu = d->flags & CNB_BN_RELU;
  CNB_REQUIRE(!add || residual, CNB_ERR_INVALID, "cnb_bn_fwd_eval: CNB_BN_ADD without residual");
  cudaStream_t st = (cudaStream_t)stream;
  const int C = d->C;
  float* coeff = (float*)workspace;
  float* scale = coeff;
  float* shift = coeff + C;
  bn_eval_coeff_k<<<(unsigned)ceil_div(C, 128), 128, 0, st>>>(gamma, beta, running_mean,
                                                              running_var, scale, shift, C, d->eps);
  CNB_TRY(check_launch("bn_eval_coeff"));
  int v = pick_vec(C, dtype_size(d->dtype), x, y, add ? residual : nullptr);
  const bool ahot = d->dtype == CNB_BF16 && v == 8;
  const int avariant = bn_fwd_variant();
  BnGeom ga = bn_geom(d->M, C, v, apply_waves(ahot, avariant));
  dim3 agrid(ga.cblocks, ga.splits), ablock(ga.cx, ga.ry);
  const int64_t arows = ga.rows_per_split;
  uint8_t* mask = nullptr;
  const float* mean_ptr = running_mean;
  APPLY_ANY();
#undef APPLY_ANY
#undef APPLY_HOT
#undef APPLY
#undef APPLY_ARGS
  return check_launch("bn_apply(eval)");
}

int cnb_bn_bwd(const cnb_bn_desc* d, const void* x, const void* y, const void* relu_mask,
               const void* dy, const float* gamma, const float* save_mean, const float* save_rstd, void* dx,
               void* d_residual, float* dgamma, float* dbeta, void* workspace,
               size_t workspace_bytes, void* stream) {
  CNB_TRY(validate_bn(d, "cnb_bn_bwd"));
  CNB_REQUIRE(x && dy && save_mean && save_rstd && dx && workspace, CNB_ERR_INVALID,
              "cnb_bn_bwd: null pointer");
  const bool relu = d->flags & CNB_BN_RELU;
  CNB_REQUIRE(!relu || y || relu_mask, CNB_ERR_INVALID,
              "cnb_bn_bwd: CNB_BN_RELU needs the forward output y or the relu_mask");
  CNB_REQUIRE(!relu_mask || (relu && d->dtype == CNB_BF16 && d->C % 8 == 0), CNB_ERR_INVALID,
              "cnb_bn_bwd: relu_mask needs CNB_BN_RELU, bf16 and C %% 8 == 0 (C=%d)", d->C);
  const uint8_t* mask = (const uint8_t*)relu_mask;
  CNB_REQUIRE(workspace_bytes >= bn_ws_bytes(d->C), CNB_ERR_WORKSPACE,
              "cnb_bn_bwd: workspace %zu < %zu", workspace_bytes, bn_ws_bytes(d->C));
  cudaStream_t st = (cudaStream_t)stream;
  const int C = d->C;
  const int64_t M = d->M;
  int v = pick_vec(C, dtype_size(d->dtype), x, dy, dx, (relu && !mask) ? y : nullptr);
  if (d_residual)
    while (v > 1 && (uintptr_t)d_residual % (v * dtype_size(d->dtype)) != 0) v >>= 1;
  CNB_REQUIRE(!mask || v == 8, CNB_ERR_INVALID, "cnb_bn_bwd: relu_mask needs 16-byte aligned tensors");
  const bool hot = d->dtype == CNB_BF16 && v == 8;   // the feature-map configuration of every zoo model
  const int variant = bn_bwd_variant();
  const int mode = relu ? (mask ? 2 : 1) : 0;
  // grids are whole multiples of the CTAs that are resident at once: 2 per SM with the (.., 2)
  // loop shapes, 3 per SM with the (.., 3) ones (80 registers)
  const bool three = hot && (variant == 1 || variant == 4);
  float* part = (float*)workspace;
  if (hot && (int64_t)2 * M * C * 2 <= bn_bwd_fused_limit()) {
    // x + dy fit L2: one cooperative kernel, one DRAM read of every operand
    bool taken = false;
    BnBwdFin fin{gamma, save_rstd, dgamma, dbeta, nullptr, nullptr, nullptr};
#define FUSED(R, D)                                                                                   \
  CNB_TRY((launch_bn_bwd_fused<R, D>((const bf16*)x, (const bf16*)y, mask, (const bf16*)dy, save_mean, \
                                     part, (bf16*)dx, (bf16*)d_residual, M, C, fin, &taken, st)))
    if (mode == 2 && d_residual) FUSED(2, true);
    else if (mode == 2) FUSED(2, false);
    else if (mode == 1 && d_residual) FUSED(1, true);
    else if (mode == 1) FUSED(1, false);
    else if (d_residual) FUSED(0, true);
    else FUSED(0, false);
#undef FUSED
    if (taken) return CNB_OK;
  }
  BnGeom g = bn_geom(M, C, v, three ? 3 : kReduceWaves, kReduceCx);
  float* c1 = part + (size_t)2 * C * g.splits;
  float* c2 = c1 + C;
  float* c3 = c2 + C;
  int* counter = bn_counters(g.cblocks);
  CNB_REQUIRE(counter, CNB_ERR_CUDA, "cnb_bn_bwd: cannot allocate the arrival counters");
  {
    dim3 grid(g.cblocks, g.splits), block(g.cx, g.ry);
    size_t smem = (size_t)g.ry * 2 * g.cx * v * sizeof(float);
    if (smem < 4096) smem = 4096;  // the merge step re-uses it for 256 x 2 doubles
    BnBwdFin fin{gamma, save_rstd, dgamma, dbeta, c1, c2, c3};
#define RED_ARGS (const T*)x, (const T*)y, mask, (const T*)dy, save_mean, part, M, C, g.rows_per_split, \
                 g.splits, counter, fin
#define RED_HOT(UU, MB)                                                                      \
  do {                                                                                       \
    using T = bf16;                                                                          \
    if (mode == 2) bn_bwd_reduce_k<T, 8, 2, UU, MB><<<grid, block, smem, st>>>(RED_ARGS);    \
    else if (mode == 1) bn_bwd_reduce_k<T, 8, 1, UU, MB><<<grid, block, smem, st>>>(RED_ARGS); \
    else bn_bwd_reduce_k<T, 8, 0, UU, MB><<<grid, block, smem, st>>>(RED_ARGS);              \
  } while (0)
    if (hot && variant == 1) RED_HOT(2, 3);
    else if (hot && variant == 2) RED_HOT(4, 2);
    else if (hot && variant == 3) RED_HOT(3, 2);
    else if (hot && variant == 4) RED_HOT(4, 3);
    else if (mode == 2)
      CNB_DISPATCH_TV(d->dtype, v, (bn_bwd_reduce_k<T, V, 2, 2, 2><<<grid, block, smem, st>>>(RED_ARGS)));
    else if (mode == 1)
      CNB_DISPATCH_TV(d->dtype, v, (bn_bwd_reduce_k<T, V, 1, 2, 2><<<grid, block, smem, st>>>(RED_ARGS)));
    else
      CNB_DISPATCH_TV(d->dtype, v, (bn_bwd_reduce_k<T, V, 0, 2, 2><<<grid, block, smem, st>>>(RED_ARGS)));
#undef RED_HOT
#undef RED_ARGS
    CNB_TRY(check_launch("bn_bwd_reduce"));
  }
  return launch_bn_bwd_dx(d, x, y, mask, dy, save_mean, c1, c2, c3, dx, d_residual, v, hot, variant, mode, three, st);
}

int cnb_bn_bwd_from_partials(const cnb_bn_desc* d, const void* x, const void* relu_mask, const void* dy,
                             const float* gamma, const float* save_mean, const float* save_rstd,
                             const float* partials, int32_t n_partials, void* dx, void* d_residual,
                             float* dgamma, float* dbeta, void* workspace, size_t workspace_bytes,
                             void* stream) {
  CNB_TRY(validate_bn(d, "cnb_bn_bwd_from_partials"));
  CNB_REQUIRE(x && dy && save_mean && save_rstd && dx && workspace && partials && n_partials > 0,
              CNB_ERR_INVALID, "cnb_bn_bwd_from_partials: null pointer / no rows");
  const bool relu = d->flags & CNB_BN_RELU;
  CNB_REQUIRE(relu && relu_mask && d->dtype == CNB_BF16 && d->C % 8 == 0, CNB_ERR_UNSUPPORTED,
              "cnb_bn_bwd_from_partials: bf16 BatchNorm + ReLU with the 1-bit mask only (C=%d)", d->C);
  CNB_REQUIRE(workspace_bytes >= bn_ws_bytes(d->C), CNB_ERR_WORKSPACE,
              "cnb_bn_bwd_from_partials: workspace %zu < %zu", workspace_bytes, bn_ws_bytes(d->C));
  cudaStream_t st = (cudaStream_t)stream;
  const int C = d->C;
  const uint8_t* mask = (const uint8_t*)relu_mask;
  int v = pick_vec(C, dtype_size(d->dtype), x, dy, dx);
  if (d_residual)
    while (v > 1 && (uintptr_t)d_residual % (v * dtype_size(d->dtype)) != 0) v >>= 1;
  CNB_REQUIRE(v == 8, CNB_ERR_INVALID, "cnb_bn_bwd_from_partials: tensors must be 16-byte aligned");
  const int variant = bn_bwd_variant();
  const bool three = variant == 1 || variant == 4;
  float* part = (float*)workspace;
  BnGeom g;
  g.V = 1; g.cx = n_partials <= 512 ? 16 : 64; g.ry = kBnThreads / g.cx;
  g.cblocks = (int)ceil_div(C, g.cx);
  g.splits = 1;
  g.rows_per_split = n_partials;
  if (n_partials > 512) {
    int64_t splits = ceil_div((int64_t)sm_count() * 2, g.cblocks);
    int64_t max_splits = ceil_div((int64_t)n_partials, (int64_t)g.ry * 4);
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
    g.rows_per_split = ceil_div((int64_t)n_partials, splits);
    g.splits = (int)ceil_div((int64_t)n_partials, g.rows_per_split);
  }
  float* c1 = part + (size_t)2 * C * g.splits;
  float* c2 = c1 + C;
  float* c3 = c2 + C;
  int* counter = bn_counters(g.cblocks);
  CNB_REQUIRE(counter, CNB_ERR_CUDA, "cnb_bn_bwd_from_partials: cannot allocate the arrival counters");
  BnBwdFin fin{gamma, save_rstd, dgamma, dbeta, c1, c2, c3};
  dim3 grid(g.cblocks, g.splits), block(g.cx, g.ry);
  bn_bwd_partials_k<<<grid, block, 4096, st>>>(partials, part, n_partials, d->M, C, g.rows_per_split, g.splits,
                                               counter, fin);
  CNB_TRY(check_launch("bn_bwd_partials"));
  return launch_bn_bwd_dx(d, x, nullptr, mask, dy, save_mean, c1, c2, c3, dx, d_residual, 8, true, variant, 2, three,
                          st);
}

// ---------------------------------------------------------------- allocation-free BatchNorm forward
// cnb_bn_fwd_train_static: the training forward of cnb_bn_fwd_train (statistics from x) for callers
// that may not allocate — under CUDA-graph capture, in a tight serving loop, in a harness that
// forbids cudaMalloc.  No counterpart in the reference.
// The scratch the two-kernel form needs (per-CTA partial sums, the per-channel scale / shift, the
// arrival counters of the "last CTA merges" step) lives in ONE file-scope device array that every
// GPU loading the library holds until exit:
//   [0, kStaticScratchBytes)            partials [splits][2][C] floats, then scale[C], shift[C]
//   [kStaticScratchBytes, + 4 KiB)      arrival counters (kCountersPerSlot ints)
// In host code the array's NAME is a host-side stand-in, not a device address: its device address is
// looked up once per device with cudaGetSymbolAddress (that device current) and passed to the
// kernels.  Every call shares the array and finds the previous call's contents in it: the counters
// are cleared on the caller's stream before the statistics kernel; partials and coefficients are
// fully rewritten before they are read.  Two calls that overlap on the GPU (different streams, no
// ordering between them) race on it.
constexpr size_t kStaticScratchBytes = (size_t)1 << 20;
constexpr int kStaticMaxChannels = 16384;
constexpr size_t kStaticTotalBytes = kStaticScratchBytes + (size_t)kCountersPerSlot * sizeof(int);
static __device__ __align__(256) unsigned char g_bn_static_scratch[kStaticTotalBytes];

static int bn_static_scratch_ptr(unsigned char** out) {
  static std::mutex mu;
  static unsigned char* addr[64] = {nullptr};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  CNB_REQUIRE(e == cudaSuccess && dev >= 0 && dev < 64, CNB_ERR_CUDA,
              "cnb_bn_fwd_train_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    unsigned char* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_bn_static_scratch);
    CNB_REQUIRE(e == cudaSuccess && p != nullptr, CNB_ERR_CUDA,
                "cnb_bn_fwd_train_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
    addr[dev] = p;
  }
  *out = addr[dev];
  return CNB_OK;
}

size_t cnb_bn_static_scratch_bytes(void) { return kStaticScratchBytes; }

int cnb_bn_fwd_train_static(const cnb_bn_desc* d, const void* x, const void* residual, const float* gamma,
                            const float* beta, float* running_mean, float* running_var, void* y,
                            float* save_mean, float* save_rstd, void* relu_mask, void* stream) {
  CNB_TRY(validate_bn(d, "cnb_bn_fwd_train_static"));
  CNB_REQUIRE(!relu_mask || ((d->flags & CNB_BN_RELU) && d->dtype == CNB_BF16 && d->C % 8 == 0),
              CNB_ERR_INVALID,
              "cnb_bn_fwd_train_static: relu_mask needs CNB_BN_RELU, bf16 and C %% 8 == 0 (C=%d)", d->C);
  CNB_REQUIRE(x && y && save_mean && save_rstd, CNB_ERR_INVALID, "cnb_bn_fwd_train_static: null pointer");
  const bool add = d->flags & CNB_BN_ADD, relu = d->flags & CNB_BN_RELU;
  CNB_REQUIRE(!add || residual, CNB_ERR_INVALID, "cnb_bn_fwd_train_static: CNB_BN_ADD without residual");
  CNB_REQUIRE(d->M > 1, CNB_ERR_INVALID,
              "cnb_bn_fwd_train_static: expected more than 1 value per channel when training (M=%lld)",
              (long long)d->M);
  CNB_REQUIRE(d->C <= kStaticMaxChannels, CNB_ERR_WORKSPACE,
              "cnb_bn_fwd_train_static: C=%d exceeds the %d channels the built-in scratch is sized for "
              "(use cnb_bn_fwd_train with a caller workspace)", d->C, kStaticMaxChannels);
  cudaStream_t st = (cudaStream_t)stream;
  const int C = d->C;
  const int64_t M = d->M;
  int v = pick_vec(C, dtype_size(d->dtype), x, y, add ? residual : nullptr);
  CNB_REQUIRE(!relu_mask || v == 8, CNB_ERR_INVALID,
              "cnb_bn_fwd_train_static: relu_mask needs 16-byte aligned tensors");
  // same kernels and launch geometry as cnb_bn_fwd_train => the same bits
  const int svariant = (d->dtype == CNB_BF16 && v == 8) ? bn_stats_variant() : 0;
  BnGeom g = bn_geom(M, C, v, svariant == 2 ? 3 : kReduceWaves, kReduceCx);
  const size_t need = ((size_t)2 * C * g.splits + (size_t)2 * C) * sizeof(float);
  CNB_REQUIRE(need <= kStaticScratchBytes && g.cblocks <= kCountersPerSlot, CNB_ERR_WORKSPACE,
              "cnb_bn_fwd_train_static: needs %zu bytes of scratch (C=%d, %d row splits), the built-in "
              "scratch holds %zu (use cnb_bn_fwd_train with a caller workspace)", need, C, g.splits,
              kStaticScratchBytes);
  unsigned char* base = nullptr;
  CNB_TRY(bn_static_scratch_ptr(&base));
  float* part = reinterpret_cast<float*>(base);
  int* counter = reinterpret_cast<int*>(base + kStaticScratchBytes);
  float* scale = part + (size_t)2 * C * g.splits;
  float* shift = scale + C;
  BnFwdFin fin{gamma, beta, running_mean, running_var, save_mean, save_rstd, scale, shift,
               d->eps, d->momentum, (long long*)d->num_batches_tracked};
  cudaError_t e = cudaMemsetAsync(counter, 0, (size_t)g.cblocks * sizeof(int), st);
  CNB_REQUIRE(e == cudaSuccess, CNB_ERR_CUDA, "cnb_bn_fwd_train_static: clearing the counters: %s",
              cudaGetErrorString(e));
  {
    dim3 grid(g.cblocks, g.splits), block(g.cx, g.ry);
    size_t smem = (size_t)g.ry * 2 * g.cx * v * sizeof(float);
    if (smem < 4096) smem = 4096;  // the merge step re-uses it for 256 x 2 doubles
    if (svariant == 1)
      bn_stats_walk_k<bf16, 8, 4, 4><<<grid, block, smem, st>>>((const bf16*)x, part, M, C, g.rows_per_split,
                                                                g.splits, counter, fin);
    else if (svariant == 2)
      bn_stats_walk_k<bf16, 8, 8, 3><<<grid, block, smem, st>>>((const bf16*)x, part, M, C, g.rows_per_split,
                                                                g.splits, counter, fin);
    else
      CNB_DISPATCH_TV(d->dtype, v, (bn_stats_k<T, V><<<grid, block, smem, st>>>(
                                       (const T*)x, part, M, C, g.rows_per_split, g.splits, counter,
                                       fin)));
    CNB_TRY(check_launch("bn_stats(static)"));
  }
  const float* mean_ptr = save_mean;
  uint8_t* mask = (uint8_t*)relu_mask;
  const bool ahot = d->dtype == CNB_BF16 && v == 8;
  const int avariant = bn_fwd_variant();
  BnGeom ga = bn_geom(M, C, v, apply_waves(ahot, avariant));
  dim3 agrid(ga.cblocks, ga.splits), ablock(ga.cx, ga.ry);
  const int64_t arows = ga.rows_per_split;
#define SAPPLY_ARGS (const T*)x, (const T*)residual, scale, shift, mean_ptr, (T*)y, mask, M, C, arows
#define SAPPLY(R, A) \
  CNB_DISPATCH_TV(d->dtype, v, (bn_apply_k<T, V, R, A, 4, 2><<<agrid, ablock, 0, st>>>(SAPPLY_ARGS)))
#define SAPPLY_HOT(UU, MB)                                                                          \
  do {                                                                                              \
    using T = bf16;                                                                                 \
    if (relu && add) bn_apply_k<T, 8, true, true, UU, MB><<<agrid, ablock, 0, st>>>(SAPPLY_ARGS);   \
    else if (relu) bn_apply_k<T, 8, true, false, UU, MB><<<agrid, ablock, 0, st>>>(SAPPLY_ARGS);    \
    else if (add) bn_apply_k<T, 8, false, true, UU, MB><<<agrid, ablock, 0, st>>>(SAPPLY_ARGS);     \
    else bn_apply_k<T, 8, false, false, UU, MB><<<agrid, ablock, 0, st>>>(SAPPLY_ARGS);             \
  } while (0)
  if (ahot && avariant == 1) SAPPLY_HOT(2, 3);
  else if (ahot && avariant == 2) SAPPLY_HOT(4, 3);
  else if (ahot && avariant == 3) SAPPLY_HOT(2, 4);
  else if (relu && add) SAPPLY(true, true);
  else if (relu) SAPPLY(true, false);
  else if (add) SAPPLY(false, true);
  else SAPPLY(false, false);
#undef SAPPLY_HOT
#undef SAPPLY
#undef SAPPLY_ARGS
  return check_launch("bn_apply(static)");
}
}  // extern "C"
