// bn.cu — BatchNorm2d forward (training / inference) and backward over NHWC tensors viewed as
// [M = N*H*W, C].  Replaces aten::native_batch_norm(_backward) reached from reference
// src/basemodel.py:1008,1061; src/densenet.py:80,117,124,182; src/mobilenet_v1.py:108,119.
//
// Structure (all HBM-bound, 128-bit accesses along C):
//   stats   : per-(row-chunk, channel) shifted sums  -> fp32 partials in the workspace
//   finalize: Chan-merge of the partials in fp64, running-stat update, scale/shift vectors
//   apply   : y = [relu](x*scale + shift [+ residual])
//   bwd     : reduce (sum dy', sum dy'*xhat) -> finalize (dgamma, dbeta, coefficients) -> dx
// Two-stage reductions (no atomics) keep results bit-reproducible run to run.
#include "common.cuh"

namespace cnb {

constexpr int kBnThreads = 256;

struct BnGeom {
  int V;          // channels per thread
  int cx;         // threads along channels
  int ry;         // threads along rows
  int cblocks;    // blocks along channels
  int splits;     // blocks along rows
  int64_t rows_per_split;
};

static BnGeom bn_geom(int64_t M, int C, int V) {
  BnGeom g;
  g.V = V;
  int cvs = C / V;
  int p = 1;
  while (p * 2 <= cvs && p * 2 <= 32) p *= 2;
  g.cx = p;
  g.ry = kBnThreads / g.cx;
  g.cblocks = (int)ceil_div(cvs, g.cx);
  int64_t target = (int64_t)sm_count() * 8;
  int64_t splits = ceil_div(target, g.cblocks);
  int64_t max_splits = ceil_div(M, (int64_t)g.ry * 4);  // at least ~4 rows per thread
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  if (splits > 4096) splits = 4096;
  g.rows_per_split = ceil_div(M, splits);
  g.splits = (int)ceil_div(M, g.rows_per_split);
  return g;
}

// workspace layout (floats): [0, 3*splits*C) partials | then up to 5*C floats of per-channel
// coefficients (scale, shift / c1, c2, c3)

// ---------------------------------------------------------------- forward statistics
// partial[split][0][c] = sum(x - K_c), partial[split][1][c] = sum((x - K_c)^2) with the pivot
// K_c = x[first row of the split][c] (kills the mean^2/var cancellation of a raw sumsq).
// pivot[split][c] is stored too.
template <typename T, int V>
__global__ void bn_stats_k(const T* __restrict__ x, float* __restrict__ part, int64_t M, int C,
                           int64_t rows_per_split) {
  extern __shared__ float red[];  // [ry][2][cx*V]
  const int cv = blockIdx.x * blockDim.x + threadIdx.x;
  const int c0 = cv * V;
  const int64_t r0 = blockIdx.y * rows_per_split;
  int64_t r1 = r0 + rows_per_split;
  if (r1 > M) r1 = M;
  float s[V], q[V], piv[V];
#pragma unroll
  for (int j = 0; j < V; ++j) s[j] = q[j] = piv[j] = 0.f;
  if (c0 < C) {
    Pack<T, V> p0 = ld_pack<T, V>(x + r0 * C + c0);
#pragma unroll
    for (int j = 0; j < V; ++j) piv[j] = to_f(p0.v[j]);
    for (int64_t r = r0 + threadIdx.y; r < r1; r += blockDim.y) {
      Pack<T, V> a = ld_pack<T, V>(x + r * C + c0);
#pragma unroll
      for (int j = 0; j < V; ++j) {
        float d = to_f(a.v[j]) - piv[j];
        s[j] += d;
        q[j] = fmaf(d, d, q[j]);
      }
    }
  }
  const int W = blockDim.x * V;
#pragma unroll
  for (int j = 0; j < V; ++j) {
    red[(threadIdx.y * 2 + 0) * W + threadIdx.x * V + j] = s[j];
    red[(threadIdx.y * 2 + 1) * W + threadIdx.x * V + j] = q[j];
  }
  __syncthreads();
  if (threadIdx.y == 0 && c0 < C) {
    float* out = part + (int64_t)blockIdx.y * 3 * C;
#pragma unroll
    for (int j = 0; j < V; ++j) {
      float ts = 0.f, tq = 0.f;
      for (int r = 0; r < (int)blockDim.y; ++r) {
        ts += red[(r * 2 + 0) * W + threadIdx.x * V + j];
        tq += red[(r * 2 + 1) * W + threadIdx.x * V + j];
      }
      out[0 * C + c0 + j] = ts;
      out[1 * C + c0 + j] = tq;
      out[2 * C + c0 + j] = piv[j];
    }
  }
}

// One thread per channel: merge split partials (Chan et al.) in fp64, emit mean / rstd, update
// the running statistics (momentum, unbiased variance) and the fused scale / shift vectors.
__global__ void bn_finalize_fwd_k(const float* __restrict__ part, int splits, int64_t M,
                                  int64_t rows_per_split, int C, const float* __restrict__ gamma,
                                  const float* __restrict__ beta, float* __restrict__ rmean,
                                  float* __restrict__ rvar, float* __restrict__ save_mean,
                                  float* __restrict__ save_rstd, float* __restrict__ scale,
                                  float* __restrict__ shift, float eps, float momentum) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double n_acc = 0.0, mean = 0.0, m2 = 0.0;
  for (int s = 0; s < splits; ++s) {
    int64_t r0 = (int64_t)s * rows_per_split;
    int64_t r1 = r0 + rows_per_split;
    if (r1 > M) r1 = M;
    double nb = (double)(r1 - r0);
    const float* p = part + (int64_t)s * 3 * C;
    double sb = p[c], qb = p[C + c], kb = p[2 * C + c];
    double mean_b = kb + sb / nb;
    double m2_b = qb - sb * sb / nb;
    if (m2_b < 0.0) m2_b = 0.0;
    double delta = mean_b - mean;
    double n_new = n_acc + nb;
    mean += delta * nb / n_new;
    m2 += m2_b + delta * delta * n_acc * nb / n_new;
    n_acc = n_new;
  }
  double var = m2 / (double)M;  // biased: used for normalisation
  float rstd = (float)(1.0 / sqrt(var + (double)eps));
  float mu = (float)mean;
  save_mean[c] = mu;
  save_rstd[c] = rstd;
  if (rmean) rmean[c] = (1.f - momentum) * rmean[c] + momentum * mu;
  if (rvar) {
    double unbiased = M > 1 ? m2 / (double)(M - 1) : var;
    rvar[c] = (1.f - momentum) * rvar[c] + momentum * (float)unbiased;
  }
  float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
  scale[c] = g * rstd;
  shift[c] = b - mu * g * rstd;
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
  shift[c] = b - rmean[c] * g * rstd;
}

// y = [relu](x*scale + shift [+ res])
template <typename T, int V, bool RELU, bool ADD>
__global__ void bn_apply_k(const T* __restrict__ x, const T* __restrict__ res,
                           const float* __restrict__ scale, const float* __restrict__ shift,
                           T* __restrict__ y, int64_t total_vec, int Cv) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total_vec;
       i += (int64_t)gridDim.x * blockDim.x) {
    int cv = (int)(i % Cv);
    Pack<T, V> a = ld_pack<T, V>(x + i * V), o;
    Pack<T, V> r;
    if (ADD) r = ld_pack<T, V>(res + i * V);
#pragma unroll
    for (int j = 0; j < V; ++j) {
      float f = fmaf(to_f(a.v[j]), scale[cv * V + j], shift[cv * V + j]);
      if (ADD) {
        // the reference materialises bn(x) in the storage dtype before `output += identity`
        f = to_f(from_f<T>(f)) + to_f(r.v[j]);
      }
      if (RELU) f = (f > 0.f || f != f) ? f : 0.f;
      o.v[j] = from_f<T>(f);
    }
    st_pack<T, V>(y + i * V, o);
  }
}

// ---------------------------------------------------------------- backward reduce
// partial[split][0][c] = sum dy', partial[split][1][c] = sum dy' * (x - mean)   (dy' = masked dy)
template <typename T, int V, bool RELU>
__global__ void bn_bwd_reduce_k(const T* __restrict__ x, const T* __restrict__ y,
                                const T* __restrict__ dy, const float* __restrict__ mean,
                                float* __restrict__ part, int64_t M, int C,
                                int64_t rows_per_split) {
  extern __shared__ float red[];
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
    for (int64_t r = r0 + threadIdx.y; r < r1; r += blockDim.y) {
      Pack<T, V> a = ld_pack<T, V>(x + r * C + c0);
      Pack<T, V> g = ld_pack<T, V>(dy + r * C + c0);
      Pack<T, V> o;
      if (RELU) o = ld_pack<T, V>(y + r * C + c0);
#pragma unroll
      for (int j = 0; j < V; ++j) {
        float gf = to_f(g.v[j]);
        if (RELU && !(to_f(o.v[j]) > 0.f)) gf = 0.f;
        s[j] += gf;
        q[j] = fmaf(gf, to_f(a.v[j]) - mu[j], q[j]);
      }
    }
  }
  const int W = blockDim.x * V;
#pragma unroll
  for (int j = 0; j < V; ++j) {
    red[(threadIdx.y * 2 + 0) * W + threadIdx.x * V + j] = s[j];
    red[(threadIdx.y * 2 + 1) * W + threadIdx.x * V + j] = q[j];
  }
  __syncthreads();
  if (threadIdx.y == 0 && c0 < C) {
    float* out = part + (int64_t)blockIdx.y * 3 * C;
#pragma unroll
    for (int j = 0; j < V; ++j) {
      float ts = 0.f, tq = 0.f;
      for (int r = 0; r < (int)blockDim.y; ++r) {
        ts += red[(r * 2 + 0) * W + threadIdx.x * V + j];
        tq += red[(r * 2 + 1) * W + threadIdx.x * V + j];
      }
      out[0 * C + c0 + j] = ts;
      out[1 * C + c0 + j] = tq;
    }
  }
}

// dbeta = sum dy' ; dgamma = rstd * sum dy'(x-mean) ;
// dx = c1 * (dy' - c2 - (x - mean) * c3), c1 = gamma*rstd, c2 = dbeta/M, c3 = rstd^2 * sumxc / M
__global__ void bn_finalize_bwd_k(const float* __restrict__ part, int splits, int64_t M, int C,
                                  const float* __restrict__ gamma, const float* __restrict__ rstd,
                                  float* __restrict__ dgamma, float* __restrict__ dbeta,
                                  float* __restrict__ c1, float* __restrict__ c2,
                                  float* __restrict__ c3) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double s = 0.0, q = 0.0;
  for (int i = 0; i < splits; ++i) {
    const float* p = part + (int64_t)i * 3 * C;
    s += (double)p[c];
    q += (double)p[C + c];
  }
  float rs = rstd[c];
  float g = gamma ? gamma[c] : 1.f;
  if (dbeta) dbeta[c] = (float)s;
  if (dgamma) dgamma[c] = (float)(q * (double)rs);
  c1[c] = g * rs;
  c2[c] = (float)(s / (double)M);
  c3[c] = (float)(q * (double)rs * (double)rs / (double)M);
}

template <typename T, int V, bool RELU, bool DRES>
__global__ void bn_bwd_dx_k(const T* __restrict__ x, const T* __restrict__ y,
                            const T* __restrict__ dy, const float* __restrict__ mean,
                            const float* __restrict__ c1, const float* __restrict__ c2,
                            const float* __restrict__ c3, T* __restrict__ dx,
                            T* __restrict__ dres, int64_t total_vec, int Cv) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total_vec;
       i += (int64_t)gridDim.x * blockDim.x) {
    int cv = (int)(i % Cv);
    Pack<T, V> a = ld_pack<T, V>(x + i * V), g = ld_pack<T, V>(dy + i * V), o, m, dr;
    if (RELU) m = ld_pack<T, V>(y + i * V);
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const int c = cv * V + j;
      float gf = to_f(g.v[j]);
      if (RELU && !(to_f(m.v[j]) > 0.f)) gf = 0.f;
      if (DRES) dr.v[j] = from_f<T>(gf);
      float xc = to_f(a.v[j]) - mean[c];
      o.v[j] = from_f<T>(c1[c] * (gf - c2[c] - xc * c3[c]));
    }
    st_pack<T, V>(dx + i * V, o);
    if (DRES) st_pack<T, V>(dres + i * V, dr);
  }
}

constexpr int kMaxSplits = 4096;
static size_t bn_ws_bytes(int C) {
  // partials: up to kMaxSplits x 3 x C floats is far too pessimistic; geometry caps
  // splits*cblocks near 8 waves, so bound by (8*SMs + 1) blocks of at least 1 channel-vector:
  // splits <= min(kMaxSplits, 8*SMs) -> 3*C*splits floats.
  int64_t splits = (int64_t)sm_count() * 8;
  if (splits > kMaxSplits) splits = kMaxSplits;
  return (size_t)(3 * (int64_t)C * splits + 5 * (int64_t)C) * sizeof(float) + 256;
}

static int validate_bn(const cnb_bn_desc* d, const char* who) {
  CNB_REQUIRE(d != nullptr, CNB_ERR_INVALID, "%s: null descriptor", who);
  CNB_REQUIRE(d->M > 0 && d->C > 0 && valid_dtype(d->dtype), CNB_ERR_INVALID,
              "%s: bad descriptor (M=%lld C=%d dtype=%d)", who, (long long)d->M, d->C, d->dtype);
  return CNB_OK;
}

}  // namespace cnb

using namespace cnb;

extern "C" {

size_t cnb_bn_workspace_bytes(const cnb_bn_desc* d) {
  if (!d || d->C <= 0) return 0;
  return bn_ws_bytes(d->C);
}

int cnb_bn_fwd_train(const cnb_bn_desc* d, const void* x, const void* residual, const float* gamma,
                     const float* beta, float* running_mean, float* running_var, void* y,
                     float* save_mean, float* save_rstd, void* workspace, size_t workspace_bytes,
                     void* stream) {
  CNB_TRY(validate_bn(d, "cnb_bn_fwd_train"));
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
  BnGeom g = bn_geom(M, C, v);
  float* part = (float*)workspace;
  float* scale = part + (size_t)3 * C * g.splits;
  float* shift = scale + C;
  dim3 grid(g.cblocks, g.splits), block(g.cx, g.ry);
  size_t smem = (size_t)g.ry * 2 * g.cx * v * sizeof(float);
  CNB_DISPATCH_TV(d->dtype, v, (bn_stats_k<T, V><<<grid, block, smem, st>>>((const T*)x, part, M, C,
                                                                           g.rows_per_split)));
  CNB_TRY(check_launch("bn_stats"));
  bn_finalize_fwd_k<<<(unsigned)ceil_div(C, 128), 128, 0, st>>>(
      part, g.splits, M, g.rows_per_split, C, gamma, beta, running_mean, running_var, save_mean,
      save_rstd, scale, shift, d->eps, d->momentum);
  CNB_TRY(check_launch("bn_finalize_fwd"));
  int64_t total = M * (C / v);
  int gr = grid_for(total, kBnThreads);
#define APPLY(R, A)                                                                              \
  CNB_DISPATCH_TV(d->dtype, v, (bn_apply_k<T, V, R, A><<<gr, kBnThreads, 0, st>>>(               \
                                   (const T*)x, (const T*)residual, scale, shift, (T*)y, total, \
                                   C / V)))
  if (relu && add) APPLY(true, true);
  else if (relu) APPLY(true, false);
  else if (add) APPLY(false, true);
  else APPLY(false, false);
  return check_launch("bn_apply");
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
  int64_t total = d->M * (C / v);
  int gr = grid_for(total, kBnThreads);
  if (relu && add) APPLY(true, true);
  else if (relu) APPLY(true, false);
  else if (add) APPLY(false, true);
  else APPLY(false, false);
#undef APPLY
  return check_launch("bn_apply(eval)");
}

int cnb_bn_bwd(const cnb_bn_desc* d, const void* x, const void* y, const void* dy,
               const float* gamma, const float* save_mean, const float* save_rstd, void* dx,
               void* d_residual, float* dgamma, float* dbeta, void* workspace,
               size_t workspace_bytes, void* stream) {
  CNB_TRY(validate_bn(d, "cnb_bn_bwd"));
  CNB_REQUIRE(x && dy && save_mean && save_rstd && dx && workspace, CNB_ERR_INVALID,
              "cnb_bn_bwd: null pointer");
  const bool relu = d->flags & CNB_BN_RELU;
  CNB_REQUIRE(!relu || y, CNB_ERR_INVALID, "cnb_bn_bwd: CNB_BN_RELU needs the forward output y");
  CNB_REQUIRE(workspace_bytes >= bn_ws_bytes(d->C), CNB_ERR_WORKSPACE,
              "cnb_bn_bwd: workspace %zu < %zu", workspace_bytes, bn_ws_bytes(d->C));
  cudaStream_t st = (cudaStream_t)stream;
  const int C = d->C;
  const int64_t M = d->M;
  int v = pick_vec(C, dtype_size(d->dtype), x, dy, dx, relu ? y : nullptr);
  if (d_residual)
    while (v > 1 && (uintptr_t)d_residual % (v * dtype_size(d->dtype)) != 0) v >>= 1;
  BnGeom g = bn_geom(M, C, v);
  float* part = (float*)workspace;
  float* c1 = part + (size_t)3 * C * g.splits;
  float* c2 = c1 + C;
  float* c3 = c2 + C;
  dim3 grid(g.cblocks, g.splits), block(g.cx, g.ry);
  size_t smem = (size_t)g.ry * 2 * g.cx * v * sizeof(float);
  if (relu)
    CNB_DISPATCH_TV(d->dtype, v, (bn_bwd_reduce_k<T, V, true><<<grid, block, smem, st>>>(
                                     (const T*)x, (const T*)y, (const T*)dy, save_mean, part, M, C,
                                     g.rows_per_split)));
  else
    CNB_DISPATCH_TV(d->dtype, v, (bn_bwd_reduce_k<T, V, false><<<grid, block, smem, st>>>(
                                     (const T*)x, (const T*)y, (const T*)dy, save_mean, part, M, C,
                                     g.rows_per_split)));
  CNB_TRY(check_launch("bn_bwd_reduce"));
  bn_finalize_bwd_k<<<(unsigned)ceil_div(C, 128), 128, 0, st>>>(part, g.splits, M, C, gamma,
                                                                save_rstd, dgamma, dbeta, c1, c2, c3);
  CNB_TRY(check_launch("bn_finalize_bwd"));
  int64_t total = M * (C / v);
  int gr = grid_for(total, kBnThreads);
#define DX(R, D)                                                                                  \
  CNB_DISPATCH_TV(d->dtype, v, (bn_bwd_dx_k<T, V, R, D><<<gr, kBnThreads, 0, st>>>(               \
                                   (const T*)x, (const T*)y, (const T*)dy, save_mean, c1, c2, c3, \
                                   (T*)dx, (T*)d_residual, total, C / V)))
  if (relu && d_residual) DX(true, true);
  else if (relu) DX(true, false);
  else if (d_residual) DX(false, true);
  else DX(false, false);
#undef DX
  return check_launch("bn_bwd_dx");
}

}  // extern "C"
