// conv_tc.cu — tcgen05 / TMEM / TMA implicit-GEMM convolution for sm_100a (bf16 in, fp32 accumulate).
// Replaces aten::convolution / convolution_backward for the dense (groups == 1) layers the
// reference builds at src/basemodel.py:974 (k = 1, 3, 5, 7; strided; dilated).
//
// Two kernels:
//  * conv_fprop_tc_k : D[M = pixels, N = out-channels] = sum over (tap, channel-block) of
//        A[128 pixels x 64 ch]  (one im2col-mode TMA load per (tap, channel block); zero-fill
//                                gives the padding)          -> K-major SW128 smem tile
//        B[BLOCK_N x 64 ch]     (weights, KRSC viewed as {C, R*S, K}) -> K-major SW128 smem tile
//    accumulated by tcgen05.mma (M=128, N=BLOCK_N, K=16) into TMEM; epilogue tcgen05.ld ->
//    (+bias, ReLU) -> bf16 -> swizzled smem -> TMA store into the NHWC output (a plain
//    [M, K] matrix).  The same kernel computes data-grad: stride-1 dgrad is a forward conv of dy
//    with the transposed filter; strided dgrad is decomposed by output parity class into stride-1
//    sub-convolutions whose epilogues scatter their rows to the class's pixels of dx.
//  * conv_wgrad_tc_k : per tap, D[M = out-channels(128), N = in-channels(NT)] += dy^T x_tap with
//    the GEMM-K dimension = output pixels; both operands are "MN-major" smem tiles
//    ([64 pixels x 64 ch] TMA boxes), one TMEM accumulator per tap, split-K over pixel ranges
//    into fp32 partials summed by a second (deterministic) kernel.
//
// Warp roles: warp 0 = TMA producer, warp 1 = TMEM allocator + single-thread MMA issuer,
// warps 2..5 = epilogue (warp_id % 4 selects the TMEM lane quarter).  mbarrier full/empty rings
// between producer and MMA; tcgen05.commit releases smem slots and signals the epilogue.
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"
#include "sm100.cuh"

namespace cnb {
using namespace sm100;

constexpr int kMaxTaps = 49;
constexpr int kTcThreads = 192;
constexpr int kPersistThreads = 320;  // persistent fprop: producer, MMA and 8 epilogue warps
constexpr int TILE_M = 128;   // pixels per CTA tile (fprop) / out-channels per CTA tile (wgrad)
constexpr int TILE_K = 64;    // bf16 elements per 128-byte swizzle row
constexpr int A_TILE_BYTES = TILE_M * TILE_K * 2;  // 16 KB

// ------------------------------------------------------------------------------------------
// driver entry points (no link-time dependency on libcuda)
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*PFN_encodeIm2col)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                     const cuuint64_t*, const cuuint64_t*, const int*, const int*,
                                     cuuint32_t, cuuint32_t, const cuuint32_t*,
                                     CUtensorMapInterleave, CUtensorMapSwizzle,
                                     CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled g_encodeTiled = nullptr;
static PFN_encodeIm2col g_encodeIm2col = nullptr;
static int g_driver_version = 0;

static int load_driver_fns() {
  if (g_encodeTiled && g_encodeIm2col) return CNB_OK;
  cudaDriverEntryPointQueryResult qres;
  void* fn = nullptr;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  CNB_REQUIRE(e == cudaSuccess && fn && qres == cudaDriverEntryPointSuccess, CNB_ERR_CUDA,
              "cuTensorMapEncodeTiled not available: %s", cudaGetErrorString(e));
  g_encodeTiled = (PFN_encodeTiled)fn;
  fn = nullptr;
  e = cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fn, cudaEnableDefault, &qres);
  CNB_REQUIRE(e == cudaSuccess && fn && qres == cudaDriverEntryPointSuccess, CNB_ERR_CUDA,
              "cuTensorMapEncodeIm2col not available: %s", cudaGetErrorString(e));
  g_encodeIm2col = (PFN_encodeIm2col)fn;
  cudaDriverGetVersion(&g_driver_version);
  return CNB_OK;
}

// The encode calls are DRIVER API: they need a context current on the calling thread.  A thread that
// has only ever made calls that do not bind one (the autograd engine's device thread can reach a
// convolution that way) gets CUDA_ERROR_INVALID_CONTEXT; a runtime call that needs the device binds
// the primary context, after which the encode is retried once.  Never taken on a thread that has
// launched anything.
template <typename F>
static CUresult encode_with_context(const void* base, F&& encode) {
  CUresult r = encode();
  if (r == CUDA_ERROR_INVALID_CONTEXT) {
    cudaPointerAttributes a;   // the device that owns the operand, not whatever device 0 is
    if (cudaPointerGetAttributes(&a, base) == cudaSuccess && a.type == cudaMemoryTypeDevice)
      cudaSetDevice(a.device);
    if (cudaFree(nullptr) == cudaSuccess) r = encode();
  }
  return r;
}

// 2-D / 3-D tiled bf16 map with 128B swizzle; dims/box innermost first; OOB reads give zeros.
static int make_tiled_map(CUtensorMap* m, const void* base, int rank, const uint64_t* dims,
                          const uint64_t* strides_bytes /* rank-1 */, const uint32_t* box,
                          const char* what) {
  cuuint64_t gdim[5], gstr[4];
  cuuint32_t b[5], es[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    b[i] = box[i];
    es[i] = 1;
  }
  for (int i = 0; i < rank - 1; ++i) gstr[i] = strides_bytes[i];
  CUresult r = encode_with_context(base, [&] {
    return g_encodeTiled(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank,
                         const_cast<void*>(base), gdim, gstr, b, es,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  });
  CNB_REQUIRE(r == CUDA_SUCCESS, CNB_ERR_CUDA, "cuTensorMapEncodeTiled(%s) failed: CUresult %d", what,
              (int)r);
  return CNB_OK;
}

// im2col-mode bf16 map over an NHWC activation tensor {C, W, H, N}.
static int make_im2col_map(CUtensorMap* m, const void* base, int N, int H, int W, int C,
                           int lower_w, int lower_h, int upper_w, int upper_h, int trav_w,
                           int trav_h, int pixels, const char* what) {
  cuuint64_t gdim[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t gstr[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  int lower[2] = {lower_w, lower_h};
  int upper[2] = {upper_w, upper_h};
  cuuint32_t es[4] = {1, (cuuint32_t)trav_w, (cuuint32_t)trav_h, 1};
  CUresult r = encode_with_context(base, [&] {
    return g_encodeIm2col(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), gdim,
                          gstr, lower, upper, (cuuint32_t)TILE_K, (cuuint32_t)pixels, es,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  });
  CNB_REQUIRE(r == CUDA_SUCCESS, CNB_ERR_CUDA,
              "cuTensorMapEncodeIm2col(%s) failed: CUresult %d (N=%d H=%d W=%d C=%d lower=%d,%d "
              "upper=%d,%d trav=%d,%d)",
              what, (int)r, N, H, W, C, lower_w, lower_h, upper_w, upper_h, trav_w, trav_h);
  // Driver workaround also applied by CUTLASS (cute/atom/copy_traits_sm90_im2col.hpp): for small
  // tensors, drivers <= 13.1 set a descriptor bit that must be cleared.
  if (g_driver_version <= 13010 && (size_t)N * H * W * C * 2 < 131072)
    reinterpret_cast<uint64_t*>(m)[1] &= ~(1ull << 21);
  return CNB_OK;
}

// ==========================================================================================
// forward / data-grad kernel
// ==========================================================================================
// internal epilogue flag (next to CNB_CONV_BIAS / CNB_CONV_RELU): out = acc + addend
constexpr int kConvAccum = 1 << 8;

struct FpropParams {
  CUtensorMap tmA;  // activations: im2col {C,W,H,N} (box 64ch x 128 pixels) or tiled {C, M}
  CUtensorMap tmB;  // weights {Cg, RS, Kout}, box {64, 1, BLOCK_N}
  CUtensorMap tmD;  // output {Kout, M}, box {64, 128}
  const float* bias;
  const bf16* addend;  // optional [M][Kout]: added to the accumulator before the bf16 rounding (kConvAccum)
  const uint8_t* addend_mask;  // EPI == 4: bit (m * Kout + c) selects which addend elements count (1-bit ReLU mask)
  // data-grad whose output is the gradient of a BatchNorm2d + ReLU output: `stats` then receives that
  // BatchNorm's BACKWARD sums per CTA row — (sum dy', sum dy' * (x - mean)) with dy' = the stored
  // output of this kernel where the ReLU bit is set — from bn_x (the BatchNorm's input, [M][Kout]),
  // bn_mask (its 1-bit ReLU mask) and bn_mean
  const bf16* bn_x;
  const uint8_t* bn_mask;
  const float* bn_mean;
  float* stats;     // optional [persistent CTAs / n_tiles][2][Kout] column sums / sums of squares of the stored output
  int M;            // output pixels in this launch
  int outP, outQ;   // output grid per image (pixel index -> (n, i, j))
  int lower_h, lower_w, trav_h, trav_w;
  int num_cblk;     // ceil(C / 64)
  int num_taps;
  int n_tiles;      // ceil(Kout / BLOCK_N)
  int m_tiles;      // ceil(M / 128)
  int grp_out, grp_in;  // grouped conv: output / reduction channels per group (0 = dense)
  int Kout;
  int flags;
  // EPI == 3 (parity class of a strided data-grad): output pixel (n, i, j) of this launch is pixel
  // (n, i * sc_sh + sc_a, j * sc_sw + sc_b) of the [N, sc_H, sc_W, Kout] tensor `scatter`
  bf16* scatter;
  int sc_H, sc_W, sc_sh, sc_sw, sc_a, sc_b;
  int16_t off_h[kMaxTaps], off_w[kMaxTaps];
  uint8_t widx[kMaxTaps];
};

// ------------------------------------------------------------------------------------------
// Persistent variant: one CTA per SM.  CTA b owns ONE n-tile (b % n_tiles) and walks the m-tiles
// b / n_tiles, + grid / n_tiles, ... — CTAs that run side by side share the activation tile through
// L2, and because a CTA's output columns never change, the BatchNorm column sums of all its tiles
// accumulate in registers and leave the kernel as ONE partial row per CTA.  The TMA producer runs
// ahead across tile boundaries (the smem ring never drains), the accumulator is double-buffered in
// TMEM (2 x BLOCK_N columns) and the epilogue of tile i — tcgen05.ld, bf16 pack, swizzled staging,
// TMA store, column sums — overlaps the MMAs of tile i+1.
//   PLAIN : no bias / ReLU in the epilogue (every conv that feeds a BatchNorm, and data-grad);
//           the generic variant reads the flags at run time.
//   grouped convolutions (p.grp_out > 0): the n-tile selects the input-channel slice it reads
//           (a_c0) and the filter operand is the block-diagonal expansion built by
//           expand_grouped_filter_k, see "grouped convolutions" below.
//   BRES  : the CTA's whole filter slice (its n-tile x every tap x every channel block, at most
//           kBresIters tiles) is loaded ONCE and stays in shared memory; the ring then carries the
//           activation tile only.  A persistent CTA never changes its n-tile, and a 64-channel
//           3x3 layer re-fetched 8 KB of filter next to every 16 KB activation tile — a third of
//           the L2 -> SM traffic that bounds these layers (~50 B / clk / SM, ncu r02c).
constexpr int kBresIters = 9;
//   BNRED : data-grad into a BatchNorm + ReLU output: the column-sum pass forms that BatchNorm's
//           backward sums (see FpropParams::bn_x) instead of forward statistics.  A template
//           parameter, not a run-time branch: the extra live values cost the N = 256 variants their
//           last registers (168-register cap at 320 threads), which the forward kernels must not pay.
template <int BLOCK_N, int STAGES, bool IM2COL, int EPI, bool BRES = false, bool BNRED = false>
__global__ void __launch_bounds__(kPersistThreads, 1)
conv_fprop_tc_persist_k(const __grid_constant__ FpropParams p) {
  // EPI: 0 = bias / ReLU by flags, 1 = plain store, 2 = plain + addend (data-grad + skip gradient),
  //      3 = plain, rows scattered to the strided pixel positions of a larger tensor (one parity
  //          class of a strided data-grad writes dx directly: no compact copy, no interleave pass)
  //      4 = 2 with the addend masked by a 1-bit mask: the skip gradient of a residual block is the
  //          block-output gradient where the output ReLU was active, so BatchNorm backward does not
  //          have to materialise it (src/resnet.py:191-194 `output += identity; F.relu(output)`)
  constexpr bool PLAIN = EPI != 0;   // rows past the end of the tensor stay zero (no bias added)
  constexpr int B_TILE_BYTES = BLOCK_N * TILE_K * 2;
  constexpr int STAGE_BYTES = BRES ? A_TILE_BYTES : A_TILE_BYTES + B_TILE_BYTES;
  constexpr int BRES_BYTES = BRES ? kBresIters * B_TILE_BYTES : 0;
  constexpr int OUT_BYTES = TILE_M * BLOCK_N * 2;
  constexpr uint32_t TMEM_COLS = 2 * BLOCK_N < 32 ? 32 : 2 * BLOCK_N;

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~(uintptr_t)1023);
  uint8_t* bres = smem + STAGES * STAGE_BYTES;   // resident filter tiles, k-block order (BRES)
  uint8_t* stage_out = bres + BRES_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(stage_out + OUT_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;   // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;   // [2]
  uint64_t* bres_bar = tmem_empty_bar + 2;        // [1]
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bres_bar + 1);

  const int warp = threadIdx.x / 32;
  const int lane = threadIdx.x % 32;
  const int num_iters = p.num_taps * p.num_cblk;
  const int n_idx = blockIdx.x % p.n_tiles;
  const int m_first = blockIdx.x / p.n_tiles;
  const int m_step = gridDim.x / p.n_tiles;
  const int n0 = n_idx * BLOCK_N;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmA);
    tma_prefetch_desc(&p.tmB);
    tma_prefetch_desc(&p.tmD);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tmem_full_bar[b], 1);
      mbar_init(&tmem_empty_bar[b], kPersistThreads - 64);
    }
    mbar_init(bres_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_ptr, TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  // Both issuing roles run their loops with the WHOLE warp (uniform control flow, every lane polls
  // the mbarriers) and hand only the TMA / tcgen05 instructions to one lane chosen by elect.sync.
  // Addresses, coordinates and descriptors are then warp-uniform values in uniform registers; under
  // `if (lane == 0)` the compiler has to treat them as per-thread values and wraps every UTMALDG /
  // UTCHMMA in an ELECT / R2UR.BROADCAST / BRA.U.ANY loop (~30 dependent instructions per MMA: the
  // issuing warp, not the tensor pipe, bounded every layer — ncu r02b: 19 % pipe-active on 64
  // -channel 3x3 layers with 84 % of the MMA warp's samples inside its own issue loop).
  if (warp == 0) {
    // ===================== TMA producer =====================
    // every per-tile / per-tap value goes through uni(): the compiler then keeps it in a uniform
    // register for the whole inner loop instead of moving it there (R2UR) once per k-block
    const int a_c0 = p.grp_out > 0 ? (n0 / p.grp_out) * p.grp_in : 0;
    const uint32_t smem0 = smem_u32(smem), full0 = smem_u32(full_bar), empty0 = smem_u32(empty_bar);
    uint32_t s = 0, ph = 1;   // ring slot and the parity a free slot shows
    if (BRES) {
      if (elect_one()) {
        const uint32_t bb = smem_u32(bres_bar);
        mbar_expect_tx_u32(bb, (uint32_t)num_iters * B_TILE_BYTES);
        uint32_t dst = smem_u32(bres);
        for (int tap = 0; tap < p.num_taps; ++tap)
          for (int cb = 0; cb < p.num_cblk; ++cb, dst += B_TILE_BYTES)
            tma_load_3d_u32(&p.tmB, bb, dst, cb * TILE_K, (int)p.widx[tap], n0);
      }
      __syncwarp();
    }
    for (int mt = m_first; mt < p.m_tiles; mt += m_step) {
      const int m0 = mt * TILE_M;
      int bn = 0, bh = 0, bw = 0;
      if (IM2COL) {
        int j = m0 % p.outQ;
        int t = m0 / p.outQ;
        int i = t % p.outP;
        bn = uni(t / p.outP);
        bh = uni(p.lower_h + i * p.trav_h);
        bw = uni(p.lower_w + j * p.trav_w);
      }
      for (int tap = 0; tap < p.num_taps; ++tap) {
        const int widx = uni((int)p.widx[tap]);
        const uint32_t offs = uni((uint32_t)(uint16_t)p.off_w[tap] | ((uint32_t)(uint16_t)p.off_h[tap] << 16));
        for (int cb = 0; cb < p.num_cblk; ++cb) {
          mbar_wait_u32(empty0 + s * 8, ph);
          if (elect_one()) {
            const uint32_t sa = smem0 + s * STAGE_BYTES;
            const uint32_t fb = full0 + s * 8;
            mbar_expect_tx_u32(fb, STAGE_BYTES);
            if (IM2COL)
              tma_load_im2col_4d_u32(&p.tmA, fb, sa, a_c0 + cb * TILE_K, bw, bh, bn, (uint16_t)(offs & 0xffff),
                                     (uint16_t)(offs >> 16));
            else
              tma_load_2d_u32(&p.tmA, fb, sa, a_c0 + cb * TILE_K, m0);
            if (!BRES) tma_load_3d_u32(&p.tmB, fb, sa + A_TILE_BYTES, cb * TILE_K, widx, n0);
          }
          __syncwarp();
          if (++s == STAGES) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc = make_idesc_bf16(TILE_M, BLOCK_N, 0, 0);
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t full0 = smem_u32(full_bar), empty0 = smem_u32(empty_bar);
    const uint32_t tfull0 = smem_u32(tmem_full_bar), tempty0 = smem_u32(tmem_empty_bar);
    // descriptor low words (start address >> 4) of slot 0's A / B tiles; + STAGE_BYTES/16 per slot,
    // + 2 per 16-element k-step.  The high word (SBO 1024, version 1, SWIZZLE_128B) is a constant.
    const uint32_t a_lo0 = (smem_u32(smem) >> 4);
    const uint32_t b_lo0 = BRES ? (smem_u32(bres) >> 4) : ((smem_u32(smem) + A_TILE_BYTES) >> 4);
    if (BRES) mbar_wait_u32(smem_u32(bres_bar), 0);   // the filter slice has landed (once per kernel)
    constexpr uint32_t DESC_HI = (1024u >> 4) | (1u << 14) | (2u << 29);
    uint32_t s = 0, ph = 0, lt = 0;
    for (int mt = m_first; mt < p.m_tiles; mt += m_step, ++lt) {
      const uint32_t buf = lt & 1;
      mbar_wait_u32(tempty0 + buf * 8, ((lt >> 1) & 1) ^ 1);  // epilogue has drained this buffer
      tc_fence_after();
      const uint32_t acc = tmem_u + buf * BLOCK_N;
      for (int k_it = 0; k_it < num_iters; ++k_it) {
        mbar_wait_u32(full0 + s * 8, ph);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t a_lo = a_lo0 + s * (STAGE_BYTES >> 4);
          const uint32_t b_lo = b_lo0 + (BRES ? (uint32_t)k_it * (B_TILE_BYTES >> 4) : s * (STAGE_BYTES >> 4));
#pragma unroll
          for (int k = 0; k < TILE_K / 16; ++k)
            umma_bf16_lohi(acc, a_lo + 2 * k, b_lo + 2 * k, DESC_HI, idesc, (k_it | k) != 0 ? 1u : 0u);
          umma_commit_u32(empty0 + s * 8);
          if (k_it == num_iters - 1) umma_commit_u32(tfull0 + buf * 8);
        }
        __syncwarp();
        if (++s == STAGES) { s = 0; ph ^= 1; }
      }
    }
  } else {
    // ===================== epilogue (warps 2..9) =====================
    // A warp can only read the TMEM lane quarter (warp % 4); two warps share each quarter and
    // take one half of the tile's columns each, so the drain keeps up with one-k-block tiles.
    const int quarter = warp % 4;
    const int chalf = (warp - 2) / 4;
    constexpr int HALF_N = BLOCK_N / 2;
    constexpr int NCH = HALF_N / 32;          // 32-column chunks per warp: 1, 2 or 4
    const int row = quarter * 32 + lane;
    // TMA stores: issued, committed and waited for by the lane elect.sync picks in warp 2 (the
    // election is deterministic for a given member mask, so it is the same thread every time)
    const bool store_warp = warp == 2;
    const uint32_t st_base = smem_u32(stage_out);
    const uint32_t row_addr = st_base + row * 128;
    const uint32_t row7 = (uint32_t)(row & 7);
    const uint32_t lane_taddr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    // BatchNorm column sums: thread (pr, rg) owns the column pair 2*pr of row group rg
    const int et = (int)threadIdx.x - 64;
    constexpr int PAIRS = BLOCK_N / 2;
    constexpr int RG = 256 / PAIRS;           // row groups: 2, 4 or 8
    constexpr int RPG = TILE_M / RG;          // rows per group: 64, 32 or 16
    const int pr = et % PAIRS, rg = et / PAIRS;
    const int sc = pr * 2;
    const uint32_t col_addr = st_base + (sc / 64) * (TILE_M * 128) + (sc % 8) * 2 + rg * RPG * 128;
    uint32_t soff[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) soff[i] = i * 128 + ((((sc % 64) / 8) ^ i) << 4);  // RPG % 8 == 0
    float as0 = 0.f, as1 = 0.f, aq0 = 0.f, aq1 = 0.f;
    const bool want_stats = p.stats != nullptr;
    float bsum[BNRED ? 8 : 1], bsq[BNRED ? 8 : 1];   // BNRED: this thread's 8 columns, all its tiles
#pragma unroll
    for (int j = 0; j < (BNRED ? 8 : 1); ++j) bsum[j] = bsq[j] = 0.f;

    uint32_t lt = 0;
    for (int mt = m_first; mt < p.m_tiles; mt += m_step, ++lt) {
      const int m0 = mt * TILE_M;
      const uint32_t buf = lt & 1;
      if (EPI == 2 || EPI == 4) {
        // the addend row segment this thread will add comes from HBM (a 100-400 MB tensor): ask for
        // it now, while the accumulator is still filling, so the loads below hit L2
        const int64_t grow = (int64_t)m0 + row;
        if (grow < p.M) {
          const bf16* ap = p.addend + grow * p.Kout + n0 + chalf * HALF_N;
#pragma unroll
          for (int b = 0; b < (HALF_N * 2 + 127) / 128; ++b)
            if (n0 + chalf * HALF_N + b * 64 < p.Kout) prefetch_l2(ap + b * 64);
        }
      }
      if (BNRED) {
        // the BatchNorm input rows of this tile come from HBM (written in the forward pass): ask for
        // their lines now — the column-sum pass below then finds them in L2.  256 threads, one
        // 128-byte line each per round: line l = (row l / segs, 64-channel segment l % segs).
        constexpr int SEGS = BLOCK_N / 64;
#pragma unroll
        for (int l = et; l < TILE_M * SEGS; l += 256) {
          const int r = l / SEGS, sg = l % SEGS;
          if (m0 + r < p.M && n0 + sg * 64 < p.Kout)
            prefetch_l2(p.bn_x + ((int64_t)m0 + r) * p.Kout + n0 + sg * 64);
        }
      }
      mbar_wait(&tmem_full_bar[buf], (lt >> 1) & 1);
      tc_fence_after();
      // the previous tile's TMA store must have finished READING the staging buffer
      if (EPI != 3 && store_warp) {
        if (elect_one()) tma_store_wait_read0();
        __syncwarp();
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      const uint32_t tcol = lane_taddr + buf * BLOCK_N + (uint32_t)(chalf * HALF_N);
      uint32_t v[2][32];
      tmem_ld_32x32b_x32(tcol, v[0]);
#pragma unroll
      for (int ch = 0; ch < NCH; ++ch) {
        tmem_ld_wait();
        if (ch + 1 < NCH) tmem_ld_32x32b_x32(tcol + (uint32_t)((ch + 1) * 32), v[(ch + 1) & 1]);
        const uint32_t(&cur)[32] = v[ch & 1];
        const int c0 = chalf * HALF_N + ch * 32;
        const uint32_t box = row_addr + (uint32_t)(c0 / 64) * (TILE_M * 128);
        const uint32_t chunk0 = (uint32_t)((c0 % 64) / 8);
        // skip-gradient accumulation: this lane's 32 columns of the addend row, 4 x 16 bytes, all
        // requested before the first use (rows past M / columns past Kout are never stored)
        uint4 add4[4];
        if (EPI == 2 || EPI == 4) {
          const int64_t grow = (int64_t)m0 + row;
          uint32_t mbits = 0xffffffffu;   // one bit per column of this 32-column chunk
          if (EPI == 4 && grow < p.M && n0 + c0 < p.Kout)
            mbits = __ldg(reinterpret_cast<const uint32_t*>(p.addend_mask + ((grow * p.Kout + n0 + c0) >> 3)));
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int col = n0 + c0 + j * 8;
            add4[j] = (grow < p.M && col < p.Kout)
                          ? __ldg(reinterpret_cast<const uint4*>(p.addend + grow * p.Kout + col))
                          : make_uint4(0u, 0u, 0u, 0u);
            if (EPI == 4) {
              const uint32_t mb = (mbits >> (8 * j)) & 0xffu;
              add4[j].x &= ((mb & 1u) ? 0x0000ffffu : 0u) | ((mb & 2u) ? 0xffff0000u : 0u);
              add4[j].y &= ((mb & 4u) ? 0x0000ffffu : 0u) | ((mb & 8u) ? 0xffff0000u : 0u);
              add4[j].z &= ((mb & 16u) ? 0x0000ffffu : 0u) | ((mb & 32u) ? 0xffff0000u : 0u);
              add4[j].w &= ((mb & 64u) ? 0x0000ffffu : 0u) | ((mb & 128u) ? 0xffff0000u : 0u);
            }
          }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint32_t packed[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            float f0 = __uint_as_float(cur[j * 8 + e * 2 + 0]);
            float f1 = __uint_as_float(cur[j * 8 + e * 2 + 1]);
            if (EPI == 2 || EPI == 4) {
              const uint32_t a2 = e == 0 ? add4[j].x : (e == 1 ? add4[j].y : (e == 2 ? add4[j].z : add4[j].w));
              f0 += __uint_as_float(a2 << 16);
              f1 += __uint_as_float(a2 & 0xffff0000u);
            }
            if (!PLAIN) {
              const int col = n0 + c0 + j * 8 + e * 2;
              if (p.flags & CNB_CONV_BIAS) {
                f0 += (col < p.Kout) ? p.bias[col] : 0.f;
                f1 += (col + 1 < p.Kout) ? p.bias[col + 1] : 0.f;
              }
              if (p.flags & CNB_CONV_RELU) {
                f0 = (f0 > 0.f || f0 != f0) ? f0 : 0.f;
                f1 = (f1 > 0.f || f1 != f1) ? f1 : 0.f;
              }
            }
            __nv_bfloat162 h = __floats2bfloat162_rn(f0, f1);
            packed[e] = *reinterpret_cast<uint32_t*>(&h);
          }
          sts_v4(box + (((chunk0 + j) ^ row7) << 4), packed[0], packed[1], packed[2], packed[3]);
        }
      }
      // accumulator buffer fully read: hand it back to the MMA warp
      tc_fence_before();
      mbar_arrive(&tmem_empty_bar[buf]);
      fence_proxy_async_smem();
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (EPI == 3) {
        // staged tile -> dx: a thread moves one 16-byte chunk of 4 pixel rows per 64-channel box; the 8
        // lanes of a pixel write its 128 contiguous bytes, a warp 4 full lines.  The next tile's
        // writers meet every reader at the bar.sync above before they touch the staging buffer.
        const int chunk = et & 7, r0 = et >> 3;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int r = r0 + 32 * q;
          const int m = m0 + r;
          if (m < p.M) {
            const int j = m % p.outQ;
            const int t = m / p.outQ;
            const int i = t % p.outP;
            const int n = t / p.outP;
            const int64_t pix = ((int64_t)n * p.sc_H + (i * p.sc_sh + p.sc_a)) * p.sc_W + (j * p.sc_sw + p.sc_b);
            bf16* dst = p.scatter + pix * p.Kout + n0 + chunk * 8;
#pragma unroll
            for (int b = 0; b < BLOCK_N / 64; ++b) {
              if (n0 + b * 64 + chunk * 8 < p.Kout) {
                const uint4 v = lds_v4(st_base + b * (TILE_M * 128) + r * 128 + ((chunk ^ (r & 7)) << 4));
                *reinterpret_cast<uint4*>(dst + b * 64) = v;
              }
            }
          }
        }
      } else if (store_warp) {
        if (elect_one()) {
#pragma unroll
          for (int b = 0; b < BLOCK_N / 64; ++b) {
            if (n0 + b * 64 < p.Kout)
              tma_store_2d_u32(&p.tmD, st_base + b * (TILE_M * 128), n0 + b * 64, m0);
          }
          tma_store_commit();
        }
        __syncwarp();
      }
      if (BNRED) {
        // backward sums of the BatchNorm whose output gradient this tile is: the staged (stored) dy,
        // the ReLU bits and the BatchNorm input x of the same elements.  A thread owns 8 adjacent
        // columns (one 16-byte vector of x, of the staged tile, one mask byte per row) and every
        // RGB-th row: 4 rows' loads are in flight together, a warp's lanes read whole rows.
        constexpr int CG = BLOCK_N / 8;        // column groups: 8, 16, 32
        constexpr int RGB = 256 / CG;          // row groups:   32, 16, 8
        constexpr int RPB = TILE_M / RGB;      // rows per thread and tile: 4, 8, 16
        const int cg = et % CG, rgb = et / CG;
        const int col8 = n0 + cg * 8;
        if (col8 < p.Kout) {
          const float4 mlo = __ldg(reinterpret_cast<const float4*>(p.bn_mean + col8));
          const float4 mhi = __ldg(reinterpret_cast<const float4*>(p.bn_mean + col8 + 4));
          const float mu[8] = {mlo.x, mlo.y, mlo.z, mlo.w, mhi.x, mhi.y, mhi.z, mhi.w};
          const uint32_t sbase = st_base + (uint32_t)(cg / 8) * (TILE_M * 128);
#pragma unroll
          for (int b4 = 0; b4 < RPB / 4; ++b4) {
            uint4 xv[4];
            uint32_t mb[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int r = rgb + RGB * (b4 * 4 + i);
              const bool ok = m0 + r < p.M;
              const int64_t e = ((int64_t)m0 + r) * p.Kout + col8;
              xv[i] = ok ? __ldg(reinterpret_cast<const uint4*>(p.bn_x + e)) : make_uint4(0u, 0u, 0u, 0u);
              mb[i] = ok ? (uint32_t)__ldg(p.bn_mask + (e >> 3)) : 0u;
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int r = rgb + RGB * (b4 * 4 + i);
              const uint4 dv = lds_v4(sbase + (uint32_t)r * 128 + ((uint32_t)((cg & 7) ^ (r & 7)) << 4));
              const uint32_t dw[4] = {dv.x, dv.y, dv.z, dv.w}, xw[4] = {xv[i].x, xv[i].y, xv[i].z, xv[i].w};
#pragma unroll
              for (int e2 = 0; e2 < 4; ++e2) {
                const float ga = (mb[i] >> (2 * e2) & 1u) ? __uint_as_float(dw[e2] << 16) : 0.f;
                const float gb = (mb[i] >> (2 * e2 + 1) & 1u) ? __uint_as_float(dw[e2] & 0xffff0000u) : 0.f;
                bsum[2 * e2] += ga;
                bsq[2 * e2] = fmaf(ga, __uint_as_float(xw[e2] << 16) - mu[2 * e2], bsq[2 * e2]);
                bsum[2 * e2 + 1] += gb;
                bsq[2 * e2 + 1] = fmaf(gb, __uint_as_float(xw[e2] & 0xffff0000u) - mu[2 * e2 + 1], bsq[2 * e2 + 1]);
              }
            }
          }
        }
      } else if (want_stats) {
        // BatchNorm statistics of this tile from the staged (bf16-rounded, i.e. as stored) values.
        // Rows past the end of the tensor are zero (TMA zero-fill x weights) unless a bias was
        // added, so only the generic variant needs the row bound.
        float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
        const int rows_left = p.M - m0 - rg * RPG;  // valid rows of this thread's group
#pragma unroll
        for (int r8 = 0; r8 < RPG / 8; ++r8) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            if (PLAIN || r8 * 8 + i < rows_left) {
              const uint32_t w2 = lds_u32(col_addr + r8 * 1024 + soff[i]);
              const float a = __uint_as_float(w2 << 16), b = __uint_as_float(w2 & 0xffff0000u);
              s0 += a; q0 = fmaf(a, a, q0);
              s1 += b; q1 = fmaf(b, b, q1);
            }
          }
        }
        as0 += s0; as1 += s1; aq0 += q0; aq1 += q1;  // two-level sum: per tile, then per CTA
      }
    }
    if (EPI != 3 && store_warp) {
      if (elect_one()) tma_store_wait0();
      __syncwarp();
    }
    if (BNRED) {
      // [RGB][2][BLOCK_N] floats in the staging buffer (16 KB), then one thread per column sums the
      // row groups in fixed order
      constexpr int CG = BLOCK_N / 8, RGB = 256 / CG;
      const int cg = et % CG, rgb = et / CG;
      asm volatile("bar.sync 1, 256;" ::: "memory");
      const uint32_t red = st_base;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        sts_f32(red + ((rgb * 2 + 0) * BLOCK_N + cg * 8 + j) * 4, bsum[j]);
        sts_f32(red + ((rgb * 2 + 1) * BLOCK_N + cg * 8 + j) * 4, bsq[j]);
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (et < BLOCK_N) {
        float S = 0.f, Qs = 0.f;
#pragma unroll
        for (int g = 0; g < RGB; ++g) {
          S += lds_f32(red + ((g * 2 + 0) * BLOCK_N + et) * 4);
          Qs += lds_f32(red + ((g * 2 + 1) * BLOCK_N + et) * 4);
        }
        const int col = n0 + et;
        if (col < p.Kout) {
          float* dst = p.stats + (size_t)m_first * 2 * p.Kout;
          dst[col] = S;
          dst[p.Kout + col] = Qs;
        }
      }
    } else if (want_stats) {
      // combine the row groups in fixed order and write this CTA's partial row
      asm volatile("bar.sync 1, 256;" ::: "memory");  // staging is idle (last store completed)
      const uint32_t red = st_base;                    // [RG][2][BLOCK_N] floats
      sts_f32(red + ((rg * 2 + 0) * BLOCK_N + sc) * 4, as0);
      sts_f32(red + ((rg * 2 + 0) * BLOCK_N + sc + 1) * 4, as1);
      sts_f32(red + ((rg * 2 + 1) * BLOCK_N + sc) * 4, aq0);
      sts_f32(red + ((rg * 2 + 1) * BLOCK_N + sc + 1) * 4, aq1);
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (et < BLOCK_N) {
        float S = 0.f, Qs = 0.f;
#pragma unroll
        for (int g = 0; g < RG; ++g) {
          S += lds_f32(red + ((g * 2 + 0) * BLOCK_N + et) * 4);
          Qs += lds_f32(red + ((g * 2 + 1) * BLOCK_N + et) * 4);
        }
        const int col = n0 + et;
        if (col < p.Kout) {
          float* dst = p.stats + (size_t)m_first * 2 * p.Kout;
          dst[col] = S;
          dst[p.Kout + col] = Qs;
        }
      }
    }
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ==========================================================================================
// weight-grad kernel
// ==========================================================================================
struct WgradParams {
  CUtensorMap tmDy;  // {K, M} tiled, box {64 out-ch, 64 pixels}
  CUtensorMap tmX;   // im2col {C,W,H,N}, box 64 ch x 64 pixels (or tiled {C, M} for 1x1 s1 p0)
  CUtensorMap tmOut; // fp32 {Cg, RS, K, splits}, box {32, 1, 128, 1}: coalesced epilogue stores
  float* out;        // [splits][K][RS][Cg] partials (or dw itself when splits == 1)
  int M;             // output pixels N*P*Q
  int outP, outQ, lower_h, lower_w, trav_h, trav_w;
  int K, Cg, RS;
  int num_taps;      // all taps
  int taps_per_cta, tap_groups;
  int cin_tiles, cout_tiles;
  int splits, pixblk_per_split;  // pixel blocks (64 pixels) per split
  int grp_out, grp_in;           // grouped conv: out / in channels per group (0 = dense)
  int16_t off_h[kMaxTaps], off_w[kMaxTaps];
};

// TS > 1 (narrow layers, NT = 64): the x tiles of all (<= TS) taps of the CTA share one ring stage,
// laid out like the 64-channel boxes of a wide tile, so ONE MMA chain of N = taps*64 per pixel block
// covers them all (tap t lands in TMEM columns [64t, 64t+64)) instead of one N = 64 chain per tap.
template <int NT, int SA, int SB, bool IM2COL, int TS = 1>
__global__ void __launch_bounds__(kTcThreads, 1)
conv_wgrad_tc_k(const __grid_constant__ WgradParams p) {
  static_assert(TS == 1 || NT == 64, "multi-tap stages are for 64-channel tiles");
  constexpr int PIX = 64;                         // GEMM-K block: pixels per TMA box
  constexpr int DY_BYTES = 2 * PIX * 128;         // two 64-channel boxes -> 128 out-channels
  constexpr int X_BYTES = TS * (NT / 64) * PIX * 128;  // NT/64 boxes per tap, TS taps
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~(uintptr_t)1023);
  uint8_t* sA = smem;
  uint8_t* sB = smem + SA * DY_BYTES;
  uint64_t* fullA = reinterpret_cast<uint64_t*>(sB + SB * X_BYTES);
  uint64_t* emptyA = fullA + SA;
  uint64_t* fullB = emptyA + SA;
  uint64_t* emptyB = fullB + SB;
  uint64_t* tmem_full_bar = emptyB + SB;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = threadIdx.x / 32;
  const int lane = threadIdx.x % 32;

  // work item decode: blockIdx.x = ((split * cout_tiles + ct) * cin_tiles + it_) * tap_groups + tg
  int w = blockIdx.x;
  const int tg = w % p.tap_groups; w /= p.tap_groups;
  const int cit = w % p.cin_tiles; w /= p.cin_tiles;
  const int ct = w % p.cout_tiles; w /= p.cout_tiles;
  const int split = w;
  const int tap0 = tg * p.taps_per_cta;
  const int ntap = min(p.taps_per_cta, p.num_taps - tap0);
  const int total_pixblk = (p.M + PIX - 1) / PIX;
  const int pb0 = split * p.pixblk_per_split;
  const int pb1 = min(pb0 + p.pixblk_per_split, total_pixblk);
  const int npb = max(pb1 - pb0, 0);
  const int cout0 = ct * TILE_M;
  const int cin0 = cit * NT;   // input-channel coordinate of the OUTPUT tile
  // input-channel coordinate of the x LOADS: grouped layers read the slice of their out-channel tile
  const int xc0 = p.grp_out > 0 ? (cout0 / p.grp_out) * p.grp_in + cin0 : cin0;

  // TMEM columns: power of two >= taps_per_cta * NT (host guarantees <= 512)
  uint32_t tmem_cols = 32;
  while ((int)tmem_cols < p.taps_per_cta * NT) tmem_cols <<= 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmDy);
    tma_prefetch_desc(&p.tmX);
    tma_prefetch_desc(&p.tmOut);
    for (int s = 0; s < SA; ++s) { mbar_init(&fullA[s], 1); mbar_init(&emptyA[s], 1); }
    for (int s = 0; s < SB; ++s) { mbar_init(&fullB[s], 1); mbar_init(&emptyB[s], 1); }
    mbar_init(tmem_full_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_ptr, tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  // issuing roles: whole-warp loops, one elected lane per TMA / tcgen05 instruction (see the note in
  // conv_fprop_tc_persist_k)
  if (warp == 0) {
    const uint32_t sA0 = smem_u32(sA), sB0 = smem_u32(sB);
    const uint32_t fullA0 = smem_u32(fullA), emptyA0 = smem_u32(emptyA);
    const uint32_t fullB0 = smem_u32(fullB), emptyB0 = smem_u32(emptyB);
    uint32_t sa = 0, pha = 1, sb = 0, phb = 1;   // ring slots and the parity their free state shows
    for (int pb = 0; pb < npb; ++pb) {
      const int pix0 = (pb0 + pb) * PIX;
      mbar_wait_u32(emptyA0 + sa * 8, pha);
      if (elect_one()) {  // dy tile: two 64-channel boxes
        const uint32_t fb = fullA0 + sa * 8, dst = sA0 + sa * DY_BYTES;
        mbar_expect_tx_u32(fb, DY_BYTES);
        tma_load_2d_u32(&p.tmDy, fb, dst, cout0, pix0);
        tma_load_2d_u32(&p.tmDy, fb, dst + PIX * 128, cout0 + 64, pix0);
      }
      __syncwarp();
      if (++sa == SA) { sa = 0; pha ^= 1; }
      int bn = 0, bh = 0, bw = 0;
      if (IM2COL) {
        int j = pix0 % p.outQ;
        int t = pix0 / p.outQ;
        int i = t % p.outP;
        bn = t / p.outP;
        bh = p.lower_h + i * p.trav_h;
        bw = p.lower_w + j * p.trav_w;
      }
      if (TS > 1) {
        mbar_wait_u32(emptyB0 + sb * 8, phb);
        if (elect_one()) {
          const uint32_t fb = fullB0 + sb * 8, dst = sB0 + sb * X_BYTES;
          mbar_expect_tx_u32(fb, (uint32_t)ntap * PIX * 128);
          for (int tp = 0; tp < ntap; ++tp)
            tma_load_im2col_4d_u32(&p.tmX, fb, dst + tp * PIX * 128, xc0, bw, bh, bn,
                                   (uint16_t)p.off_w[tap0 + tp], (uint16_t)p.off_h[tap0 + tp]);
        }
        __syncwarp();
        if (++sb == SB) { sb = 0; phb ^= 1; }
        continue;
      }
      for (int tp = 0; tp < ntap; ++tp) {
        mbar_wait_u32(emptyB0 + sb * 8, phb);
        if (elect_one()) {
          const uint32_t fb = fullB0 + sb * 8, dst = sB0 + sb * X_BYTES;
          mbar_expect_tx_u32(fb, X_BYTES);
#pragma unroll
          for (int b = 0; b < NT / 64; ++b) {
            if (IM2COL)
              tma_load_im2col_4d_u32(&p.tmX, fb, dst + b * PIX * 128, xc0 + b * 64, bw, bh, bn,
                                     (uint16_t)p.off_w[tap0 + tp], (uint16_t)p.off_h[tap0 + tp]);
            else
              tma_load_2d_u32(&p.tmX, fb, dst + b * PIX * 128, xc0 + b * 64, pix0);
          }
        }
        __syncwarp();
        if (++sb == SB) { sb = 0; phb ^= 1; }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = make_idesc_bf16(TILE_M, NT, 1, 1);  // both operands MN-major
    const uint32_t idesc_mt = make_idesc_bf16(TILE_M, (uint32_t)ntap * 64u, 1, 1);
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t fullA0 = smem_u32(fullA), emptyA0 = smem_u32(emptyA);
    const uint32_t fullB0 = smem_u32(fullB), emptyB0 = smem_u32(emptyB);
    const uint32_t tfull = smem_u32(tmem_full_bar);
    // MN-major SW128 descriptors: 16 pixel rows = 2 swizzle atoms of 1024 B (SBO); the next
    // 64-channel box sits PIX*128 bytes further (LBO).  Low word = address >> 4 | LBO >> 4 << 16.
    constexpr uint32_t DESC_HI = (1024u >> 4) | (1u << 14) | (2u << 29);
    constexpr uint32_t LBO_BITS = ((uint32_t)(PIX * 128) >> 4) << 16;
    const uint32_t a_lo0 = (smem_u32(sA) >> 4) | LBO_BITS, b_lo0 = (smem_u32(sB) >> 4) | LBO_BITS;
    uint32_t sa = 0, pha = 0, sb = 0, phb = 0;
    for (int pb = 0; pb < npb; ++pb) {
      mbar_wait_u32(fullA0 + sa * 8, pha);
      const uint32_t a_lo = a_lo0 + sa * (DY_BYTES >> 4);
      if (TS > 1) {
        mbar_wait_u32(fullB0 + sb * 8, phb);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t b_lo = b_lo0 + sb * (X_BYTES >> 4);
#pragma unroll
          for (int k = 0; k < PIX / 16; ++k)
            umma_bf16_lohi(tmem_u, a_lo + k * (2048 >> 4), b_lo + k * (2048 >> 4), DESC_HI, idesc_mt,
                           (pb | k) != 0 ? 1u : 0u);
          umma_commit_u32(emptyB0 + sb * 8);
          umma_commit_u32(emptyA0 + sa * 8);
          if (pb == npb - 1) umma_commit_u32(tfull);
        }
        __syncwarp();
        if (++sb == SB) { sb = 0; phb ^= 1; }
        if (++sa == SA) { sa = 0; pha ^= 1; }
        continue;
      }
      for (int tp = 0; tp < ntap; ++tp) {
        mbar_wait_u32(fullB0 + sb * 8, phb);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t b_lo = b_lo0 + sb * (X_BYTES >> 4);
#pragma unroll
          for (int k = 0; k < PIX / 16; ++k)
            umma_bf16_lohi(tmem_u + (uint32_t)(tp * NT), a_lo + k * (2048 >> 4), b_lo + k * (2048 >> 4),
                           DESC_HI, idesc, (pb | k) != 0 ? 1u : 0u);
          umma_commit_u32(emptyB0 + sb * 8);
          if (tp == ntap - 1) umma_commit_u32(emptyA0 + sa * 8);
          if (pb == npb - 1 && tp == ntap - 1) umma_commit_u32(tfull);
        }
        __syncwarp();
        if (++sb == SB) { sb = 0; phb ^= 1; }
      }
      if (++sa == SA) { sa = 0; pha ^= 1; }
    }
  } else {
    // Epilogue: TMEM -> registers -> swizzled fp32 staging (the operand rings are idle by now:
    // every load has landed and every MMA that read them has completed) -> TMA stores.  One
    // [128 out-ch x 32 in-ch] box per store: full 128-byte rows instead of the 4-byte scattered
    // writes a row-per-thread store pattern would issue; rows >= K and columns >= Cg are clipped
    // by the tensor map.
    const int quarter = warp % 4;
    const int row = quarter * 32 + lane;  // out-channel within the tile
    const bool store_warp = warp == 2;
    if (npb > 0) {
      mbar_wait(tmem_full_bar, 0);
      tc_fence_after();
    }
    uint8_t* staging = smem;
    for (int tp = 0; tp < ntap; ++tp) {
      if (store_warp) {
        if (elect_one()) tma_store_wait_read0();
        __syncwarp();
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
#pragma unroll 1
      for (int c0 = 0; c0 < NT; c0 += 32) {
        uint32_t v[32];
        if (npb > 0) {
          tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(tp * NT + c0), v);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0u;
        }
        uint8_t* box = staging + (c0 / 32) * (TILE_M * 128);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          uint4* dst = reinterpret_cast<uint4*>(box + row * 128 + ((j ^ (row & 7)) * 16));
          *dst = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        }
      }
      fence_proxy_async_smem();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (store_warp) {
        if (elect_one()) {
          for (int c0 = 0; c0 < NT; c0 += 32)
            if (cin0 + c0 < p.Cg)
              tma_store_4d(&p.tmOut, staging + (c0 / 32) * (TILE_M * 128), cin0 + c0, tap0 + tp, cout0,
                           split);
          tma_store_commit();
        }
        __syncwarp();
      }
    }
    if (store_warp) {
      if (elect_one()) tma_store_wait0();
      __syncwarp();
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, tmem_cols);
  }
}

// ==========================================================================================
// small helper kernels
// ==========================================================================================
// wt[c][rs][k] = w[k][rs][c]   (dgrad operand: the filter with in/out channels exchanged)
__global__ void transpose_filter_k(const bf16* __restrict__ w, bf16* __restrict__ wt, int K, int RS,
                                   int C) {
  int64_t total = (int64_t)K * RS * C;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    int k = (int)(i % K);
    int64_t t = i / K;
    int rs = (int)(t % RS);
    int c = (int)(t / RS);
    wt[i] = w[((int64_t)k * RS + rs) * C + c];
  }
}

// Strided data-grad: the parity classes that receive no filter tap (1x1 stride 2: three of four)
// are zero; every other pixel of dx is written by its class's convolution (EPI == 3).
struct ZeroClassP {
  int N, H, W, C, sh, sw;
  uint32_t zero_mask;   // bit (a * sw + b): class (a, b) is all zeros
};
__global__ void dgrad_zero_classes_k(ZeroClassP p, bf16* __restrict__ dx) {
  const int Cv = p.C / 8;
  const int64_t total = (int64_t)p.N * p.H * p.W * Cv;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t pix = i / Cv;
    const int w = (int)(pix % p.W);
    const int h = (int)((pix / p.W) % p.H);
    const int cls = (h % p.sh) * p.sw + (w % p.sw);
    if ((p.zero_mask >> cls) & 1u) *reinterpret_cast<uint4*>(dx + i * 8) = make_uint4(0, 0, 0, 0);
  }
}

__global__ void sum_splits_tc_k(const float* __restrict__ part, float* __restrict__ out, int splits,
                                int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    float s = 0.f;
    for (int k = 0; k < splits; ++k) s += part[(int64_t)k * n + i];
    out[i] = s;
  }
}

// ==========================================================================================
// host side
// ==========================================================================================
struct FpropLaunch {
  const void* a;                 // activation tensor fed to the loads, NHWC bf16
  int aN, aH, aW, aC;
  int lower_h, lower_w, upper_h, upper_w, trav_h, trav_w;
  int outP, outQ;                // output grid per image
  int ntaps;
  int16_t off_h[kMaxTaps], off_w[kMaxTaps];
  uint8_t widx[kMaxTaps];
  const void* w;                 // weights [Kout][RS][aC] bf16
  int Kout, RS;
  void* out;                     // [aN*outP*outQ, Kout] bf16
  const float* bias;
  const void* addend;            // optional [M][Kout] bf16 summed into the output (data-grad + skip gradient)
  const void* addend_mask;       // optional 1-bit mask over the addend (M*Kout/8 bytes, Kout % 32 == 0)
  const void* bn_x;              // data-grad into a BatchNorm + ReLU output: that layer's input, mask, mean
  const void* bn_mask;
  const float* bn_mean;
  float* stats;                  // optional BatchNorm partials (forward only)
  int flags;
  bool tiled2d;                  // pure GEMM: A is the [M, C] matrix itself
  int grp_out, grp_in, wC;       // grouped: out / reduction channels per group, expanded filter width
  void* scatter;                 // strided data-grad class: the full dx tensor (EPI 3), else nullptr
  int sc_H, sc_W, sc_sh, sc_sw, sc_a, sc_b;
};

// Persistent grid: a multiple of n_tiles (every CTA keeps one n-tile), at most one CTA per SM, and
// no CTA without work.  grid / n_tiles is also the number of BatchNorm partial rows.
static int persist_grid(int64_t m_tiles, int n_tiles) {
  int64_t per = sm_count() / n_tiles;
  if (per < 1) per = 1;
  if (per > m_tiles) per = m_tiles;
  return (int)(per * n_tiles);
}

static int fprop_block_n(int Kout, bool grouped) {
  if (grouped) return 64;
  return Kout <= 64 ? 64 : (Kout <= 128 ? 128 : 256);
}

template <int BLOCK_N, int STAGES, bool IM2COL, int EPI, bool BRES = false, bool BNRED = false>
static int launch_fprop_persist_inst(const FpropParams& prm, cudaStream_t st) {
  constexpr int STAGE_BYTES = BRES ? A_TILE_BYTES : A_TILE_BYTES + BLOCK_N * TILE_K * 2;
  constexpr size_t smem = (size_t)STAGES * STAGE_BYTES + (BRES ? (size_t)kBresIters * BLOCK_N * TILE_K * 2 : 0) +
                          (size_t)TILE_M * BLOCK_N * 2 + (2 * STAGES + 5) * 8 + 16 + 1024;
  static_assert(smem <= 232448, "exceeds the 227 KB dynamic shared memory limit");
  static bool attr_done = false;
  auto kern = conv_fprop_tc_persist_k<BLOCK_N, STAGES, IM2COL, EPI, BRES, BNRED>;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    CNB_REQUIRE(e == cudaSuccess, CNB_ERR_CUDA, "cudaFuncSetAttribute(fprop persistent): %s",
                cudaGetErrorString(e));
    attr_done = true;
  }
  const int grid = persist_grid(prm.m_tiles, prm.n_tiles);
  kern<<<grid, kPersistThreads, smem, st>>>(prm);
  return check_launch("conv_fprop_tc_persist");
}

// CNB_BRES=0 keeps the filter tiles in the ring (the r02c kernels) for A/B measurements
static bool bres_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("CNB_BRES");
    v = (e && atoi(e) == 0) ? 0 : 1;
  }
  return v == 1;
}

template <bool IM2COL, int EPI, bool BNRED = false>
static int launch_fprop_bn(const FpropParams& prm, int BLOCK_N, cudaStream_t st) {
  if (BLOCK_N == 64 && prm.num_taps * prm.num_cblk <= kBresIters && bres_enabled())
    return launch_fprop_persist_inst<64, 8, IM2COL, EPI, true, BNRED>(prm, st);
  if (BLOCK_N == 64) return launch_fprop_persist_inst<64, 8, IM2COL, EPI, false, BNRED>(prm, st);
  if (BLOCK_N == 128) return launch_fprop_persist_inst<128, 5, IM2COL, EPI, false, BNRED>(prm, st);
  return launch_fprop_persist_inst<256, 3, IM2COL, EPI, false, BNRED>(prm, st);
}

static int launch_fprop(const FpropLaunch& L, cudaStream_t st) {
  CNB_TRY(load_driver_fns());
  FpropParams prm;
  memset(&prm, 0, sizeof(prm));
  const int64_t M64 = (int64_t)L.aN * L.outP * L.outQ;
  CNB_REQUIRE(M64 < (1ll << 31) - 256, CNB_ERR_UNSUPPORTED, "conv_tc: too many output pixels");
  const int M = (int)M64;
  if (L.tiled2d) {
    uint64_t dims[2] = {(uint64_t)L.aC, (uint64_t)M};
    uint64_t str[1] = {(uint64_t)L.aC * 2};
    uint32_t box[2] = {TILE_K, TILE_M};
    CNB_TRY(make_tiled_map(&prm.tmA, L.a, 2, dims, str, box, "A(2d)"));
  } else {
    CNB_TRY(make_im2col_map(&prm.tmA, L.a, L.aN, L.aH, L.aW, L.aC, L.lower_w, L.lower_h, L.upper_w,
                            L.upper_h, L.trav_w, L.trav_h, TILE_M, "A(im2col)"));
  }
  const bool grouped = L.grp_out > 0;
  const int BLOCK_N = fprop_block_n(L.Kout, grouped);
  const int wC = grouped ? L.wC : L.aC;   // reduction width of the filter operand
  {
    uint64_t dims[3] = {(uint64_t)wC, (uint64_t)L.RS, (uint64_t)L.Kout};
    uint64_t str[2] = {(uint64_t)wC * 2, (uint64_t)wC * L.RS * 2};
    uint32_t box[3] = {TILE_K, 1, (uint32_t)BLOCK_N};
    CNB_TRY(make_tiled_map(&prm.tmB, L.w, 3, dims, str, box, "B(weights)"));
  }
  if (!L.scatter) {
    uint64_t dims[2] = {(uint64_t)L.Kout, (uint64_t)M};
    uint64_t str[1] = {(uint64_t)L.Kout * 2};
    uint32_t box[2] = {64, TILE_M};
    CNB_TRY(make_tiled_map(&prm.tmD, L.out, 2, dims, str, box, "D(output)"));
  } else {
    prm.tmD = prm.tmB;   // not used by the scattering epilogue; keep the prefetch target valid
    prm.scatter = reinterpret_cast<bf16*>(L.scatter);
    prm.sc_H = L.sc_H; prm.sc_W = L.sc_W; prm.sc_sh = L.sc_sh; prm.sc_sw = L.sc_sw;
    prm.sc_a = L.sc_a; prm.sc_b = L.sc_b;
  }
  prm.bias = L.bias;
  prm.stats = L.stats;
  prm.M = M;
  prm.outP = L.outP; prm.outQ = L.outQ;
  prm.lower_h = L.lower_h; prm.lower_w = L.lower_w;
  prm.trav_h = L.trav_h; prm.trav_w = L.trav_w;
  prm.num_cblk = (wC + TILE_K - 1) / TILE_K;
  prm.num_taps = L.ntaps;
  prm.n_tiles = (L.Kout + BLOCK_N - 1) / BLOCK_N;
  prm.m_tiles = (int)ceil_div(M, TILE_M);
  prm.grp_out = L.grp_out;
  prm.grp_in = L.grp_in;
  prm.Kout = L.Kout;
  prm.flags = L.flags & ((L.bias ? CNB_CONV_BIAS : 0) | CNB_CONV_RELU);
  prm.addend = reinterpret_cast<const bf16*>(L.addend);
  prm.addend_mask = reinterpret_cast<const uint8_t*>(L.addend_mask);
  prm.bn_x = reinterpret_cast<const bf16*>(L.bn_x);
  prm.bn_mask = reinterpret_cast<const uint8_t*>(L.bn_mask);
  prm.bn_mean = L.bn_mean;
  CNB_REQUIRE(!L.bn_x || (L.stats && L.bn_mask && L.bn_mean && L.Kout % 8 == 0), CNB_ERR_INVALID,
              "conv_tc: BatchNorm backward sums need the partial table, the bit mask and the mean");
  if (L.addend) prm.flags |= kConvAccum;
  CNB_REQUIRE(!L.addend_mask || (L.addend && L.Kout % 32 == 0), CNB_ERR_UNSUPPORTED,
              "conv_tc: a masked addend needs the addend and Kout %% 32 == 0 (Kout=%d)", L.Kout);
  for (int t = 0; t < L.ntaps; ++t) {
    prm.off_h[t] = L.off_h[t];
    prm.off_w[t] = L.off_w[t];
    prm.widx[t] = L.widx[t];
  }
  // epilogue variant: 1 = plain store, 2 = plain + addend, 0 = bias / ReLU by flags
  const int epi = prm.flags == 0 ? 1 : (prm.flags == kConvAccum ? 2 : 0);
  CNB_REQUIRE(epi != 0 || !(prm.flags & kConvAccum), CNB_ERR_UNSUPPORTED,
              "conv_tc: the accumulating epilogue does not combine with bias / ReLU");
  if (L.scatter) {
    CNB_REQUIRE(epi == 1 && !L.tiled2d && !L.stats, CNB_ERR_UNSUPPORTED,
                "conv_tc: the scattering epilogue is a plain im2col store");
    return launch_fprop_bn<true, 3>(prm, BLOCK_N, st);
  }
  if (prm.bn_x) {   // data-grad + BatchNorm backward sums
    CNB_REQUIRE(epi != 0, CNB_ERR_UNSUPPORTED, "conv_tc: BatchNorm backward sums do not combine with bias / ReLU");
    if (L.tiled2d) {
      if (epi == 1) return launch_fprop_bn<false, 1, true>(prm, BLOCK_N, st);
      if (L.addend_mask) return launch_fprop_bn<false, 4, true>(prm, BLOCK_N, st);
      return launch_fprop_bn<false, 2, true>(prm, BLOCK_N, st);
    }
    if (epi == 1) return launch_fprop_bn<true, 1, true>(prm, BLOCK_N, st);
    if (L.addend_mask) return launch_fprop_bn<true, 4, true>(prm, BLOCK_N, st);
    return launch_fprop_bn<true, 2, true>(prm, BLOCK_N, st);
  }
  if (L.tiled2d) {
    if (epi == 1) return launch_fprop_bn<false, 1>(prm, BLOCK_N, st);
    if (epi == 2 && L.addend_mask) return launch_fprop_bn<false, 4>(prm, BLOCK_N, st);
    if (epi == 2) return launch_fprop_bn<false, 2>(prm, BLOCK_N, st);
    return launch_fprop_bn<false, 0>(prm, BLOCK_N, st);
  }
  if (epi == 1) return launch_fprop_bn<true, 1>(prm, BLOCK_N, st);
  if (epi == 2 && L.addend_mask) return launch_fprop_bn<true, 4>(prm, BLOCK_N, st);
  if (epi == 2) return launch_fprop_bn<true, 2>(prm, BLOCK_N, st);
  return launch_fprop_bn<true, 0>(prm, BLOCK_N, st);
}

// ==========================================================================================
// grouped convolutions (ResNeXt / SENet / SKNet 3x3 with 4..32 channels per group, ShuffleNet 1x1)
// ==========================================================================================
// GEMM-N side = "out" channels (out_per_g per group), reduction side = "in" channels (in_per_g).
// An n-tile [n0, n0 + tile) of out channels belongs to groups g_lo..g_hi and therefore reads the
// input-channel slice [g_lo * in_per_g, (g_hi + 1) * in_per_g): its widest extent is the `span`.
// The filter operand is EXPANDED to width wC = round_up(span, 64): row k holds weight (k, c) at
// column (g(k) - g_lo) * in_per_g + c and zeros elsewhere, i.e. the dense kernels run unchanged on
// a block-diagonal filter restricted to the slice (the A operand's channel coordinate starts at
// g_lo * in_per_g).  The padded multiplications by zero are not counted as algorithmic FLOPs;
// these layers are HBM-bound (AI 5..60 FLOP/B) and the tensor pipe has the slack.
static int grouped_span(int out_total, int out_per_g, int in_per_g, int tile) {
  int mx = 0;
  for (int n0 = 0; n0 < out_total; n0 += tile) {
    const int g_lo = n0 / out_per_g;
    const int last = (n0 + tile < out_total ? n0 + tile : out_total) - 1;
    const int g_hi = last / out_per_g;
    const int sp = (g_hi - g_lo + 1) * in_per_g;
    mx = sp > mx ? sp : mx;
  }
  return mx;
}

// The A operand's channel coordinate of an n-tile starts at g_lo * in_per_g.  TMA needs the first
// byte of every box row 16-byte aligned, i.e. that coordinate a multiple of 8 bf16 elements
// (ShuffleNet-v1 g = 4: 68 channels per group -> 136-byte starts -> "illegal instruction").
static bool grouped_starts_aligned(int out_total, int out_per_g, int in_per_g, int tile) {
  for (int n0 = 0; n0 < out_total; n0 += tile)
    if (((n0 / out_per_g) * in_per_g) % 8 != 0) return false;
  return true;
}

constexpr int kGroupTileN = 64;     // fprop / dgrad n-tile of the grouped path
constexpr int kGroupMaxWidth = 1024; // widest expanded reduction slice

// out[o][rs][cc], cc in [0, wC).  transposed == 0 (forward operand): o = k, reduction = c.
// transposed == 1 (data-grad operand): o = c, reduction = k.  w is the plain [K][RS][Cg] filter.
__global__ void expand_grouped_filter_k(const bf16* __restrict__ w, bf16* __restrict__ out, int K,
                                        int RS, int Cg, int Kg, int wC, int tile, int transposed) {
  const int C = (K / Kg) * Cg;
  const int O = transposed ? C : K;
  const int out_per_g = transposed ? Cg : Kg, in_per_g = transposed ? Kg : Cg;
  const int64_t total = (int64_t)O * RS * wC;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int cc = (int)(i % wC);
    const int64_t t = i / wC;
    const int rs = (int)(t % RS);
    const int o = (int)(t / RS);
    const int g_lo = ((o / tile) * tile) / out_per_g;
    const int g = o / out_per_g;
    const int rl = cc - (g - g_lo) * in_per_g;   // reduction channel within the group
    bf16 v = __float2bfloat16_rn(0.f);
    if (rl >= 0 && rl < in_per_g) {
      const int k = transposed ? g * Kg + rl : o;
      const int c = transposed ? o - g * Cg : rl;
      v = w[((int64_t)k * RS + rs) * Cg + c];
    }
    out[i] = v;
  }
}

// dw[k][rs][c] = sum over splits of part[split][k][rs][(g(k) - g_lo(k-tile)) * Cg + c]: picks the
// block diagonal out of the dense [128 out-ch x NT] tiles the grouped weight-grad produced.
__global__ void extract_grouped_wgrad_k(const float* __restrict__ part, float* __restrict__ dw,
                                        int splits, int K, int RS, int Cg, int Kg, int NT) {
  const int64_t total = (int64_t)K * RS * Cg;
  const int64_t psz = (int64_t)K * RS * NT;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % Cg);
    const int64_t t = i / Cg;
    const int k = (int)(t / RS);
    const int g_lo = ((k / TILE_M) * TILE_M) / Kg;
    const int64_t src = t * NT + (int64_t)(k / Kg - g_lo) * Cg + c;
    float sum = 0.f;
    for (int sp = 0; sp < splits; ++sp) sum += part[(int64_t)sp * psz + src];
    dw[i] = sum;
  }
}

// expanded reduction width (multiple of 64) for the forward (dir 0) / data-grad (dir 1) operand
static int grouped_width(const cnb_conv_desc& d, int transposed) {
  const int Cg = d.C / d.groups, Kg = d.K / d.groups;
  const int span = transposed ? grouped_span(d.C, Cg, Kg, kGroupTileN)
                              : grouped_span(d.K, Kg, Cg, kGroupTileN);
  return (span + 63) / 64 * 64;
}
// widest input-channel slice of a 128-out-channel tile of the grouped weight-grad
static int grouped_wgrad_span(const cnb_conv_desc& d) {
  const int Cg = d.C / d.groups, Kg = d.K / d.groups;
  return grouped_span(d.K, Kg, Cg, TILE_M);
}

static bool corners_ok(int v) { return v >= -128 && v <= 127; }

bool conv_tc_supported(const cnb_conv_desc& d, int direction, const char** why) {
  static const char* dummy;
  if (!why) why = &dummy;
  if (d.dtype != CNB_BF16) { *why = "tcgen05 path is bf16-only (fp32-IO mode uses CUDA cores)"; return false; }
  if (d.C % 8 != 0 || d.K % 8 != 0) { *why = "C and K must be multiples of 8 (16-byte TMA strides)"; return false; }
  if (d.groups != 1) {
    if (direction != CNB_WGRAD && grouped_width(d, direction == CNB_DGRAD) > kGroupMaxWidth) {
      *why = "grouped convolution: channel slice wider than 1024"; return false;
    }
    const int Cg = d.C / d.groups, Kg = d.K / d.groups;
    const bool ok = direction == CNB_FWD     ? grouped_starts_aligned(d.K, Kg, Cg, kGroupTileN)
                    : direction == CNB_DGRAD ? grouped_starts_aligned(d.C, Cg, Kg, kGroupTileN)
                                             : grouped_starts_aligned(d.K, Kg, Cg, TILE_M);
    if (!ok) { *why = "grouped convolution: a tile's channel slice does not start on a 16-byte boundary"; return false; }
  }
  if (d.R * d.S > kMaxTaps) { *why = "more than 49 filter taps"; return false; }
  int P, Q;
  conv_out_dims(d, P, Q);
  if (P <= 0 || Q <= 0) { *why = "empty output"; return false; }
  const int eh = (d.R - 1) * d.dil_h, ew = (d.S - 1) * d.dil_w;
  if (eh > 255 || ew > 255) { *why = "dilated filter extent > 255"; return false; }
  if (!corners_ok(-d.pad_h) || !corners_ok(-d.pad_w) || !corners_ok(d.pad_h - eh) ||
      !corners_ok(d.pad_w - ew)) { *why = "padding outside the im2col corner range"; return false; }
  if (d.stride_h > 8 || d.stride_w > 8) { *why = "stride > 8"; return false; }
  if (direction == CNB_DGRAD) {
    if (d.stride_h * d.stride_w > 16) { *why = "strided dgrad with more than 16 parity classes"; return false; }
    if (d.R > 16 || d.S > 16) { *why = "filter larger than 16 in one dimension"; return false; }
  }
  if ((int64_t)d.N * d.H * d.W >= (1ll << 31) - 256 || (int64_t)d.N * P * Q >= (1ll << 31) - 256) {
    *why = "more than 2^31 pixels"; return false;
  }
  return true;
}

static void wgrad_geometry(const cnb_conv_desc& d, int& NT, int& taps_per_cta, int& tap_groups,
                           int& cin_tiles, int& cout_tiles, int& splits, int& pixblk_per_split) {
  const bool grouped = d.groups > 1;
  const int Cg = grouped ? grouped_wgrad_span(d) : d.C;   // grouped: the slice of one out-channel tile
  NT = Cg >= 256 ? 256 : (Cg >= 128 ? 128 : 64);
  const int taps = d.R * d.S;
  int maxt = 512 / NT;
  if (NT == 64 && taps > 1) maxt = 4;  // multi-tap stages: one N = taps*64 MMA chain per pixel block
  tap_groups = (taps + maxt - 1) / maxt;
  taps_per_cta = (taps + tap_groups - 1) / tap_groups;
  tap_groups = (taps + taps_per_cta - 1) / taps_per_cta;
  cin_tiles = (Cg + NT - 1) / NT;
  cout_tiles = (d.K + TILE_M - 1) / TILE_M;
  int P, Q;
  conv_out_dims(d, P, Q);
  const int64_t M = (int64_t)d.N * P * Q;
  const int64_t total_pb = ceil_div(M, 64);
  const int64_t tiles = (int64_t)tap_groups * cin_tiles * cout_tiles;
  // One CTA is resident per SM (the operand rings take > 114 KB), so the grid runs in waves of
  // sm_count CTAs.  Pick the split count that minimises waves x per-CTA time: per-CTA time ~
  // (pixel blocks per split) x (taps per CTA) MMA groups + a fixed prologue/epilogue cost; every
  // extra split also adds one row to the second-stage sum.  At least 8 pixel blocks per split.
  const int64_t sms = sm_count();
  int64_t maxs = ceil_div(total_pb, 8);
  if (maxs > 512) maxs = 512;
  if (maxs < 1) maxs = 1;
  int64_t best_s = 1;
  double best_cost = 1e300;
  for (int64_t sp = 1; sp <= maxs; ++sp) {
    const int64_t waves = ceil_div(sp * tiles, sms);
    const double per_cta = (double)ceil_div(total_pb, sp) * taps_per_cta + 12.0 + 2.0 * taps_per_cta;
    const double cost = (double)waves * per_cta + 0.05 * (double)sp;
    if (cost < best_cost * 0.999) {
      best_cost = cost;
      best_s = sp;
    }
  }
  pixblk_per_split = (int)ceil_div(total_pb, best_s);
  splits = (int)ceil_div(total_pb, pixblk_per_split);
}

static size_t dgrad_filter_bytes(const cnb_conv_desc& d) {
  if (d.groups > 1) return align_up((size_t)d.C * d.R * d.S * grouped_width(d, 1) * 2, 1024);
  return align_up((size_t)d.K * d.R * d.S * d.C * 2, 1024);
}

size_t conv_tc_workspace(const cnb_conv_desc& d, int direction) {
  if (direction == CNB_FWD)
    return d.groups > 1 ? align_up((size_t)d.K * d.R * d.S * grouped_width(d, 0) * 2, 1024) + 1024 : 0;
  if (direction == CNB_DGRAD) {
    size_t wt = dgrad_filter_bytes(d);
    const size_t tmp = 0;   // strided data-grad writes dx from the class convolutions' epilogues
    return wt + tmp + 1024;
  }
  int NT, tpc, tg, cit, cot, splits, pps;
  wgrad_geometry(d, NT, tpc, tg, cit, cot, splits, pps);
  if (d.groups > 1)  // dense [K][RS][NT] tiles per split; the block diagonal is extracted afterwards
    return (size_t)splits * d.K * d.R * d.S * (cit * NT) * sizeof(float) + 1024;
  return (size_t)(splits > 1 ? splits : 0) * d.K * d.R * d.S * d.C * sizeof(float) + 1024;
}

int conv_tc_stats_rows(const cnb_conv_desc& d) {
  int P, Q;
  conv_out_dims(d, P, Q);
  const int BLOCK_N = fprop_block_n(d.K, d.groups > 1);
  const int n_tiles = (d.K + BLOCK_N - 1) / BLOCK_N;
  // one row per persistent CTA of an n-tile column (conv_fprop_tc_persist_k)
  return persist_grid(ceil_div((int64_t)d.N * P * Q, TILE_M), n_tiles) / n_tiles;
}

// rows of the table a stride-1 data-grad writes when it also forms BatchNorm backward sums
int conv_tc_dgrad_stats_rows(const cnb_conv_desc& d) {
  const int BLOCK_N = fprop_block_n(d.C, d.groups > 1);
  const int n_tiles = (d.C + BLOCK_N - 1) / BLOCK_N;
  return persist_grid(ceil_div((int64_t)d.N * d.H * d.W, TILE_M), n_tiles) / n_tiles;
}

int conv_tc_fwd(const cnb_conv_desc& d, const void* x, const void* w, const float* bias, void* y,
                float* stats, void* ws, size_t ws_bytes, cudaStream_t st) {
  FpropLaunch L;
  memset(&L, 0, sizeof(L));
  int P, Q;
  conv_out_dims(d, P, Q);
  if (d.groups > 1) {
    // block-diagonal expansion of the filter, rebuilt per call (a few hundred KB, one tiny kernel)
    CNB_REQUIRE(ws && ws_bytes >= conv_tc_workspace(d, CNB_FWD), CNB_ERR_WORKSPACE,
                "conv_tc_fwd(grouped): workspace %zu < %zu", ws_bytes, conv_tc_workspace(d, CNB_FWD));
    const int Cg = d.C / d.groups, Kg = d.K / d.groups, wC = grouped_width(d, 0);
    bf16* wexp = reinterpret_cast<bf16*>((reinterpret_cast<uintptr_t>(ws) + 1023) & ~(uintptr_t)1023);
    const int64_t total = (int64_t)d.K * d.R * d.S * wC;
    expand_grouped_filter_k<<<grid_for(total, 256), 256, 0, st>>>((const bf16*)w, wexp, d.K, d.R * d.S,
                                                                 Cg, Kg, wC, kGroupTileN, 0);
    CNB_TRY(check_launch("expand_grouped_filter"));
    w = wexp;
    L.grp_out = Kg; L.grp_in = Cg; L.wC = wC;
  }
  L.a = x; L.aN = d.N; L.aH = d.H; L.aW = d.W; L.aC = d.C;
  L.lower_h = -d.pad_h; L.lower_w = -d.pad_w;
  L.upper_h = d.pad_h - (d.R - 1) * d.dil_h;
  L.upper_w = d.pad_w - (d.S - 1) * d.dil_w;
  L.trav_h = d.stride_h; L.trav_w = d.stride_w;
  L.outP = P; L.outQ = Q;
  L.ntaps = d.R * d.S;
  for (int r = 0; r < d.R; ++r)
    for (int s = 0; s < d.S; ++s) {
      int t = r * d.S + s;
      L.off_h[t] = (int16_t)(r * d.dil_h);
      L.off_w[t] = (int16_t)(s * d.dil_w);
      L.widx[t] = (uint8_t)t;
    }
  L.w = w; L.Kout = d.K; L.RS = d.R * d.S;
  L.out = y; L.bias = bias; L.flags = d.flags;
  L.stats = stats;
  L.tiled2d = (d.R == 1 && d.S == 1 && d.stride_h == 1 && d.stride_w == 1 && d.pad_h == 0 &&
               d.pad_w == 0);
  return launch_fprop(L, st);
}

int conv_tc_dgrad(const cnb_conv_desc& d, const void* dy, const void* w, const void* w_t, void* dx,
                  void* ws, size_t ws_bytes, cudaStream_t st, const void* addend, const void* addend_mask,
                  const void* bn_x, const void* bn_mask, const float* bn_mean, float* bn_partials) {
  CNB_REQUIRE(!addend || (d.stride_h == 1 && d.stride_w == 1), CNB_ERR_UNSUPPORTED,
              "conv_tc_dgrad: the accumulating epilogue exists for stride-1 data-grad only");
  CNB_REQUIRE(ws && ws_bytes >= conv_tc_workspace(d, CNB_DGRAD), CNB_ERR_WORKSPACE,
              "conv_tc_dgrad: workspace %zu < %zu", ws_bytes, conv_tc_workspace(d, CNB_DGRAD));
  int P, Q;
  conv_out_dims(d, P, Q);
  const int RS = d.R * d.S;
  // transposed filter wt[C][RS][K]
  bf16* wt_ws = reinterpret_cast<bf16*>((reinterpret_cast<uintptr_t>(ws) + 1023) & ~(uintptr_t)1023);
  const bf16* wt = reinterpret_cast<const bf16*>(w_t);
  const bool grouped = d.groups > 1;
  const int Cg = d.C / d.groups, Kg = d.K / d.groups;
  const int wCd = grouped ? grouped_width(d, 1) : 0;
  if (grouped) {
    // block-diagonal expansion of the transposed filter: [C][RS][wCd] (w_t is not used)
    const int64_t total = (int64_t)d.C * RS * wCd;
    expand_grouped_filter_k<<<grid_for(total, 256), 256, 0, st>>>((const bf16*)w, wt_ws, d.K, RS, Cg, Kg,
                                                                 wCd, kGroupTileN, 1);
    CNB_TRY(check_launch("expand_grouped_filter(dgrad)"));
    wt = wt_ws;
  } else if (!wt) {  // no pre-transposed copy from cnb_pack_filter: build it in the workspace
    int64_t total = (int64_t)d.K * RS * d.C;
    transpose_filter_k<<<grid_for(total, 256), 256, 0, st>>>((const bf16*)w, wt_ws, d.K, RS, d.C);
    CNB_TRY(check_launch("transpose_filter"));
    wt = wt_ws;
  }

  if (d.stride_h == 1 && d.stride_w == 1) {
    FpropLaunch L;
    memset(&L, 0, sizeof(L));
    L.a = dy; L.aN = d.N; L.aH = P; L.aW = Q; L.aC = d.K;
    L.lower_h = d.pad_h - (d.R - 1) * d.dil_h;
    L.lower_w = d.pad_w - (d.S - 1) * d.dil_w;
    L.upper_h = L.lower_h + d.H - P;
    L.upper_w = L.lower_w + d.W - Q;
    CNB_REQUIRE(corners_ok(L.lower_h) && corners_ok(L.lower_w) && corners_ok(L.upper_h) &&
                    corners_ok(L.upper_w),
                CNB_ERR_UNSUPPORTED, "conv_tc_dgrad: corners outside the im2col range");
    L.trav_h = 1; L.trav_w = 1;
    L.outP = d.H; L.outQ = d.W;
    L.ntaps = RS;
    for (int r = 0; r < d.R; ++r)
      for (int s = 0; s < d.S; ++s) {
        int t = r * d.S + s;
        L.off_h[t] = (int16_t)((d.R - 1 - r) * d.dil_h);
        L.off_w[t] = (int16_t)((d.S - 1 - s) * d.dil_w);
        L.widx[t] = (uint8_t)t;
      }
    L.w = wt; L.Kout = d.C; L.RS = RS;
    if (grouped) { L.grp_out = Cg; L.grp_in = Kg; L.wC = wCd; }
    L.out = dx; L.bias = nullptr; L.flags = 0;
    L.addend = addend;
    L.addend_mask = addend_mask;
    L.bn_x = bn_x; L.bn_mask = bn_mask; L.bn_mean = bn_mean;
    L.stats = bn_x ? bn_partials : nullptr;
    L.tiled2d = (RS == 1 && d.pad_h == 0 && d.pad_w == 0);
    return launch_fprop(L, st);
  }

  // ---- strided: one stride-1 sub-convolution per output parity class (a, b), each writing its
  // pixels of dx directly; classes without a tap are zero-filled by one small kernel
  CNB_REQUIRE(!addend && !bn_x, CNB_ERR_UNSUPPORTED,
              "conv_tc_dgrad: the addend / BatchNorm-sum epilogues are stride-1 only");
  uint32_t zero_mask = 0;
  for (int a = 0; a < d.stride_h; ++a)
    for (int b = 0; b < d.stride_w; ++b) {
      const int cls = a * d.stride_w + b;
      const int Ha = (d.H - a + d.stride_h - 1) / d.stride_h;
      const int Wb = (d.W - b + d.stride_w - 1) / d.stride_w;
      if (Ha <= 0 || Wb <= 0) continue;
      // taps of this class: (a + pad_h - r*dil_h) % sh == 0, offset dr = (...) / sh
      int drs[16], rr[16], nr = 0, dss[16], ss[16], ns = 0;
      for (int r = 0; r < d.R; ++r) {
        int v = a + d.pad_h - r * d.dil_h;
        if (((v % d.stride_h) + d.stride_h) % d.stride_h == 0 && nr < 16) {
          drs[nr] = (v >= 0) ? v / d.stride_h : -((-v) / d.stride_h);
          rr[nr++] = r;
        }
      }
      for (int s = 0; s < d.S; ++s) {
        int v = b + d.pad_w - s * d.dil_w;
        if (((v % d.stride_w) + d.stride_w) % d.stride_w == 0 && ns < 16) {
          dss[ns] = (v >= 0) ? v / d.stride_w : -((-v) / d.stride_w);
          ss[ns++] = s;
        }
      }
      if (nr == 0 || ns == 0) { zero_mask |= 1u << cls; continue; }  // no contribution -> zeros
      int lo_h = drs[0], lo_w = dss[0];
      for (int i = 1; i < nr; ++i) lo_h = drs[i] < lo_h ? drs[i] : lo_h;
      for (int i = 1; i < ns; ++i) lo_w = dss[i] < lo_w ? dss[i] : lo_w;
      FpropLaunch L;
      memset(&L, 0, sizeof(L));
      L.a = dy; L.aN = d.N; L.aH = P; L.aW = Q; L.aC = d.K;
      L.lower_h = lo_h; L.lower_w = lo_w;
      L.upper_h = lo_h + Ha - P; L.upper_w = lo_w + Wb - Q;
      CNB_REQUIRE(corners_ok(L.lower_h) && corners_ok(L.lower_w) && corners_ok(L.upper_h) &&
                      corners_ok(L.upper_w),
                  CNB_ERR_UNSUPPORTED, "conv_tc_dgrad(strided): corners outside the im2col range");
      L.trav_h = 1; L.trav_w = 1;
      L.outP = Ha; L.outQ = Wb;
      L.ntaps = 0;
      for (int i = 0; i < nr; ++i)
        for (int j = 0; j < ns; ++j) {
          int t = L.ntaps++;
          L.off_h[t] = (int16_t)(drs[i] - lo_h);
          L.off_w[t] = (int16_t)(dss[j] - lo_w);
          L.widx[t] = (uint8_t)(rr[i] * d.S + ss[j]);
        }
      L.w = wt; L.Kout = d.C; L.RS = RS;
      if (grouped) { L.grp_out = Cg; L.grp_in = Kg; L.wC = wCd; }
      L.out = nullptr; L.bias = nullptr; L.flags = 0;
      L.tiled2d = false;
      L.scatter = dx;
      L.sc_H = d.H; L.sc_W = d.W; L.sc_sh = d.stride_h; L.sc_sw = d.stride_w; L.sc_a = a; L.sc_b = b;
      CNB_TRY(launch_fprop(L, st));
    }
  if (zero_mask) {
    ZeroClassP zp;
    zp.N = d.N; zp.H = d.H; zp.W = d.W; zp.C = d.C; zp.sh = d.stride_h; zp.sw = d.stride_w;
    zp.zero_mask = zero_mask;
    int64_t total = (int64_t)d.N * d.H * d.W * (d.C / 8);
    dgrad_zero_classes_k<<<grid_for(total, 256), 256, 0, st>>>(zp, (bf16*)dx);
    return check_launch("dgrad_zero_classes");
  }
  return CNB_OK;
}

template <int NT, int SA, int SB, bool IM2COL, int TS = 1>
static int launch_wgrad_inst(const WgradParams& prm, int grid, cudaStream_t st) {
  constexpr size_t smem = (size_t)SA * 2 * 64 * 128 + (size_t)SB * TS * (NT / 64) * 64 * 128 +
                          (2 * SA + 2 * SB + 1) * 8 + 16 + 1024;
  static bool attr_done = false;
  auto kern = conv_wgrad_tc_k<NT, SA, SB, IM2COL, TS>;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    CNB_REQUIRE(e == cudaSuccess, CNB_ERR_CUDA, "cudaFuncSetAttribute(wgrad): %s",
                cudaGetErrorString(e));
    attr_done = true;
  }
  kern<<<grid, kTcThreads, smem, st>>>(prm);
  return check_launch("conv_wgrad_tc");
}

int conv_tc_wgrad(const cnb_conv_desc& d, const void* x, const void* dy, float* dw, void* ws,
                  size_t ws_bytes, cudaStream_t st) {
  CNB_TRY(load_driver_fns());
  int P, Q;
  conv_out_dims(d, P, Q);
  const int RS = d.R * d.S;
  const int M = d.N * P * Q;
  WgradParams prm;
  memset(&prm, 0, sizeof(prm));
  int NT;
  wgrad_geometry(d, NT, prm.taps_per_cta, prm.tap_groups, prm.cin_tiles, prm.cout_tiles, prm.splits,
                 prm.pixblk_per_split);
  const bool grouped = d.groups > 1;
  const int CW = grouped ? prm.cin_tiles * NT : d.C;   // innermost extent of the kernel's output tensor
  const int64_t wsz = (int64_t)d.K * RS * CW;
  float* part = dw;
  if (prm.splits > 1 || grouped) {
    CNB_REQUIRE(ws && ws_bytes >= (size_t)prm.splits * wsz * sizeof(float) + 1024, CNB_ERR_WORKSPACE,
                "conv_tc_wgrad: workspace %zu < %zu", ws_bytes,
                (size_t)prm.splits * wsz * sizeof(float) + 1024);
    part = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(ws) + 255) & ~(uintptr_t)255);
  }
  const bool tiled2d = (RS == 1 && d.stride_h == 1 && d.stride_w == 1 && d.pad_h == 0 && d.pad_w == 0);
  {
    uint64_t dims[2] = {(uint64_t)d.K, (uint64_t)M};
    uint64_t str[1] = {(uint64_t)d.K * 2};
    uint32_t box[2] = {64, 64};
    CNB_TRY(make_tiled_map(&prm.tmDy, dy, 2, dims, str, box, "dy"));
  }
  if (tiled2d) {
    uint64_t dims[2] = {(uint64_t)d.C, (uint64_t)M};
    uint64_t str[1] = {(uint64_t)d.C * 2};
    uint32_t box[2] = {64, 64};
    CNB_TRY(make_tiled_map(&prm.tmX, x, 2, dims, str, box, "x(2d)"));
  } else {
    int upper_h = d.pad_h - (d.R - 1) * d.dil_h;
    int upper_w = d.pad_w - (d.S - 1) * d.dil_w;
    CNB_TRY(make_im2col_map(&prm.tmX, x, d.N, d.H, d.W, d.C, -d.pad_w, -d.pad_h, upper_w, upper_h,
                            d.stride_w, d.stride_h, 64, "x(im2col)"));
  }
  prm.out = part;
  {
    cuuint64_t gdim[4] = {(cuuint64_t)CW, (cuuint64_t)RS, (cuuint64_t)d.K,
                          (cuuint64_t)(prm.splits > 1 ? prm.splits : 1)};
    cuuint64_t gstr[3] = {(cuuint64_t)CW * 4, (cuuint64_t)RS * CW * 4, (cuuint64_t)wsz * 4};
    cuuint32_t box[4] = {32, 1, (cuuint32_t)TILE_M, 1};
    cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = encode_with_context(part, [&] {
      return g_encodeTiled(&prm.tmOut, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, part, gdim, gstr, box, es,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    });
    CNB_REQUIRE(r == CUDA_SUCCESS, CNB_ERR_CUDA, "cuTensorMapEncodeTiled(wgrad out) failed: CUresult %d",
                (int)r);
  }
  prm.M = M;
  prm.outP = P; prm.outQ = Q;
  prm.lower_h = -d.pad_h; prm.lower_w = -d.pad_w;
  prm.trav_h = d.stride_h; prm.trav_w = d.stride_w;
  prm.K = d.K; prm.Cg = CW; prm.RS = RS;
  prm.grp_out = grouped ? d.K / d.groups : 0;
  prm.grp_in = grouped ? d.C / d.groups : 0;
  prm.num_taps = RS;
  for (int r = 0; r < d.R; ++r)
    for (int s = 0; s < d.S; ++s) {
      prm.off_h[r * d.S + s] = (int16_t)(r * d.dil_h);
      prm.off_w[r * d.S + s] = (int16_t)(s * d.dil_w);
    }
  const int64_t grid64 = (int64_t)prm.splits * prm.cout_tiles * prm.cin_tiles * prm.tap_groups;
  CNB_REQUIRE(grid64 < (1ll << 31), CNB_ERR_UNSUPPORTED, "conv_tc_wgrad: grid too large");
  const int grid = (int)grid64;
  int s;
  // Ring depths <dy stages, x stages>.  The producer is one thread walking pixel blocks in order
  // (dy tile, then one x tile per tap), so it runs ahead by min(dy stages x taps per CTA, x stages)
  // MMA groups: 1x1 layers (one tap per block) need as many dy stages as x stages.
  if (tiled2d) {
    if (NT == 64) s = launch_wgrad_inst<64, 8, 8, false>(prm, grid, st);
    else if (NT == 128) s = launch_wgrad_inst<128, 6, 6, false>(prm, grid, st);
    else s = launch_wgrad_inst<256, 4, 4, false>(prm, grid, st);
  } else {
    if (NT == 64) s = launch_wgrad_inst<64, 3, 5, true, 4>(prm, grid, st);
    else if (NT == 128) s = launch_wgrad_inst<128, 3, 10, true>(prm, grid, st);
    else s = launch_wgrad_inst<256, 3, 5, true>(prm, grid, st);
  }
  CNB_TRY(s);
  if (grouped) {
    const int64_t n = (int64_t)d.K * RS * (d.C / d.groups);
    extract_grouped_wgrad_k<<<grid_for(n, 256), 256, 0, st>>>(part, dw, prm.splits, d.K, RS,
                                                             d.C / d.groups, d.K / d.groups, CW);
    return check_launch("extract_grouped_wgrad");
  }
  if (prm.splits > 1) {
    sum_splits_tc_k<<<grid_for(wsz, 256), 256, 0, st>>>(part, dw, prm.splits, wsz);
    CNB_TRY(check_launch("wgrad_sum_splits"));
  }
  return CNB_OK;
}

}  // namespace cnb
