// pair_tc.cu — two convolutions of a NonBt1D block in ONE kernel (reference models/erfnet.py:65-71, 77-81):
//
//     T = relu(bn1(conv3x1(X)))            (ConvBNAct(c, c, (3, 1)))
//     Y = act(bn2(conv1x3(T) [+ R]))       (ConvBNAct(c, c, (1, 3)), or the block tail: Conv2d(1, 3) -> += x -> BN -> act)
//
// The intermediate T never leaves the SM: it is written by the epilogue warps into shared memory in exactly the
// swizzled K-major layout the second convolution's tcgen05.mma reads as its A operand.  HBM traffic of the pair drops
// from (read X, write T, read T, write Y) to (read X, write Y) — the layer-wise ceiling SURVEY §8d computes no longer
// binds.  Undilated pairs of 16- and 64-channel blocks (all of ERFNet's layer3_7, layer18_19, layer21_22).
//
// Geometry.  An output tile is TH rows x TW columns; every tile in shared memory uses one pixel pitch P = TW + 2
// (a power of two), so that with the flattened pixel index g = row * P + column
//     stage 1 (vertical taps):    T[g] = sum_t W1[t] . X[g + t*P]          X tile = (TH + 2) x P pixels, one TMA box
//     stage 2 (horizontal taps):  Y[g] = sum_t W2[t] . T[g + t - 1]
// are both "1-D convolutions over g": each 128-pixel chunk of g is one M = 128 accumulator and every tap is the same
// shared-memory tile read through a descriptor whose start address is shifted by t*P (resp. t - 1) pixel rows.  (A
// start address that is not a multiple of the swizzle atom is fine: the tensor core derives the swizzle phase from
// the shared-memory address itself — DESIGN.md §3, the fact conv_tc.cu's halo modes rest on.)  The first and last
// column of a tile row are halo: stage 2 produces garbage there, which is never stored.  T outside the image is
// forced to zero by the stage-1 epilogue (the second convolution pads T, not X).
//
// BatchNorm without epilogue arithmetic.  When the caller hands over weights that already carry the BatchNorm scale
// (w' = bf16(scale[co] * w), flag bits of `w_scaled`), the shift enters the accumulator through the tensor core as
// well: the first MMA of every chunk multiplies a constant [128 x 16] tile whose first two columns are 1 by a
// [c x 16] tile whose first two columns hold the bf16 pair (hi, lo) with hi + lo = shift[co] (16 mantissa bits), and
// the convolution's MMAs accumulate on top.  The epilogue of such a stage is tcgen05.ld -> cvt.rn.relu.bf16x2 ->
// st.shared: no shared-memory reads of scale / shift, no FMAs — these epilogues are issue- and shared-memory-pipe
// bound, not bandwidth bound.  A PRE residual still needs its scale: v = acc + scale[co] * res.
//
// Roles (576 threads, one persistent CTA per SM): warp 0 streams X tiles through an mbarrier ring with TMA; warp 1
// issues both stages' MMAs (stage 1 of tile i+1 ahead of stage 2 of tile i); warps 2-17 run the two epilogues
// interleaved — A(i): stage-1 accumulator -> BN/ReLU -> T in shared memory; B(i-1): stage-2 accumulator (+ residual,
// which a TMA load has placed in the staging buffer) -> BN/act -> staging -> TMA store.  Accumulators are double
// buffered in TMEM for both stages.
#include <cuda.h>

#include <algorithm>
#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "ptx.cuh"

namespace rtseg {

namespace {

constexpr int kPairEpiWarps = 16;   // four warps per TMEM lane quarter: the epilogues are issue-bound, not bandwidth-bound
constexpr int kPairEpiThreads = 32 * kPairEpiWarps;
constexpr int kPairThreads = 64 + kPairEpiThreads;
constexpr int kPairMaxNx = 6;
constexpr uint32_t kPairSmemBudget = 218 * 1024;

struct PairParams {
  int n, h, w;
  int th, tw, p, p_shift;   // tile rows, valid columns (p - 2), pixel pitch (power of two)
  int mc;                   // 128-pixel chunks per stage = th * p / 128
  int tiles_w, tiles_h;
  int num_tiles;
  int c;                    // channels = K block
  int x_coff, y_coff, r_coff;
  int nx;                   // X ring depth
  int has_res, out_bufs;
  int fold1, fold2;         // stage 1 / 2: weights carry the BatchNorm scale and the shift enters through the bias MMA
  uint32_t x_stage_bytes, t1_bytes, out_buf_bytes, w_block_bytes, x_tx_bytes, o_tx_bytes;
  uint32_t tmem_cols;
  float lo1, hi1, lo2, hi2;
};

struct PairMaps {
  CUtensorMap x, w1, w2, y, r;
};

#ifdef RTSEG_PAIR_TRACE
// Development aid (tools/pair_trace.py builds with -DRTSEG_PAIR_TRACE): CTA 0 records clock64() at the hand-over
// points of its first tiles.  Not compiled into the product library.
__device__ long long g_pair_trace[64][16];
#define PAIR_TRACE(tile_, ev_) do { if (blockIdx.x == 0 && (tile_) < 64) g_pair_trace[(tile_)][(ev_)] = clock64(); } while (0)
#else
#define PAIR_TRACE(tile_, ev_) do { } while (0)
#endif

__device__ __forceinline__ void pair_bar_sync(int id) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "n"(kPairEpiThreads) : "memory");
}

template <int KB>
__device__ __forceinline__ uint64_t pair_desc(uint32_t addr) {
  constexpr uint64_t layout = KB == 64 ? 2 : (KB == 32 ? 4 : 6);  // SWIZZLE_128B / 64B / 32B
  constexpr uint64_t sbo = (8 * KB * 2) >> 4;                      // 8 consecutive pixel rows per core-matrix group
  return (uint64_t)((addr & 0x3ffffu) >> 4) | (1ull << 16) | (sbo << 32) | (1ull << 46) | (layout << 61);
}

// 16 accumulator columns (+ residual) -> folded BatchNorm -> clamp -> 8 packed bf16 pairs.  RELU: the clamp is
// [0, inf) and rides in the conversion; otherwise NaN-propagating max / min with (lo, hi).
// FOLD: the accumulator already holds scale * conv + shift (pre-scaled weights, bias MMA); a residual then enters as
// acc + scale * res.
template <bool RELU, bool RES, bool FOLD>
__device__ __forceinline__ void affine_pack16(const uint32_t (&v)[16], const float (&rv)[16], uint32_t s_sc, uint32_t s_sh,
                                              float lo, float hi, uint32_t (&pk)[8]) {
#pragma unroll
  for (int e4 = 0; e4 < 16; e4 += 4) {
    float o[4] = {__uint_as_float(v[e4]), __uint_as_float(v[e4 + 1]), __uint_as_float(v[e4 + 2]), __uint_as_float(v[e4 + 3])};
    if constexpr (FOLD) {
      if constexpr (RES) {
        const float4 sc = lds128(s_sc + (uint32_t)e4 * 4);
        o[0] = fmaf(rv[e4], sc.x, o[0]), o[1] = fmaf(rv[e4 + 1], sc.y, o[1]);
        o[2] = fmaf(rv[e4 + 2], sc.z, o[2]), o[3] = fmaf(rv[e4 + 3], sc.w, o[3]);
      }
    } else {
      const float4 sc = lds128(s_sc + (uint32_t)e4 * 4), sh = lds128(s_sh + (uint32_t)e4 * 4);
      if constexpr (RES) {
#pragma unroll
        for (int u = 0; u < 4; ++u) o[u] += rv[e4 + u];   // += x before the BatchNorm (reference models/erfnet.py:79-80)
      }
      o[0] = fmaf(o[0], sc.x, sh.x), o[1] = fmaf(o[1], sc.y, sh.y), o[2] = fmaf(o[2], sc.z, sh.z), o[3] = fmaf(o[3], sc.w, sh.w);
    }
    if constexpr (RELU) {
      pk[(e4 >> 1) + 0] = pack_relu_bf16(o[0], o[1]);
      pk[(e4 >> 1) + 1] = pack_relu_bf16(o[2], o[3]);
    } else {
#pragma unroll
      for (int u = 0; u < 4; ++u) o[u] = fmin_nan(fmax_nan(o[u], lo), hi);
      pk[(e4 >> 1) + 0] = pack_bf16(o[0], o[1]);
      pk[(e4 >> 1) + 1] = pack_bf16(o[2], o[3]);
    }
  }
}

__device__ __forceinline__ float g_aff_src(const float *v, int i) { return v ? v[i] : 0.f; }

// 32-byte-row K-major operand (K = 16), SWIZZLE_32B: the bias MMA's two tiles.
__device__ __forceinline__ uint64_t bias_desc(uint32_t addr) {
  return (uint64_t)((addr & 0x3ffffu) >> 4) | (1ull << 16) | ((uint64_t)(256 >> 4) << 32) | (1ull << 46) | (6ull << 61);
}

template <int KB, bool RES>
__global__ void __launch_bounds__(kPairThreads, 1) conv_pair_tc_kernel(const __grid_constant__ PairMaps maps,
                                                                       const PairParams p, const float *__restrict__ scale1,
                                                                       const float *__restrict__ shift1,
                                                                       const float *__restrict__ scale2,
                                                                       const float *__restrict__ shift2) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  constexpr uint32_t kRowB = KB * 2;                           // bytes of one pixel row of a tile
  constexpr uint32_t kSwMask = KB == 64 ? 7u : (KB == 32 ? 3u : 1u);
  constexpr int kGroups = KB / 16;                             // 16-column groups per chunk
  const uint32_t s_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t s_w = s_base;                                 // 6 weight blocks: W1[0..2], W2[0..2]
  const uint32_t s_x = s_w + 6 * p.w_block_bytes;
  const uint32_t s_t1 = s_x + (uint32_t)p.nx * p.x_stage_bytes;
  const uint32_t s_out = s_t1 + 2 * p.t1_bytes;                // T is double buffered: A(i+1) never waits for stage 2 of tile i
  const uint32_t s_aff = s_out + (uint32_t)p.out_bufs * p.out_buf_bytes;  // scale1, shift1, scale2, shift2
  const uint32_t s_ones = (s_aff + 4u * KB * 4 + 255u) & ~255u;            // [128 x 16] bf16, SWIZZLE_32B: columns 0, 1 = 1
  const uint32_t s_bias = s_ones + 128u * 32u;                              // 2 x [KB x 16] bf16: columns 0, 1 = (hi, lo) of shift
  const uint32_t s_bars = s_bias + 2u * KB * 32u;
  const uint32_t full_x = s_bars, empty_x = full_x + 8 * kPairMaxNx;
  const uint32_t w_bar = empty_x + 8 * kPairMaxNx;
  const uint32_t s1_full = w_bar + 8, s1_empty = s1_full + 16, s2_full = s1_empty + 16, s2_empty = s2_full + 16;
  const uint32_t t1_full = s2_empty + 16, res_full = t1_full + 16;  // t1_full[2]: one per T buffer; res_full[3]: one per staging buffer
  uint8_t *gen = smem_raw + (s_base - smem_u32(smem_raw));
  float *g_aff = reinterpret_cast<float *>(gen + (s_aff - s_base));
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(gen + (res_full + 24 - s_base));
  volatile uint32_t *dead = tmem_slot + 1;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.nx; ++s) mbar_init(full_x + 8 * s, 1), mbar_init(empty_x + 8 * s, 1);
    mbar_init(w_bar, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(s1_full + 8 * s, 1), mbar_init(s1_empty + 8 * s, kPairEpiWarps);
      mbar_init(s2_full + 8 * s, 1), mbar_init(s2_empty + 8 * s, kPairEpiWarps);
    }
    for (int s = 0; s < 3; ++s) mbar_init(res_full + 8 * s, 1);
    // one barrier per T buffer: a warp that runs one tile ahead arrives on the other barrier, never on this one early
    for (int s = 0; s < 2; ++s) mbar_init(t1_full + 8 * s, kPairEpiWarps);
    *dead = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int i = threadIdx.x; i < KB; i += blockDim.x) {
    g_aff[i] = scale1 ? scale1[i] : 1.f;
    g_aff[KB + i] = shift1 ? shift1[i] : 0.f;
    g_aff[2 * KB + i] = scale2 ? scale2[i] : 1.f;
    g_aff[3 * KB + i] = shift2 ? shift2[i] : 0.f;
  }
  // constant operand tiles of the bias MMA (K-major, 32-byte rows, SWIZZLE_32B: 16-byte chunk j of row r sits at
  // r*32 + ((j ^ ((r >> 2) & 1)) << 4))
  for (int r = threadIdx.x; r < 128 + 2 * KB; r += blockDim.x) {
    uint32_t w0 = 0x3f803f80u;   // (1.0, 1.0)
    uint32_t base = s_ones;
    int rr = r;
    if (r >= 128) {
      rr = r - 128;
      const int st = rr / KB, co = rr - st * KB;
      const float sh = g_aff_src(st == 0 ? shift1 : shift2, co);
      const __nv_bfloat16 hi = __float2bfloat16_rn(sh);
      const __nv_bfloat16 lo = __float2bfloat16_rn(sh - __bfloat162float(hi));
      w0 = (uint32_t)__bfloat16_as_ushort(hi) | ((uint32_t)__bfloat16_as_ushort(lo) << 16);
      base = s_bias;
    }
    const uint32_t a = base + (uint32_t)rr * 32u;
    const uint32_t c0 = ((uint32_t)(rr >> 2) & 1u) << 4;
    sts128(a + c0, w0, 0u, 0u, 0u);
    sts128(a + (c0 ^ 16u), 0u, 0u, 0u, 0u);
  }
  fence_async_smem();
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(p.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t acc_cols = (uint32_t)(p.mc * KB);   // TMEM columns of one accumulator buffer (mc chunks x KB columns)
  const int n_my = (p.num_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;   // tiles of this CTA (>= 1)
  auto tile_coords = [&](int j, int &w0, int &h0, int &img) {   // j-th tile of this CTA
    const int t = (int)blockIdx.x + j * (int)gridDim.x;
    const int twi = t % p.tiles_w, r = t / p.tiles_w;
    w0 = twi * p.tw, h0 = (r % p.tiles_h) * p.th, img = r / p.tiles_h;
  };

  if (warp == 0) {
    // ===================================================== TMA producer ====
    if (lane == 0) {
      prefetch_tmap(&maps.x);
      prefetch_tmap(&maps.y);
      mbar_expect_tx(w_bar, 6u * (uint32_t)(KB * KB * 2));
      for (int t = 0; t < 3; ++t) {
        tma_load_2d(s_w + (uint32_t)t * p.w_block_bytes, &maps.w1, w_bar, 0, t * KB);
        tma_load_2d(s_w + (uint32_t)(3 + t) * p.w_block_bytes, &maps.w2, w_bar, 0, t * KB);
      }
      uint32_t stage = 0, phase = 0;
      for (int j = 0; j < n_my; ++j) {
        int w0, h0, img;
        tile_coords(j, w0, h0, img);
        mbar_wait(empty_x + 8 * stage, phase ^ 1, dead);
        mbar_expect_tx(full_x + 8 * stage, p.x_tx_bytes);
        tma_load_4d(s_x + stage * p.x_stage_bytes, &maps.x, full_x + 8 * stage, p.x_coff, w0 - 1, h0 - 1, img);
        PAIR_TRACE(j, 0);
        if (++stage == (uint32_t)p.nx) stage = 0, phase ^= 1;
      }
    }
  } else if (warp == 1) {
    // ======================================================= MMA issuer ====
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(KB >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint64_t w_desc0 = pair_desc<KB>(s_w);
    const uint32_t wblk16 = p.w_block_bytes >> 4;
    const uint32_t chunk16 = (128u * kRowB) >> 4;          // one 128-pixel chunk of a tile
    const uint32_t vtap16 = ((uint32_t)p.p * kRowB) >> 4;  // one tile row down
    const uint64_t ones_d = bias_desc(s_ones), bias_d = bias_desc(s_bias);
    mbar_wait(w_bar, 0, dead);
    uint32_t xs = 0, xph = 0;
    auto issue_s1 = [&](int j) {
      const uint32_t b = (uint32_t)j & 1u, use = (uint32_t)j >> 1;
      mbar_wait(full_x + 8 * xs, xph, dead);
      if (lane == 0) PAIR_TRACE(j, 1);
      mbar_wait(s1_empty + 8 * b, (use & 1u) ^ 1u, dead);
      tc_fence_after();
      if (lane == 0) PAIR_TRACE(j, 2);
      if (elect_one()) {
        const uint64_t a0 = pair_desc<KB>(s_x + xs * p.x_stage_bytes);
        for (int m = 0; m < p.mc; ++m) {
          const uint32_t d = tmem_base + b * acc_cols + (uint32_t)(m * KB);
          if (p.fold1) umma_bf16(d, ones_d, bias_d, idesc, 0u);   // accumulator := shift1 (exact to 16 mantissa bits)
#pragma unroll
          for (int t = 0; t < 3; ++t) {
            const uint64_t a_d = a0 + (uint64_t)((uint32_t)m * chunk16 + (uint32_t)t * vtap16);
            const uint64_t b_d = w_desc0 + (uint64_t)((uint32_t)t * wblk16);
#pragma unroll
            for (int k = 0; k < KB / 16; ++k) umma_bf16(d, a_d + 2 * k, b_d + 2 * k, idesc, (t | k | p.fold1) ? 1u : 0u);
          }
        }
        umma_commit(empty_x + 8 * xs);      // the X tile is free once these MMAs have read it
        umma_commit(s1_full + 8 * b);       // ... and the stage-1 accumulators are complete
      }
      __syncwarp();
      if (++xs == (uint32_t)p.nx) xs = 0, xph ^= 1;
    };
    auto issue_s2 = [&](int i) {
      const uint32_t b = (uint32_t)i & 1u, use = (uint32_t)i >> 1;
      mbar_wait(t1_full + 8 * b, use & 1u, dead);               // every epilogue warp has written (and fenced) its part of T[b]
      mbar_wait(s2_empty + 8 * b, (use & 1u) ^ 1u, dead);
      tc_fence_after();
      if (lane == 0) PAIR_TRACE(i, 3);
      if (elect_one()) {
        const uint64_t a0 = pair_desc<KB>(s_t1 + b * p.t1_bytes) - (uint64_t)(kRowB >> 4);   // tap 0 starts one pixel row before g
        for (int m = 0; m < p.mc; ++m) {
          const uint32_t d = tmem_base + (2u + b) * acc_cols + (uint32_t)(m * KB);
          if (p.fold2) umma_bf16(d, ones_d, bias_d + (uint64_t)((KB * 32) >> 4), idesc, 0u);   // accumulator := shift2
#pragma unroll
          for (int t = 0; t < 3; ++t) {
            const uint64_t a_d = a0 + (uint64_t)((uint32_t)m * chunk16 + (uint32_t)t * (kRowB >> 4));
            const uint64_t b_d = w_desc0 + (uint64_t)((uint32_t)(3 + t) * wblk16);
#pragma unroll
            for (int k = 0; k < KB / 16; ++k) umma_bf16(d, a_d + 2 * k, b_d + 2 * k, idesc, (t | k | p.fold2) ? 1u : 0u);
          }
        }
        umma_commit(s2_full + 8 * b);       // stage-2 accumulators complete; T may be overwritten
      }
      __syncwarp();
    };
    issue_s1(0);
    for (int i = 0; i < n_my; ++i) {
      if (i + 1 < n_my) issue_s1(i + 1);
      issue_s2(i);
    }
  } else {
    // ========================================================= epilogue ====
    const int e = warp - 2;
    const int q = warp & 3;          // TMEM lane quarter of this warp
    const int sub = e >> 2;          // which of the four warps of the quarter: items sub, sub + 4, ...
    const int row = q * 32 + lane;   // pixel inside a chunk
    const bool elected = threadIdx.x == 64;
    const int items = p.mc * kGroups;                 // (chunk, 16-column group) pairs per stage
    const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
    const uint32_t s_sc1 = s_aff, s_sh1 = s_aff + KB * 4, s_sc2 = s_aff + 2 * KB * 4, s_sh2 = s_aff + 3 * KB * 4;
    const bool relu1 = p.lo1 == 0.f && p.hi1 > 3.0e38f, relu2 = p.lo2 == 0.f && p.hi2 > 3.0e38f;   // warp-uniform

    if (RES && elected) {  // residual tiles of the first units: straight into the staging buffers
      prefetch_tmap(&maps.r);
      for (int j = 0; j < p.out_bufs && j < n_my; ++j) {
        int w0, h0, img;
        tile_coords(j, w0, h0, img);
        mbar_expect_tx(res_full + 8 * j, p.o_tx_bytes);
        tma_load_4d(s_out + (uint32_t)j * p.out_buf_bytes, &maps.r, res_full + 8 * j, p.r_coff, w0, h0, img);
      }
    }
    // tile -> (image, tile row, tile column), advanced by the decomposed grid stride: no division per tile
    int twi, thi, img, step_w, step_h, step_n;
    {
      const int r0 = (int)blockIdx.x / p.tiles_w, rs = (int)gridDim.x / p.tiles_w;
      twi = (int)blockIdx.x % p.tiles_w, thi = r0 % p.tiles_h, img = r0 / p.tiles_h;
      step_w = (int)gridDim.x % p.tiles_w, step_h = rs % p.tiles_h, step_n = rs / p.tiles_h;
    }
    int pw0 = 0, ph0 = 0, pimg = 0;   // coordinates of the previous tile (stage B works one tile behind stage A)
    uint32_t ob = 0, ob_use = 0;      // staging buffer of the tile stage B handles, and how often it has been used
    float no_res[16];                 // (placeholder operand of the residual-free instantiations)
#pragma unroll
    for (int u = 0; u < 16; ++u) no_res[u] = 0.f;

    for (int i = 0; i <= n_my; ++i) {
      const int w0 = twi * p.tw, h0 = thi * p.th, img0 = img;
      if (i < n_my) {
        // ---- A(i): stage-1 accumulator -> BN + clamp -> T (bf16, swizzled K-major) ----
        const uint32_t b = (uint32_t)i & 1u, use = (uint32_t)i >> 1;
        mbar_wait(s1_full + 8 * b, use & 1u, dead);
        if (elected) PAIR_TRACE(i, 4);
        // T[b] is free: stage 2 of tile i-2 read it, and B(i-2) — earlier in this warp's program order — waited for
        // that stage's accumulator barrier
        tc_fence_after();
        const uint32_t t_buf = s_t1 + b * p.t1_bytes;
        const bool border = w0 == 0 || w0 + p.tw >= p.w;    // only these tiles hold T columns outside the image
        for (int k = sub; k < items; k += 4) {
          const int m = k / kGroups, c0 = (k - m * kGroups) * 16;
          uint32_t v[16], pk[8];
          tmem_ld16(lane_base + b * acc_cols + (uint32_t)(m * KB + c0), v);
          tmem_ld_wait();
          const uint32_t sc_a = s_sc1 + (uint32_t)c0 * 4, sh_a = s_sh1 + (uint32_t)c0 * 4;
          if (p.fold1) {
            if (relu1) affine_pack16<true, false, true>(v, no_res, sc_a, sh_a, p.lo1, p.hi1, pk);
            else affine_pack16<false, false, true>(v, no_res, sc_a, sh_a, p.lo1, p.hi1, pk);
          } else {
            if (relu1) affine_pack16<true, false, false>(v, no_res, sc_a, sh_a, p.lo1, p.hi1, pk);
            else affine_pack16<false, false, false>(v, no_res, sc_a, sh_a, p.lo1, p.hi1, pk);
          }
          const int g = m * 128 + row;
          if (border) {   // T outside the image is the second conv's zero padding
            const int cimg = w0 - 1 + (g & (p.p - 1));
            if (cimg < 0 || cimg >= p.w) {
#pragma unroll
              for (int u = 0; u < 8; ++u) pk[u] = 0u;
            }
          }
          const uint32_t off0 = (uint32_t)g * kRowB + (uint32_t)c0 * 2;
          const uint32_t a = t_buf + (off0 ^ (((off0 >> 7) & kSwMask) << 4));
          sts128(a, pk[0], pk[1], pk[2], pk[3]);
          sts128(a ^ 16u, pk[4], pk[5], pk[6], pk[7]);
        }
        fence_async_smem();     // T is read by the tensor core (async proxy)
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(s1_empty + 8 * b);
          mbar_arrive(t1_full + 8 * b);
        }
        if (elected) PAIR_TRACE(i, 5);
      }
      if (i >= 1) {
        // ---- B(j): stage-2 accumulator (+ residual) -> BN + clamp -> staging (compact TH x TW) -> TMA store ----
        const int j = i - 1;
        const uint32_t b = (uint32_t)j & 1u, use = (uint32_t)j >> 1;
        const uint32_t buf = s_out + ob * p.out_buf_bytes;
        mbar_wait(s2_full + 8 * b, use & 1u, dead);
        if (elected) PAIR_TRACE(j, 6);
        if constexpr (RES) mbar_wait(res_full + 8 * ob, ob_use & 1u, dead);
        tc_fence_after();
        if constexpr (!RES) pair_bar_sync(2);   // the elected thread has seen the store before last read this buffer out
        if (elected) PAIR_TRACE(j, 7);
        for (int k = sub; k < items; k += 4) {
          const int m = k / kGroups, c0 = (k - m * kGroups) * 16;
          uint32_t v[16];
          tmem_ld16(lane_base + (2u + b) * acc_cols + (uint32_t)(m * KB + c0), v);
          tmem_ld_wait();
          const int g = m * 128 + row;
          const int cc = g & (p.p - 1), r = g >> p.p_shift;
          if (cc >= 1 && cc <= p.tw) {        // halo columns carry garbage: not stored
            const uint32_t off0 = (uint32_t)(r * p.tw + cc - 1) * kRowB + (uint32_t)c0 * 2;
            const uint32_t a = buf + (off0 ^ (((off0 >> 7) & kSwMask) << 4));
            float rv[16];
            if constexpr (RES) {
              const float4 f0 = lds128(a), f1 = lds128(a ^ 16u);
              const uint32_t w8[8] = {__float_as_uint(f0.x), __float_as_uint(f0.y), __float_as_uint(f0.z), __float_as_uint(f0.w),
                                      __float_as_uint(f1.x), __float_as_uint(f1.y), __float_as_uint(f1.z), __float_as_uint(f1.w)};
#pragma unroll
              for (int u = 0; u < 8; ++u) rv[2 * u] = bf16_lo(w8[u]), rv[2 * u + 1] = bf16_hi(w8[u]);
            }
            uint32_t pk[8];
            const uint32_t sc_a = s_sc2 + (uint32_t)c0 * 4, sh_a = s_sh2 + (uint32_t)c0 * 4;
            if (p.fold2) {
              if (relu2) affine_pack16<true, RES, true>(v, RES ? rv : no_res, sc_a, sh_a, p.lo2, p.hi2, pk);
              else affine_pack16<false, RES, true>(v, RES ? rv : no_res, sc_a, sh_a, p.lo2, p.hi2, pk);
            } else {
              if (relu2) affine_pack16<true, RES, false>(v, RES ? rv : no_res, sc_a, sh_a, p.lo2, p.hi2, pk);
              else affine_pack16<false, RES, false>(v, RES ? rv : no_res, sc_a, sh_a, p.lo2, p.hi2, pk);
            }
            sts128(a, pk[0], pk[1], pk[2], pk[3]);
            sts128(a ^ 16u, pk[4], pk[5], pk[6], pk[7]);
          }
        }
        fence_async_smem();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(s2_empty + 8 * b);
        if (elected) PAIR_TRACE(j, 8);
        pair_bar_sync(1);        // every warp has written its part of the staging tile
        if (elected) {
          PAIR_TRACE(j, 9);
          tma_store_4d(&maps.y, buf, p.y_coff, pw0, ph0, pimg);
          bulk_commit();
          bulk_wait_read<1>();   // the store BEFORE this one has read its buffer out: that buffer may be written again
          PAIR_TRACE(j, 10);
          if (RES && j >= 1 && j + 2 < n_my) {   // ... first by the residual tile of the unit that uses it next
            int w2, h2, img2;
            tile_coords(j + 2, w2, h2, img2);
            const uint32_t nb = ob == 0 ? (uint32_t)p.out_bufs - 1u : ob - 1u;   // buffer of tile j-1 == buffer of tile j+2
            mbar_expect_tx(res_full + 8 * nb, p.o_tx_bytes);
            tma_load_4d(s_out + nb * p.out_buf_bytes, &maps.r, res_full + 8 * nb, p.r_coff, w2, h2, img2);
          }
        }
        if (++ob == (uint32_t)p.out_bufs) ob = 0, ++ob_use;
      }
      pw0 = w0, ph0 = h0, pimg = img0;
      // next tile of this CTA
      twi += step_w;
      if (twi >= p.tiles_w) twi -= p.tiles_w, ++thi;
      thi += step_h;
      if (thi >= p.tiles_h) thi -= p.tiles_h, ++img;
      img += step_n;
    }
    if (elected) bulk_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
  }
}

typedef CUresult (*EncodeTiledFnPair)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                      const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                      CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFnPair pair_encode_fn() {
  static EncodeTiledFnPair fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void *ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFnPair>(ptr);
  });
  return fn;
}

CUtensorMapSwizzle pair_swizzle(int kb) {
  return kb == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : (kb == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
}

// 4-D bf16 map over channels [0, cdim) of an NHWC buffer with `ct` channels per pixel; box = kb x bw x bh x 1.
int pair_nhwc_map(CUtensorMap *m, const void *base, int cdim, int ct, int w, int h, int n, int kb, int bw, int bh) {
  EncodeTiledFnPair enc = pair_encode_fn();
  if (!enc) RTSEG_FAIL(RTSEG_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  const cuuint64_t dims[4] = {(cuuint64_t)cdim, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)n};
  const cuuint64_t strides[3] = {(cuuint64_t)ct * 2, (cuuint64_t)ct * w * 2, (cuuint64_t)ct * w * h * 2};
  const cuuint32_t box[4] = {(cuuint32_t)kb, (cuuint32_t)bw, (cuuint32_t)bh, 1};
  const cuuint32_t es[4] = {1, 1, 1, 1};
  const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void *>(base), dims, strides, box, es,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, pair_swizzle(kb), CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) RTSEG_FAIL(RTSEG_ERR_CUDA, "cuTensorMapEncodeTiled(pair, NHWC view) failed with CUresult %d", (int)r);
  return RTSEG_OK;
}

int pair_weight_map(CUtensorMap *m, const void *w, int c) {   // [3][c][c] bf16, K-major per tap; box = one tap
  EncodeTiledFnPair enc = pair_encode_fn();
  if (!enc) RTSEG_FAIL(RTSEG_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  const cuuint64_t dims[2] = {(cuuint64_t)c, (cuuint64_t)(3 * c)};
  const cuuint64_t strides[1] = {(cuuint64_t)c * 2};
  const cuuint32_t box[2] = {(cuuint32_t)c, (cuuint32_t)c};
  const cuuint32_t es[2] = {1, 1};
  const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(w), dims, strides, box, es,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, pair_swizzle(c), CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) RTSEG_FAIL(RTSEG_ERR_CUDA, "cuTensorMapEncodeTiled(pair, weights) failed with CUresult %d", (int)r);
  return RTSEG_OK;
}

const char *pair_why(const rtseg_conv_desc *d, int act1) {
  if (d->dtype != RTSEG_BF16) return "dtype must be bf16";
  if (d->cin != d->cout || (d->cin != 16 && d->cin != 64)) return "channels must be 16 or 64 (cin == cout)";
  if (d->stride != 1 || d->dil != 1 || d->groups != 1) return "stride 1, dilation 1, dense only";
  if (d->kh != 3 || d->kw != 1 || d->pad_h != 1 || d->pad_w != 0) return "the descriptor describes the first conv: 3x1, padding (1, 0)";
  if (d->x_ctotal % 8 || d->x_coff % 8 || d->y_ctotal % 8 || d->y_coff % 8) return "channel slices must be 16-byte aligned";
  if (d->res_mode != RTSEG_RES_NONE && d->res_mode != RTSEG_RES_PRE) return "residual: none or PRE (added before the BatchNorm)";
  if (d->res_mode != RTSEG_RES_NONE && (d->r_ctotal % 8 || d->r_coff % 8)) return "residual slice must be 16-byte aligned";
  float lo, hi, lo2, hi2;
  if (act1 != RTSEG_ACT_NONE && act1 != RTSEG_ACT_RELU && act1 != RTSEG_ACT_RELU6) return "first activation must be none / relu / relu6";
  if (epilogue_mode(act1, 0, RTSEG_RES_NONE, nullptr, 0.f, 0.f, lo, hi, lo2, hi2) != 0 ||
      epilogue_mode(d->act, 0, RTSEG_RES_NONE, nullptr, d->act_p0, d->act_p1, lo, hi, lo2, hi2) != 0)
    return "activations must be none / relu / relu6 / hardtanh";
  if (d->w < 8 || d->h < 2) return "image too small";
  return nullptr;
}

}  // namespace

}  // namespace rtseg

using namespace rtseg;

#ifdef RTSEG_PAIR_TRACE
extern "C" int rtseg_pair_trace_read(long long *out_host) {
  return cudaMemcpyFromSymbol(out_host, g_pair_trace, sizeof(long long) * 64 * 16) == cudaSuccess ? 0 : -3;
}
#endif

extern "C" int rtseg_conv_pair_tc_supported(const rtseg_conv_desc *d, int32_t act1) {
  static const int knob = getenv("RTSEG_PAIR_TC") ? atoi(getenv("RTSEG_PAIR_TC")) : 1;
  return d && knob && pair_why(d, act1) == nullptr ? 1 : 0;
}

extern "C" int rtseg_conv_pair_tc(const rtseg_conv_desc *d, const void *x, const void *w1_bf16, const float *scale1,
                                  const float *shift1, int32_t act1, const void *w2_bf16, const float *scale2,
                                  const float *shift2, int32_t w_scaled, const void *res, void *y, rtseg_stream_t stream) {
  RTSEG_REQUIRE(d && x && w1_bf16 && w2_bf16 && y, "NULL argument");
  RTSEG_REQUIRE(d->n > 0 && d->h > 0 && d->w > 0, "non-positive extent");
  if (const char *why = pair_why(d, act1)) RTSEG_FAIL(RTSEG_ERR_UNSUPPORTED, "rtseg_conv_pair_tc: %s", why);
  const bool has_res = d->res_mode == RTSEG_RES_PRE;
  RTSEG_REQUIRE(!has_res || res, "residual mode set but res is NULL");
  RTSEG_REQUIRE(reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(y) % 16 == 0 &&
                    reinterpret_cast<uintptr_t>(res) % 16 == 0 && reinterpret_cast<uintptr_t>(w1_bf16) % 16 == 0 &&
                    reinterpret_cast<uintptr_t>(w2_bf16) % 16 == 0, "tensor pointers must be 16-byte aligned");
  RTSEG_REQUIRE(d->x_coff + d->cin <= d->x_ctotal && d->y_coff + d->cout <= d->y_ctotal, "channel slice outside its buffer");
  const int c = d->cin;
  PairParams p{};
  p.n = d->n, p.h = d->h, p.w = d->w, p.c = c;
  // 64 channels: 128-byte pixel rows, small tiles and a deep X ring; 16 channels: 32-byte rows, large tiles
  if (c == 64) p.th = 4, p.p = 32; else p.th = 8, p.p = 64;
  static const int th_knob = getenv("RTSEG_PAIR_TH") ? atoi(getenv("RTSEG_PAIR_TH")) : 0;
  if (th_knob > 0 && (th_knob * p.p) % 128 == 0) p.th = th_knob;
  p.tw = p.p - 2;
  p.p_shift = p.p == 32 ? 5 : 6;
  p.mc = p.th * p.p / 128;
  p.tiles_w = (int)cdiv(p.w, p.tw), p.tiles_h = (int)cdiv(p.h, p.th);
  const int64_t tiles = (int64_t)p.n * p.tiles_w * p.tiles_h;
  RTSEG_REQUIRE(tiles < (1ll << 30), "too many tiles");
  p.num_tiles = (int)tiles;
  p.x_coff = d->x_coff, p.y_coff = d->y_coff, p.r_coff = d->r_coff;
  p.has_res = has_res ? 1 : 0;
  p.fold1 = (w_scaled & 1) ? 1 : 0, p.fold2 = (w_scaled & 2) ? 1 : 0;
  RTSEG_REQUIRE(!(p.fold2 && has_res) || scale2, "pre-scaled second-stage weights with a residual still need scale2 (v = acc + scale2 * res)");
  p.out_bufs = has_res ? 3 : 2;   // stores never block the epilogue; a residual tile lands two units ahead
  const uint32_t row_b = (uint32_t)c * 2;
  auto up1k = [](uint32_t v) { return (v + 1023u) / 1024u * 1024u; };
  p.x_tx_bytes = (uint32_t)(p.th + 2) * p.p * row_b;
  p.x_stage_bytes = up1k(p.x_tx_bytes);
  p.t1_bytes = up1k((uint32_t)p.th * p.p * row_b);
  p.o_tx_bytes = (uint32_t)p.th * p.tw * row_b;
  p.out_buf_bytes = up1k(p.o_tx_bytes + row_b);   // (the pixel row behind T that tap 2 of the last pixel reads lies in here)
  p.w_block_bytes = up1k((uint32_t)c * row_b);
  const uint32_t misc = 4u * c * 4 + 256 + 128 * 32 + 2u * c * 32 + (2 * kPairMaxNx + 14) * 8 + 64;
  const uint32_t fixed = 6 * p.w_block_bytes + 2 * p.t1_bytes + (uint32_t)p.out_bufs * p.out_buf_bytes + misc;
  RTSEG_REQUIRE(fixed + 2 * p.x_stage_bytes <= kPairSmemBudget, "pair plan: tiles do not fit in shared memory");
  p.nx = (int)std::min<uint32_t>(kPairMaxNx, (kPairSmemBudget - fixed) / p.x_stage_bytes);
  const uint32_t smem = fixed + (uint32_t)p.nx * p.x_stage_bytes + 1024;
  uint32_t cols = 32;
  while (cols < (uint32_t)(4 * p.mc * c)) cols <<= 1;
  RTSEG_REQUIRE(cols <= 512, "pair plan: accumulators exceed tensor memory");
  p.tmem_cols = cols;
  float lo, hi;
  epilogue_mode(act1, 0, RTSEG_RES_NONE, nullptr, 0.f, 0.f, p.lo1, p.hi1, lo, hi);
  epilogue_mode(d->act, 0, RTSEG_RES_NONE, nullptr, d->act_p0, d->act_p1, p.lo2, p.hi2, lo, hi);
  RTSEG_REQUIRE(p.p <= 256 && p.th + 2 <= 256 && p.tw >= 1, "pair plan: box exceeds the TMA limits");

  PairMaps maps;
  if (int rc = pair_nhwc_map(&maps.x, x, d->x_coff + c, d->x_ctotal, d->w, d->h, d->n, c, p.p, p.th + 2)) return rc;
  if (int rc = pair_nhwc_map(&maps.y, y, d->y_coff + c, d->y_ctotal, d->w, d->h, d->n, c, p.tw, p.th)) return rc;
  if (has_res) {
    RTSEG_REQUIRE(d->r_coff + c <= d->r_ctotal, "residual slice outside its buffer");
    if (int rc = pair_nhwc_map(&maps.r, res, d->r_coff + c, d->r_ctotal, d->w, d->h, d->n, c, p.tw, p.th)) return rc;
  } else {
    maps.r = maps.y;
  }
  if (int rc = pair_weight_map(&maps.w1, w1_bf16, c)) return rc;
  if (int rc = pair_weight_map(&maps.w2, w2_bf16, c)) return rc;

  const unsigned grid = (unsigned)std::min<int64_t>(tiles, num_sms());
  cudaStream_t st = as_stream(stream);
#define RTSEG_PAIR(KB_, RES_)                                                                                          \
  do {                                                                                                                 \
    static bool attr_set[64] = {false};                                                                                \
    int dev = 0;                                                                                                       \
    RTSEG_CUDA(cudaGetDevice(&dev));                                                                                   \
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {                                                                      \
      RTSEG_CUDA(cudaFuncSetAttribute(conv_pair_tc_kernel<KB_, RES_>, cudaFuncAttributeMaxDynamicSharedMemorySize,     \
                                      (int)kPairSmemBudget + 2048));                                                   \
      attr_set[dev] = true;                                                                                            \
    }                                                                                                                  \
    conv_pair_tc_kernel<KB_, RES_><<<grid, kPairThreads, smem, st>>>(maps, p, scale1, shift1, scale2, shift2);         \
  } while (0)
  if (c == 64) { if (has_res) RTSEG_PAIR(64, true); else RTSEG_PAIR(64, false); }
  else { if (has_res) RTSEG_PAIR(16, true); else RTSEG_PAIR(16, false); }
#undef RTSEG_PAIR
  RTSEG_LAUNCH_CHECK("conv_pair_tc_kernel");
  return RTSEG_OK;
}
