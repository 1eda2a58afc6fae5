// dwpw_tc.cu — a depthwise 3x3 convolution and the pointwise convolution behind it in ONE kernel:
//
//     T = act1(bn1(dwconv3x3(X)))                      DWConvBNAct        (reference models/modules.py:46-59)
//     Y = act2(bn2(conv1x1(T) [+ R])) [+ R]            PWConvBNAct / conv1x1 + BN (reference models/modules.py:63-69)
//
// i.e. DSConvBNAct (models/modules.py:36-42: Fast-SCNN's classifier), the depthwise + linear-projection tail of
// InvertedResidual (models/fastscnn.py:113-121) and the low-resolution branch of FeatureFusionModule (:80-92).
// The reference (and the two-launch path) writes T — for InvertedResidual the 6x-expanded tensor — to HBM and reads
// it straight back.  Here T exists only as a [128 pixel x 64 channel] bf16 tile in shared memory, written by CUDA
// cores in exactly the swizzled K-major layout the pointwise convolution's tcgen05.mma reads as its A operand.
//
// Roles (one persistent CTA per SM, 448 threads):
//   warp 0       TMA producer: per tile and 64-channel block one box of (TH + 2) x (TW + 2) input pixels (zero fill
//                outside the image = the depthwise convolution's padding) through an mbarrier ring;
//   warp 1       MMA issuer: four K = 16 MMAs per block into a double-buffered TMEM accumulator [128 x N];
//   warps 2-5    epilogue: tcgen05.ld -> BatchNorm / residual / clamp -> swizzled staging -> TMA store;
//   warps 6-13   depthwise workers: a thread owns 4 neighbouring pixels x 8 channels, reads its 3 x 6 window as 16-byte
//                vectors (a quarter warp covers one 128-byte pixel row: conflict-free), 9 taps of packed fp32 FMAs,
//                + shift, clamp, bf16, four 16-byte stores into the A tile.
// The pointwise weights (N x C bf16, <= 64 KB) stay resident in shared memory.
#include <cuda.h>

#include <algorithm>
#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "ptx.cuh"

namespace rtseg {

namespace {

constexpr int kDpTH = 8, kDpTW = 16;                 // output tile (128 pixels = one accumulator)
constexpr int kDpBH = kDpTH + 2, kDpBW = kDpTW + 2;  // input box
constexpr int kDpDwWarps = 8, kDpEpiWarps = 4;
constexpr int kDpThreads = 64 + 32 * (kDpEpiWarps + kDpDwWarps);
constexpr int kDpMaxStages = 6;
constexpr int kDpABufs = 2;
constexpr uint32_t kDpSmemBudget = 220 * 1024;
constexpr uint32_t kDpXStage = ((uint32_t)kDpBH * kDpBW * 128u + 1023u) / 1024u * 1024u;

struct DpParams {
  int n, h, w;
  int c, nkb;        // depthwise channels, 64-channel blocks
  int nout, nhalf;   // pointwise outputs (padded to 16), 64-channel halves of the staging tile
  int tiles_w, tiles_h, num_tiles;
  int x_coff, y_coff, r_ct, r_coff;
  int stages;
  int res_mode;
  uint32_t wpw_block_bytes, out_buf_bytes, tmem_cols;
  float lo1, hi1, lo, hi, lo2, hi2;
};

struct DpMaps {
  CUtensorMap x, w, y;
};

__device__ __forceinline__ void dp_epi_sync(int id) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "n"(32 * kDpEpiWarps) : "memory");
}

__device__ __forceinline__ uint64_t dp_desc(uint32_t addr) {   // K-major, 128-byte rows, SWIZZLE_128B, 8-row groups 1024 B apart
  return (uint64_t)((addr & 0x3ffffu) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}

__device__ __forceinline__ void unpack8(const float4 &raw, float (&v)[8]) {
  const uint32_t w[4] = {__float_as_uint(raw.x), __float_as_uint(raw.y), __float_as_uint(raw.z), __float_as_uint(raw.w)};
#pragma unroll
  for (int i = 0; i < 4; ++i) v[2 * i] = bf16_lo(w[i]), v[2 * i + 1] = bf16_hi(w[i]);
}

template <int RESMODE>
__global__ void __launch_bounds__(kDpThreads, 1) dwpw_tc_kernel(const __grid_constant__ DpMaps maps, const DpParams p,
                                                                const float *__restrict__ w_dw, const float *__restrict__ scale_dw,
                                                                const float *__restrict__ shift_dw,
                                                                const float *__restrict__ scale_pw,
                                                                const float *__restrict__ shift_pw,
                                                                const __nv_bfloat16 *__restrict__ res) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t s_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t s_w = s_base;                                              // nkb blocks [nout x 64] bf16
  const uint32_t s_x = s_w + (uint32_t)p.nkb * p.wpw_block_bytes;           // input ring
  const uint32_t s_a = s_x + (uint32_t)p.stages * kDpXStage;               // A tiles [128 x 64] bf16
  const uint32_t s_out = s_a + kDpABufs * 16384u;                           // staging: 2 x nhalf x [128 x 64] bf16
  const uint32_t s_dw = s_out + 2u * p.out_buf_bytes;                       // fp32: 9 x C scaled taps, C shifts
  const uint32_t s_aff = s_dw + 10u * (uint32_t)p.c * 4u;                   // fp32: nout scales, nout shifts
  const uint32_t s_bars = (s_aff + 2u * (uint32_t)p.nout * 4u + 15u) & ~15u;
  const uint32_t x_full = s_bars, x_empty = x_full + 8 * kDpMaxStages;
  const uint32_t a_full = x_empty + 8 * kDpMaxStages, a_empty = a_full + 8 * kDpABufs;
  const uint32_t t_full = a_empty + 8 * kDpABufs, t_empty = t_full + 16, w_bar = t_empty + 16;
  uint8_t *gen = smem_raw + (s_base - smem_u32(smem_raw));
  float *g_dw = reinterpret_cast<float *>(gen + (s_dw - s_base));
  float *g_aff = reinterpret_cast<float *>(gen + (s_aff - s_base));
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(gen + (w_bar + 8 - s_base));
  volatile uint32_t *dead = tmem_slot + 1;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) mbar_init(x_full + 8 * s, 1), mbar_init(x_empty + 8 * s, kDpDwWarps);
    for (int s = 0; s < kDpABufs; ++s) mbar_init(a_full + 8 * s, kDpDwWarps), mbar_init(a_empty + 8 * s, 1);
    for (int s = 0; s < 2; ++s) mbar_init(t_full + 8 * s, 1), mbar_init(t_empty + 8 * s, kDpEpiWarps);
    mbar_init(w_bar, 1);
    *dead = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  // depthwise taps with the BatchNorm scale folded in (fp32), shifts, pointwise scale / shift
  // layout: [tap 0..8, shift][64-channel block][half 0 / 1][8-channel vector j][4 floats]: the eight vectors j of one
  // (tap, block, half) are 128 contiguous bytes, so a quarter warp (j = 0..7) reads them without bank conflicts
  for (int i = threadIdx.x; i < 10 * p.c; i += blockDim.x) {
    const int tp = i / p.c, ch = i - tp * p.c;
    const int kb = ch >> 6, j = (ch >> 3) & 7, e = ch & 7;
    const float v = tp < 9 ? w_dw[i] * (scale_dw ? scale_dw[ch] : 1.f) : (shift_dw ? shift_dw[ch] : 0.f);
    g_dw[(tp * p.nkb + kb) * 64 + (e >> 2) * 32 + j * 4 + (e & 3)] = v;
  }
  for (int i = threadIdx.x; i < p.nout; i += blockDim.x) {
    g_aff[i] = scale_pw ? scale_pw[i] : 1.f;
    g_aff[p.nout + i] = shift_pw ? shift_pw[i] : 0.f;
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(p.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int n_my = (p.num_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  auto tile_coords = [&](int j, int &w0, int &h0, int &img) {
    const int t = (int)blockIdx.x + j * (int)gridDim.x;
    const int twi = t % p.tiles_w, r = t / p.tiles_w;
    w0 = twi * kDpTW, h0 = (r % p.tiles_h) * kDpTH, img = r / p.tiles_h;
  };

  if (warp == 0) {
    // ===================================================== TMA producer ====
    if (lane == 0) {
      prefetch_tmap(&maps.x);
      prefetch_tmap(&maps.y);
      mbar_expect_tx(w_bar, (uint32_t)p.nkb * (uint32_t)p.nout * 128u);
      for (int kb = 0; kb < p.nkb; ++kb) tma_load_2d(s_w + (uint32_t)kb * p.wpw_block_bytes, &maps.w, w_bar, kb * 64, 0);
      uint32_t stage = 0, phase = 0;
      for (int j = 0; j < n_my; ++j) {
        int w0, h0, img;
        tile_coords(j, w0, h0, img);
        for (int kb = 0; kb < p.nkb; ++kb) {
          mbar_wait_relaxed(x_empty + 8 * stage, phase ^ 1, dead);
          mbar_expect_tx(x_full + 8 * stage, (uint32_t)kDpBH * kDpBW * 128u);
          tma_load_4d(s_x + stage * kDpXStage, &maps.x, x_full + 8 * stage, p.x_coff + kb * 64, w0 - 1, h0 - 1, img);
          if (++stage == (uint32_t)p.stages) stage = 0, phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ======================================================= MMA issuer ====
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.nout >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    mbar_wait(w_bar, 0, dead);
    uint32_t ab = 0, aph = 0;
    for (int j = 0; j < n_my; ++j) {
      const uint32_t acc = (uint32_t)j & 1u, use = (uint32_t)j >> 1;
      mbar_wait_relaxed(t_empty + 8 * acc, (use & 1u) ^ 1u, dead);
      for (int kb = 0; kb < p.nkb; ++kb) {
        mbar_wait_relaxed(a_full + 8 * ab, aph, dead);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t a_d = dp_desc(s_a + ab * 16384u), b_d = dp_desc(s_w + (uint32_t)kb * p.wpw_block_bytes);
          const uint32_t d = tmem_base + acc * (uint32_t)p.nout;
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16(d, a_d + 2 * k, b_d + 2 * k, idesc, (kb | k) ? 1u : 0u);
          umma_commit(a_empty + 8 * ab);                       // the A tile may be overwritten once these MMAs have read it
          if (kb == p.nkb - 1) umma_commit(t_full + 8 * acc);  // ... and the tile's accumulator is complete
        }
        __syncwarp();
        if (++ab == kDpABufs) ab = 0, aph ^= 1;
      }
    }
  } else if (warp < 2 + kDpEpiWarps) {
    // ========================================================= epilogue ====
    const int q = warp & 3;                    // TMEM lane quarter of this warp (warps 2..5 -> 2, 3, 0, 1)
    const int m = q * 32 + lane;               // pixel of the tile: m = ty * TW + tx
    const int ty = m / kDpTW, tx = m % kDpTW;
    const bool elected = threadIdx.x == 64;
    const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
    const int groups = p.nout / 16;
    uint32_t ob = 0;
    for (int j = 0; j < n_my; ++j) {
      int w0, h0, img;
      tile_coords(j, w0, h0, img);
      const uint32_t acc = (uint32_t)j & 1u, use = (uint32_t)j >> 1;
      const uint32_t buf = s_out + ob * p.out_buf_bytes;
      // the residual row of this thread's pixel is requested BEFORE the wait for the accumulator: the loads fly while
      // the depthwise workers and the MMAs are still busy with this tile (the epilogue warps would only be waiting)
      const int gy = h0 + ty, gx = w0 + tx;
      uint4 rres[RESMODE != 0 ? 16 : 1];
      if (RESMODE != 0) {
        const bool inside = gy < p.h && gx < p.w;
        const __nv_bfloat16 *rp = res + (((int64_t)img * p.h + (inside ? gy : 0)) * p.w + (inside ? gx : 0)) * p.r_ct + p.r_coff;
#pragma unroll
        for (int u = 0; u < 16; ++u)
          rres[u] = (inside && u * 8 < p.nout) ? __ldg(reinterpret_cast<const uint4 *>(rp + u * 8)) : make_uint4(0u, 0u, 0u, 0u);
        // ... and the NEXT tile's row is pulled into L2 now: when the epilogue is the slowest stage the accumulator
        // is ready at once and the loads above would wait a full DRAM latency per tile (ncu: the residual unpack held
        // 8 % of the stall samples)
        if (j + 1 < n_my) {
          int w1, h1, img1;
          tile_coords(j + 1, w1, h1, img1);
          const int ny = h1 + ty, nx = w1 + tx;
          if (ny < p.h && nx < p.w) {
            const __nv_bfloat16 *np = res + (((int64_t)img1 * p.h + ny) * p.w + nx) * p.r_ct + p.r_coff;
            for (int b = 0; b < p.nout * 2; b += 128)
              asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char *>(np) + b));
          }
        }
      }
      mbar_wait_relaxed(t_full + 8 * acc, use & 1u, dead);
      tc_fence_after();
      dp_epi_sync(2);     // the elected thread has seen the store before last read this buffer out
#pragma unroll
      for (int g = 0; g < 8; ++g) {
        if (g >= groups) break;
        const int c0 = g * 16;
        uint32_t v[16], pk[8];
        tmem_ld16(lane_base + acc * (uint32_t)p.nout + (uint32_t)c0, v);
        float rv[16];
        if (RESMODE != 0) {
          const uint4 r0 = rres[RESMODE != 0 ? 2 * g : 0], r1 = rres[RESMODE != 0 ? 2 * g + 1 : 0];
          const uint32_t w8[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
          for (int u = 0; u < 8; ++u) rv[2 * u] = bf16_lo(w8[u]), rv[2 * u + 1] = bf16_hi(w8[u]);
        }
        tmem_ld_wait();
#pragma unroll
        for (int e4 = 0; e4 < 16; e4 += 4) {
          const float4 sc = lds128(s_aff + (uint32_t)(c0 + e4) * 4), sh = lds128(s_aff + (uint32_t)(p.nout + c0 + e4) * 4);
          float o[4] = {__uint_as_float(v[e4]), __uint_as_float(v[e4 + 1]), __uint_as_float(v[e4 + 2]), __uint_as_float(v[e4 + 3])};
          if (RESMODE == RTSEG_RES_PRE) {
#pragma unroll
            for (int u = 0; u < 4; ++u) o[u] += rv[e4 + u];
          }
          o[0] = fmaf(o[0], sc.x, sh.x), o[1] = fmaf(o[1], sc.y, sh.y), o[2] = fmaf(o[2], sc.z, sh.z), o[3] = fmaf(o[3], sc.w, sh.w);
#pragma unroll
          for (int u = 0; u < 4; ++u) o[u] = fmin_nan(fmax_nan(o[u], p.lo), p.hi);
          if (RESMODE == RTSEG_RES_POST) {
#pragma unroll
            for (int u = 0; u < 4; ++u) o[u] = fmin_nan(fmax_nan(o[u] + rv[e4 + u], p.lo2), p.hi2);
          }
          pk[(e4 >> 1) + 0] = pack_bf16(o[0], o[1]);
          pk[(e4 >> 1) + 1] = pack_bf16(o[2], o[3]);
        }
        // staging: per 64-channel half a [128 x 128 B] tile, 16-byte chunk index xor (row & 7)
        const uint32_t half = (uint32_t)c0 >> 6, cin_half = (uint32_t)c0 & 63u;
        const uint32_t off0 = (uint32_t)m * 128u + cin_half * 2u;
        const uint32_t a = buf + half * 16384u + (off0 ^ (((off0 >> 7) & 7u) << 4));
        sts128(a, pk[0], pk[1], pk[2], pk[3]);
        sts128(a ^ 16u, pk[4], pk[5], pk[6], pk[7]);
      }
      fence_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(t_empty + 8 * acc);
      dp_epi_sync(1);        // every warp has written its part of the staging tile
      if (elected) {
        for (int hf = 0; hf < p.nhalf; ++hf) tma_store_4d(&maps.y, buf + (uint32_t)hf * 16384u, p.y_coff + hf * 64, w0, h0, img);
        bulk_commit();
        bulk_wait_read<1>();   // the store BEFORE this one has read its buffer out: that buffer may be written again
      }
      ob ^= 1u;
    }
    if (elected) bulk_wait_all();
  } else {
    // ================================================= depthwise workers ====
    const int t = threadIdx.x - 32 * (2 + kDpEpiWarps);   // 0 .. 255
    const int j8 = t & 7;                                  // which 8-channel vector of the 64-channel block
    const int strip = t >> 3;                              // 0 .. 31: four neighbouring pixels of one tile row
    const int ty = strip >> 2, tx0 = (strip & 3) * 4;
    uint32_t xs = 0, xph = 0, ab = 0, aph = 0;
    const bool relu1 = p.lo1 == 0.f && p.hi1 > 3.0e38f;   // warp-uniform
    for (int j = 0; j < n_my; ++j) {
      for (int kb = 0; kb < p.nkb; ++kb) {
        // This thread's taps (BatchNorm scale folded in) and shift for channels [kb*64 + j8*8, +8) are read from shared
        // memory one filter row at a time, next to that row's six input vectors: 12 loads issued back to back, then the
        // row's 48 packed FMAs.  (All nine taps in registers — 72 of them — left no room to have loads in flight, and
        // every input vector was consumed the moment it was requested.)
        const uint32_t wbase = s_dw + (uint32_t)(kb * 64 + j8 * 4) * 4;
        float2 acc[4][4];
        {
          const float4 lo4 = lds128(wbase + (uint32_t)(9 * p.nkb * 64) * 4), hi4 = lds128(wbase + (uint32_t)(9 * p.nkb * 64 + 32) * 4);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            acc[i][0] = make_float2(lo4.x, lo4.y), acc[i][1] = make_float2(lo4.z, lo4.w);
            acc[i][2] = make_float2(hi4.x, hi4.y), acc[i][3] = make_float2(hi4.z, hi4.w);
          }
        }
        mbar_wait(x_full + 8 * xs, xph, dead);
        const uint32_t xb = s_x + xs * kDpXStage;
#pragma unroll
        for (int dy = 0; dy < 3; ++dy) {
          float4 raw[6];
#pragma unroll
          for (int cx = 0; cx < 6; ++cx) {
            const uint32_t pix = (uint32_t)((ty + dy) * kDpBW + tx0 + cx);            // pixel row of the box
            raw[cx] = lds128(xb + pix * 128u + (((uint32_t)j8 ^ (pix & 7u)) << 4));
          }
          float2 wt[3][4];
#pragma unroll
          for (int dx = 0; dx < 3; ++dx) {
            const int tp = dy * 3 + dx;
            const float4 lo4 = lds128(wbase + (uint32_t)(tp * p.nkb * 64) * 4), hi4 = lds128(wbase + (uint32_t)(tp * p.nkb * 64 + 32) * 4);
            wt[dx][0] = make_float2(lo4.x, lo4.y), wt[dx][1] = make_float2(lo4.z, lo4.w);
            wt[dx][2] = make_float2(hi4.x, hi4.y), wt[dx][3] = make_float2(hi4.z, hi4.w);
          }
#pragma unroll
          for (int cx = 0; cx < 6; ++cx) {
            const uint32_t rw[4] = {__float_as_uint(raw[cx].x), __float_as_uint(raw[cx].y), __float_as_uint(raw[cx].z),
                                    __float_as_uint(raw[cx].w)};
            float2 xv[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) xv[e] = make_float2(bf16_lo(rw[e]), bf16_hi(rw[e]));
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int dx = cx - i;                                                    // tap column for output pixel i
              if (dx >= 0 && dx < 3) {
#pragma unroll
                for (int e = 0; e < 4; ++e) acc[i][e] = __ffma2_rn(xv[e], wt[dx][e], acc[i][e]);   // packed fp32 FMA
              }
            }
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(x_empty + 8 * xs);            // this warp has read the input box out
        if (++xs == (uint32_t)p.stages) xs = 0, xph ^= 1;
        mbar_wait(a_empty + 8 * ab, aph ^ 1, dead);              // the MMAs that read this A tile last time are done
        const uint32_t abuf = s_a + ab * 16384u;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint32_t pk[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            if (relu1) {
              pk[e] = pack_relu_bf16(acc[i][e].x, acc[i][e].y);       // the clamp [0, inf) rides in the conversion
            } else {
              pk[e] = pack_bf16(fmin_nan(fmax_nan(acc[i][e].x, p.lo1), p.hi1), fmin_nan(fmax_nan(acc[i][e].y, p.lo1), p.hi1));
            }
          }
          const uint32_t mrow = (uint32_t)(ty * kDpTW + tx0 + i);
          sts128(abuf + mrow * 128u + (((uint32_t)j8 ^ (mrow & 7u)) << 4), pk[0], pk[1], pk[2], pk[3]);
        }
        fence_async_smem();      // the A tile is read by the tensor core (async proxy)
        __syncwarp();
        if (lane == 0) mbar_arrive(a_full + 8 * ab);
        if (++ab == kDpABufs) ab = 0, aph ^= 1;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
  }
}

// ---- stride-2 variant for narrow layers (Fast-SCNN's learning-to-downsample DSConvBNAct 32 -> 48 and 48 -> 64,
// reference models/fastscnn.py:38-40): channels in 16-wide blocks (32-byte pixel rows, SWIZZLE_32B), one box of
// (2 TH + 1) x (2 TW + 1) input pixels per tile and block, a depthwise thread owns ONE output pixel x 8 channels (nine
// 16-byte window reads), one K = 16 MMA per block; staging rows are N x 2 bytes, unswizzled.
constexpr int kD2BH = 2 * kDpTH + 1, kD2BW = 2 * kDpTW + 1;
// The input box arrives as TWO boxes, its even and its odd columns (TMA traversal stride 2 along w): a tap then reads
// NEIGHBOURING 32-byte pixel slots for neighbouring output pixels — a quarter warp covers 128 contiguous bytes.  With
// all 33 columns in one box the taps read every other slot and every shared load ran into a 2-way bank conflict
// (ncu: 27 M of 80 M wavefronts, LSU data pipe 88 % busy).
constexpr int kD2HW = kDpTW + 1;                                             // pixels per row of a half box
constexpr uint32_t kD2Half = ((uint32_t)kD2BH * kD2HW * 32u + 127u) / 128u * 128u;
constexpr uint32_t kD2XStage = (2 * kD2Half + 1023u) / 1024u * 1024u;   // the A tiles behind the ring stay 1 KB aligned
constexpr int kD2MaxStages = 8;

__device__ __forceinline__ uint64_t d2_desc(uint32_t addr) {   // K-major, 32-byte rows, SWIZZLE_32B, 8-row groups 256 B apart
  return (uint64_t)((addr & 0x3ffffu) >> 4) | (1ull << 16) | ((uint64_t)(256 >> 4) << 32) | (1ull << 46) | (6ull << 61);
}

__global__ void __launch_bounds__(kDpThreads, 1) dwpw_s2_tc_kernel(const __grid_constant__ DpMaps maps, const DpParams p,
                                                                   const float *__restrict__ w_dw, const float *__restrict__ scale_dw,
                                                                   const float *__restrict__ shift_dw,
                                                                   const float *__restrict__ scale_pw,
                                                                   const float *__restrict__ shift_pw) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t s_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t s_w = s_base;                                              // nkb blocks [nout x 16] bf16 (32-byte rows)
  const uint32_t s_x = s_w + (uint32_t)p.nkb * p.wpw_block_bytes;           // input ring
  const uint32_t s_a = s_x + (uint32_t)p.stages * kD2XStage;               // A tiles [128 x 16] bf16
  const uint32_t s_out = s_a + kDpABufs * 4096u;                            // staging: 2 x [128 x nout] bf16, unswizzled
  const uint32_t s_dw = s_out + 2u * p.out_buf_bytes;                       // fp32: [tap 0..8, shift][block][16]
  const uint32_t s_aff = s_dw + 10u * (uint32_t)p.c * 4u;
  const uint32_t s_bars = (s_aff + 2u * (uint32_t)p.nout * 4u + 15u) & ~15u;
  const uint32_t x_full = s_bars, x_empty = x_full + 8 * kD2MaxStages;
  const uint32_t a_full = x_empty + 8 * kD2MaxStages, a_empty = a_full + 8 * kDpABufs;
  const uint32_t t_full = a_empty + 8 * kDpABufs, t_empty = t_full + 16, w_bar = t_empty + 16;
  uint8_t *gen = smem_raw + (s_base - smem_u32(smem_raw));
  float *g_dw = reinterpret_cast<float *>(gen + (s_dw - s_base));
  float *g_aff = reinterpret_cast<float *>(gen + (s_aff - s_base));
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(gen + (w_bar + 8 - s_base));
  volatile uint32_t *dead = tmem_slot + 1;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) mbar_init(x_full + 8 * s, 1), mbar_init(x_empty + 8 * s, kDpDwWarps);
    for (int s = 0; s < kDpABufs; ++s) mbar_init(a_full + 8 * s, kDpDwWarps), mbar_init(a_empty + 8 * s, 1);
    for (int s = 0; s < 2; ++s) mbar_init(t_full + 8 * s, 1), mbar_init(t_empty + 8 * s, kDpEpiWarps);
    mbar_init(w_bar, 1);
    *dead = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  // depthwise taps (BatchNorm scale folded in) and shifts: [tap][channel], channel-contiguous (16 per block)
  for (int i = threadIdx.x; i < 10 * p.c; i += blockDim.x) {
    const int tp = i / p.c, ch = i - tp * p.c;
    g_dw[i] = tp < 9 ? w_dw[i] * (scale_dw ? scale_dw[ch] : 1.f) : (shift_dw ? shift_dw[ch] : 0.f);
  }
  for (int i = threadIdx.x; i < p.nout; i += blockDim.x) {
    g_aff[i] = scale_pw ? scale_pw[i] : 1.f;
    g_aff[p.nout + i] = shift_pw ? shift_pw[i] : 0.f;
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(p.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int n_my = (p.num_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  auto tile_coords = [&](int j, int &w0, int &h0, int &img) {   // OUTPUT coordinates of the j-th tile of this CTA
    const int t = (int)blockIdx.x + j * (int)gridDim.x;
    const int twi = t % p.tiles_w, r = t / p.tiles_w;
    w0 = twi * kDpTW, h0 = (r % p.tiles_h) * kDpTH, img = r / p.tiles_h;
  };

  if (warp == 0) {
    if (lane == 0) {
      prefetch_tmap(&maps.x);
      prefetch_tmap(&maps.y);
      mbar_expect_tx(w_bar, (uint32_t)p.nkb * (uint32_t)p.nout * 32u);
      for (int kb = 0; kb < p.nkb; ++kb) tma_load_2d(s_w + (uint32_t)kb * p.wpw_block_bytes, &maps.w, w_bar, kb * 16, 0);
      uint32_t stage = 0, phase = 0;
      for (int j = 0; j < n_my; ++j) {
        int w0, h0, img;
        tile_coords(j, w0, h0, img);
        for (int kb = 0; kb < p.nkb; ++kb) {
          mbar_wait_relaxed(x_empty + 8 * stage, phase ^ 1, dead);
          mbar_expect_tx(x_full + 8 * stage, 2u * (uint32_t)kD2BH * kD2HW * 32u);
          tma_load_4d(s_x + stage * kD2XStage, &maps.x, x_full + 8 * stage, p.x_coff + kb * 16, 2 * w0 - 1, 2 * h0 - 1, img);
          tma_load_4d(s_x + stage * kD2XStage + kD2Half, &maps.x, x_full + 8 * stage, p.x_coff + kb * 16, 2 * w0, 2 * h0 - 1, img);
          if (++stage == (uint32_t)p.stages) stage = 0, phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.nout >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    mbar_wait(w_bar, 0, dead);
    uint32_t ab = 0, aph = 0;
    for (int j = 0; j < n_my; ++j) {
      const uint32_t acc = (uint32_t)j & 1u, use = (uint32_t)j >> 1;
      mbar_wait_relaxed(t_empty + 8 * acc, (use & 1u) ^ 1u, dead);
      for (int kb = 0; kb < p.nkb; ++kb) {
        mbar_wait_relaxed(a_full + 8 * ab, aph, dead);
        tc_fence_after();
        if (elect_one()) {
          umma_bf16(tmem_base + acc * (uint32_t)p.nout, d2_desc(s_a + ab * 4096u), d2_desc(s_w + (uint32_t)kb * p.wpw_block_bytes),
                    idesc, kb ? 1u : 0u);
          umma_commit(a_empty + 8 * ab);
          if (kb == p.nkb - 1) umma_commit(t_full + 8 * acc);
        }
        __syncwarp();
        if (++ab == kDpABufs) ab = 0, aph ^= 1;
      }
    }
  } else if (warp < 2 + kDpEpiWarps) {
    const int q = warp & 3;
    const int m = q * 32 + lane;
    const bool elected = threadIdx.x == 64;
    const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
    const int groups = p.nout / 16;
    const uint32_t row_b = (uint32_t)p.nout * 2u;
    uint32_t ob = 0;
    for (int j = 0; j < n_my; ++j) {
      int w0, h0, img;
      tile_coords(j, w0, h0, img);
      const uint32_t acc = (uint32_t)j & 1u, use = (uint32_t)j >> 1;
      const uint32_t buf = s_out + ob * p.out_buf_bytes;
      mbar_wait_relaxed(t_full + 8 * acc, use & 1u, dead);
      tc_fence_after();
      dp_epi_sync(2);
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        if (g >= groups) break;
        const int c0 = g * 16;
        uint32_t v[16], pk[8];
        tmem_ld16(lane_base + acc * (uint32_t)p.nout + (uint32_t)c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int e4 = 0; e4 < 16; e4 += 4) {
          const float4 sc = lds128(s_aff + (uint32_t)(c0 + e4) * 4), sh = lds128(s_aff + (uint32_t)(p.nout + c0 + e4) * 4);
          float o[4] = {__uint_as_float(v[e4]), __uint_as_float(v[e4 + 1]), __uint_as_float(v[e4 + 2]), __uint_as_float(v[e4 + 3])};
          o[0] = fmaf(o[0], sc.x, sh.x), o[1] = fmaf(o[1], sc.y, sh.y), o[2] = fmaf(o[2], sc.z, sh.z), o[3] = fmaf(o[3], sc.w, sh.w);
#pragma unroll
          for (int u = 0; u < 4; ++u) o[u] = fmin_nan(fmax_nan(o[u], p.lo), p.hi);
          pk[(e4 >> 1) + 0] = pack_bf16(o[0], o[1]);
          pk[(e4 >> 1) + 1] = pack_bf16(o[2], o[3]);
        }
        const uint32_t a = buf + (uint32_t)m * row_b + (uint32_t)c0 * 2u;
        sts128(a, pk[0], pk[1], pk[2], pk[3]);
        sts128(a + 16u, pk[4], pk[5], pk[6], pk[7]);
      }
      fence_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(t_empty + 8 * acc);
      dp_epi_sync(1);
      if (elected) {
        tma_store_4d(&maps.y, buf, p.y_coff, w0, h0, img);
        bulk_commit();
        bulk_wait_read<1>();
      }
      ob ^= 1u;
    }
    if (elected) bulk_wait_all();
  } else {
    // depthwise workers: thread = one output pixel x 8 channels of the 16-channel block
    const int t = threadIdx.x - 32 * (2 + kDpEpiWarps);   // 0 .. 255
    const int j2 = t & 1;                                  // which half (8 channels) of the block
    const int m = t >> 1;                                  // output pixel of the tile
    const int ty = m / kDpTW, tx = m % kDpTW;
    uint32_t xs = 0, xph = 0, ab = 0, aph = 0;
    const bool relu1 = p.lo1 == 0.f && p.hi1 > 3.0e38f;
    for (int j = 0; j < n_my; ++j) {
      for (int kb = 0; kb < p.nkb; ++kb) {
        const uint32_t wbase = s_dw + (uint32_t)(kb * 16 + j2 * 8) * 4;
        float2 acc[4];
        {
          const float4 lo4 = lds128(wbase + (uint32_t)(9 * p.c) * 4), hi4 = lds128(wbase + (uint32_t)(9 * p.c + 4) * 4);
          acc[0] = make_float2(lo4.x, lo4.y), acc[1] = make_float2(lo4.z, lo4.w);
          acc[2] = make_float2(hi4.x, hi4.y), acc[3] = make_float2(hi4.z, hi4.w);
        }
        mbar_wait(x_full + 8 * xs, xph, dead);
        const uint32_t xb = s_x + xs * kD2XStage;
#pragma unroll
        for (int dy = 0; dy < 3; ++dy) {
#pragma unroll
          for (int dx = 0; dx < 3; ++dx) {
            // dx = 0 / 2: even columns tx / tx + 1; dx = 1: odd column tx
            const uint32_t pix = (uint32_t)((2 * ty + dy) * kD2HW + tx + (dx >> 1));
            const float4 raw = lds128(xb + (dx == 1 ? kD2Half : 0u) + pix * 32u + (uint32_t)j2 * 16u);
            const uint32_t rw[4] = {__float_as_uint(raw.x), __float_as_uint(raw.y), __float_as_uint(raw.z), __float_as_uint(raw.w)};
            const float4 wl = lds128(wbase + (uint32_t)((dy * 3 + dx) * p.c) * 4), wh = lds128(wbase + (uint32_t)((dy * 3 + dx) * p.c + 4) * 4);
            acc[0] = __ffma2_rn(make_float2(bf16_lo(rw[0]), bf16_hi(rw[0])), make_float2(wl.x, wl.y), acc[0]);
            acc[1] = __ffma2_rn(make_float2(bf16_lo(rw[1]), bf16_hi(rw[1])), make_float2(wl.z, wl.w), acc[1]);
            acc[2] = __ffma2_rn(make_float2(bf16_lo(rw[2]), bf16_hi(rw[2])), make_float2(wh.x, wh.y), acc[2]);
            acc[3] = __ffma2_rn(make_float2(bf16_lo(rw[3]), bf16_hi(rw[3])), make_float2(wh.z, wh.w), acc[3]);
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(x_empty + 8 * xs);
        if (++xs == (uint32_t)p.stages) xs = 0, xph ^= 1;
        mbar_wait(a_empty + 8 * ab, aph ^ 1, dead);
        uint32_t pk[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          if (relu1) pk[e] = pack_relu_bf16(acc[e].x, acc[e].y);
          else pk[e] = pack_bf16(fmin_nan(fmax_nan(acc[e].x, p.lo1), p.hi1), fmin_nan(fmax_nan(acc[e].y, p.lo1), p.hi1));
        }
        sts128(s_a + ab * 4096u + (uint32_t)m * 32u + (((uint32_t)j2 ^ (((uint32_t)m >> 2) & 1u)) << 4), pk[0], pk[1], pk[2], pk[3]);
        fence_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(a_full + 8 * ab);
        if (++ab == kDpABufs) ab = 0, aph ^= 1;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
  }
}

typedef CUresult (*EncodeTiledFnDp)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFnDp dp_encode_fn() {
  static EncodeTiledFnDp fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void *ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFnDp>(ptr);
  });
  return fn;
}

// 4-D bf16 map over channels [0, cdim) of an NHWC buffer with `ct` channels per pixel; box = 64 x bw x bh x 1, SWIZZLE_128B
// es_w: traversal stride along w (2 = every other column of the bw-wide box: (bw + 1) / 2 pixels per row arrive)
int dp_nhwc_map(CUtensorMap *m, const void *base, int cdim, int ct, int w, int h, int n, int bw, int bh, int bc = 64,
                CUtensorMapSwizzle sw = CU_TENSOR_MAP_SWIZZLE_128B, int es_w = 1) {
  EncodeTiledFnDp enc = dp_encode_fn();
  if (!enc) RTSEG_FAIL(RTSEG_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  const cuuint64_t dims[4] = {(cuuint64_t)cdim, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)n};
  const cuuint64_t strides[3] = {(cuuint64_t)ct * 2, (cuuint64_t)ct * w * 2, (cuuint64_t)ct * w * h * 2};
  const cuuint32_t box[4] = {(cuuint32_t)bc, (cuuint32_t)bw, (cuuint32_t)bh, 1};
  const cuuint32_t es[4] = {1, (cuuint32_t)es_w, 1, 1};
  const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void *>(base), dims, strides, box, es,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) RTSEG_FAIL(RTSEG_ERR_CUDA, "cuTensorMapEncodeTiled(dwpw, NHWC view) failed with CUresult %d", (int)r);
  return RTSEG_OK;
}

// [nout][c] bf16 K-major; box = bc channels x nout rows
int dp_weight_map(CUtensorMap *m, const void *w, int c, int nout, int bc = 64, CUtensorMapSwizzle sw = CU_TENSOR_MAP_SWIZZLE_128B) {
  EncodeTiledFnDp enc = dp_encode_fn();
  if (!enc) RTSEG_FAIL(RTSEG_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  const cuuint64_t dims[2] = {(cuuint64_t)c, (cuuint64_t)nout};
  const cuuint64_t strides[1] = {(cuuint64_t)c * 2};
  const cuuint32_t box[2] = {(cuuint32_t)bc, (cuuint32_t)nout};
  const cuuint32_t es[2] = {1, 1};
  const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(w), dims, strides, box, es,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) RTSEG_FAIL(RTSEG_ERR_CUDA, "cuTensorMapEncodeTiled(dwpw, weights) failed with CUresult %d", (int)r);
  return RTSEG_OK;
}

// the stride-2 narrow-layer variant (dwpw_s2_tc_kernel)
const char *dp_why_s2(const rtseg_conv_desc *dw, const rtseg_conv_desc *pw) {
  if (dw->dtype != RTSEG_BF16 || pw->dtype != RTSEG_BF16) return "dtype must be bf16";
  if (dw->groups != dw->cin || dw->cout != dw->cin) return "first op: depthwise with channel multiplier 1";
  if (dw->kh != 3 || dw->kw != 3 || dw->stride != 2 || dw->dil != 1 || dw->pad_h != 1 || dw->pad_w != 1) return "first op: 3x3, stride 2, padding 1";
  if (dw->cin % 16 || dw->cin > 64) return "depthwise channels must be a multiple of 16 (<= 64)";
  if ((dw->h & 1) || (dw->w & 1)) return "even input size";
  if (pw->groups != 1 || pw->kh != 1 || pw->kw != 1 || pw->stride != 1 || pw->pad_h != 0 || pw->pad_w != 0) return "second op: dense 1x1, stride 1";
  if (pw->cin != dw->cout || pw->n != dw->n || pw->h != dw->h / 2 || pw->w != dw->w / 2) return "the two ops do not chain";
  if (pw->cout % 16 || pw->cout > 64) return "pointwise outputs must be a multiple of 16 (<= 64)";
  if (dw->x_ctotal % 8 || dw->x_coff % 8 || pw->y_ctotal % 8 || pw->y_coff % 8) return "channel slices must be 16-byte aligned";
  if (dw->res_mode != RTSEG_RES_NONE || pw->res_mode != RTSEG_RES_NONE) return "no residual in the stride-2 variant";
  float lo, hi, lo2, hi2;
  if (epilogue_mode(dw->act, 0, RTSEG_RES_NONE, nullptr, dw->act_p0, dw->act_p1, lo, hi, lo2, hi2) != 0) return "first activation must be none / relu / relu6 / hardtanh";
  if (epilogue_mode(pw->act, 0, RTSEG_RES_NONE, nullptr, pw->act_p0, pw->act_p1, lo, hi, lo2, hi2) != 0) return "second activation must be none / relu / relu6 / hardtanh";
  if (dw->w < 16 || dw->h < 8) return "image too small";
  return nullptr;
}

const char *dp_why(const rtseg_conv_desc *dw, const rtseg_conv_desc *pw) {
  if (dw->dtype != RTSEG_BF16 || pw->dtype != RTSEG_BF16) return "dtype must be bf16";
  if (dw->groups != dw->cin || dw->cout != dw->cin) return "first op: depthwise with channel multiplier 1";
  if (dw->kh != 3 || dw->kw != 3 || dw->stride != 1 || dw->dil != 1 || dw->pad_h != 1 || dw->pad_w != 1)
    return "first op: 3x3, stride 1, dilation 1, padding 1";
  if (dw->cin % 64 || dw->cin > 512) return "depthwise channels must be a multiple of 64 (<= 512)";
  if (pw->groups != 1 || pw->kh != 1 || pw->kw != 1 || pw->stride != 1 || pw->pad_h != 0 || pw->pad_w != 0) return "second op: dense 1x1, stride 1";
  if (pw->cin != dw->cout || pw->n != dw->n || pw->h != dw->h || pw->w != dw->w) return "the two ops do not chain";
  if (pw->cout % 64 || pw->cout > 128) return "pointwise outputs must be 64 or 128";
  if ((size_t)pw->cout * dw->cin * 2 > 64 * 1024) return "pointwise weights exceed the resident 64 KB";
  if (dw->x_ctotal % 8 || dw->x_coff % 8 || pw->y_ctotal % 8 || pw->y_coff % 8) return "channel slices must be 16-byte aligned";
  if (dw->res_mode != RTSEG_RES_NONE) return "the depthwise op takes no residual";
  if (pw->res_mode != RTSEG_RES_NONE && (pw->r_ctotal % 8 || pw->r_coff % 8)) return "residual slice must be 16-byte aligned";
  float lo, hi, lo2, hi2;
  if (epilogue_mode(dw->act, 0, RTSEG_RES_NONE, nullptr, dw->act_p0, dw->act_p1, lo, hi, lo2, hi2) != 0) return "first activation must be none / relu / relu6 / hardtanh";
  if (epilogue_mode(pw->act, pw->act2, pw->res_mode, nullptr, pw->act_p0, pw->act_p1, lo, hi, lo2, hi2) != 0) return "second activation must be none / relu / relu6 / hardtanh";
  if (dw->w < 8 || dw->h < 4) return "image too small";
  return nullptr;
}

}  // namespace

}  // namespace rtseg

using namespace rtseg;

extern "C" int rtseg_dwpw_tc_supported(const rtseg_conv_desc *dw, const rtseg_conv_desc *pw) {
  static const int knob = getenv("RTSEG_DWPW_TC") ? atoi(getenv("RTSEG_DWPW_TC")) : 1;
  static const int knob2 = getenv("RTSEG_DWPW_S2") ? atoi(getenv("RTSEG_DWPW_S2")) : 1;
  return dw && pw && knob && (dp_why(dw, pw) == nullptr || (knob2 && dp_why_s2(dw, pw) == nullptr)) ? 1 : 0;
}

static int launch_dwpw_s2(const rtseg_conv_desc *dw, const void *x, const float *w_dw, const float *scale_dw, const float *shift_dw,
                          const rtseg_conv_desc *pw, const void *w_pw_bf16, const float *scale_pw, const float *shift_pw, void *y,
                          cudaStream_t st) {
  DpParams p{};
  p.n = dw->n, p.h = pw->h, p.w = pw->w, p.c = dw->cin, p.nkb = dw->cin / 16;     // h, w: the OUTPUT grid
  p.nout = pw->cout, p.nhalf = 1;
  p.tiles_w = (int)cdiv(p.w, kDpTW), p.tiles_h = (int)cdiv(p.h, kDpTH);
  const int64_t tiles = (int64_t)p.n * p.tiles_w * p.tiles_h;
  RTSEG_REQUIRE(tiles < (1ll << 30), "too many tiles");
  p.num_tiles = (int)tiles;
  p.x_coff = dw->x_coff, p.y_coff = pw->y_coff;
  p.wpw_block_bytes = ((uint32_t)p.nout * 32u + 1023u) / 1024u * 1024u;
  p.out_buf_bytes = ((uint32_t)p.nout * 2u * 128u + 1023u) / 1024u * 1024u;
  const uint32_t misc = 10u * p.c * 4 + 2u * p.nout * 4 + (2 * kD2MaxStages + 2 * kDpABufs + 5) * 8 + 64 + 16;
  const uint32_t fixed = (uint32_t)p.nkb * p.wpw_block_bytes + kDpABufs * 4096u + 2 * p.out_buf_bytes + misc;
  p.stages = (int)std::min<uint32_t>(kD2MaxStages, (kDpSmemBudget - fixed) / kD2XStage);
  RTSEG_REQUIRE(p.stages >= 2, "dwpw stride-2 plan: tiles do not fit in shared memory");
  const uint32_t smem = fixed + (uint32_t)p.stages * kD2XStage + 1024;
  uint32_t cols = 32;
  while (cols < (uint32_t)(2 * p.nout)) cols <<= 1;
  p.tmem_cols = cols;
  float lo2, hi2;
  epilogue_mode(dw->act, 0, RTSEG_RES_NONE, nullptr, dw->act_p0, dw->act_p1, p.lo1, p.hi1, lo2, hi2);
  epilogue_mode(pw->act, 0, RTSEG_RES_NONE, nullptr, pw->act_p0, pw->act_p1, p.lo, p.hi, p.lo2, p.hi2);
  DpMaps maps;
  if (int rc = dp_nhwc_map(&maps.x, x, dw->x_coff + p.c, dw->x_ctotal, dw->w, dw->h, p.n, kD2BW, kD2BH, 16, CU_TENSOR_MAP_SWIZZLE_NONE, 2)) return rc;
  if (int rc = dp_nhwc_map(&maps.y, y, pw->y_coff + p.nout, pw->y_ctotal, p.w, p.h, p.n, kDpTW, kDpTH, p.nout, CU_TENSOR_MAP_SWIZZLE_NONE)) return rc;
  if (int rc = dp_weight_map(&maps.w, w_pw_bf16, p.c, p.nout, 16, CU_TENSOR_MAP_SWIZZLE_32B)) return rc;
  const unsigned grid = (unsigned)std::min<int64_t>(tiles, num_sms());
  static bool attr_set[64] = {false};
  int dev = 0;
  RTSEG_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    RTSEG_CUDA(cudaFuncSetAttribute(dwpw_s2_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kDpSmemBudget + 2048));
    attr_set[dev] = true;
  }
  dwpw_s2_tc_kernel<<<grid, kDpThreads, smem, st>>>(maps, p, w_dw, scale_dw, shift_dw, scale_pw, shift_pw);
  RTSEG_LAUNCH_CHECK("dwpw_s2_tc_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_dwpw_tc(const rtseg_conv_desc *dw, const void *x, const float *w_dw, const float *scale_dw,
                             const float *shift_dw, const rtseg_conv_desc *pw, const void *w_pw_bf16, const float *scale_pw,
                             const float *shift_pw, const void *res, void *y, rtseg_stream_t stream) {
  RTSEG_REQUIRE(dw && pw && x && w_dw && w_pw_bf16 && y, "NULL argument");
  RTSEG_REQUIRE(dw->n > 0 && dw->h > 0 && dw->w > 0, "non-positive extent");
  if (dw->stride == 2) {
    if (const char *why = dp_why_s2(dw, pw)) RTSEG_FAIL(RTSEG_ERR_UNSUPPORTED, "rtseg_dwpw_tc: %s", why);
    RTSEG_REQUIRE(reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(y) % 16 == 0 &&
                      reinterpret_cast<uintptr_t>(w_pw_bf16) % 16 == 0, "tensor pointers must be 16-byte aligned");
    RTSEG_REQUIRE(dw->x_coff + dw->cin <= dw->x_ctotal && pw->y_coff + pw->cout <= pw->y_ctotal, "channel slice outside its buffer");
    return launch_dwpw_s2(dw, x, w_dw, scale_dw, shift_dw, pw, w_pw_bf16, scale_pw, shift_pw, y, as_stream(stream));
  }
  if (const char *why = dp_why(dw, pw)) RTSEG_FAIL(RTSEG_ERR_UNSUPPORTED, "rtseg_dwpw_tc: %s", why);
  const bool has_res = pw->res_mode != RTSEG_RES_NONE;
  RTSEG_REQUIRE(!has_res || res, "residual mode set but res is NULL");
  RTSEG_REQUIRE(reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(y) % 16 == 0 &&
                    reinterpret_cast<uintptr_t>(res) % 16 == 0 && reinterpret_cast<uintptr_t>(w_pw_bf16) % 16 == 0 &&
                    reinterpret_cast<uintptr_t>(w_dw) % 4 == 0, "tensor pointers must be 16-byte aligned");
  RTSEG_REQUIRE(dw->x_coff + dw->cin <= dw->x_ctotal && pw->y_coff + pw->cout <= pw->y_ctotal, "channel slice outside its buffer");
  RTSEG_REQUIRE(!has_res || pw->r_coff + pw->cout <= pw->r_ctotal, "residual slice outside its buffer");
  DpParams p{};
  p.n = dw->n, p.h = dw->h, p.w = dw->w, p.c = dw->cin, p.nkb = dw->cin / 64;
  p.nout = pw->cout, p.nhalf = pw->cout / 64;
  p.tiles_w = (int)cdiv(p.w, kDpTW), p.tiles_h = (int)cdiv(p.h, kDpTH);
  const int64_t tiles = (int64_t)p.n * p.tiles_w * p.tiles_h;
  RTSEG_REQUIRE(tiles < (1ll << 30), "too many tiles");
  p.num_tiles = (int)tiles;
  p.x_coff = dw->x_coff, p.y_coff = pw->y_coff, p.r_ct = pw->r_ctotal, p.r_coff = pw->r_coff;
  p.res_mode = pw->res_mode;
  p.wpw_block_bytes = (uint32_t)p.nout * 128u;     // a multiple of 1024 (nout is a multiple of 64)
  p.out_buf_bytes = (uint32_t)p.nhalf * 16384u;
  const uint32_t misc = 10u * p.c * 4 + 2u * p.nout * 4 + (2 * kDpMaxStages + 2 * kDpABufs + 5) * 8 + 64 + 16;
  const uint32_t fixed = (uint32_t)p.nkb * p.wpw_block_bytes + kDpABufs * 16384u + 2 * p.out_buf_bytes + misc;
  RTSEG_REQUIRE(fixed + 2 * kDpXStage <= kDpSmemBudget, "dwpw plan: tiles do not fit in shared memory");
  p.stages = (int)std::min<uint32_t>(kDpMaxStages, (kDpSmemBudget - fixed) / kDpXStage);
  const uint32_t smem = fixed + (uint32_t)p.stages * kDpXStage + 1024;
  uint32_t cols = 32;
  while (cols < (uint32_t)(2 * p.nout)) cols <<= 1;
  p.tmem_cols = cols;
  float lo2, hi2;
  epilogue_mode(dw->act, 0, RTSEG_RES_NONE, nullptr, dw->act_p0, dw->act_p1, p.lo1, p.hi1, lo2, hi2);
  epilogue_mode(pw->act, pw->act2, pw->res_mode, nullptr, pw->act_p0, pw->act_p1, p.lo, p.hi, p.lo2, p.hi2);

  DpMaps maps;
  if (int rc = dp_nhwc_map(&maps.x, x, dw->x_coff + p.c, dw->x_ctotal, p.w, p.h, p.n, kDpBW, kDpBH)) return rc;
  if (int rc = dp_nhwc_map(&maps.y, y, pw->y_coff + p.nout, pw->y_ctotal, p.w, p.h, p.n, kDpTW, kDpTH)) return rc;
  if (int rc = dp_weight_map(&maps.w, w_pw_bf16, p.c, p.nout)) return rc;

  const unsigned grid = (unsigned)std::min<int64_t>(tiles, num_sms());
  cudaStream_t st = as_stream(stream);
  const __nv_bfloat16 *rp = static_cast<const __nv_bfloat16 *>(res);
#define RTSEG_DWPW(RM_)                                                                                             \
  do {                                                                                                              \
    static bool attr_set[64] = {false};                                                                             \
    int dev = 0;                                                                                                    \
    RTSEG_CUDA(cudaGetDevice(&dev));                                                                                \
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {                                                                   \
      RTSEG_CUDA(cudaFuncSetAttribute(dwpw_tc_kernel<RM_>, cudaFuncAttributeMaxDynamicSharedMemorySize,             \
                                      (int)kDpSmemBudget + 2048));                                                  \
      attr_set[dev] = true;                                                                                         \
    }                                                                                                               \
    dwpw_tc_kernel<RM_><<<grid, kDpThreads, smem, st>>>(maps, p, w_dw, scale_dw, shift_dw, scale_pw, shift_pw, rp); \
  } while (0)
  if (p.res_mode == RTSEG_RES_NONE) RTSEG_DWPW(0);
  else if (p.res_mode == RTSEG_RES_PRE) RTSEG_DWPW(RTSEG_RES_PRE);
  else RTSEG_DWPW(RTSEG_RES_POST);
#undef RTSEG_DWPW
  RTSEG_LAUNCH_CHECK("dwpw_tc_kernel");
  return RTSEG_OK;
}
