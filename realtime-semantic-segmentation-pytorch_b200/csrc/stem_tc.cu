// stem_tc.cu — the stride-2 3x3 image stem on the tensor cores (bf16 activations).
//
// Same operator as stem.cu (InitialBlock, reference models/enet.py:38-48, and the plain
// ConvBNAct(3, C, 3, 2) stems of models/fastscnn.py:37 / models/dabnet.py:19), different engine.
// The CUDA-core stem spends ~1200 (16 channels) to ~2400 (32 channels) issue slots per output pixel on
// FMAs and broadcast weight reads and is issue-bound at a quarter of the HBM roofline.  Here a thread
// still owns one output pixel and reads its 3x3xcin window straight from the caller's NCHW (or NHWC)
// image, but only to convert it to bf16 and drop it as one 64-byte im2col row (K = 9*cin <= 27, padded to
// 32) into a 128-row shared-memory tile in the UMMA K-major SWIZZLE_64B layout.  One elected thread then
// issues two tcgen05.mma (M = 128 pixels, N = 16 or 32 channels, K = 2 x 16) against the resident weight
// tile; the accumulator comes back from TMEM one row per thread, gets BN + activation, the max-pool
// channels (the window maximum, already in registers) are appended and the pixel is stored as 16-byte
// chunks.  ~150 issue slots per pixel; the kernel becomes a streaming pass over the image.
//
// Operands are rounded to bf16 (image and weights), accumulation is fp32 — the same contract as every
// other tensor-core layer of the bf16 path.  fp32 activations keep the CUDA-core stem.
#include <cfloat>

#include "common.cuh"
#include "ptx.cuh"

namespace rtseg {

namespace {

constexpr int kRows = 128;      // pixels per tile = MMA M
constexpr int kRowBytes = 64;   // K = 32 bf16 per im2col row

// K-major SWIZZLE_64B operand tile: rows of 64 bytes, 8-row groups 512 bytes apart.
__device__ __forceinline__ uint64_t desc_sw64(uint32_t addr) {
  return (uint64_t)((addr & 0x3ffffu) >> 4) | (1ull << 16) | ((uint64_t)(512 >> 4) << 32) | (1ull << 46) | (4ull << 61);
}

// byte offset of 16-byte chunk `c` of row `r` inside such a tile
__device__ __forceinline__ uint32_t sw64_off(int r, int c) {
  return (uint32_t)r * kRowBytes + ((uint32_t)(c ^ ((r >> 1) & 3)) << 4);
}

template <typename TI, int CIN, int N, int MODE, bool NCHW>
__global__ void __launch_bounds__(kRows, 5) stem_tc_kernel(StemParams p, const TI *__restrict__ x,
                                                           const float *__restrict__ wgt, Epilogue ep,
                                                           __nv_bfloat16 *__restrict__ y, int num_tiles) {
  constexpr int K = 9 * CIN;  // <= 27
  __shared__ __align__(1024) uint8_t s_a_buf[kRows * kRowBytes];
  __shared__ __align__(1024) uint8_t s_b_buf[N * kRowBytes];
  __shared__ __align__(16) float s_aff[3][32];
  __shared__ __align__(8) uint64_t s_bar_word;
  __shared__ uint32_t s_tmem;
  __shared__ uint32_t s_dead;
  const int tid = threadIdx.x, warp = tid >> 5;
  const uint32_t s_a = smem_u32(s_a_buf), s_b = smem_u32(s_b_buf), bar = smem_u32(&s_bar_word);
  volatile uint32_t *dead = &s_dead;

  // ---- one-time setup: weights -> bf16 K-major tile, affine vectors, TMEM, barrier ----------------
  if (tid < N) {  // row n of B = output channel n: k = (ky*3 + kx)*CIN + ci, source [3][3][CIN][cconv]
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      uint32_t pk[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int k0 = c * 8 + e * 2;
        const float a = (k0 < K && tid < p.cconv) ? __ldg(wgt + (int64_t)k0 * p.cconv + tid) : 0.f;
        const float b = (k0 + 1 < K && tid < p.cconv) ? __ldg(wgt + (int64_t)(k0 + 1) * p.cconv + tid) : 0.f;
        pk[e] = pack_bf16(a, b);
      }
      sts128(s_b + sw64_off(tid, c), pk[0], pk[1], pk[2], pk[3]);
    }
  }
  if (tid < 32) {
    const bool ok = tid < p.cconv;
    s_aff[0][tid] = (ok && ep.scale) ? ep.scale[tid] : 1.f;
    s_aff[1][tid] = (ok && ep.shift) ? ep.shift[tid] : 0.f;
    s_aff[2][tid] = (ok && ep.slope) ? ep.slope[tid] : 0.f;
  }
  if (tid == 0) {
    s_dead = 0;
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(32)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = s_tmem;
  const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(kRows >> 4) << 24);
  const uint64_t a_desc = desc_sw64(s_a), b_desc = desc_sw64(s_b);
  const int total = p.n * p.ho * p.wo;  // < 2^31 (checked by the launcher)
  const int ctot = p.cconv + (p.pool ? CIN : 0);
  uint32_t phase = 0;
  const bool relu_only = ep.lo == 0.f && ep.hi == INFINITY;

  // Index arithmetic is what this kernel would otherwise spend its issue slots on (the first version executed
  // ~690 instructions per pixel, 28 of them loads), hence: 32-bit pixel decomposition, element offsets of the
  // window's (channel, row) pairs hoisted out of the tile loop, an unpredicated path for interior pixels.
  const int kxs = NCHW ? 1 : (int)p.sw;  // element stride between horizontally adjacent taps
  int off[CIN][3];
#pragma unroll
  for (int ci = 0; ci < CIN; ++ci)
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) off[ci][ky] = (int)(ci * p.sc + ky * p.sh);
  const bool rows_aligned = (p.wo % kRows) == 0;  // then a tile never straddles an output row
  const int tiles_per_row = p.wo / kRows;

  // Loads the 3x3xCIN window of this thread's pixel of `tile` (loads only: they stay in flight across whatever
  // the caller does next).  okmask: bit t set when tap t lies inside the image (0x1ff for interior pixels).
  // Aligned rows: the tile's (image, row, tile-in-row) is carried from tile to tile (a tile advances by gridDim.x, i.e.
  // by a fixed number of rows and tiles-in-row) instead of being divided out of the tile index every time.
  int nx_b = 0, nx_oy = 0, nx_tx = 0, step_row = 0, step_tx = 0;
  if (rows_aligned) {
    const int row0 = (int)blockIdx.x / tiles_per_row;
    nx_tx = (int)blockIdx.x - row0 * tiles_per_row;
    nx_b = row0 / p.ho, nx_oy = row0 - nx_b * p.ho;
    step_row = (int)gridDim.x / tiles_per_row, step_tx = (int)gridDim.x - step_row * tiles_per_row;
  }
  auto load_window = [&](int tile, float (&v)[K], int &pixel, int &okmask) {
    int row, ox, b, oy;
    if (rows_aligned) {
      b = nx_b, oy = nx_oy, ox = nx_tx * kRows + tid;
      row = b * p.ho + oy;
      nx_tx += step_tx, nx_oy += step_row;
      if (nx_tx >= tiles_per_row) nx_tx -= tiles_per_row, ++nx_oy;
      while (nx_oy >= p.ho) nx_oy -= p.ho, ++nx_b;
    } else {
      const unsigned q = (unsigned)tile * kRows + tid;
      row = (int)(q / (unsigned)p.wo);
      ox = (int)(q - (unsigned)row * (unsigned)p.wo);
      b = row / p.ho, oy = row - b * p.ho;
    }
    pixel = row * p.wo + ox;
    const bool valid = tile < num_tiles && pixel < total;
    const int iy0 = oy * 2 - 1, ix0 = ox * 2 - 1;
    const TI *base = x + ((int64_t)b * p.sn + (int64_t)iy0 * p.sh + (int64_t)ix0 * kxs);
    const bool interior = valid && iy0 >= 0 && iy0 + 2 < p.h && ix0 >= 0 && ix0 + 2 < p.w;
    if (interior && sizeof(TI) == 1 && !NCHW && CIN == 3 && p.u8_words) {
      // decoded uint8 HWC image: a window row is 9 contiguous bytes.  Three aligned 32-bit loads cover them wherever
      // they start (the launcher checked 4-byte aligned rows and a 4-byte multiple of total bytes, so no word
      // reaches past the buffer); two funnel shifts line the bytes up.  9 loads per pixel instead of 27.
      okmask = 0x1ff;
#pragma unroll
      for (int ky = 0; ky < 3; ++ky) {
        const uintptr_t a = reinterpret_cast<uintptr_t>(base + off[0][ky]);
        const uint32_t r8 = (uint32_t)(a & 3u) * 8u;
        const uint32_t *wp = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t)3);
        const uint32_t w0 = __ldg(wp), w1 = __ldg(wp + 1), w2 = __ldg(wp + 2);
        const uint32_t s[3] = {__funnelshift_r(w0, w1, r8), __funnelshift_r(w1, w2, r8), w2 >> r8};
#pragma unroll
        for (int j = 0; j < 9; ++j) {
          const uint32_t byte = (s[j >> 2] >> (8 * (j & 3))) & 0xffu;
          v[ky * 9 + j] = __fmul_rn(__fsub_rn((float)byte, p.nz.m255[j % 3]), p.nz.rden[j % 3]);   // v[(ky*3 + kx)*3 + ci], j = kx*3 + ci
        }
      }
    } else if (interior) {
      okmask = 0x1ff;
#pragma unroll
      for (int ci = 0; ci < CIN; ++ci)
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
          const TI *rp = base + off[ci][ky];
#pragma unroll
          for (int kx = 0; kx < 3; ++kx) v[(ky * 3 + kx) * CIN + ci] = ld_img<TI>(rp + kx * kxs, p.nz, ci);
        }
    } else {
      okmask = 0;
#pragma unroll
      for (int ky = 0; ky < 3; ++ky)
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          const bool ok = valid && iy0 + ky >= 0 && iy0 + ky < p.h && ix0 + kx >= 0 && ix0 + kx < p.w;
          okmask |= ok ? 1 << (ky * 3 + kx) : 0;
#pragma unroll
          for (int ci = 0; ci < CIN; ++ci)
            v[(ky * 3 + kx) * CIN + ci] = ok ? ld_img<TI>(base + off[ci][ky] + kx * kxs, p.nz, ci) : 0.f;
        }
      if (!valid) pixel = -1;
    }
  };
  // MaxPool2d(3, 2, 1) per channel from the same window (padding never wins: the centre tap is always inside).
  auto window_max = [&](const float (&v)[K], int okmask, float (&mx)[CIN]) {
    if (okmask == 0x1ff) {
#pragma unroll
      for (int ci = 0; ci < CIN; ++ci) {
        float m = v[ci];
#pragma unroll
        for (int t = 1; t < 9; ++t) m = fmax_nan(m, v[t * CIN + ci]);
        mx[ci] = m;
      }
    } else {
#pragma unroll
      for (int ci = 0; ci < CIN; ++ci) {
        float m = v[4 * CIN + ci];
#pragma unroll
        for (int t = 0; t < 9; ++t)
          if ((okmask >> t) & 1) m = fmax_nan(m, v[t * CIN + ci]);
        mx[ci] = m;
      }
    }
  };

  float v[K];
  int pixel, okmask;
  load_window(blockIdx.x, v, pixel, okmask);
  for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
    float cur_mx[CIN] = {};
    if (p.pool) window_max(v, okmask, cur_mx);
    // ---- im2col row -> shared memory (bf16, swizzled) ----------------------------------------------
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      uint32_t pk[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int k0 = c * 8 + e * 2;
        pk[e] = pack_bf16(k0 < K ? v[k0 < K ? k0 : 0] : 0.f, k0 + 1 < K ? v[k0 + 1 < K ? k0 + 1 : 0] : 0.f);
      }
      sts128(s_a + sw64_off(tid, c), pk[0], pk[1], pk[2], pk[3]);
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();  // rows complete; every thread has also finished reading the previous accumulator
    if (tid == 0) {
      tc_fence_after();
      umma_bf16(tmem_base, a_desc, b_desc, idesc, 0u);
      umma_bf16(tmem_base, a_desc + 2, b_desc + 2, idesc, 1u);
      umma_commit(bar);
    }
    // the next tile's window is in flight while this tile's MMA runs and its accumulator is written out
    const int cur_pixel = pixel;
    load_window(tile + (int)gridDim.x, v, pixel, okmask);
    mbar_wait(bar, phase, dead);
    phase ^= 1;
    tc_fence_after();
    // ---- accumulator row -> BN + act -> NHWC ---------------------------------------------------------
    uint32_t acc[N];
    {
      uint32_t t16[16];
      tmem_ld16(tmem_base + ((uint32_t)(warp * 32) << 16), t16);
#pragma unroll
      for (int j = 0; j < 16; ++j) acc[j] = t16[j];
      if constexpr (N == 32) {
        tmem_ld16(tmem_base + ((uint32_t)(warp * 32) << 16) + 16, t16);
#pragma unroll
        for (int j = 0; j < 16; ++j) acc[16 + j] = t16[j];
      }
      tmem_ld_wait();
    }
    tc_fence_before();
    __nv_bfloat16 *yp = y + (int64_t)cur_pixel * p.y_ct + p.y_co;
    // 16-byte chunks; the pooled channels are the last CIN channels of the last chunk (ctot % 8 == 0)
#pragma unroll
    for (int c8 = 0; c8 < N + 8; c8 += 8) {
      if (MODE == 0 && relu_only && c8 < N && !(p.pool && c8 + 8 == ctot)) {
        // plain ReLU: the clamp rides on the bf16 conversion (cvt.rn.relu), two channels per instruction
        uint32_t pk[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int c = (c8 + 2 * e) < N ? c8 + 2 * e : 0;
          pk[e] = pack_relu_bf16(fmaf(__uint_as_float(acc[c]), s_aff[0][c], s_aff[1][c]),
                                 fmaf(__uint_as_float(acc[c + 1]), s_aff[0][c + 1], s_aff[1][c + 1]));
        }
        if (cur_pixel >= 0 && c8 < ctot) *reinterpret_cast<uint4 *>(yp + c8) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        continue;
      }
      float o[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int c = c8 + e;
        if (c < N) {
          const float t = fmaf(__uint_as_float(acc[c < N ? c : 0]), s_aff[0][c < N ? c : 0], s_aff[1][c < N ? c : 0]);
          if constexpr (MODE == 1) o[e] = fmaf(s_aff[2][c < N ? c : 0], fmin_nan(t, 0.f), fmax_nan(t, 0.f));
          else o[e] = fmin_nan(fmax_nan(t, ep.lo), ep.hi);
        } else {
          o[e] = 0.f;
        }
      }
      if (p.pool && c8 + 8 == ctot) {
#pragma unroll
        for (int ci = 0; ci < CIN; ++ci) o[8 - CIN + ci] = cur_mx[ci];
      }
      if (cur_pixel >= 0 && c8 < ctot)
        *reinterpret_cast<uint4 *>(yp + c8) =
            make_uint4(pack_bf16(o[0], o[1]), pack_bf16(o[2], o[3]), pack_bf16(o[4], o[5]), pack_bf16(o[6], o[7]));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(32) : "memory");
  }
}

template <typename TI, int CIN, int N>
int launch_mode(const StemParams &p, const TI *x, const float *w, const Epilogue &ep, __nv_bfloat16 *y, cudaStream_t st) {
  const int64_t total = (int64_t)p.n * p.ho * p.wo;
  const int64_t tiles = cdiv(total, kRows);
  // 32-bit index arithmetic inside the kernel: pixel counts and per-image element offsets must fit
  if (total + kRows >= INT32_MAX || (int64_t)CIN * p.sc + 3 * p.sh + 3 * p.sw >= INT32_MAX) return 1;
  const unsigned grid = (unsigned)std::min<int64_t>(tiles, (int64_t)num_sms() * 5);
  const bool nchw = p.sw == 1;
#define RTSEG_STEM_TC(MODE_, NCHW_) \
  stem_tc_kernel<TI, CIN, N, MODE_, NCHW_><<<grid, kRows, 0, st>>>(p, x, w, ep, y, (int)tiles)
  if (ep.mode == 0) { if (nchw) RTSEG_STEM_TC(0, true); else RTSEG_STEM_TC(0, false); }
  else { if (nchw) RTSEG_STEM_TC(1, true); else RTSEG_STEM_TC(1, false); }
#undef RTSEG_STEM_TC
  RTSEG_LAUNCH_CHECK("stem_tc_kernel");
  return RTSEG_OK;
}

template <typename TI>
int launch_in(const StemParams &p, const TI *x, const float *w, const Epilogue &ep, __nv_bfloat16 *y, cudaStream_t st) {
  const bool n16 = p.cconv <= 16;
  switch (p.cin) {
    case 1: return n16 ? launch_mode<TI, 1, 16>(p, x, w, ep, y, st) : launch_mode<TI, 1, 32>(p, x, w, ep, y, st);
    case 2: return n16 ? launch_mode<TI, 2, 16>(p, x, w, ep, y, st) : launch_mode<TI, 2, 32>(p, x, w, ep, y, st);
    case 3: return n16 ? launch_mode<TI, 3, 16>(p, x, w, ep, y, st) : launch_mode<TI, 3, 32>(p, x, w, ep, y, st);
    default: return 1;
  }
}

}  // namespace

// RTSEG_OK: launched.  1: not this kernel's shape (caller falls back to the CUDA-core stem).  < 0: error.
int stem_tc_try(const StemParams &p, int x_dtype, const void *x, const float *w, const Epilogue &ep, void *y,
                cudaStream_t st) {
  static const int knob = getenv("RTSEG_STEM_TC") ? atoi(getenv("RTSEG_STEM_TC")) : 1;
  if (!knob) return 1;
  const int ctot = p.cconv + (p.pool ? p.cin : 0);
  if (p.cin < 1 || p.cin > 3 || p.cconv > 32 || ep.mode > 1) return 1;
  if (ctot % 8 || p.y_ct % 8 || p.y_co % 8 || reinterpret_cast<uintptr_t>(y) % 16) return 1;  // 16-byte pixel stores
  if (ctot > (p.cconv <= 16 ? 16 : 32) + 8) return 1;
  __nv_bfloat16 *yo = static_cast<__nv_bfloat16 *>(y);
  if (x_dtype == RTSEG_F32) return launch_in<float>(p, static_cast<const float *>(x), w, ep, yo, st);
  if (x_dtype == RTSEG_BF16) return launch_in<__nv_bfloat16>(p, static_cast<const __nv_bfloat16 *>(x), w, ep, yo, st);
  if (x_dtype == RTSEG_U8) {
    StemParams q = p;
    static const int words_knob = getenv("RTSEG_STEM_U8_WORDS") ? atoi(getenv("RTSEG_STEM_U8_WORDS")) : 1;
    q.u8_words = words_knob && p.cin == 3 && p.sc == 1 && p.sw == 3 && p.sh % 4 == 0 && p.sn % 4 == 0 &&
                 reinterpret_cast<uintptr_t>(x) % 4 == 0;
    return launch_in<uint8_t>(q, static_cast<const uint8_t *>(x), w, ep, yo, st);
  }
  return 1;
}

int stem_tc_timeout_flag() { return read_clear_timeout_flag(); }

}  // namespace rtseg
