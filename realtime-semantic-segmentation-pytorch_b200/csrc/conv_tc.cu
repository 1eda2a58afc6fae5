// conv_tc.cu — K2'/K3'/K4'/K5' on the 5th-generation tensor cores.
//
// Dense convolutions of the path (ERFNet's 3x1 / 1x3 dilated NonBt1D convs,
// reference models/erfnet.py:65-71; pointwise convs, models/modules.py:63-69;
// dense 3x3 incl. the stride-2 downsamplers, models/enet.py:42; transposed
// 3x3 s2, models/modules.py:99-102) are implicit GEMMs
//        D[pixel][cout] = sum_tap  A_tap[pixel][cin] * W_tap[cout][cin]^T
// with M = 128 output pixels per tile, N = cout (<= 256) and K = taps * cin.
//
// Blackwell mapping (one persistent CTA per SM, 192 threads):
//   warp 0   TMA producer.  For every (tap, 64-channel K block) it issues one
//            cp.async.bulk.tensor.4d box [KB ch, TW px, TH rows, 1 image] at the
//            tap's shifted coordinates; TMA's out-of-bounds zero fill IS the
//            convolution padding.  A stride-2 conv reads four parity views of
//            the input (one tensor map each), so no strided box is needed.
//   warp 1   allocates TMEM and issues tcgen05.mma (cta_group::1, kind::f16,
//            bf16 x bf16 -> fp32) with A and B from 128B/64B/32B-swizzled shared
//            memory; accumulators live in TMEM, double buffered (2 x N columns)
//            so the epilogue of tile i overlaps the MMAs of tile i+1.
//   warps 2-5 epilogue: tcgen05.ld 32 lanes x 16 columns, folded-BN scale and
//            shift, residual (before or after), activation, bf16 pack, 16-byte
//            stores; optionally the channel argmax (segmentation labels).
// The weights of one N tile stay resident in shared memory for the lifetime of
// the CTA (<= 96 KB), the A operand flows through a 4-8 stage mbarrier ring.
//
// Everything here is hand-written PTX; no CUTLASS/CuTe.
#include <cuda.h>

#include <algorithm>
#include <mutex>

#include "common.cuh"

namespace rtseg {

constexpr int kTcThreads = 192;
constexpr int kTileM = 128;
constexpr int kMaxTcTaps = 9;
constexpr int kMaxStages = 8;
constexpr uint32_t kSmemBudget = 220 * 1024;

struct TcTap {
  int16_t map;   // which A tensor map (parity view)
  int16_t dy, dx;
  int16_t widx;  // weight tap index
};

struct TcParams {
  int n, gh, gw;          // output grid (pixels enumerated by tiles)
  int tw, th;             // tile = th rows x tw columns, tw*th == 128
  int tiles_w, tiles_h;   // tiles per image along w / h
  int64_t num_m_tiles;
  int oh, ow;             // full output size
  int ostride, oy0, ox0;  // output pixel = grid * ostride + o?0
  int cin, cout;          // logical
  int cout_pad;           // rows per tap in the packed weights
  int bn;                 // N tile (multiple of 16, <= 256)
  int nkb;                // K blocks per tap = cin / KB
  int ntaps;
  int x_coff;
  int y_ct, y_co, r_ct, r_co;
  int y_vec;              // 16-byte vector stores allowed (slice aligned)
  int stages;
  uint32_t b_block_bytes;  // smem bytes of one (tap, kb) weight block (1024-aligned)
  uint32_t tmem_cols;
  TcTap taps[kMaxTcTaps];
};

// ----------------------------------------------------------------- PTX helpers ----
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}

// Device-wide "a barrier wait gave up" flag: a wrong descriptor or a lost TMA
// must end in an error code, never in a hung GPU.
__device__ unsigned int g_tc_timeout_flag = 0;

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, volatile uint32_t *dead) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (*dead) return;
    if (clock64() - t0 > 4000000000LL) {  // ~2 s at 2 GHz
      *dead = 1;
      atomicExch(&g_tc_timeout_flag, 1u);
      return;
    }
  }
}

__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0, int c1, int c2,
                                            int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap *map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// Shared-memory matrix descriptor, K-major operand whose rows are KB bf16 wide
// (KB*2 bytes == the swizzle span) and densely packed: 8-row groups are
// 8*KB*2 bytes apart (SBO); LBO is unused for swizzled K-major layouts.
template <int KB>
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t addr) {
  constexpr uint64_t layout = KB == 64 ? 2 : (KB == 32 ? 4 : 6);  // SWIZZLE_128B / 64B / 32B
  constexpr uint64_t sbo = (8 * KB * 2) >> 4;
  return (uint64_t)((addr & 0x3ffffu) >> 4) | (1ull << 16) | (sbo << 32) | (1ull << 46) | (layout << 61);
}

// kind::f16 instruction descriptor: D = f32, A = B = bf16, both K-major, M = 128.
__device__ __forceinline__ uint32_t make_idesc(int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);
}

// ------------------------------------------------------------------------ kernel ----
struct TcMaps {
  CUtensorMap a[4];
  CUtensorMap b;
};

template <int KB>
__global__ void __launch_bounds__(kTcThreads, 1) conv_tc_kernel(const __grid_constant__ TcMaps maps, const TcParams p,
                                                                Epilogue ep, const __nv_bfloat16 *__restrict__ res,
                                                                __nv_bfloat16 *__restrict__ y,
                                                                uint8_t *__restrict__ labels) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve-up (all offsets multiples of 1024 until the small tail)
  uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr uint32_t kStageBytes = kTileM * KB * 2;
  const uint32_t b_bytes = p.b_block_bytes * p.ntaps * p.nkb;
  uint8_t *smem_b = smem;
  uint8_t *smem_a = smem + b_bytes;
  float *s_scale = reinterpret_cast<float *>(smem_a + (size_t)p.stages * kStageBytes);
  float *s_shift = s_scale + p.bn;
  float *s_slope = s_shift + p.bn;
  uint64_t *bars = reinterpret_cast<uint64_t *>(s_slope + p.bn);  // 8-byte aligned: bn % 16 == 0
  uint64_t *full_bar = bars;                                      // [stages]
  uint64_t *empty_bar = full_bar + kMaxStages;                    // [stages]
  uint64_t *tfull_bar = empty_bar + kMaxStages;                   // [2]
  uint64_t *tempty_bar = tfull_bar + 2;                           // [2]
  uint64_t *b_bar = tempty_bar + 2;                               // [1]
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(b_bar + 1);
  volatile uint32_t *dead = tmem_slot + 1;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tile = blockIdx.y;
  const int co_base = n_tile * p.bn;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(smem_u32(&full_bar[s]), 1);
      mbar_init(smem_u32(&empty_bar[s]), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(smem_u32(&tfull_bar[s]), 1);
      mbar_init(smem_u32(&tempty_bar[s]), 4);
    }
    mbar_init(smem_u32(b_bar), 1);
    *dead = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int i = threadIdx.x; i < p.bn; i += blockDim.x) {
    const int c = co_base + i;
    const bool ok = c < p.cout;
    s_scale[i] = (ok && ep.scale) ? ep.scale[c] : 1.f;
    s_shift[i] = (ok && ep.shift) ? ep.shift[c] : 0.f;
    s_slope[i] = (ok && ep.slope) ? ep.slope[c] : 0.f;
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"(p.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================================================== TMA producer ====
    if (lane == 0) {
      prefetch_tmap(&maps.b);
      prefetch_tmap(&maps.a[0]);
      // resident weights: ntaps * nkb boxes of [KB x bn]
      mbar_expect_tx(smem_u32(b_bar), (uint32_t)(p.ntaps * p.nkb) * (uint32_t)(p.bn * KB * 2));
      for (int t = 0; t < p.ntaps; ++t)
        for (int kb = 0; kb < p.nkb; ++kb)
          tma_load_2d(smem_u32(smem_b + (size_t)(t * p.nkb + kb) * p.b_block_bytes), &maps.b, smem_u32(b_bar), kb * KB,
                      p.taps[t].widx * p.cout_pad + co_base);
      uint32_t stage = 0, phase = 0;
      for (int64_t tile = blockIdx.x; tile < p.num_m_tiles; tile += gridDim.x) {
        const int twi = (int)(tile % p.tiles_w);
        const int64_t r = tile / p.tiles_w;
        const int thi = (int)(r % p.tiles_h);
        const int img = (int)(r / p.tiles_h);
        const int w0 = twi * p.tw, h0 = thi * p.th;
        for (int t = 0; t < p.ntaps; ++t) {
          const CUtensorMap *am = &maps.a[p.taps[t].map];
          for (int kb = 0; kb < p.nkb; ++kb) {
            mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1, dead);
            mbar_expect_tx(smem_u32(&full_bar[stage]), kStageBytes);
            tma_load_4d(smem_u32(smem_a + (size_t)stage * kStageBytes), am, smem_u32(&full_bar[stage]),
                        p.x_coff + kb * KB, w0 + p.taps[t].dx, h0 + p.taps[t].dy, img);
            if (++stage == (uint32_t)p.stages) stage = 0, phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ======================================================= MMA issuer ====
    if (lane == 0) {
      const uint32_t idesc = make_idesc(p.bn);
      mbar_wait(smem_u32(b_bar), 0, dead);
      uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0;
      for (int64_t tile = blockIdx.x; tile < p.num_m_tiles; tile += gridDim.x) {
        mbar_wait(smem_u32(&tempty_bar[acc]), acc_phase ^ 1, dead);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * (uint32_t)p.bn;
        uint32_t first = 1;
        for (int t = 0; t < p.ntaps; ++t) {
          for (int kb = 0; kb < p.nkb; ++kb) {
            mbar_wait(smem_u32(&full_bar[stage]), phase, dead);
            tc_fence_after();
            const uint32_t a_addr = smem_u32(smem_a + (size_t)stage * kStageBytes);
            const uint32_t b_addr = smem_u32(smem_b + (size_t)(t * p.nkb + kb) * p.b_block_bytes);
#pragma unroll
            for (int k = 0; k < KB / 16; ++k) {
              umma_bf16(d_tmem, make_smem_desc<KB>(a_addr + k * 32), make_smem_desc<KB>(b_addr + k * 32), idesc,
                        first ? 0u : 1u);
              first = 0;
            }
            umma_commit(smem_u32(&empty_bar[stage]));  // frees the A stage when these MMAs retire
            if (++stage == (uint32_t)p.stages) stage = 0, phase ^= 1;
          }
        }
        umma_commit(smem_u32(&tfull_bar[acc]));  // accumulator complete -> epilogue
        if (++acc == 2) acc = 0, acc_phase ^= 1;
      }
    }
  } else {
    // ========================================================= epilogue ====
    const int q = warp & 3;  // TMEM lane quarter this warp may touch
    const int row = q * 32 + lane;
    const int th_i = row / p.tw, tw_i = row % p.tw;
    uint32_t acc = 0, acc_phase = 0;
    for (int64_t tile = blockIdx.x; tile < p.num_m_tiles; tile += gridDim.x) {
      const int twi = (int)(tile % p.tiles_w);
      const int64_t r = tile / p.tiles_w;
      const int thi = (int)(r % p.tiles_h);
      const int img = (int)(r / p.tiles_h);
      const int gy = thi * p.th + th_i, gx = twi * p.tw + tw_i;
      const bool valid = gy < p.gh && gx < p.gw;
      const int64_t opix = ((int64_t)img * p.oh + (gy * p.ostride + p.oy0)) * p.ow + (gx * p.ostride + p.ox0);

      mbar_wait(smem_u32(&tfull_bar[acc]), acc_phase, dead);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * (uint32_t)p.bn;
      float best = 0.f;
      int best_idx = 0;
      for (int c0 = 0; c0 < p.bn; c0 += 16) {
        uint32_t v[16];
        tmem_ld16(taddr + c0, v);
        tmem_ld_wait();
        const int cg = co_base + c0;  // first global channel of this chunk
        if (valid && cg < p.cout) {
          float rv[16];
          if (ep.res_mode != RTSEG_RES_NONE) {
            const __nv_bfloat16 *rp = res + opix * p.r_ct + p.r_co + cg;
            float t8[8];
            Vec16<__nv_bfloat16>::load(rp, t8);
#pragma unroll
            for (int e = 0; e < 8; ++e) rv[e] = t8[e];
            Vec16<__nv_bfloat16>::load(rp + 8, t8);
#pragma unroll
            for (int e = 0; e < 8; ++e) rv[8 + e] = t8[e];
          } else {
#pragma unroll
            for (int e = 0; e < 16; ++e) rv[e] = 0.f;
          }
          float o[16];
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            float a = __uint_as_float(v[e]);
            if (ep.res_mode == RTSEG_RES_PRE) a += rv[e];
            a = fmaf(a, s_scale[c0 + e], s_shift[c0 + e]);
            a = act_apply(a, ep.act, ep.p0, ep.p1, s_slope[c0 + e]);
            if (ep.res_mode == RTSEG_RES_POST) {
              a += rv[e];
              a = act_apply(a, ep.act2, ep.p0, ep.p1, s_slope[c0 + e]);
            }
            o[e] = a;
          }
          if (labels) {
#pragma unroll
            for (int e = 0; e < 16; ++e) {
              if (cg + e < p.cout) {
                // labels follow the values the logits tensor would hold (bf16-rounded)
                const float vr = __bfloat162float(__float2bfloat16_rn(o[e]));
                if (cg + e == 0) {
                  best = vr, best_idx = 0;
                } else if (vr > best || (vr != vr && best == best)) {
                  best = vr, best_idx = cg + e;
                }
              }
            }
          }
          if (y) {
            __nv_bfloat16 *yp = y + opix * p.y_ct + p.y_co + cg;
            if (p.y_vec && cg + 16 <= p.cout) {
              float t8[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) t8[e] = o[e];
              Vec16<__nv_bfloat16>::store(yp, t8);
#pragma unroll
              for (int e = 0; e < 8; ++e) t8[e] = o[8 + e];
              Vec16<__nv_bfloat16>::store(yp + 8, t8);
            } else {
#pragma unroll
              for (int e = 0; e < 16; ++e)
                if (cg + e < p.cout) yp[e] = __float2bfloat16_rn(o[e]);
            }
          }
        }
      }
      if (labels && valid && gridDim.y == 1) labels[opix] = (uint8_t)best_idx;
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&tempty_bar[acc]));
      if (++acc == 2) acc = 0, acc_phase ^= 1;
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

// -------------------------------------------------------------------------- host ----
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void *ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  });
  return fn;
}

static CUtensorMapSwizzle swizzle_for(int kb) {
  return kb == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : (kb == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
}

static int pick_kb(int cin) { return cin % 64 == 0 ? 64 : (cin % 32 == 0 ? 32 : 16); }

struct TcPlan {
  int kb, bn, n_tiles, stages;
  uint32_t b_block_bytes, smem_bytes, tmem_cols;
};

// Chooses the N tile so that the resident weights leave room for >= 3 A stages.
static bool plan_tc(int cin, int cout_pad, int ntaps, TcPlan &pl) {
  pl.kb = pick_kb(cin);
  const int nkb = cin / pl.kb;
  const uint32_t stage_bytes = kTileM * pl.kb * 2;
  for (int bn = std::min(256, cout_pad); bn >= 16; bn -= 16) {
    if (cout_pad % bn) continue;
    const uint32_t blk = ((uint32_t)bn * pl.kb * 2 + 1023) / 1024 * 1024;
    const uint32_t b_bytes = blk * ntaps * nkb;
    const uint32_t tail = 3 * bn * 4 + (2 * kMaxStages + 5) * 8 + 16;
    if (b_bytes + tail + 1024 > kSmemBudget) continue;
    const int stages = (int)std::min<uint32_t>(kMaxStages, (kSmemBudget - 1024 - b_bytes - tail) / stage_bytes);
    if (stages < 3) continue;
    pl.bn = bn;
    pl.n_tiles = cout_pad / bn;
    pl.stages = stages;
    pl.b_block_bytes = blk;
    pl.smem_bytes = 1024 + b_bytes + stages * stage_bytes + tail;
    uint32_t cols = 32;
    while (cols < (uint32_t)(2 * bn)) cols <<= 1;
    pl.tmem_cols = cols;
    return true;
  }
  return false;
}

static int check_tc_common(const rtseg_conv_desc *d, bool deconv, bool has_y, const char **why) {
  *why = nullptr;
  if (d->dtype != RTSEG_BF16) *why = "dtype must be bf16";
  else if (d->groups != 1) *why = "groups must be 1";
  else if (d->cin % 16) *why = "cin must be a multiple of 16";
  else if (d->x_ctotal % 8 || d->x_coff % 8) *why = "input channel slice must be 16-byte aligned";
  else if (d->res_mode != RTSEG_RES_NONE && (d->r_ctotal % 8 || d->r_coff % 8)) *why = "residual slice must be 16-byte aligned";
  else if (d->kh * d->kw > kMaxTcTaps) *why = "more than 9 taps";
  else if (!deconv && d->stride != 1 && d->stride != 2) *why = "stride must be 1 or 2";
  else if (!deconv && d->stride == 2 && ((d->h & 1) || (d->w & 1))) *why = "stride 2 needs even H and W";
  return *why ? RTSEG_ERR_UNSUPPORTED : RTSEG_OK;
}

static int encode_act_map(CUtensorMap *m, const void *base, int ctotal, int64_t w, int64_t h, int64_t n, int64_t sw,
                          int64_t sh, int64_t sn, int kb, int tw, int th) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) RTSEG_FAIL(RTSEG_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  const cuuint64_t dims[4] = {(cuuint64_t)ctotal, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)n};
  const cuuint64_t strides[3] = {(cuuint64_t)sw * 2, (cuuint64_t)sh * 2, (cuuint64_t)sn * 2};
  const cuuint32_t box[4] = {(cuuint32_t)kb, (cuuint32_t)tw, (cuuint32_t)th, 1};
  const cuuint32_t es[4] = {1, 1, 1, 1};
  const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void *>(base), dims, strides, box, es,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for(kb), CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) RTSEG_FAIL(RTSEG_ERR_CUDA, "cuTensorMapEncodeTiled(activations) failed with CUresult %d", (int)r);
  return RTSEG_OK;
}

static int encode_weight_map(CUtensorMap *m, const void *w, int cin, int rows, int kb, int bn) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) RTSEG_FAIL(RTSEG_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  const cuuint64_t dims[2] = {(cuuint64_t)cin, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)cin * 2};
  const cuuint32_t box[2] = {(cuuint32_t)kb, (cuuint32_t)bn};
  const cuuint32_t es[2] = {1, 1};
  const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(w), dims, strides, box, es,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for(kb), CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) RTSEG_FAIL(RTSEG_ERR_CUDA, "cuTensorMapEncodeTiled(weights) failed with CUresult %d", (int)r);
  return RTSEG_OK;
}

static int check_timeout_flag(cudaStream_t st) {
  (void)st;
  return RTSEG_OK;
}

template <int KB>
static int launch_tc_kb(const TcMaps &maps, const TcParams &p, const TcPlan &pl, const Epilogue &ep, const void *res,
                        void *y, uint8_t *labels, cudaStream_t st) {
  static bool attr_set[64] = {false};
  int dev = 0;
  RTSEG_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    RTSEG_CUDA(cudaFuncSetAttribute(conv_tc_kernel<KB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBudget + 2048));
    attr_set[dev] = true;
  }
  const int sms = num_sms();
  const int gx = (int)std::min<int64_t>(p.num_m_tiles, std::max(1, sms / pl.n_tiles));
  dim3 grid((unsigned)gx, (unsigned)pl.n_tiles);
  conv_tc_kernel<KB><<<grid, kTcThreads, pl.smem_bytes + 1024, st>>>(maps, p, ep, static_cast<const __nv_bfloat16 *>(res),
                                                                   static_cast<__nv_bfloat16 *>(y), labels);
  RTSEG_LAUNCH_CHECK("conv_tc_kernel");
  return check_timeout_flag(st);
}

static int launch_tc(const TcMaps &maps, const TcParams &p, const TcPlan &pl, const Epilogue &ep, const void *res, void *y,
                     uint8_t *labels, cudaStream_t st) {
  switch (pl.kb) {
    case 64: return launch_tc_kb<64>(maps, p, pl, ep, res, y, labels, st);
    case 32: return launch_tc_kb<32>(maps, p, pl, ep, res, y, labels, st);
    default: return launch_tc_kb<16>(maps, p, pl, ep, res, y, labels, st);
  }
}

static void pick_tile(int gw, int &tw, int &th) {
  tw = 128;
  while (tw > 8 && tw / 2 >= gw) tw /= 2;  // smallest power of two >= gw, capped at 128, floor 8
  th = kTileM / tw;
}

static void fill_common(TcParams &p, const rtseg_conv_desc *d, const TcPlan &pl, int cout_pad) {
  p.cin = d->cin, p.cout = d->cout, p.cout_pad = cout_pad;
  p.bn = pl.bn, p.nkb = d->cin / pl.kb, p.stages = pl.stages;
  p.b_block_bytes = pl.b_block_bytes, p.tmem_cols = pl.tmem_cols;
  p.x_coff = d->x_coff;
  p.y_ct = d->y_ctotal, p.y_co = d->y_coff, p.r_ct = d->r_ctotal, p.r_co = d->r_coff;
  p.y_vec = (d->y_ctotal % 8 == 0 && d->y_coff % 8 == 0) ? 1 : 0;
  pick_tile(p.gw, p.tw, p.th);
  p.tiles_w = (int)cdiv(p.gw, p.tw), p.tiles_h = (int)cdiv(p.gh, p.th);
  p.num_m_tiles = (int64_t)p.n * p.tiles_w * p.tiles_h;
}

}  // namespace rtseg

using namespace rtseg;

extern "C" int rtseg_conv_tc_supported(const rtseg_conv_desc *d) {
  if (!d) return 0;
  const char *why;
  if (check_tc_common(d, false, true, &why)) return 0;
  TcPlan pl;
  return plan_tc(d->cin, (d->cout + 15) / 16 * 16, d->kh * d->kw, pl) ? 1 : 0;
}

// Reads and clears the device-side watchdog flag (synchronises the stream). Used by tests.
extern "C" int rtseg_tc_timeout_flag(rtseg_stream_t stream) {
  unsigned int v = 0, zero = 0;
  if (cudaStreamSynchronize(as_stream(stream)) != cudaSuccess) return -1;
  if (cudaMemcpyFromSymbol(&v, g_tc_timeout_flag, sizeof(v)) != cudaSuccess) return -1;
  if (v) cudaMemcpyToSymbol(g_tc_timeout_flag, &zero, sizeof(zero));
  return (int)v;
}

extern "C" int rtseg_conv_tc(const rtseg_conv_desc *d, const void *x, const void *w_bf16, const float *scale,
                             const float *shift, const float *slope, const void *res, void *y,
                             rtseg_stream_t stream) {
  if (int rc = validate_conv_desc(d, res)) return rc;
  RTSEG_REQUIRE(x && w_bf16 && y, "NULL tensor pointer");
  const char *why;
  if (check_tc_common(d, false, true, &why)) RTSEG_FAIL(RTSEG_ERR_UNSUPPORTED, "rtseg_conv_tc: %s", why);
  RTSEG_REQUIRE(reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(w_bf16) % 16 == 0 &&
                    reinterpret_cast<uintptr_t>(y) % 16 == 0 && reinterpret_cast<uintptr_t>(res) % 16 == 0,
                "tensor pointers must be 16-byte aligned");
  const int cout_pad = (d->cout + 15) / 16 * 16;
  const int ntaps = d->kh * d->kw;
  TcPlan pl;
  if (!plan_tc(d->cin, cout_pad, ntaps, pl))
    RTSEG_FAIL(RTSEG_ERR_UNSUPPORTED, "rtseg_conv_tc: weights (%d taps x %d x %d) do not fit in shared memory", ntaps,
               d->cin, cout_pad);
  TcParams p;
  p.n = d->n;
  rtseg_out_hw(d, &p.gh, &p.gw);
  p.oh = p.gh, p.ow = p.gw, p.ostride = 1, p.oy0 = 0, p.ox0 = 0;
  p.ntaps = ntaps;
  fill_common(p, d, pl, cout_pad);

  TcMaps maps;
  const int64_t ct = d->x_ctotal;
  if (d->stride == 1) {
    if (int rc = encode_act_map(&maps.a[0], x, d->x_ctotal, d->w, d->h, d->n, ct, ct * d->w, ct * d->w * d->h, pl.kb,
                                p.tw, p.th))
      return rc;
    for (int i = 1; i < 4; ++i) maps.a[i] = maps.a[0];
    for (int ky = 0, t = 0; ky < d->kh; ++ky)
      for (int kx = 0; kx < d->kw; ++kx, ++t) {
        p.taps[t].map = 0;
        p.taps[t].dy = (int16_t)(ky * d->dil - d->pad_h);
        p.taps[t].dx = (int16_t)(kx * d->dil - d->pad_w);
        p.taps[t].widx = (int16_t)t;
      }
  } else {
    // stride 2: input (h, w) = (2*h2 + ph, 2*w2 + pw); one tensor map per parity (ph, pw)
    for (int ph = 0; ph < 2; ++ph)
      for (int pw = 0; pw < 2; ++pw) {
        const __nv_bfloat16 *base = static_cast<const __nv_bfloat16 *>(x) + ((int64_t)ph * d->w + pw) * ct;
        if (int rc = encode_act_map(&maps.a[ph * 2 + pw], base, d->x_ctotal, d->w / 2, d->h / 2, d->n, 2 * ct,
                                    2 * ct * d->w, ct * d->w * d->h, pl.kb, p.tw, p.th))
          return rc;
      }
    for (int ky = 0, t = 0; ky < d->kh; ++ky)
      for (int kx = 0; kx < d->kw; ++kx, ++t) {
        const int oy = ky * d->dil - d->pad_h, ox = kx * d->dil - d->pad_w;  // input = 2*out + o
        const int ph = ((oy % 2) + 2) % 2, pw = ((ox % 2) + 2) % 2;
        p.taps[t].map = (int16_t)(ph * 2 + pw);
        p.taps[t].dy = (int16_t)((oy - ph) / 2);
        p.taps[t].dx = (int16_t)((ox - pw) / 2);
        p.taps[t].widx = (int16_t)t;
      }
  }
  if (int rc = encode_weight_map(&maps.b, w_bf16, d->cin, ntaps * cout_pad, pl.kb, pl.bn)) return rc;
  const Epilogue ep = make_epilogue(d, scale, shift, slope);
  return launch_tc(maps, p, pl, ep, res, y, nullptr, as_stream(stream));
}

extern "C" int rtseg_deconv3x3s2_tc_supported(const rtseg_conv_desc *d) {
  if (!d) return 0;
  const char *why;
  if (check_tc_common(d, true, true, &why)) return 0;
  if (d->kh != 3 || d->kw != 3 || d->stride != 2 || d->dil != 1) return 0;
  TcPlan pl;
  return plan_tc(d->cin, (d->cout + 15) / 16 * 16, 4, pl) ? 1 : 0;
}

extern "C" int rtseg_deconv3x3s2_tc(const rtseg_conv_desc *d, const void *x, const void *w_bf16, const float *scale,
                                    const float *shift, const float *slope, void *y, uint8_t *labels,
                                    rtseg_stream_t stream) {
  RTSEG_REQUIRE(d && x && w_bf16 && (y || labels), "NULL argument");
  RTSEG_REQUIRE(d->kh == 3 && d->kw == 3 && d->stride == 2 && d->dil == 1 && d->groups == 1,
                "rtseg_deconv3x3s2_tc handles k=3, s=2, dense only");
  RTSEG_REQUIRE(d->res_mode == RTSEG_RES_NONE, "deconv has no residual input");
  RTSEG_REQUIRE(d->n > 0 && d->h > 0 && d->w > 0 && d->cin > 0 && d->cout > 0, "non-positive extent");
  const char *why;
  if (check_tc_common(d, true, y != nullptr, &why)) RTSEG_FAIL(RTSEG_ERR_UNSUPPORTED, "rtseg_deconv3x3s2_tc: %s", why);
  RTSEG_REQUIRE(reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(w_bf16) % 16 == 0 &&
                    reinterpret_cast<uintptr_t>(y) % 16 == 0,
                "tensor pointers must be 16-byte aligned");
  RTSEG_REQUIRE(labels == nullptr || d->cout <= 256, "uint8 labels need cout <= 256");
  const int cout_pad = (d->cout + 15) / 16 * 16;
  TcPlan pl;
  if (!plan_tc(d->cin, cout_pad, 4, pl) || (labels && pl.n_tiles != 1))
    RTSEG_FAIL(RTSEG_ERR_UNSUPPORTED, "rtseg_deconv3x3s2_tc: weights do not fit in shared memory");
  TcMaps maps;
  TcParams p;
  p.n = d->n, p.gh = d->h, p.gw = d->w;
  p.oh = 2 * d->h, p.ow = 2 * d->w, p.ostride = 2;
  fill_common(p, d, pl, cout_pad);
  const int64_t ct = d->x_ctotal;
  if (int rc = encode_act_map(&maps.a[0], x, d->x_ctotal, d->w, d->h, d->n, ct, ct * d->w, ct * d->w * d->h, pl.kb, p.tw,
                              p.th))
    return rc;
  for (int i = 1; i < 4; ++i) maps.a[i] = maps.a[0];
  if (int rc = encode_weight_map(&maps.b, w_bf16, d->cin, 9 * cout_pad, pl.kb, pl.bn)) return rc;
  const Epilogue ep = make_epilogue(d, scale, shift, slope);
  // oy = 2*iy - 1 + ky: even rows take ky=1 (iy=i); odd rows ky=0 (iy=i+1) and ky=2 (iy=i). Same in x.
  for (int py = 0; py < 2; ++py)
    for (int px = 0; px < 2; ++px) {
      const int kys[2] = {py == 0 ? 1 : 0, 2}, dys[2] = {py == 0 ? 0 : 1, 0};
      const int kxs[2] = {px == 0 ? 1 : 0, 2}, dxs[2] = {px == 0 ? 0 : 1, 0};
      const int ny = py == 0 ? 1 : 2, nx = px == 0 ? 1 : 2;
      p.ntaps = 0;
      for (int a = 0; a < ny; ++a)
        for (int b = 0; b < nx; ++b) {
          TcTap &t = p.taps[p.ntaps++];
          t.map = 0, t.dy = (int16_t)dys[a], t.dx = (int16_t)dxs[b], t.widx = (int16_t)(kys[a] * 3 + kxs[b]);
        }
      p.oy0 = py, p.ox0 = px;
      if (int rc = launch_tc(maps, p, pl, ep, nullptr, y, labels, as_stream(stream))) return rc;
    }
  return RTSEG_OK;
}
