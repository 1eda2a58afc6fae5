// conv_tc.cu — K2'/K3'/K4'/K5' on the 5th-generation tensor cores.
//
// Dense convolutions of the path (ERFNet's 3x1 / 1x3 dilated NonBt1D convs,
// reference models/erfnet.py:65-71; pointwise convs, models/modules.py:63-69;
// dense 3x3 incl. the stride-2 downsamplers, models/enet.py:42; transposed
// 3x3 s2, models/modules.py:99-102) are implicit GEMMs
//        D[pixel][cout] = sum_tap  A_tap[pixel][cin] * W_tap[cout][cin]^T
// with M = 128 output pixels per tile, N = cout (<= 256) and K = taps * cin.
//
// Blackwell mapping (one persistent CTA per SM, 320 threads):
//   warp 0    TMA producer.  For every (tap, 64-channel K block) it issues one
//             cp.async.bulk.tensor.4d box [KB ch, TW px, TH rows, 1 image] at the
//             tap's shifted coordinates; TMA's out-of-bounds zero fill IS the
//             convolution padding.  A stride-2 conv reads four parity views of
//             the input (one tensor map each), so no strided box is needed.
//   warp 1    allocates TMEM and issues tcgen05.mma (cta_group::1, kind::f16,
//             bf16 x bf16 -> fp32), A and B from 128B/64B/32B-swizzled shared
//             memory.  Accumulators live in TMEM, double buffered, so the
//             epilogue of tile i overlaps the MMAs of tile i+1.  A transposed
//             conv keeps FOUR accumulators (one per output parity): each of its
//             four shifted input tiles is loaded once and feeds up to four MMAs.
//   warps 2-9 epilogue (two warps per TMEM lane quarter, alternating 16-column
//             chunks): tcgen05.ld, folded-BN scale/shift from shared memory,
//             residual (prefetched into registers before the accumulator is
//             ready), activation, bf16 pack, swizzled st.shared into a staging
//             tile, then ONE elected thread issues cp.async.bulk.tensor stores —
//             full 128-byte lines leave the SM, partial tiles are clipped by TMA.
//             Optional channel argmax (segmentation labels) in the same pass.
// The weights of one N tile stay resident in shared memory for the lifetime of
// the CTA, the A operand flows through a 3-8 stage mbarrier ring.
//
// Everything here is hand-written PTX; no CUTLASS/CuTe.
#include <cuda.h>

#include <algorithm>
#include <cfloat>
#include <type_traits>
#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "ptx.cuh"

namespace rtseg {

int dw_tma_timeout_flag();   // dw_tma.cu
int stem_tc_timeout_flag();  // stem_tc.cu

constexpr int kEpiWarps = 8;
constexpr int kEpiThreads = 32 * kEpiWarps;
constexpr int kTcThreads = 64 + kEpiThreads;
constexpr int kTileM = 128;
constexpr int kMaxTcTaps = 9;
constexpr int kMaxStages = 8;
constexpr int kMaxAcc = 4;
constexpr int kOutBufs = 2;
constexpr int kMaxMyChunks = 8;  // 16-column chunks one epilogue warp may own (bn <= 256 split over two warps)
constexpr uint32_t kSmemBudget = 224 * 1024;

struct TcTap {
  int16_t map;  // which A tensor map (parity view)
  int16_t dy, dx;
  int16_t nops;      // MMAs fed by this A tile
  int16_t widx[4];   // weight tap of each MMA
  int16_t acc[4];    // accumulator of each MMA
};

struct TcParams {
  int n, gh, gw;          // output grid (pixels enumerated by tiles)
  int tw, th;             // MMA sub-tile = th rows x tw columns, tw*th == 128
  int ms;                 // sub-tiles per unit tile
  int wstack;             // 0: sub-tiles stacked along h (unit = ms*th rows x tw cols); 1: along w (th rows x ms*tw cols)
  int unit_w, unit_h;     // unit tile extent in output-grid pixels
  // A-operand sharing ("halo tiles"): 0 = one TMA box per (tap, kb); 1 = ONE box per kb that contains every
  // tap's rows, each tap addressing it through a shifted UMMA descriptor.
  int a_share;
  int a_dy, a_dx;          // shared box origin relative to the unit tile origin
  uint32_t stage_bytes;    // bytes of one pipeline stage (1024-aligned)
  uint32_t a_tx_bytes;     // bytes one A box delivers (what the full barrier expects)
  uint32_t a_sub16;        // (byte offset between consecutive sub-tiles inside a stage) >> 4
  uint32_t a_sbo16;        // (byte stride between 8-row groups of the A tile) >> 4
  int a_boff;              // 1: start addresses are not 8-row aligned -> fill the descriptor's base-offset field
  int box_w, box_h;        // spatial extent of one A box (host side: tensor-map box)
  int res_boxes;           // residual sub-boxes (bn / ob) that travel through the A ring per unit, 0 = no residual
  int a_stages_per_unit;   // ring stages the MMA consumes per unit
  uint32_t tap_off16[kMaxTcTaps];  // a_share: (byte offset of tap t's first row inside the stage) >> 4
  int tiles_w, tiles_h;   // tiles per image along w / h
  int64_t num_m_tiles;
  int oh, ow;             // full output size
  int ostride;            // output pixel = grid * ostride + (oy0, ox0)[acc]
  int oy0[kMaxAcc], ox0[kMaxAcc];
  int nacc;
  int cin, cout;          // logical
  int cout_pad;           // rows per weight tap in the packed weights
  int bn;                 // N tile (multiple of 16, <= 256)
  int nkb;                // K blocks per tap = cin / KB
  int ntaps;              // A tiles per output tile
  int x_coff;
  int y_ct, y_co, r_ct, r_co;
  int y_mode;             // 0: no logits (labels only), 1: direct global stores, 2: staged TMA stores
  int ob;                 // channels per staged sub-box (64 / 32 / 16)
  int stages;
  uint32_t b_block_bytes;  // smem bytes of one (weight tap, kb) block (1024-aligned)
  uint32_t b_bytes;        // all resident weight blocks
  uint32_t out_buf_bytes;  // one staging buffer = ms * 128 * bn * 2
  int out_bufs;            // 2 (store of unit u overlaps the epilogue of u+1) or 1 when shared memory is tight
  int split_out;           // 1: a staging buffer holds ONE ob-channel sub-box; the epilogue runs bn / ob passes per unit
  uint32_t tmem_cols;
  float lo, hi;            // fast epilogue: clamp(v, lo, hi) covers none / relu / relu6 / hardtanh
  float lo2, hi2;          // ... and clamp(v + res, lo2, hi2) the activation after a POST residual
  TcTap taps[kMaxTcTaps];
};

#ifdef RTSEG_TC_TRACE
// Development aid (tools/tc_trace.py builds with -DRTSEG_TC_TRACE): CTA 0 records clock64() at the hand-over points of
// its first tiles.  Not compiled into the product library.
__device__ long long g_tc_trace[64][16];
#define TC_TRACE(tile_, ev_) do { if (blockIdx.x == 0 && blockIdx.y == 0 && (tile_) < 64) g_tc_trace[(tile_)][(ev_)] = clock64(); } while (0)
#else
#define TC_TRACE(tile_, ev_) do { } while (0)
#endif

__device__ __forceinline__ void epi_bar_sync(int id) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "n"(kEpiThreads) : "memory");
}

// Shared-memory matrix descriptor, K-major operand whose rows are KB bf16 wide
// (KB*2 bytes == the swizzle span) and densely packed: 8-row groups are
// 8*KB*2 bytes apart (SBO); LBO is unused for swizzled K-major layouts.
template <int KB>
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t addr) {
  constexpr uint64_t layout = KB == 64 ? 2 : (KB == 32 ? 4 : 6);  // SWIZZLE_128B / 64B / 32B
  constexpr uint64_t sbo = (8 * KB * 2) >> 4;
  return (uint64_t)((addr & 0x3ffffu) >> 4) | (1ull << 16) | (sbo << 32) | (1ull << 46) | (layout << 61);
}

template <int KB>
__device__ __forceinline__ uint64_t make_smem_desc_sbo(uint32_t addr, uint32_t sbo16) {
  constexpr uint64_t layout = KB == 64 ? 2 : (KB == 32 ? 4 : 6);
  return (uint64_t)((addr & 0x3ffffu) >> 4) | (1ull << 16) | ((uint64_t)sbo16 << 32) | (1ull << 46) | (layout << 61);
}

// kind::f16 instruction descriptor: D = f32, A = B = bf16, both K-major, M = 128.
__device__ __forceinline__ uint32_t make_idesc(int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);
}

// ------------------------------------------------------------------------ kernel ----
struct TcMaps {
  CUtensorMap a[4];        // input (parity views for stride 2)
  CUtensorMap b;           // packed weights
  CUtensorMap c[kMaxAcc];  // output (parity views for the transposed conv)
  CUtensorMap r;           // residual (same geometry as c[0])
};

// MODE 0: clamp (none / relu / relu6 / hardtanh).  MODE 1: max(v,0) + slope*min(v,0) (prelu).
template <int MODE>
__device__ __forceinline__ float fast_act(float v, float slope, float lo, float hi) {
  if constexpr (MODE == 1) return fmaf(slope, fmin_nan(v, 0.f), fmax_nan(v, 0.f));
  return fmin_nan(fmax_nan(v, lo), hi);
}

// {hi = bf16(a), lo = bf16(b)}, round to nearest even, NaN -> 0x7fff.
__device__ __forceinline__ uint32_t bf16x2_bits(float a, float b) {
  uint32_t d;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(a), "f"(b));
  return d;
}

__device__ __forceinline__ uint32_t bf16x2_relu_bits(float a, float b) {   // {hi = bf16(max(a, 0)), lo = bf16(max(b, 0))}
  uint32_t d;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(a), "f"(b));
  return d;
}

// The bf16 rounding the labels head compares in (value in the upper 16 bits).
__device__ __forceinline__ uint32_t bf16_hi_bits(float t) {
  uint32_t d;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(t), "f"(0.f));  // {hi = bf16(t), lo = bf16(0) = 0}
  return d;
}

__device__ __noinline__ float act_generic(float v, int act, float p0, float p1, float slope) {
  return act_apply(v, act, p0, p1, slope);
}

// MODE 0: clamp epilogue (none / relu / relu6 / hardtanh), staged output.
// MODE 1: slope epilogue (leakyrelu / prelu), staged output.
// MODE 2: everything else (any activation, fused argmax, unaligned output, labels only).
// MODE 3: transposed conv with labels only (clamp activation): argmax straight out of TMEM.
// The labels head (MODE 3) is bound by its epilogue warps' issue slots: there the producer and the MMA issuer — which
// wait most of the time — use the suspend-hinted wait (they were 30 % of the issued instructions).  Everywhere else
// these two warps sit on the critical path and keep the tight poll (hinted waits cost 5-8 % on the 128-channel layers).
template <int MODE>
__device__ __forceinline__ void tc_wait(uint32_t bar, uint32_t parity, volatile uint32_t *dead) {
  if constexpr (MODE == 3) mbar_wait_relaxed(bar, parity, dead);
  else mbar_wait(bar, parity, dead);
}

template <int KB, int MODE, bool RES, bool DIRECT = false>
__global__ void __launch_bounds__(kTcThreads, 1) conv_tc_kernel(const __grid_constant__ TcMaps maps, const TcParams p,
                                                                Epilogue ep, const __nv_bfloat16 *__restrict__ res,
                                                                __nv_bfloat16 *__restrict__ y,
                                                                uint8_t *__restrict__ labels) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t kStageBytes = p.stage_bytes;        // one pipeline stage = the A data of one (tap, kb) or one kb
  // carve-up in the shared address space (all offsets multiples of 1024 until the small tail)
  const uint32_t s_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t s_b = s_base;
  const uint32_t s_a = s_b + p.b_bytes;
  const uint32_t s_out = s_a + (uint32_t)p.stages * kStageBytes;
  const uint32_t s_scale = s_out + (uint32_t)p.out_bufs * p.out_buf_bytes;
  const uint32_t s_shift = s_scale + (uint32_t)p.bn * 4;
  const uint32_t s_slope = s_shift + (uint32_t)p.bn * 4;
  const uint32_t s_bars = s_slope + (uint32_t)p.bn * 4;  // 8-byte aligned: bn % 16 == 0
  const uint32_t full_bar = s_bars;                       // [kMaxStages]
  const uint32_t empty_bar = full_bar + 8 * kMaxStages;   // [kMaxStages]
  const uint32_t tfull_bar = empty_bar + 8 * kMaxStages;  // [2]
  const uint32_t tempty_bar = tfull_bar + 16;             // [2]
  const uint32_t b_bar = tempty_bar + 16;                 // [1]
  uint8_t *gen = smem_raw + (s_base - smem_u32(smem_raw));  // generic view of the same carve-up
  float *g_scale = reinterpret_cast<float *>(gen + (s_scale - s_base));
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(gen + (b_bar + 8 - s_base));
  volatile uint32_t *dead = tmem_slot + 1;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tile = blockIdx.y;
  const int co_base = n_tile * p.bn;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(full_bar + 8 * s, 1);
      mbar_init(empty_bar + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar + 8 * s, 1);
      mbar_init(tempty_bar + 8 * s, kEpiWarps);
    }
    mbar_init(b_bar, 1);
    *dead = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int i = threadIdx.x; i < p.bn; i += blockDim.x) {
    const int c = co_base + i;
    const bool ok = c < p.cout;
    g_scale[i] = (ok && ep.scale) ? ep.scale[c] : 1.f;
    // (labels-only mode: padded classes get shift = -inf so that they evaluate to the clamp's lower bound)
    g_scale[p.bn + i] = ok ? (ep.shift ? ep.shift[c] : 0.f) : (MODE == 3 ? -INFINITY : 0.f);
    g_scale[2 * p.bn + i] = (ok && ep.slope) ? ep.slope[c] : 0.f;
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
  const uint32_t acc_cols = (uint32_t)(p.ms * p.nacc * p.bn);  // TMEM columns of one accumulator stage
  const int unit_h = p.unit_h, unit_w = p.unit_w;

  if (warp == 0) {
    // ===================================================== TMA producer ====
    if (lane == 0) {
      prefetch_tmap(&maps.b);
      prefetch_tmap(&maps.a[0]);
      const int nwt = (int)(p.b_bytes / p.b_block_bytes) / p.nkb;  // resident weight taps
      mbar_expect_tx(b_bar, (uint32_t)(nwt * p.nkb) * (uint32_t)(p.bn * KB * 2));
      for (int t = 0; t < nwt; ++t)
        for (int kb = 0; kb < p.nkb; ++kb)
          tma_load_2d(s_b + (uint32_t)(t * p.nkb + kb) * p.b_block_bytes, &maps.b, b_bar, kb * KB,
                      t * p.cout_pad + co_base);
      uint32_t stage = 0, phase = 0;
      // tile -> (image, tile row, tile column), advanced by the decomposed grid stride: no division per tile
      const int64_t r0 = (int64_t)blockIdx.x / p.tiles_w, rs = (int64_t)gridDim.x / p.tiles_w;
      int twi = (int)((int64_t)blockIdx.x % p.tiles_w), thi = (int)(r0 % p.tiles_h), img = (int)(r0 / p.tiles_h);
      const int step_w = (int)((int64_t)gridDim.x % p.tiles_w), step_h = (int)(rs % p.tiles_h), step_n = (int)(rs / p.tiles_h);
      auto next_tile = [&]() {
        twi += step_w;
        if (twi >= p.tiles_w) twi -= p.tiles_w, ++thi;
        thi += step_h;
        if (thi >= p.tiles_h) thi -= p.tiles_h, ++img;
        img += step_n;
      };
      for (int64_t tile = blockIdx.x; tile < p.num_m_tiles; tile += gridDim.x, next_tile()) {
        const int w0 = twi * unit_w, h0 = thi * unit_h;
        if (p.a_share) {
          for (int kb = 0; kb < p.nkb; ++kb) {
            tc_wait<MODE>(empty_bar + 8 * stage, phase ^ 1, dead);
            TC_TRACE((int)((tile - blockIdx.x) / gridDim.x), kb == 0 ? 0 : 1);
            mbar_expect_tx(full_bar + 8 * stage, p.a_tx_bytes);
            tma_load_4d(s_a + stage * kStageBytes, &maps.a[0], full_bar + 8 * stage, p.x_coff + kb * KB, w0 + p.a_dx,
                        h0 + p.a_dy, img);
            if (++stage == (uint32_t)p.stages) stage = 0, phase ^= 1;
          }
        } else {
          for (int t = 0; t < p.ntaps; ++t) {
            const CUtensorMap *am = &maps.a[p.taps[t].map];
            for (int kb = 0; kb < p.nkb; ++kb) {
              tc_wait<MODE>(empty_bar + 8 * stage, phase ^ 1, dead);
              mbar_expect_tx(full_bar + 8 * stage, p.a_tx_bytes);
              tma_load_4d(s_a + stage * kStageBytes, am, full_bar + 8 * stage, p.x_coff + kb * KB, w0 + p.taps[t].dx,
                          h0 + p.taps[t].dy, img);
              if (++stage == (uint32_t)p.stages) stage = 0, phase ^= 1;
            }
          }
        }
        // The residual tile rides the same ring (consumed by the epilogue warps, not by the MMA): it is
        // in shared memory, swizzled like the output staging tile, long before its accumulator is ready.
        for (int sb = 0; sb < p.res_boxes; ++sb) {
          tc_wait<MODE>(empty_bar + 8 * stage, phase ^ 1, dead);
          mbar_expect_tx(full_bar + 8 * stage, (uint32_t)p.ms * kTileM * (uint32_t)p.ob * 2);
          tma_load_4d(s_a + stage * kStageBytes, &maps.r, full_bar + 8 * stage, p.r_co + co_base + sb * p.ob, w0, h0, img);
          if (++stage == (uint32_t)p.stages) stage = 0, phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ======================================================= MMA issuer ====
    // The whole warp walks the loop (uniform control flow, uniform registers); one elected lane issues.
    {
      const uint32_t idesc = make_idesc(p.bn);
      // Descriptors differ only in their 14-bit start-address field: build them once and add
      // (byte offset >> 4) — the issuing thread is a single lane, every instruction counts.
      const uint64_t a_desc0 = make_smem_desc_sbo<KB>(s_a, p.a_sbo16);
      const uint64_t b_desc0 = make_smem_desc<KB>(s_b);
      const uint32_t stage16 = kStageBytes >> 4, sub16 = p.a_sub16, blk16 = p.b_block_bytes >> 4;
      mbar_wait(b_bar, 0, dead);
      uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0;
      for (int64_t tile = blockIdx.x; tile < p.num_m_tiles; tile += gridDim.x) {
        tc_wait<MODE>(tempty_bar + 8 * acc, acc_phase ^ 1, dead);
        tc_fence_after();
        if (lane == 0) TC_TRACE((int)((tile - blockIdx.x) / gridDim.x), 2);
        const uint32_t d_tmem = tmem_base + acc * acc_cols;
        uint32_t started = 0;  // bit a: accumulator a already holds a partial sum
        if (p.a_share) {
          // one stage per kb holds the rows of every tap; tap t starts tap_off16[t] into it
          for (int kb = 0; kb < p.nkb; ++kb) {
            tc_wait<MODE>(full_bar + 8 * stage, phase, dead);
            tc_fence_after();
            if (lane == 0) TC_TRACE((int)((tile - blockIdx.x) / gridDim.x), kb == 0 ? 3 : 4);
            if (elect_one()) {
            if (p.nacc == 1) {
              // forward conv: one accumulator, one weight tap per A tap — the leanest possible issue loop (this single
              // lane feeds the tensor core; every extra instruction here shows up in the C=128 layers)
              for (int t = 0; t < p.ntaps; ++t) {
                const uint64_t b_d = b_desc0 + (uint64_t)((uint32_t)(p.taps[t].widx[0] * p.nkb + kb) * blk16);
                const uint32_t fresh = (started & 1u) | (t > 0 ? 1u : 0u);
                for (int sub = 0; sub < p.ms; ++sub) {
                  const uint32_t off16 = stage * stage16 + p.tap_off16[t] + (uint32_t)sub * sub16;
                  uint64_t a_d = a_desc0 + (uint64_t)off16;
                  // rows not aligned to the 8-row swizzle atom: tell the MMA the phase of the first row
                  if (p.a_boff) a_d |= (uint64_t)((((s_a >> 4) + off16) >> 3) & 7u) << 49;
                  const uint32_t d_col = d_tmem + (uint32_t)sub * (uint32_t)p.bn;
                  umma_bf16(d_col, a_d, b_d, idesc, fresh);
#pragma unroll
                  for (int k = 1; k < KB / 16; ++k) umma_bf16(d_col, a_d + 2 * k, b_d + 2 * k, idesc, 1u);
                }
              }
            } else {
              uint32_t st = started;  // bit a: accumulator a already holds a partial sum (this lane's running copy)
              for (int t = 0; t < p.ntaps; ++t) {
                const TcTap &tap = p.taps[t];
                for (int o = 0; o < tap.nops; ++o) {  // a transposed conv feeds up to four accumulators from one tap
                  const uint64_t b_d = b_desc0 + (uint64_t)((uint32_t)(tap.widx[o] * p.nkb + kb) * blk16);
                  const uint32_t a_acc = (uint32_t)tap.acc[o];
                  const uint32_t fresh = (st >> a_acc) & 1u;
                  st |= 1u << a_acc;
                  for (int sub = 0; sub < p.ms; ++sub) {
                    const uint64_t a_d = a_desc0 + (uint64_t)(stage * stage16 + p.tap_off16[t] + (uint32_t)sub * sub16);
                    const uint32_t d_col = d_tmem + (uint32_t)(sub * p.nacc + (int)a_acc) * (uint32_t)p.bn;
                    umma_bf16(d_col, a_d, b_d, idesc, fresh);
#pragma unroll
                    for (int k = 1; k < KB / 16; ++k) umma_bf16(d_col, a_d + 2 * k, b_d + 2 * k, idesc, 1u);
                  }
                }
              }
            }
            umma_commit(empty_bar + 8 * stage);
            }
            __syncwarp();
            started = 0xfu;  // every accumulator of the unit has received this K block's contribution
            if (++stage == (uint32_t)p.stages) stage = 0, phase ^= 1;
          }
        } else
        for (int t = 0; t < p.ntaps; ++t) {
          const TcTap &tap = p.taps[t];
          for (int kb = 0; kb < p.nkb; ++kb) {
            mbar_wait(full_bar + 8 * stage, phase, dead);
            tc_fence_after();
            const uint64_t a_st = a_desc0 + (uint64_t)(stage * stage16);
            if (elect_one()) {
              for (int o = 0; o < tap.nops; ++o) {
                const uint64_t b_d = b_desc0 + (uint64_t)((uint32_t)(tap.widx[o] * p.nkb + kb) * blk16);
                const uint32_t a_acc = (uint32_t)tap.acc[o];
                const uint32_t fresh = (started >> a_acc) & 1u;  // same for every sub-tile of this accumulator
                for (int sub = 0; sub < p.ms; ++sub) {
                  const uint64_t a_d = a_st + (uint64_t)((uint32_t)sub * sub16);
                  const uint32_t d_col = d_tmem + (uint32_t)(sub * p.nacc + (int)a_acc) * (uint32_t)p.bn;
                  umma_bf16(d_col, a_d, b_d, idesc, fresh);
#pragma unroll
                  for (int k = 1; k < KB / 16; ++k) umma_bf16(d_col, a_d + 2 * k, b_d + 2 * k, idesc, 1u);
                }
              }
              umma_commit(empty_bar + 8 * stage);  // frees the A stage when these MMAs retire
            }
            __syncwarp();
            for (int o = 0; o < tap.nops; ++o) started |= 1u << (uint32_t)tap.acc[o];
            if (++stage == (uint32_t)p.stages) stage = 0, phase ^= 1;
          }
        }
        if (elect_one()) umma_commit(tfull_bar + 8 * acc);  // accumulators complete -> epilogue
        __syncwarp();
        if (lane == 0) TC_TRACE((int)((tile - blockIdx.x) / gridDim.x), 5);
        // The residual stages belong to the epilogue, but this warp must still SEE each of their fills complete
        // before it moves on: a parity wait can only tell "this lap" from "the previous lap".  If the ring is
        // shorter than one unit's boxes (stages < A boxes + residual boxes), the slot of a later A tile held a
        // residual box one lap earlier; skipping that fill unobserved let the wait for the A tile succeed on the
        // lap before it (stale operand, then a second release of the slot and barrier phases out of step).
        for (int sb = 0; sb < p.res_boxes; ++sb) {
          mbar_wait(full_bar + 8 * stage, phase, dead);
          if (++stage == (uint32_t)p.stages) stage = 0, phase ^= 1;
        }
        if (++acc == 2) acc = 0, acc_phase ^= 1;
      }
    }
  } else {
    // ========================================================= epilogue ====
    const int e = warp - 2;
    const int q = warp & 3;      // TMEM lane quarter this warp may touch
    const int hsel = e >> 2;     // which of the two warps of that quarter
    const int row = q * 32 + lane;
    const bool elected = threadIdx.x == 64;
    const int nchunks = p.bn >> 4;
    const bool want_labels = labels != nullptr;
    const uint32_t sw_mask = p.ob == 64 ? 7u : (p.ob == 32 ? 3u : 1u);
    const uint32_t ob_bytes = (uint32_t)p.ob * 2;
    const uint32_t box_bytes = (uint32_t)p.ms * kTileM * ob_bytes;  // one staged channel group of the unit tile
    const int total_chunks = p.ms * nchunks;
    const bool has_res = ep.res_mode != RTSEG_RES_NONE;
    const float pre_w = ep.res_mode == RTSEG_RES_PRE ? 1.f : 0.f, post_w = ep.res_mode == RTSEG_RES_POST ? 1.f : 0.f;
    // Which (sub-tile, 16-column chunk) pairs this warp owns never changes: hoist all index math
    // (divisions, swizzle) out of the tile loop.  Flattened chunk k = hsel + 2*ci.
    // row of this thread's pixel (sub-tile `sub`) inside the staged output box [unit_h][unit_w]
    auto stage_row = [&](int sub) -> uint32_t {
      return p.wstack ? (uint32_t)((row / p.tw) * p.unit_w + sub * p.tw + row % p.tw) : (uint32_t)(sub * kTileM + row);
    };
    constexpr int kMy = kMaxMyChunks;
    uint32_t ck_meta[kMy];  // c0 | sub << 16
    uint32_t ck_tcol[kMy], ck_st0[kMy];
    int my_chunks = 0;
#pragma unroll
    for (int ci = 0; ci < kMy; ++ci) {
      const int k = hsel + 2 * ci;
      const int sub = k / nchunks, c0 = (k - sub * nchunks) * 16;
      if (k < total_chunks) my_chunks = ci + 1;
      ck_meta[ci] = (uint32_t)c0 | ((uint32_t)sub << 16) | ((uint32_t)(c0 / p.ob) << 24);  // c0 | sub | sub-box
      ck_tcol[ci] = (uint32_t)(sub * p.nacc * p.bn + c0);
      // 32-byte-aligned offset of the chunk in the staged box; its second 16-byte half is `^ 16`
      const uint32_t off0 = stage_row(sub) * ob_bytes + (uint32_t)(c0 % p.ob) * 2;
      // (split staging: the buffer holds this chunk's sub-box only)
      ck_st0[ci] = (p.split_out ? 0u : (uint32_t)(c0 / p.ob) * box_bytes) + (off0 ^ (((off0 >> 7) & sw_mask) << 4));
    }
    const uint32_t st_sb_bytes = p.split_out ? 0u : box_bytes;   // distance of consecutive sub-boxes inside a staging buffer
    uint32_t acc = 0, acc_phase = 0, unit = 0;
    uint32_t ring = 0, ring_phase = 0;  // position in the A ring (the residual stages live there)
    auto ring_advance = [&](int nsteps) {
      for (int i = 0; i < nsteps; ++i)
        if (++ring == (uint32_t)p.stages) ring = 0, ring_phase ^= 1;
    };
    // Tile -> (image, tile row, tile column) is carried along incrementally (the grid stride decomposed once),
    // and the thread's position inside a sub-tile is fixed: no division inside the tile loop.
    const int row_y = row / p.tw, row_x = row % p.tw;  // tw is a power of two dividing kTileM
    const int sub_dy = p.wstack ? 0 : kTileM / p.tw, sub_dx = p.wstack ? p.tw : 0;
    int twi, thi, img, step_w, step_h, step_n;
    {
      const int64_t r0 = (int64_t)blockIdx.x / p.tiles_w, rs = (int64_t)gridDim.x / p.tiles_w;
      twi = (int)((int64_t)blockIdx.x % p.tiles_w), thi = (int)(r0 % p.tiles_h), img = (int)(r0 / p.tiles_h);
      step_w = (int)((int64_t)gridDim.x % p.tiles_w), step_h = (int)(rs % p.tiles_h), step_n = (int)(rs / p.tiles_h);
    }
    auto next_tile = [&]() {
      twi += step_w;
      if (twi >= p.tiles_w) twi -= p.tiles_w, ++thi;
      thi += step_h;
      if (thi >= p.tiles_h) thi -= p.tiles_h, ++img;
      img += step_n;
    };
    for (int64_t tile = blockIdx.x; tile < p.num_m_tiles; tile += gridDim.x, next_tile()) {
      // grid coordinates of this thread's pixel in sub-tile `sub`
      auto pixel = [&](int sub, int &gy, int &gx) -> bool {
        gy = thi * unit_h + sub * sub_dy + row_y;
        gx = twi * unit_w + sub * sub_dx + row_x;
        return gy < p.gh && gx < p.gw;
      };

      if constexpr (MODE == 3) {
        // ---- labels-only transposed conv (the segmentation head): the logits never leave TMEM.
        // The two warps of a lane quarter split the output ROW parity; each thread reduces both column
        // parities of its input pixel and writes two adjacent uint8 labels with one store.  The argmax
        // runs as four independent integer-max chains (2 output pixels x 2 interleaved channel groups);
        // padded classes are skipped.
        mbar_wait(tfull_bar + 8 * acc, acc_phase, dead);
        tc_fence_after();
#pragma unroll 1
        for (int sub = 0; sub < p.ms; ++sub) {
          const uint32_t t0 =
              tmem_base + ((uint32_t)(q * 32) << 16) + acc * acc_cols + (uint32_t)((sub * p.nacc + hsel * 2) * p.bn);
          // key = ordered(bf16(logit)) << 16 | bf16(256 - class): "largest value, first index on ties" becomes ONE
          // unsigned max per class.  Non-negative logits (ReLU-family head, lo >= 0): the bf16 bit pattern is already
          // monotonic and a single two-value conversion builds the key.  Otherwise the usual sign fix (negative:
          // flip all bits, positive: set the top bit) after folding -0 into +0, which torch.max treats as equal.
          // NaN-propagating clamps keep a NaN logit (0x7fff after conversion) above +inf (0x7f80), and the index
          // half makes the first NaN win, like torch.max.
          // No divergent code anywhere near tcgen05.ld (.sync.aligned needs the whole warp).
          uint32_t key[2][2] = {{0u, 0u}, {0u, 0u}};
          const bool has_hi = p.hi < 3.0e38f;
          auto reduce = [&](auto nonneg_tag, auto relu_tag) {
            constexpr bool kNonNeg = decltype(nonneg_tag)::value;
            constexpr bool kRelu = decltype(relu_tag)::value;   // plain ReLU: the clamp rides in the key's conversion
#pragma unroll
            for (int ci = 0; ci < kMaxMyChunks; ++ci) {
              if (ci < nchunks) {  // warp-uniform
                const int c0 = ci * 16;
                uint32_t v[2][16];
                tmem_ld16(t0 + c0, v[0]);
                tmem_ld16(t0 + (uint32_t)p.bn + c0, v[1]);
                tmem_ld_wait();
#pragma unroll
                for (int e4 = 0; e4 < 16; e4 += 4) {
                  if (c0 + e4 < p.cout) {  // whole groups of padded classes: skip (warp-uniform)
                    const float4 sc = lds128(s_scale + (uint32_t)(c0 + e4) * 4);
                    const float4 sh = lds128(s_shift + (uint32_t)(c0 + e4) * 4);
                    const float scv[4] = {sc.x, sc.y, sc.z, sc.w}, shv[4] = {sh.x, sh.y, sh.z, sh.w};
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                      // the last group's padded classes (warp-uniform): 0 * inf in their accumulators must not win.
                      // (A branch-free variant with straight-line full groups measured slower: 0.485 vs 0.460 ms.)
                      if (c0 + e4 + j >= p.cout) continue;
#pragma unroll
                      for (int px = 0; px < 2; ++px) {
                        float t = fmaf(__uint_as_float(v[px][e4 + j]), scv[j], shv[j]);
                        uint32_t k;
                        if constexpr (kRelu) {
                          // cvt.rn.relu.bf16x2: {bf16(max(logit, 0)), bf16(256 - class)} in ONE instruction (a NaN
                          // stays a NaN and wins, as with the NaN-propagating clamp below)
                          k = bf16x2_relu_bits(t, (float)(256 - (c0 + e4 + j)));
                          key[px][j & 1] = max(key[px][j & 1], k);
                          continue;
                        }
                        t = fmax_nan(t, p.lo);
                        if (has_hi) t = fmin_nan(t, p.hi);
                        if constexpr (kNonNeg) {
                          // one conversion builds the whole key: {bf16(logit), bf16(256 - class)}; 256 - class is
                          // an exact, positive bf16 for all 256 classes the tensor-core path takes
                          k = bf16x2_bits(t, (float)(256 - (c0 + e4 + j)));
                        } else {
                          uint32_t b = bf16_hi_bits(t);
                          b = b == 0x80000000u ? 0u : b;
                          b = (b ^ ((uint32_t)((int32_t)b >> 31) | 0x80000000u)) & 0xffff0000u;
                          k = b | (bf16x2_bits(0.f, (float)(256 - (c0 + e4 + j))) & 0xffffu);
                        }
                        key[px][j & 1] = max(key[px][j & 1], k);
                      }
                    }
                  }
                }
              }
            }
          };
          if (p.lo == 0.f && !has_hi) reduce(std::true_type{}, std::true_type{});
          else if (p.lo >= 0.f) reduce(std::true_type{}, std::false_type{});
          else reduce(std::false_type{}, std::false_type{});
          const uint32_t k0 = max(key[0][0], key[0][1]), k1 = max(key[1][0], key[1][1]);
          // class = 256 - value of the low bf16
          const uint32_t lab2 = (uint32_t)(256 - (int)__uint_as_float(k0 << 16)) | ((uint32_t)(256 - (int)__uint_as_float(k1 << 16)) << 8);
          int gy, gx;
          if (pixel(sub, gy, gx)) {
            const int64_t opix = ((int64_t)img * p.oh + (gy * 2 + hsel)) * p.ow + gx * 2;
            *reinterpret_cast<uint16_t *>(labels + opix) = (uint16_t)lab2;
          }
        }
      } else if constexpr (MODE != 2) {
        // ---- fast path: staged TMA-store output, branch-free activation, residual prefetched into
        // registers while the MMAs of this tile are still running, TMEM loads one chunk ahead.
        // Residual sub-boxes of this unit: ring stages right behind the unit's A stages.
        uint32_t res_first = 0;  // ring slot of residual sub-box 0 (sub-box sb sits sb slots further, modulo)
        // Order matters: the accumulator barrier first.  Once it has fired every A fill of this unit has completed,
        // so the slot of each residual box is at most one lap behind and the parity wait below cannot succeed on
        // an older lap (it could when the ring is shorter than one unit's boxes and the slot still held an A tile
        // of the same unit).
        mbar_wait(tfull_bar + 8 * acc, acc_phase, dead);
        tc_fence_after();
        if (elected) TC_TRACE((int)((tile - blockIdx.x) / gridDim.x), 6);
        if constexpr (RES) {
          ring_advance(p.a_stages_per_unit);
          res_first = ring;
          uint32_t rs = ring, rph = ring_phase;
#pragma unroll
          for (int sb = 0; sb < 4; ++sb) {
            if (sb < p.res_boxes) {
              mbar_wait(full_bar + 8 * rs, rph, dead);
              if (++rs == (uint32_t)p.stages) rs = 0, rph ^= 1;
            }
          }
          ring_advance(p.res_boxes);
        }
        for (int a = 0; a < p.nacc; ++a) {
          const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + acc * acc_cols + (uint32_t)(a * p.bn);
          uint32_t v[2][16];
          if (my_chunks > 0) tmem_ld16(tbase + ck_tcol[0], v[0]);   // (the prefetch chain runs on across passes)
          // Split staging (wide layers, ms == 1): one pass per ob-channel sub-box through half-size staging buffers —
          // the shared memory this frees goes to the A ring, and bytes in flight are what bound these layers.
          const int npass = p.split_out ? p.bn / p.ob : 1;
          for (int pass = 0; pass < npass; ++pass) {
          const uint32_t buf = s_out + (p.out_bufs == 2 ? (unit & 1u) : 0u) * p.out_buf_bytes;
          constexpr bool direct = DIRECT;   // no staging buffers in this plan: rows go straight to global memory
                                            // (a template flag: as a run-time branch it slowed the staged layers down)
          if constexpr (!direct) epi_bar_sync(2);  // the staging buffer's previous TMA store has been read out
#pragma unroll
          for (int ci = 0; ci < kMy; ++ci) {
            if (ci < my_chunks && (!p.split_out || (int)(ck_meta[ci] >> 24) == pass)) {  // warp-uniform
              tmem_ld_wait();
              if (ci + 1 < kMy && ci + 1 < my_chunks) tmem_ld16(tbase + ck_tcol[ci + 1], v[(ci + 1) & 1]);
              const int c0 = (int)(ck_meta[ci] & 0xffffu);
              uint4 rres0 = make_uint4(0, 0, 0, 0), rres1 = rres0;
              if constexpr (RES) {  // residual chunk: same swizzled position as the output chunk, in its ring stage
                const uint32_t sb = ck_meta[ci] >> 24;
                uint32_t rslot = res_first + sb;
                if (rslot >= (uint32_t)p.stages) rslot -= (uint32_t)p.stages;
                const uint32_t ra = s_a + rslot * kStageBytes + (ck_st0[ci] - sb * st_sb_bytes);
                const float4 f0 = lds128(ra), f1 = lds128(ra ^ 16u);
                rres0 = make_uint4(__float_as_uint(f0.x), __float_as_uint(f0.y), __float_as_uint(f0.z), __float_as_uint(f0.w));
                rres1 = make_uint4(__float_as_uint(f1.x), __float_as_uint(f1.y), __float_as_uint(f1.z), __float_as_uint(f1.w));
              }
              uint32_t pk[8];
#pragma unroll
              for (int e4 = 0; e4 < 16; e4 += 4) {
                const float4 sc = lds128(s_scale + (uint32_t)(c0 + e4) * 4);
                const float4 sh = lds128(s_shift + (uint32_t)(c0 + e4) * 4);
                const float scv[4] = {sc.x, sc.y, sc.z, sc.w}, shv[4] = {sh.x, sh.y, sh.z, sh.w};
                float slv[4] = {0.f, 0.f, 0.f, 0.f};
                if constexpr (MODE == 1) {
                  const float4 sl = lds128(s_slope + (uint32_t)(c0 + e4) * 4);
                  slv[0] = sl.x, slv[1] = sl.y, slv[2] = sl.z, slv[3] = sl.w;
                }
                float o[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) o[j] = __uint_as_float(v[ci & 1][e4 + j]);
                if constexpr (!RES) {
#pragma unroll
                  for (int j = 0; j < 4; ++j) o[j] = fast_act<MODE>(fmaf(o[j], scv[j], shv[j]), slv[j], p.lo, p.hi);
                } else {
                  const uint4 w4 = (e4 & 8) ? rres1 : rres0;
                  const uint32_t lo2 = (e4 & 4) ? w4.z : w4.x, hi2 = (e4 & 4) ? w4.w : w4.y;
                  const float rv[4] = {bf16_lo(lo2), bf16_hi(lo2), bf16_lo(hi2), bf16_hi(hi2)};
                  if (ep.res_mode == RTSEG_RES_PRE) {
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                      o[j] = fast_act<MODE>(fmaf(o[j] + rv[j], scv[j], shv[j]), slv[j], p.lo, p.hi);
                  } else {
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                      o[j] = fmin_nan(fmax_nan(fast_act<MODE>(fmaf(o[j], scv[j], shv[j]), slv[j], p.lo, p.hi) + rv[j], p.lo2), p.hi2);
                  }
                }
                pk[(e4 >> 1) + 0] = pack_bf16(o[0], o[1]);
                pk[(e4 >> 1) + 1] = pack_bf16(o[2], o[3]);
              }
              if constexpr (direct) {
                int gy, gx;
                if (pixel((int)((ck_meta[ci] >> 16) & 0xffu), gy, gx)) {
                  const int64_t opix = ((int64_t)img * p.oh + (gy * p.ostride + p.oy0[a])) * p.ow + (gx * p.ostride + p.ox0[a]);
                  uint4 *yp = reinterpret_cast<uint4 *>(y + opix * p.y_ct + p.y_co + co_base + c0);
                  yp[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                  yp[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                }
              } else {
                // staged, swizzled exactly like the TMA store's tensor map expects (offsets hoisted)
                sts128(buf + ck_st0[ci], pk[0], pk[1], pk[2], pk[3]);
                sts128(buf + (ck_st0[ci] ^ 16u), pk[4], pk[5], pk[6], pk[7]);
              }
            }
          }
          if constexpr (direct) { ++unit; continue; }
          fence_async_smem();
          epi_bar_sync(1);  // every warp has written its part of the staging tile
          if constexpr (RES) {
            if (elected && pass == npass - 1) {  // ... and has read its residual chunks: hand the ring stages back to the producer
              uint32_t rs = res_first;
              for (int sb = 0; sb < p.res_boxes; ++sb) {
                mbar_arrive(empty_bar + 8 * rs);
                if (++rs == (uint32_t)p.stages) rs = 0;
              }
            }
          }
          if (elected) {
            if (p.split_out) {
              tma_store_4d(&maps.c[a], buf, p.y_co + co_base + pass * p.ob, twi * unit_w, thi * unit_h, img);
            } else {
              const int nbox = p.bn / p.ob;
              for (int sb = 0; sb < nbox; ++sb)
                tma_store_4d(&maps.c[a], buf + (uint32_t)sb * box_bytes, p.y_co + co_base + sb * p.ob, twi * unit_w,
                             thi * unit_h, img);
            }
            bulk_commit();
            if (p.out_bufs == 2) bulk_wait_read<1>(); else bulk_wait_read<0>();  // the OTHER buffer is free again for the next unit
          }
          ++unit;
          }
        }
      } else {
        // ---- general path: any activation, fused argmax, unaligned channel counts.  Compact code
        // (rolled chunk loop, out-of-line activation) — correctness first, used off the hot path.
        mbar_wait(tfull_bar + 8 * acc, acc_phase, dead);
        tc_fence_after();
        // direct stores go out as 16-byte vectors when every 16-channel chunk is whole and 16-byte aligned
        const bool vec_out = p.y_mode == 1 && p.cout % 16 == 0 && p.y_ct % 8 == 0 && p.y_co % 8 == 0 &&
                             (reinterpret_cast<uintptr_t>(y) & 15) == 0;
        for (int a = 0; a < p.nacc; ++a) {
          const uint32_t buf = s_out + (p.out_bufs == 2 ? (unit & 1u) : 0u) * p.out_buf_bytes;
          if (p.y_mode == 2) epi_bar_sync(2);
#pragma unroll 1
          for (int sub = 0; sub < p.ms; ++sub) {
          int gy, gx;
          const bool valid = pixel(sub, gy, gx);
          const uint32_t srow = stage_row(sub);
          const int64_t opix = ((int64_t)img * p.oh + (gy * p.ostride + p.oy0[a])) * p.ow + (gx * p.ostride + p.ox0[a]);
          const uint32_t taddr =
              tmem_base + ((uint32_t)(q * 32) << 16) + acc * acc_cols + (uint32_t)((sub * p.nacc + a) * p.bn);
          float best = -FLT_MAX;
          int best_idx = 0;
          // with labels the first warp of each quarter walks all chunks in order (argmax needs every channel)
          const int c_step = want_labels ? 1 : 2;
          const int c_begin = want_labels ? (hsel == 0 ? 0 : nchunks) : hsel;
#pragma unroll 1
          for (int c = c_begin; c < nchunks; c += c_step) {
            const int c0 = c * 16;
            const int cg = co_base + c0;
            uint32_t v[16];
            tmem_ld16(taddr + c0, v);
            tmem_ld_wait();
            uint32_t pk[8];
            float prev = 0.f;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float sc = g_scale[c0 + j], sh = g_scale[p.bn + c0 + j], sl = g_scale[2 * p.bn + c0 + j];
              float rv = 0.f;
              if (has_res && valid && cg + j < p.cout) rv = __bfloat162float(res[opix * p.r_ct + p.r_co + cg + j]);
              float t = fmaf(pre_w, rv, __uint_as_float(v[j]));
              t = act_generic(fmaf(t, sc, sh), ep.act, ep.p0, ep.p1, sl);
              if (ep.res_mode == RTSEG_RES_POST) t = act_generic(t + rv, ep.act2, ep.p0, ep.p1, sl);
              if (j & 1) pk[j >> 1] = pack_bf16(prev, t);
              prev = t;
              if (want_labels && cg + j < p.cout) {
                // labels follow the values the logits tensor would hold (bf16-rounded)
                const float vr = __bfloat162float(__float2bfloat16_rn(t));
                if (cg + j == 0) {
                  best = vr, best_idx = 0;
                } else if (vr > best || (vr != vr && best == best)) {
                  best = vr, best_idx = cg + j;
                }
              }
              if (p.y_mode == 1 && !vec_out && valid && cg + j < p.cout)
                y[opix * p.y_ct + p.y_co + cg + j] = __float2bfloat16_rn(t);
            }
            if (p.y_mode == 1 && vec_out && valid && cg < p.cout) {   // the chunk's 32 bytes as two 16-byte stores
              uint4 *yp = reinterpret_cast<uint4 *>(y + opix * p.y_ct + p.y_co + cg);
              yp[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
              yp[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
            }
            if (p.y_mode == 2) {
              const uint32_t box = buf + (uint32_t)(c0 / p.ob) * box_bytes;
              const uint32_t off0 = srow * ob_bytes + (uint32_t)(c0 % p.ob) * 2;
              const uint32_t off1 = off0 + 16;
              sts128(box + (off0 ^ (((off0 >> 7) & sw_mask) << 4)), pk[0], pk[1], pk[2], pk[3]);
              sts128(box + (off1 ^ (((off1 >> 7) & sw_mask) << 4)), pk[4], pk[5], pk[6], pk[7]);
            }
          }
          if (want_labels && hsel == 0 && valid && gridDim.y == 1) labels[opix] = (uint8_t)best_idx;
          }
          if (p.y_mode == 2) {
            fence_async_smem();
            epi_bar_sync(1);
            if (elected) {
              const int nbox = p.bn / p.ob;
              for (int sb = 0; sb < nbox; ++sb)
                tma_store_4d(&maps.c[a], buf + (uint32_t)sb * box_bytes, p.y_co + co_base + sb * p.ob, twi * unit_w,
                             thi * unit_h, img);
              bulk_commit();
              if (p.out_bufs == 2) bulk_wait_read<1>(); else bulk_wait_read<0>();
            }
          }
          ++unit;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar + 8 * acc);
      if (elected) TC_TRACE((int)((tile - blockIdx.x) / gridDim.x), 7);
      if (++acc == 2) acc = 0, acc_phase ^= 1;
    }
    if (elected && p.y_mode == 2) bulk_wait_all();
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

static int pick_kb(int c) { return c % 64 == 0 ? 64 : (c % 32 == 0 ? 32 : 16); }

struct TcPlan {
  int kb, bn, n_tiles, stages, ob, ms, out_bufs, split_out;
  uint32_t b_block_bytes, b_bytes, out_buf_bytes, smem_bytes, tmem_cols, stage_bytes;
};

// Chooses the N tile so that resident weights + output staging leave room for >= 3 A stages.
// Narrow layers (small N) stack `ms` 128-pixel sub-tiles per unit so that the fixed per-unit costs
// (barrier hand-offs, TMA issue, TMEM round trip) are spread over ms times more pixels.
// share: 0 = one A box per (tap, kb); 1 = vertical halo tile (8x16-pixel sub-tiles side by side);
// 2 = horizontal halo tile (128-pixel rows, taps are row-shifted descriptors); 3 = two-dimensional halo tile for
// square k x k kernels (8x16-pixel sub-tiles side by side, every tap a shifted descriptor).  halo = (k-1)*dilation.
static uint32_t tc_stage_bytes(int share, int halo, int ms, int kb) {
  uint32_t b;
  if (share == 1) b = (uint32_t)(16 + halo) * 8 * ms * kb * 2;
  else if (share == 2) b = (uint32_t)ms * (kTileM + halo) * kb * 2;
  else if (share == 3) b = (uint32_t)(16 + halo) * (8 * ms + halo) * kb * 2;
  else b = (uint32_t)ms * kTileM * kb * 2;
  return (b + 1023) / 1024 * 1024;
}

static bool plan_tc(int cin, int cout_pad, int nwtaps, int nacc, bool staged_out, int max_ms, bool has_res,
                    TcPlan &pl, int share = 0, int halo = 0, bool allow_split = false) {
  const int chunk_cap = 2 * kMaxMyChunks;  // 16-column chunks per unit (two warps share)
  pl.kb = pick_kb(cin);
  const int nkb = cin / pl.kb;
  for (int bn = std::min(256, cout_pad); bn >= 16; bn -= 16) {
    if (cout_pad % bn) continue;
    if (2 * nacc * bn > 512) continue;  // TMEM: two accumulator stages
    if (bn / 16 > chunk_cap) continue;
    const uint32_t blk = ((uint32_t)bn * pl.kb * 2 + 1023) / 1024 * 1024;
    const uint32_t b_bytes = blk * nwtaps * nkb;
    int ms = std::max(1, max_ms);
    int out_bufs = kOutBufs;
    uint32_t stage_bytes = 0, out_buf = 0, fixed = 0;
    const uint32_t misc = 3 * bn * 4 + (2 * kMaxStages + 5) * 8 + 16 + 1024;
    for (; ms >= 1; ms >>= 1) {
      stage_bytes = tc_stage_bytes(share, halo, ms, pl.kb);
      if (has_res)  // residual sub-boxes [ob ch x unit] travel through the same ring
        stage_bytes = std::max(stage_bytes, ((uint32_t)ms * kTileM * pick_kb(bn) * 2 + 1023) / 1024 * 1024);
      out_buf = staged_out ? (uint32_t)ms * kTileM * bn * 2 : 0;
      fixed = b_bytes + kOutBufs * out_buf + misc;
      // a shared halo stage feeds a whole K block (all taps), so three of them already cover three units
      const bool fits = 2 * ms * nacc * bn <= 512 && ms * (bn / 16) <= chunk_cap &&
                        fixed + (share ? 3 : 4) * stage_bytes <= kSmemBudget;
      if (fits) break;
    }
    if (ms < 1) {
      ms = 1;
      stage_bytes = tc_stage_bytes(share, halo, 1, pl.kb);
      if (has_res) stage_bytes = std::max(stage_bytes, ((uint32_t)kTileM * pick_kb(bn) * 2 + 1023) / 1024 * 1024);
      out_buf = staged_out ? (uint32_t)kTileM * bn * 2 : 0;
      fixed = b_bytes + kOutBufs * out_buf + misc;
      // Fewer than five A stages in flight starve the MMAs (each tile needs taps*nkb of them):
      // give up the second staging buffer instead.
      static const int min_stages_knob = getenv("RTSEG_TC_MINSTAGES") ? atoi(getenv("RTSEG_TC_MINSTAGES")) : 5;
      if (fixed + (uint32_t)min_stages_knob * stage_bytes > kSmemBudget && out_buf) {
        out_bufs = 1;
        fixed = b_bytes + out_buf + misc;
      }
    }
    if (fixed + 3 * stage_bytes > kSmemBudget) continue;
    // Residual layers push their residual boxes through the same ring.  When the ring cannot hold one unit's worth
    // (K blocks + residual boxes; the C=128 layers: 96 KB of weights + two 32 KB staging buffers leave three stages
    // for 2 + 2 boxes) the producer stalls on its own unit: trade the second output staging buffer for stages.
    // Measured: C=128 +res 0.51 -> 0.57 of peak; layers whose ring is deep enough keep both buffers (a single
    // staging buffer costs the non-residual C=128 layers 13 %).  RTSEG_TC_DEEP=0 disables.
    static const int deep = getenv("RTSEG_TC_DEEP") ? atoi(getenv("RTSEG_TC_DEEP")) : 1;
    const int unit_boxes = (share ? nkb : nwtaps * nkb) + bn / pick_kb(bn);
    if (deep && has_res && (share || nwtaps == 1) && out_bufs == 2 && out_buf &&
        (int)((kSmemBudget - fixed) / stage_bytes) < unit_boxes + 1) {
      out_bufs = 1;
      fixed -= out_buf;
    }
    // Split staging: stage one ob-channel sub-box at a time through two half-size buffers when that buys ring stages.
    pl.split_out = 0;
    // Measured (round 2, B200): five ring stages instead of three did NOT speed the 128-channel layers up (0.0735 vs
    // 0.0698 ms per layer, Fast-SCNN's wide pointwise layers 4 % slower) — those layers are bound by the shared-memory
    // data pipe (tensor-core operand fetch 37 % + LSU 22 % of its cycles under ncu), not by bytes in flight, and the
    // second bar.sync pair per unit costs more than the stages give.  Off by default; RTSEG_TC_SPLIT=1 enables it.
    static const int split_knob = getenv("RTSEG_TC_SPLIT") ? atoi(getenv("RTSEG_TC_SPLIT")) : 0;
    if (split_knob && allow_split && out_buf && ms == 1 && nacc == 1 && bn / pick_kb(bn) >= 2) {
      const uint32_t half = (uint32_t)kTileM * pick_kb(bn) * 2;
      const uint32_t fixed_split = b_bytes + 2 * half + misc;
      if ((kSmemBudget - fixed_split) / stage_bytes > (kSmemBudget - fixed) / stage_bytes) {
        pl.split_out = 1, out_bufs = 2, out_buf = half, fixed = fixed_split;
      }
    }
    pl.out_bufs = out_bufs;
    pl.stage_bytes = stage_bytes;
    pl.ms = ms;
    pl.bn = bn;
    pl.n_tiles = cout_pad / bn;
    pl.stages = (int)std::min<uint32_t>(kMaxStages, (kSmemBudget - fixed) / stage_bytes);
    pl.b_block_bytes = blk;
    pl.b_bytes = b_bytes;
    pl.out_buf_bytes = out_buf;
    pl.ob = pick_kb(bn);
    pl.smem_bytes = fixed + pl.stages * stage_bytes;
    uint32_t cols = 32;
    while (cols < (uint32_t)(2 * ms * nacc * bn)) cols <<= 1;
    pl.tmem_cols = cols;
    return true;
  }
  return false;
}

static int check_tc_common(const rtseg_conv_desc *d, bool deconv, const char **why) {
  *why = nullptr;
  if (d->dtype != RTSEG_BF16) *why = "dtype must be bf16";
  else if (d->groups != 1) *why = "groups must be 1";
  else if (deconv && d->cin % 16) *why = "cin must be a multiple of 16";
  else if (d->x_ctotal % 8 || d->x_coff % 8) *why = "input channel slice must be 16-byte aligned";
  else if (d->res_mode != RTSEG_RES_NONE && (d->r_ctotal % 8 || d->r_coff % 8)) *why = "residual slice must be 16-byte aligned";
  else if (d->kh * d->kw > kMaxTcTaps) *why = "more than 9 taps";
  else if (!deconv && d->stride != 1 && d->stride != 2) *why = "stride must be 1 or 2";
  else if (!deconv && d->stride == 2 && ((d->h & 1) || (d->w & 1))) *why = "stride 2 needs even H and W";
  return *why ? RTSEG_ERR_UNSUPPORTED : RTSEG_OK;
}

// 4-D bf16 tensor map over an NHWC view: dims {c, w, h, n}, element strides {1, sw, sh, sn}.
static int encode_nhwc_map(CUtensorMap *m, const void *base, int cdim, int64_t w, int64_t h, int64_t n, int64_t sw,
                           int64_t sh, int64_t sn, int cbox, int tw, int th) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) RTSEG_FAIL(RTSEG_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  const cuuint64_t dims[4] = {(cuuint64_t)cdim, (cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)n};
  const cuuint64_t strides[3] = {(cuuint64_t)sw * 2, (cuuint64_t)sh * 2, (cuuint64_t)sn * 2};
  const cuuint32_t box[4] = {(cuuint32_t)cbox, (cuuint32_t)tw, (cuuint32_t)th, 1};
  const cuuint32_t es[4] = {1, 1, 1, 1};
  const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void *>(base), dims, strides, box, es,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for(cbox), CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) RTSEG_FAIL(RTSEG_ERR_CUDA, "cuTensorMapEncodeTiled(NHWC view) failed with CUresult %d", (int)r);
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

template <int KB, int MODE, bool RES, bool DIRECT = false>
static int launch_tc_inst(const TcMaps &maps, const TcParams &p, const TcPlan &pl, const Epilogue &ep, const void *res,
                          void *y, uint8_t *labels, cudaStream_t st) {
  static bool attr_set[64] = {false};
  int dev = 0;
  RTSEG_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    RTSEG_CUDA(cudaFuncSetAttribute(conv_tc_kernel<KB, MODE, RES, DIRECT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)kSmemBudget + 2048));
    attr_set[dev] = true;
  }
  // The labels-only head is epilogue-bound and light on registers / shared memory / TMEM:
  // two CTAs per SM double the number of epilogue warps.
  const int ctas_per_sm = (MODE == 3 && pl.tmem_cols <= 256 && pl.smem_bytes + 1024 <= 110 * 1024) ? 2 : 1;
  const int sms = num_sms() * ctas_per_sm;
  const int gx = (int)std::min<int64_t>(p.num_m_tiles, std::max(1, sms / pl.n_tiles));
  dim3 grid((unsigned)gx, (unsigned)pl.n_tiles);
  conv_tc_kernel<KB, MODE, RES, DIRECT><<<grid, kTcThreads, pl.smem_bytes + 1024, st>>>(
      maps, p, ep, static_cast<const __nv_bfloat16 *>(res), static_cast<__nv_bfloat16 *>(y), labels);
  RTSEG_LAUNCH_CHECK("conv_tc_kernel");
  return RTSEG_OK;
}

static int launch_tc(const TcMaps &maps, const TcParams &p, const TcPlan &pl, const Epilogue &ep, int mode,
                     const void *res, void *y, uint8_t *labels, cudaStream_t st) {
  // mode 3 writes two adjacent labels with one 16-bit store: the class map must be 2-byte aligned (rows are: ow even)
  // direct (unstaged) stores stay on the fast epilogue when every 16-channel chunk is whole and 16-byte aligned
  const bool direct_fast = p.y_mode == 1 && mode <= 1 && p.res_boxes == 0 && !pl.split_out && p.cout % 16 == 0 &&
                           p.y_ct % 8 == 0 && p.y_co % 8 == 0 && reinterpret_cast<uintptr_t>(y) % 16 == 0;
  if (labels != nullptr && p.y_mode == 0 && p.nacc == 4 && mode == 0 && pl.n_tiles == 1 && (p.ow & 1) == 0 &&
      reinterpret_cast<uintptr_t>(labels) % 2 == 0)
    mode = 3;
  else if (labels != nullptr || (p.y_mode != 2 && !direct_fast))
    mode = 2;
  const bool has_res = p.res_boxes > 0;
#define RTSEG_TC(KB_)                                                                                          \
  if (direct_fast && mode <= 1)                                                                                \
    return mode == 0 ? launch_tc_inst<KB_, 0, false, true>(maps, p, pl, ep, res, y, labels, st)                \
                     : launch_tc_inst<KB_, 1, false, true>(maps, p, pl, ep, res, y, labels, st);               \
  return mode == 3   ? launch_tc_inst<KB_, 3, false>(maps, p, pl, ep, res, y, labels, st)                      \
         : mode == 2 ? launch_tc_inst<KB_, 2, false>(maps, p, pl, ep, res, y, labels, st)                      \
         : mode == 0 ? (has_res ? launch_tc_inst<KB_, 0, true>(maps, p, pl, ep, res, y, labels, st)            \
                                : launch_tc_inst<KB_, 0, false>(maps, p, pl, ep, res, y, labels, st))          \
                     : (has_res ? launch_tc_inst<KB_, 1, true>(maps, p, pl, ep, res, y, labels, st)            \
                                : launch_tc_inst<KB_, 1, false>(maps, p, pl, ep, res, y, labels, st))
  switch (pl.kb) {
    case 64: RTSEG_TC(64);
    case 32: RTSEG_TC(32);
    default: RTSEG_TC(16);
  }
#undef RTSEG_TC
}

static void pick_tile(int gw, int &tw, int &th) {
  tw = 128;
  while (tw > 8 && tw / 2 >= gw) tw /= 2;  // smallest power of two >= gw, capped at 128, floor 8
  th = kTileM / tw;
}

// Epilogue mode: 0 = clamp (none / relu / relu6 / hardtanh), 1 = slope form (prelu with a slope
// vector), 2 = general.
static int epilogue_mode(const rtseg_conv_desc *d, const float *slope, TcParams &p) {
  return rtseg::epilogue_mode(d->act, d->act2, d->res_mode, slope, d->act_p0, d->act_p1, p.lo, p.hi, p.lo2, p.hi2);
}

// Largest sub-tile stack that still leaves every SM at least two unit tiles.
static int max_stack(int n, int gh, int gw, int share = 0) {
  int tw, th;
  pick_tile(gw, tw, th);
  const bool wide = share == 1 || share == 3;  // sub-tiles stacked along w
  if (wide) tw = 8, th = 16;
  if (share == 2) tw = kTileM, th = 1;
  int ms = 8;
  auto units = [&](int m) {
    return wide ? (int64_t)n * cdiv(gw, tw * m) * cdiv(gh, th) : (int64_t)n * cdiv(gw, tw) * cdiv(gh, th * m);
  };
  while (ms > 1 && ((wide ? tw * ms > gw : th * ms > gh) || units(ms) < 2 * (int64_t)num_sms())) ms >>= 1;
  return ms;
}

// Which A-sharing mode a forward conv can use (0 = none).  RTSEG_TC_SHARE=0 disables, =1 vertical only.
static int pick_share(const rtseg_conv_desc *d, int gh, int gw, int kb) {
  static const int knob = getenv("RTSEG_TC_SHARE") ? atoi(getenv("RTSEG_TC_SHARE")) : 15;  // bit 3: transposed conv
  if (d->stride != 1) return 0;
  if ((knob & 1) && d->kw == 1 && d->kh >= 2 && (d->kh - 1) * d->dil < (d->kh - 1) * 16 && d->pad_w == 0 && gw >= 8 &&
      gh >= 16)
    return 1;
  if ((knob & 4) && d->kh == d->kw && d->kh >= 2 && (d->kh - 1) * d->dil <= 8 && gw >= 8 && gh >= 16) return 3;
  static const int h_any_kb = getenv("RTSEG_TC_HANYKB") ? atoi(getenv("RTSEG_TC_HANYKB")) : 1;
  if ((knob & 2) && d->kh == 1 && d->kw >= 2 && (kb == 64 || h_any_kb) && (d->kw - 1) * d->dil <= 64 && d->pad_h == 0 &&
      gw > 64)
    return 2;
  return 0;
}

static void fill_common(TcParams &p, const rtseg_conv_desc *d, const TcPlan &pl, int cout_pad, int share = 0) {
  p.cin = d->cin, p.cout = d->cout, p.cout_pad = cout_pad;
  p.bn = pl.bn, p.nkb = d->cin / pl.kb, p.stages = pl.stages, p.ob = pl.ob;
  p.b_block_bytes = pl.b_block_bytes, p.b_bytes = pl.b_bytes, p.out_buf_bytes = pl.out_buf_bytes;
  p.out_bufs = pl.out_bufs;
  p.split_out = pl.split_out;
  p.tmem_cols = pl.tmem_cols;
  p.x_coff = d->x_coff;
  p.y_ct = d->y_ctotal, p.y_co = d->y_coff, p.r_ct = d->r_ctotal, p.r_co = d->r_coff;
  p.ms = pl.ms;
  const int halo = (share == 1 || share == 3) ? (d->kh - 1) * d->dil : (share == 2 ? (d->kw - 1) * d->dil : 0);
  if (share == 1) {  // vertical taps: 8 wide x 16 tall sub-tiles, stacked along w, one box holds all taps' rows
    p.tw = 8, p.th = 16, p.wstack = 1;
    p.unit_w = 8 * p.ms, p.unit_h = 16;
    p.a_share = 1, p.a_dy = -d->pad_h, p.a_dx = 0, p.a_boff = 0;
    p.box_w = p.unit_w, p.box_h = 16 + halo;
    p.a_sub16 = (8u * pl.kb * 2) >> 4;
    p.a_sbo16 = ((uint32_t)p.unit_w * pl.kb * 2) >> 4;
    for (int t = 0; t < d->kh; ++t) p.tap_off16[t] = ((uint32_t)(t * d->dil) * p.unit_w * pl.kb * 2) >> 4;
  } else if (share == 3) {
    // k x k taps from ONE box per K block: (16 + halo) rows of (8*ms + halo) pixels.  An 8-pixel core-matrix group of
    // tap (ky, kx) starts (ky*dil) box rows and (kx*dil) pixels into the box, and consecutive groups are one box row
    // apart (the descriptor's stride-byte-offset).  Neither is a multiple of the swizzle atom: as in the horizontal
    // mode this relies on the MMA taking the swizzle phase from the shared-memory address itself.
    p.tw = 8, p.th = 16, p.wstack = 1;
    p.unit_w = 8 * p.ms, p.unit_h = 16;
    p.a_share = 1, p.a_dy = -d->pad_h, p.a_dx = -d->pad_w, p.a_boff = 0;
    p.box_w = p.unit_w + halo, p.box_h = 16 + halo;
    p.a_sub16 = (8u * pl.kb * 2) >> 4;
    p.a_sbo16 = ((uint32_t)p.box_w * pl.kb * 2) >> 4;
    for (int ky = 0, t = 0; ky < d->kh; ++ky)
      for (int kx = 0; kx < d->kw; ++kx, ++t)
        p.tap_off16[t] = ((uint32_t)(ky * d->dil * p.box_w + kx * d->dil) * pl.kb * 2) >> 4;
  } else if (share == 2) {  // horizontal taps: 128-pixel rows (+halo), stacked along h, taps = row-shifted descriptors
    p.tw = kTileM, p.th = 1, p.wstack = 0;
    p.unit_w = kTileM, p.unit_h = p.ms;
    // Measured on B200: with the base-offset field left at 0 a descriptor whose start address is shifted by
    // d rows (d*128 bytes, not a multiple of the 1024-byte swizzle atom) reads exactly the rows TMA wrote —
    // the MMA derives the 128B-swizzle phase from the absolute shared-memory address bits [7,10).  Setting
    // base_offset = (addr >> 7) & 7 instead produces wrong results (RTSEG_TC_BOFF=1 reproduces that).
    static const int boff_knob = getenv("RTSEG_TC_BOFF") ? atoi(getenv("RTSEG_TC_BOFF")) : 0;
    p.a_share = 1, p.a_dy = 0, p.a_dx = -d->pad_w, p.a_boff = boff_knob;
    p.box_w = kTileM + halo, p.box_h = p.ms;
    p.a_sub16 = ((uint32_t)(kTileM + halo) * pl.kb * 2) >> 4;
    p.a_sbo16 = (8u * pl.kb * 2) >> 4;
    for (int t = 0; t < d->kw; ++t) p.tap_off16[t] = ((uint32_t)(t * d->dil) * pl.kb * 2) >> 4;
  } else {
    pick_tile(p.gw, p.tw, p.th);
    p.wstack = 0;
    p.unit_w = p.tw, p.unit_h = p.th * p.ms;
    p.a_share = 0, p.a_dy = 0, p.a_dx = 0, p.a_boff = 0;
    p.box_w = p.tw, p.box_h = p.unit_h;
    p.a_sub16 = ((uint32_t)kTileM * pl.kb * 2) >> 4;
    p.a_sbo16 = (8u * pl.kb * 2) >> 4;
  }
  p.stage_bytes = pl.stage_bytes;
  p.a_stages_per_unit = 0;  // set by the caller
  p.res_boxes = 0;
  p.a_tx_bytes = (uint32_t)p.box_w * p.box_h * pl.kb * 2;
  p.tiles_w = (int)cdiv(p.gw, p.unit_w), p.tiles_h = (int)cdiv(p.gh, p.unit_h);
  p.num_m_tiles = (int64_t)p.n * p.tiles_w * p.tiles_h;
}

// TMA stores need 16-byte global strides: the per-pixel channel count must be a multiple of 8.
static bool can_stage_output(const rtseg_conv_desc *d, const void *y) {
  return y != nullptr && d->y_ctotal % 8 == 0 && d->y_coff % 8 == 0 && reinterpret_cast<uintptr_t>(y) % 16 == 0;
}

// ======================================================================= weight gradient on tcgen05 ====
// dW[tap][ci][co] = sum over pixels of x[pixel + tap][ci] * dy[pixel][co] for the stride-1 dense convs with
// cin = cout in {64, 128} (60 of ERFNet's 74 convolutions).  The reduction dimension K is the PIXEL index, so both
// operands are used exactly as TMA lands them for the forward conv — rows of 64 channels (128 bytes), one row per
// pixel — but read as MN-major UMMA operands (instruction-descriptor bits 15/16): M = input channel, N = output
// channel, K = pixel, 16 pixels (two 8-row swizzle atoms, SBO = 1024 bytes) per tcgen05.mma.  A 128-row M is two
// 64-channel boxes LBO bytes apart: the two K blocks of a 128-channel input, or — for 64 channels — the boxes of TWO
// taps, so that every MMA runs at M = 128.  The accumulators (one per tap or tap pair, fp32 in TMEM) live for the
// whole kernel: a CTA streams its share of the pixel tiles through a TMA ring, never leaves the MMA loop, and adds
// its TMEM contents to the caller-zeroed fp32 gradient once at the end (vector red.global).
struct WgParams {
  int units, tiles_w, tiles_h, tw, th;
  int ngroups;               // work items (MMA groups) per pixel tile, <= 8
  int b_boxes;               // dy boxes per item
  int n;                     // MMA N = cout
  int cin, cout;
  int stages, stage_bytes;
  int tmem_cols;
  // group g = up to 128 / bs_a boxes of x, one per (tap, input-channel block); box b = {tensor map, coordinates}
  int a_n[8];
  short a_map[8][8], a_c[8][8], a_dx[8][8], a_dy[8][8];
  short tap_of_blk[8][8];    // which tap block b (bs_a accumulator rows) of group g holds (-1: ignore)
  short ci_of_blk[8][8];     // first input channel of that block
  int a_lbo16[8];            // 0 = a single box, read twice
  int g_c0;                  // first channel of the dy slice
  int bs_a, bs_b;            // channels per box: 64 (SWIZZLE_128B) or 16 (SWIZZLE_32B), per operand
  int box_a, box_b, a_region;
  int kstep_a, sbo_a, layout_a, kstep_b, sbo_b, layout_b;
};

struct WgMaps {
  CUtensorMap x[4];          // input (parity views for stride 2)
  CUtensorMap g;
};

constexpr int kWgThreads = 192;  // warp 0 TMA, warp 1 MMA, warps 2-5 final read-out

__device__ __forceinline__ uint64_t wg_desc(uint32_t addr, uint32_t lbo16, uint32_t sbo16, uint32_t layout) {
  // MN-major: leading-byte-offset = distance of the next block (64 channels with SWIZZLE_128B, 16 with SWIZZLE_32B)
  // along M/N, stride-byte-offset = distance of the next 8-row group along K (rows are contiguous: 8 x row bytes)
  return (uint64_t)((addr & 0x3ffffu) >> 4) | ((uint64_t)(lbo16 & 0x3fffu) << 16) | ((uint64_t)sbo16 << 32) | (1ull << 46) |
         ((uint64_t)layout << 61);
}

__global__ void __launch_bounds__(kWgThreads, 1) conv_wgrad_tc_kernel(const __grid_constant__ WgMaps maps, const WgParams p,
                                                                      float *__restrict__ dw) {
  extern __shared__ __align__(1024) uint8_t wg_smem_raw[];
  const uint32_t s_base = (smem_u32(wg_smem_raw) + 1023u) & ~1023u;
  __shared__ __align__(8) uint64_t s_bars[2 * kMaxStages + 1];
  __shared__ uint32_t s_tmem, s_dead_word;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t full_bar = smem_u32(s_bars), empty_bar = full_bar + 8 * kMaxStages, done_bar = empty_bar + 8 * kMaxStages;
  volatile uint32_t *dead = &s_dead_word;
  if (threadIdx.x == 0) {
    s_dead_word = 0;
    for (int s = 0; s < p.stages; ++s) mbar_init(full_bar + 8 * s, 1), mbar_init(empty_bar + 8 * s, 1);
    mbar_init(done_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(p.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = s_tmem;
  const uint32_t box_a = (uint32_t)p.box_a, box_b = (uint32_t)p.box_b;  // 128 pixels x bs channels x 2 bytes
  const uint32_t b_off = (uint32_t)p.a_region;

  if (warp == 0) {
    if (lane == 0) {
      prefetch_tmap(&maps.x[0]);
      prefetch_tmap(&maps.g);
      uint32_t stage = 0, phase = 0;
      for (int u = blockIdx.x; u < p.units; u += gridDim.x) {
        const int twi = u % p.tiles_w, r = u / p.tiles_w, thi = r % p.tiles_h, img = r / p.tiles_h;
        const int w0 = twi * p.tw, h0 = thi * p.th;
        for (int g = 0; g < p.ngroups; ++g) {
          mbar_wait(empty_bar + 8 * stage, phase ^ 1, dead);
          const uint32_t dst = s_base + stage * (uint32_t)p.stage_bytes;
          const int na = p.a_n[g];
          mbar_expect_tx(full_bar + 8 * stage, (uint32_t)na * box_a + (uint32_t)p.b_boxes * box_b);
          for (int b = 0; b < na; ++b)
            tma_load_4d(dst + b * box_a, &maps.x[p.a_map[g][b]], full_bar + 8 * stage, p.a_c[g][b], w0 + p.a_dx[g][b],
                        h0 + p.a_dy[g][b], img);
          for (int b = 0; b < p.b_boxes; ++b)
            tma_load_4d(dst + b_off + b * box_b, &maps.g, full_bar + 8 * stage, p.g_c0 + b * p.bs_b, w0, h0, img);
          if (++stage == (uint32_t)p.stages) stage = 0, phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // kind::f16, D = f32, A = B = bf16, both MN-major (bits 15, 16), N at bit 17, M = 128 at bit 24
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(p.n >> 3) << 17) |
                           ((uint32_t)(kTileM >> 4) << 24);
    const uint32_t b_lbo16 = box_b >> 4;
    uint32_t stage = 0, phase = 0;
    bool first = true;
    for (int u = blockIdx.x; u < p.units; u += gridDim.x) {
      for (int g = 0; g < p.ngroups; ++g) {
        mbar_wait(full_bar + 8 * stage, phase, dead);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t a0 = s_base + stage * (uint32_t)p.stage_bytes;
          const uint64_t a_d = wg_desc(a0, (uint32_t)p.a_lbo16[g], (uint32_t)p.sbo_a, (uint32_t)p.layout_a);
          const uint64_t b_d = wg_desc(a0 + b_off, b_lbo16, (uint32_t)p.sbo_b, (uint32_t)p.layout_b);
          const uint32_t d_col = tmem_base + (uint32_t)(g * p.n);
          const uint64_t ka = (uint64_t)p.kstep_a, kb = (uint64_t)p.kstep_b;
#pragma unroll
          for (int k = 0; k < kTileM / 16; ++k)   // 16 pixels per step
            umma_bf16(d_col, a_d + ka * k, b_d + kb * k, idesc, (first && k == 0) ? 0u : 1u);
          umma_commit(empty_bar + 8 * stage);
        }
        __syncwarp();
        if (++stage == (uint32_t)p.stages) stage = 0, phase ^= 1;
      }
      first = false;
    }
    if (elect_one()) umma_commit(done_bar);
    __syncwarp();
  } else {
    // ---- final read-out: lane quarter q of TMEM belongs to warp (q + 2) % 4 ... any warp w reads quarter w % 4
    const bool any_work = (int)blockIdx.x < p.units;
    if (any_work) {
      mbar_wait(done_bar, 0, dead);
      tc_fence_after();
      const int q = warp & 3;
      const int row = q * 32 + lane, blk = row / p.bs_a, rin = row % p.bs_a;
      for (int g = 0; g < p.ngroups; ++g) {
        const int tap = p.tap_of_blk[g][blk];
        const int ci = p.ci_of_blk[g][blk] + rin;
        float *dst = dw + ((int64_t)(tap < 0 ? 0 : tap) * p.cin + ci) * p.cout;
        for (int c0 = 0; c0 < p.n; c0 += 16) {
          uint32_t v[16];
          tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(g * p.n + c0), v);
          tmem_ld_wait();
          if (tap >= 0) {
#pragma unroll
            for (int j = 0; j < 16; j += 4)
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + c0 + j), "f"(__uint_as_float(v[j])),
                           "f"(__uint_as_float(v[j + 1])), "f"(__uint_as_float(v[j + 2])), "f"(__uint_as_float(v[j + 3]))
                           : "memory");
          }
        }
      }
      tc_fence_before();
    }
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
  }
}

}  // namespace rtseg

using namespace rtseg;

#ifdef RTSEG_TC_TRACE
extern "C" int rtseg_tc_trace_read(long long *out_host) {
  return cudaMemcpyFromSymbol(out_host, g_tc_trace, sizeof(long long) * 64 * 16) == cudaSuccess ? 0 : -3;
}
#endif

extern "C" int rtseg_conv_tc_supported(const rtseg_conv_desc *d) {
  if (!d) return 0;
  const char *why;
  if (check_tc_common(d, false, &why)) return 0;
  TcPlan pl;
  return plan_tc((d->cin + 15) / 16 * 16, (d->cout + 15) / 16 * 16, d->kh * d->kw, 1, true, 1,
                 d->res_mode != RTSEG_RES_NONE, pl) ? 1 : 0;
}

// Reads and clears the device-side watchdog flag (synchronises the stream). Used by tests.
extern "C" int rtseg_tc_timeout_flag(rtseg_stream_t stream) {
  if (cudaStreamSynchronize(as_stream(stream)) != cudaSuccess) return -1;
  const int a = read_clear_timeout_flag(), b = dw_tma_timeout_flag(), c = stem_tc_timeout_flag();
  return (a < 0 || b < 0 || c < 0) ? -1 : (a | b | c);
}

// Invariants of a plan the kernel relies on but cannot check cheaply on the device: the shared-memory carve-up fits
// the launch's allocation, every TMA box fits its ring stage and the tensor-map limits, every shifted UMMA descriptor
// stays inside the bytes its TMA box delivered, tensor memory holds both accumulator stages, and one unit's residual
// boxes fit in the ring (more residual boxes than stages would wait on their own slots).
static int check_plan(const TcParams &p, const TcPlan &pl) {
  const uint32_t row_b = (uint32_t)pl.kb * 2;
  const uint32_t carve = p.b_bytes + (uint32_t)p.stages * p.stage_bytes + (uint32_t)p.out_bufs * p.out_buf_bytes +
                         3u * (uint32_t)p.bn * 4 + (2 * kMaxStages + 5) * 8 + 8;
  RTSEG_REQUIRE(carve <= pl.smem_bytes && pl.smem_bytes + 1024 <= kSmemBudget + 2048, "conv_tc plan: shared-memory carve-up exceeds the allocation");
  RTSEG_REQUIRE(p.stages >= 2 && p.stages <= kMaxStages, "conv_tc plan: ring depth out of range");
  RTSEG_REQUIRE(p.box_w >= 1 && p.box_w <= 256 && p.box_h >= 1 && p.box_h <= 256, "conv_tc plan: A box exceeds the TMA box limit");
  RTSEG_REQUIRE(p.unit_w <= 256 && p.unit_h <= 256, "conv_tc plan: output box exceeds the TMA box limit");
  RTSEG_REQUIRE(p.a_tx_bytes == (uint32_t)p.box_w * p.box_h * row_b && p.a_tx_bytes <= p.stage_bytes,
                "conv_tc plan: A box does not fit its ring stage");
  RTSEG_REQUIRE(p.stage_bytes % 1024 == 0 && p.b_block_bytes % 1024 == 0, "conv_tc plan: stage / weight block not swizzle-atom aligned");
  RTSEG_REQUIRE((uint32_t)p.unit_w * p.unit_h == (uint32_t)p.ms * kTileM,
                "conv_tc plan: unit tile is not ms x 128 pixels");
  if (p.a_share) {
    uint32_t last = 0;
    const int ntap = p.ntaps;
    for (int t = 0; t < ntap; ++t) last = std::max(last, p.tap_off16[t]);
    const uint32_t end = (last + (uint32_t)(p.ms - 1) * p.a_sub16 + 15u * p.a_sbo16) * 16u + 8u * row_b;
    RTSEG_REQUIRE(end <= p.a_tx_bytes, "conv_tc plan: a shifted tap descriptor reads past the halo box");
  } else {
    RTSEG_REQUIRE(((uint32_t)(p.ms - 1) * p.a_sub16 + 15u * p.a_sbo16) * 16u + 8u * row_b <= p.a_tx_bytes,
                  "conv_tc plan: a sub-tile descriptor reads past the A box");
  }
  RTSEG_REQUIRE((uint32_t)(2 * p.ms * p.nacc * p.bn) <= p.tmem_cols && p.tmem_cols <= 512, "conv_tc plan: accumulators exceed tensor memory");
  if (p.res_boxes) {
    RTSEG_REQUIRE(p.res_boxes <= 4 && p.res_boxes <= p.stages, "conv_tc plan: residual boxes of one unit do not fit in the ring");
    RTSEG_REQUIRE((uint32_t)p.ms * kTileM * (uint32_t)p.ob * 2 <= p.stage_bytes, "conv_tc plan: residual box does not fit its ring stage");
  }
  if (p.y_mode == 2)
    RTSEG_REQUIRE((uint32_t)p.ms * kTileM * (uint32_t)(p.split_out ? p.ob : p.bn) * 2 == p.out_buf_bytes &&
                      (!p.split_out || (p.ms == 1 && p.out_bufs == 2 && p.nacc == 1)),
                  "conv_tc plan: staging buffer size mismatch");
  RTSEG_REQUIRE(p.ms * (p.bn / 16) <= 2 * kMaxMyChunks, "conv_tc plan: more 16-column chunks than the epilogue warps own");
  return RTSEG_OK;
}

// Plans a forward conv: tile shape, A-sharing mode, stack depth, ring depth, epilogue mode.  `d` is the K-padded
// descriptor.  Shared by rtseg_conv_tc and rtseg_conv_tc_plan_info so that tests can see the plan a launch would use.
static int plan_forward(const rtseg_conv_desc *d, bool staged, const float *slope, TcPlan &pl, TcParams &p, int &share,
                        int &mode, bool may_split = true, bool may_unstage = true) {
  const int cout_pad = (d->cout + 15) / 16 * 16;
  const int ntaps = d->kh * d->kw;
  p.n = d->n;
  rtseg_out_hw(d, &p.gh, &p.gw);
  share = pick_share(d, p.gh, p.gw, pick_kb(d->cin));
  const int halo = (share == 1 || share == 3) ? (d->kh - 1) * d->dil : (share == 2 ? (d->kw - 1) * d->dil : 0);
  // split staging lives in the fast epilogue (clamp / slope activations); the general path stages whole units
  // (a residual that cannot ride the ring — more than four sub-boxes — also takes the general path)
  const int nbox_max = cout_pad / 16;
  const bool fast_epi = may_split && epilogue_mode(d, slope, p) <= 1 &&
                        (d->res_mode == RTSEG_RES_NONE || nbox_max <= 4 || cout_pad % 64 == 0);
  if (!plan_tc(d->cin, cout_pad, ntaps, 1, staged, max_stack(p.n, p.gh, p.gw, share), d->res_mode != RTSEG_RES_NONE, pl,
               share, halo, fast_epi)) {
    share = 0;
    if (!plan_tc(d->cin, cout_pad, ntaps, 1, staged, max_stack(p.n, p.gh, p.gw), d->res_mode != RTSEG_RES_NONE, pl, 0, 0,
                 fast_epi))
      RTSEG_FAIL(RTSEG_ERR_UNSUPPORTED, "rtseg_conv_tc: weights (%d taps x %d x %d) do not fit in shared memory", ntaps,
                 d->cin, cout_pad);
  }
  // Dense k x k layers whose resident weights leave no room for the output staging buffers fall back to narrow N
  // tiles (128 -> 128 3x3: N = 32, every input box fetched four times, MMAs a quarter wide).  Without staging (the
  // epilogue stores its rows straight to global memory, 2 * N contiguous bytes per thread) the same layer runs at
  // N = 64: 0.40 -> 0.22 ms at 16 x 128 x 256 (BiSeNetv2 7.78 -> 6.94 ms, DDRNet 4.85 -> 4.60 ms per step).  Only taken
  // when it widens the N tile and no residual rides the ring.
  static const int unstage_knob = getenv("RTSEG_TC_UNSTAGE") ? atoi(getenv("RTSEG_TC_UNSTAGE")) : 1;
  if (unstage_knob && may_unstage && staged && pl.n_tiles > 1 && ntaps > 3 && d->res_mode == RTSEG_RES_NONE &&
      d->cout % 16 == 0) {   // (whole 16-channel chunks: the fast epilogue's direct stores are 16-byte vectors)
    TcPlan alt;
    bool ok = plan_tc(d->cin, cout_pad, ntaps, 1, false, max_stack(p.n, p.gh, p.gw, share), false, alt, share, halo, false);
    if (!ok) ok = plan_tc(d->cin, cout_pad, ntaps, 1, false, max_stack(p.n, p.gh, p.gw), false, alt, 0, 0, false);
    if (ok && alt.bn > pl.bn) return plan_forward(d, false, slope, pl, p, share, mode, may_split, false);
  }
  p.oh = p.gh, p.ow = p.gw, p.ostride = 1, p.nacc = 1;
  for (int a = 0; a < kMaxAcc; ++a) p.oy0[a] = p.ox0[a] = 0;
  p.ntaps = ntaps;
  p.y_mode = staged ? 2 : 1;
  fill_common(p, d, pl, cout_pad, share);
  mode = epilogue_mode(d, slope, p);
  p.a_stages_per_unit = share ? p.nkb : ntaps * p.nkb;
  const bool has_res = d->res_mode != RTSEG_RES_NONE;
  const bool ring_res = has_res && staged && mode <= 1 && pl.bn / pl.ob <= 4;
  if (has_res && !ring_res) mode = 2;  // general path reads the residual with plain loads
  if (ring_res) p.res_boxes = pl.bn / pl.ob;
  if (pl.split_out && mode == 2)   // the general epilogue stages whole units: plan again without split staging
    return plan_forward(d, staged, slope, pl, p, share, mode, false);
  return RTSEG_OK;
}

// A vertically dilated k x 1 convolution (same padding, stride 1) with dilation d = f * d' is the SAME convolution
// with dilation d' on the same memory read as h / f rows of f * w pixels: row r of that view is the real rows
// r*f .. r*f + f-1 side by side, so the real row y +- d is the view row r +- d' at the same column, and the zero rows
// above / below the image stay where they are.  The vertical halo of a 16-row tile then costs 2 d' view rows instead
// of 2 d real rows (ERFNet's 3x1 d = 4 / 8 / 16 layers, reference models/erfnet.py:69).  Input, output and residual
// share the geometry, so all three are reinterpreted.  f = the largest divisor of d that leaves >= 16 view rows.
static void fold_vertical_dilation(rtseg_conv_desc &d) {
  static const int knob = getenv("RTSEG_TC_FOLD_DIL") ? atoi(getenv("RTSEG_TC_FOLD_DIL")) : 1;
  if (!knob || d.kw != 1 || d.kh < 2 || d.dil < 2 || d.stride != 1) return;
  if (d.pad_w != 0 || d.pad_h != d.dil * (d.kh - 1) / 2) return;
  int f = 1;
  for (int c = 2; c <= d.dil; ++c)
    if (d.dil % c == 0 && d.h % c == 0 && d.h / c >= 16 && (int64_t)d.w * c <= (1 << 24)) f = c;
  if (f == 1) return;
  d.h /= f, d.w *= f, d.dil /= f, d.pad_h = d.dil * (d.kh - 1) / 2;
}

// Test / diagnostics hook: the execution plan rtseg_conv_tc would use for `d` with a 16-byte-aligned output.
// info = {kb, bn, n_tiles, ms, stages, out_bufs, share, units, grid_x, smem_bytes, tmem_cols, res_boxes}.
extern "C" int rtseg_conv_tc_plan_info(const rtseg_conv_desc *d, int32_t info[12]) {
  RTSEG_REQUIRE(d && info, "NULL argument");
  const char *why;
  if (check_tc_common(d, false, &why)) RTSEG_FAIL(RTSEG_ERR_UNSUPPORTED, "rtseg_conv_tc: %s", why);
  rtseg_conv_desc padded = *d;
  padded.cin = (d->cin + 15) / 16 * 16;
  fold_vertical_dilation(padded);
  TcPlan pl;
  TcParams p;
  int share = 0, mode = 0;
  const bool staged = d->y_ctotal % 8 == 0 && d->y_coff % 8 == 0;
  if (int rc = plan_forward(&padded, staged, nullptr, pl, p, share, mode)) return rc;
  const int gx = (int)std::min<int64_t>(p.num_m_tiles, std::max(1, num_sms() / pl.n_tiles));
  const int32_t v[12] = {pl.kb, pl.bn, pl.n_tiles, pl.ms, pl.stages, pl.out_bufs, share, (int32_t)p.num_m_tiles, gx,
                         (int32_t)pl.smem_bytes, (int32_t)pl.tmem_cols, p.res_boxes};
  for (int i = 0; i < 12; ++i) info[i] = v[i];
  return RTSEG_OK;
}

extern "C" int rtseg_conv_tc(const rtseg_conv_desc *d, const void *x, const void *w_bf16, const float *scale,
                             const float *shift, const float *slope, const void *res, void *y,
                             rtseg_stream_t stream) {
  if (int rc = validate_conv_desc(d, res)) return rc;
  RTSEG_REQUIRE(x && w_bf16 && y, "NULL tensor pointer");
  const char *why;
  if (check_tc_common(d, false, &why)) RTSEG_FAIL(RTSEG_ERR_UNSUPPORTED, "rtseg_conv_tc: %s", why);
  RTSEG_REQUIRE(reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(w_bf16) % 16 == 0 &&
                    reinterpret_cast<uintptr_t>(res) % 16 == 0,
                "tensor pointers must be 16-byte aligned");
  // K is padded to a multiple of 16 with zero weights; the input tensor map ends at the last REAL channel, so the
  // TMA unit zero-fills the tail of the last K block (no padded copy of the input, nothing read past the slice).
  const int cin_real = d->cin;
  rtseg_conv_desc padded = *d;
  padded.cin = (d->cin + 15) / 16 * 16;
  fold_vertical_dilation(padded);
  d = &padded;
  const int cout_pad = (d->cout + 15) / 16 * 16;
  const int ntaps = d->kh * d->kw;
  const bool staged = can_stage_output(d, y);
  TcPlan pl;
  TcParams p;
  int share = 0, mode = 0;
  if (int rc = plan_forward(d, staged, slope, pl, p, share, mode)) return rc;
  const bool ring_res = p.res_boxes > 0;
  if (int rc = check_plan(p, pl)) return rc;

  TcMaps maps;
  const int64_t ct = d->x_ctotal;
  auto set_tap = [&](int t, int map, int dy, int dx) {
    TcTap &tp = p.taps[t];
    tp.map = (int16_t)map, tp.dy = (int16_t)dy, tp.dx = (int16_t)dx, tp.nops = 1;
    tp.widx[0] = (int16_t)t, tp.acc[0] = 0;
  };
  if (d->stride == 1) {
    if (int rc = encode_nhwc_map(&maps.a[0], x, d->x_coff + cin_real, d->w, d->h, d->n, ct, ct * d->w, ct * d->w * d->h, pl.kb,
                                 p.box_w, p.box_h))
      return rc;
    for (int i = 1; i < 4; ++i) maps.a[i] = maps.a[0];
    for (int ky = 0, t = 0; ky < d->kh; ++ky)
      for (int kx = 0; kx < d->kw; ++kx, ++t) set_tap(t, 0, ky * d->dil - d->pad_h, kx * d->dil - d->pad_w);
  } else {
    // stride 2: input (h, w) = (2*h2 + ph, 2*w2 + pw); one tensor map per parity (ph, pw)
    for (int ph = 0; ph < 2; ++ph)
      for (int pw = 0; pw < 2; ++pw) {
        const __nv_bfloat16 *base = static_cast<const __nv_bfloat16 *>(x) + ((int64_t)ph * d->w + pw) * ct;
        if (int rc = encode_nhwc_map(&maps.a[ph * 2 + pw], base, d->x_coff + cin_real, d->w / 2, d->h / 2, d->n, 2 * ct,
                                     2 * ct * d->w, ct * d->w * d->h, pl.kb, p.box_w, p.box_h))
          return rc;
      }
    for (int ky = 0, t = 0; ky < d->kh; ++ky)
      for (int kx = 0; kx < d->kw; ++kx, ++t) {
        const int oy = ky * d->dil - d->pad_h, ox = kx * d->dil - d->pad_w;  // input = 2*out + o
        const int ph = ((oy % 2) + 2) % 2, pw = ((ox % 2) + 2) % 2;
        set_tap(t, ph * 2 + pw, (oy - ph) / 2, (ox - pw) / 2);
      }
  }
  if (int rc = encode_weight_map(&maps.b, w_bf16, d->cin, ntaps * cout_pad, pl.kb, pl.bn)) return rc;
  if (staged) {
    const int64_t yc = d->y_ctotal;
    if (int rc = encode_nhwc_map(&maps.c[0], y, d->y_coff + d->cout, p.ow, p.oh, d->n, yc, yc * p.ow, yc * p.ow * p.oh,
                                 pl.ob, p.unit_w, p.unit_h))
      return rc;
  } else {
    maps.c[0] = maps.a[0];
  }
  for (int a = 1; a < kMaxAcc; ++a) maps.c[a] = maps.c[0];
  if (ring_res) {
    const int64_t rc = d->r_ctotal;
    if (int e = encode_nhwc_map(&maps.r, res, d->r_coff + d->cout, p.ow, p.oh, d->n, rc, rc * p.ow, rc * p.ow * p.oh,
                                pl.ob, p.unit_w, p.unit_h))
      return e;
  } else {
    maps.r = maps.a[0];
  }
  const Epilogue ep = make_epilogue(d, scale, shift, slope);
  return launch_tc(maps, p, pl, ep, mode, res, y, nullptr, as_stream(stream));
}

extern "C" int rtseg_deconv3x3s2_tc_supported(const rtseg_conv_desc *d) {
  if (!d) return 0;
  const char *why;
  if (check_tc_common(d, true, &why)) return 0;
  if (d->kh != 3 || d->kw != 3 || d->stride != 2 || d->dil != 1) return 0;
  TcPlan pl;
  return plan_tc(d->cin, (d->cout + 15) / 16 * 16, 9, 4, true, 1, false, pl) ? 1 : 0;
}

// Transposed conv k=3, s=2, p=1, output_padding=1: oy = 2*iy - 1 + ky.  With A(dy,dx) the input tile
// shifted by (dy,dx) and W[ky][kx] the weight taps, the four output parities of input pixel (i,j) are
//   (2i  ,2j  ) = A00*W11
//   (2i  ,2j+1) = A01*W10 + A00*W12
//   (2i+1,2j  ) = A10*W01 + A00*W21
//   (2i+1,2j+1) = A11*W00 + A10*W02 + A01*W20 + A00*W22
// so one launch loads each of the four shifted tiles once and runs nine MMAs into four accumulators.
extern "C" int rtseg_deconv3x3s2_tc(const rtseg_conv_desc *d, const void *x, const void *w_bf16, const float *scale,
                                    const float *shift, const float *slope, void *y, uint8_t *labels,
                                    rtseg_stream_t stream) {
  RTSEG_REQUIRE(d && x && w_bf16 && (y || labels), "NULL argument");
  RTSEG_REQUIRE(d->kh == 3 && d->kw == 3 && d->stride == 2 && d->dil == 1 && d->groups == 1,
                "rtseg_deconv3x3s2_tc handles k=3, s=2, dense only");
  RTSEG_REQUIRE(d->res_mode == RTSEG_RES_NONE, "deconv has no residual input");
  RTSEG_REQUIRE(d->n > 0 && d->h > 0 && d->w > 0 && d->cin > 0 && d->cout > 0, "non-positive extent");
  const char *why;
  if (check_tc_common(d, true, &why)) RTSEG_FAIL(RTSEG_ERR_UNSUPPORTED, "rtseg_deconv3x3s2_tc: %s", why);
  RTSEG_REQUIRE(reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(w_bf16) % 16 == 0,
                "tensor pointers must be 16-byte aligned");
  RTSEG_REQUIRE(labels == nullptr || d->cout <= 128, "fused argmax supports at most 128 classes");
  const int cout_pad = (d->cout + 15) / 16 * 16;
  const bool staged = can_stage_output(d, y);
  TcPlan pl;
  // One input tile with a one-pixel halo to the right and below (17 rows of 8*ms + 1 pixels) feeds all four shifted
  // operands (x, x+1, y+1, both): a quarter of the shared-memory fill traffic of four separate boxes.
  static const int share_knob = getenv("RTSEG_TC_SHARE") ? atoi(getenv("RTSEG_TC_SHARE")) : 15;
  bool share = (share_knob & 8) && d->w >= 8 && d->h >= 16;
  int stack = (labels && !y) ? 1 : max_stack(d->n, d->h, d->w, share ? 3 : 0);  // labels-only: small TMEM footprint, 2 CTAs/SM
  if (share && !(plan_tc(d->cin, cout_pad, 9, 4, staged, stack, false, pl, 3, 1) && (!labels || pl.n_tiles == 1))) share = false;
  if (!share) {
    stack = (labels && !y) ? 1 : max_stack(d->n, d->h, d->w);
    if (!plan_tc(d->cin, cout_pad, 9, 4, staged, stack, false, pl) || (labels && pl.n_tiles != 1))
      RTSEG_FAIL(RTSEG_ERR_UNSUPPORTED, "rtseg_deconv3x3s2_tc: weights do not fit in shared memory");
  }
  TcMaps maps;
  TcParams p;
  p.n = d->n, p.gh = d->h, p.gw = d->w;
  p.oh = 2 * d->h, p.ow = 2 * d->w, p.ostride = 2, p.nacc = 4;
  for (int a = 0; a < 4; ++a) p.oy0[a] = a >> 1, p.ox0[a] = a & 1;
  p.y_mode = y == nullptr ? 0 : (staged ? 2 : 1);
  fill_common(p, d, pl, cout_pad);
  const int mode = epilogue_mode(d, slope, p);
  p.a_stages_per_unit = 4 * p.nkb;
  if (share) {
    const uint32_t row_b = (uint32_t)pl.kb * 2;
    p.tw = 8, p.th = 16, p.wstack = 1;
    p.unit_w = 8 * p.ms, p.unit_h = 16;
    p.a_share = 1, p.a_dy = 0, p.a_dx = 0, p.a_boff = 0;
    p.box_w = p.unit_w + 1, p.box_h = 17;
    p.a_sub16 = (8u * row_b) >> 4;
    p.a_sbo16 = ((uint32_t)p.box_w * row_b) >> 4;
    for (int t = 0; t < 4; ++t) p.tap_off16[t] = ((uint32_t)((t >> 1) * p.box_w + (t & 1)) * row_b) >> 4;
    p.a_tx_bytes = (uint32_t)p.box_w * p.box_h * row_b;
    p.tiles_w = (int)cdiv(p.gw, p.unit_w), p.tiles_h = (int)cdiv(p.gh, p.unit_h);
    p.num_m_tiles = (int64_t)p.n * p.tiles_w * p.tiles_h;
    p.a_stages_per_unit = p.nkb;
  }
  // four shifted input tiles and the MMAs each one feeds: {ky, kx, accumulator}
  static const int kOps[4][4][3] = {
      {{1, 1, 0}, {1, 2, 1}, {2, 1, 2}, {2, 2, 3}},      // A00
      {{1, 0, 1}, {2, 0, 3}, {-1, 0, 0}, {-1, 0, 0}},    // A01 (dx = 1)
      {{0, 1, 2}, {0, 2, 3}, {-1, 0, 0}, {-1, 0, 0}},    // A10 (dy = 1)
      {{0, 0, 3}, {-1, 0, 0}, {-1, 0, 0}, {-1, 0, 0}}};  // A11
  p.ntaps = 4;
  for (int t = 0; t < 4; ++t) {
    TcTap &tp = p.taps[t];
    tp.map = 0, tp.dy = (int16_t)(t >> 1), tp.dx = (int16_t)(t & 1), tp.nops = 0;
    for (int o = 0; o < 4 && kOps[t][o][0] >= 0; ++o) {
      tp.widx[tp.nops] = (int16_t)(kOps[t][o][0] * 3 + kOps[t][o][1]);
      tp.acc[tp.nops] = (int16_t)kOps[t][o][2];
      ++tp.nops;
    }
  }
  const int64_t ct = d->x_ctotal;
  if (int rc = encode_nhwc_map(&maps.a[0], x, d->x_ctotal, d->w, d->h, d->n, ct, ct * d->w, ct * d->w * d->h, pl.kb, p.box_w,
                               p.box_h))
    return rc;
  for (int i = 1; i < 4; ++i) maps.a[i] = maps.a[0];
  if (int rc = encode_weight_map(&maps.b, w_bf16, d->cin, 9 * cout_pad, pl.kb, pl.bn)) return rc;
  for (int a = 0; a < 4; ++a) {
    if (staged) {
      const int64_t yc = d->y_ctotal;
      const __nv_bfloat16 *base = static_cast<const __nv_bfloat16 *>(y) + ((int64_t)p.oy0[a] * p.ow + p.ox0[a]) * yc;
      if (int rc = encode_nhwc_map(&maps.c[a], base, d->y_coff + d->cout, d->w, d->h, d->n, 2 * yc, 2 * yc * p.ow,
                                   yc * p.ow * p.oh, pl.ob, p.unit_w, p.unit_h))
        return rc;
    } else {
      maps.c[a] = maps.a[0];
    }
  }
  maps.r = maps.a[0];
  if (int rc = check_plan(p, pl)) return rc;
  const Epilogue ep = make_epilogue(d, scale, shift, slope);
  return launch_tc(maps, p, pl, ep, mode, nullptr, y, labels, as_stream(stream));
}


// ---- weight gradient on tcgen05 (see conv_wgrad_tc_kernel) ----------------------------------------------------
static const char *wgrad_tc_why(const rtseg_conv_desc *d) {
  if (d->dtype != RTSEG_BF16) return "dtype must be bf16";
  if (d->groups != 1) return "dense convolutions only";
  if (d->kh * d->kw > 9) return "at most nine taps";
  if (d->cin % 16 || d->cout % 16 || d->cout > 256) return "cin and cout must be multiples of 16, cout <= 256";
  if (d->x_ctotal % 8 || d->x_coff % 8 || d->y_ctotal % 8 || d->y_coff % 8) return "channel slices must be 16-byte aligned";
  int gh, gw;
  rtseg_out_hw(d, &gh, &gw);
  if (d->stride == 1) {
    if (gh != d->h || gw != d->w) return "stride 1 needs a same-size convolution";
  } else if (d->stride == 2) {
    if ((d->h & 1) || (d->w & 1) || gh * 2 != d->h || gw * 2 != d->w) return "stride 2 needs even sizes and an exact halving";
  } else {
    return "stride must be 1 or 2";
  }
  return nullptr;
}

extern "C" int rtseg_conv_wgrad_tc_supported(const rtseg_conv_desc *d) { return d && wgrad_tc_why(d) == nullptr ? 1 : 0; }

extern "C" int rtseg_conv_wgrad_tc(const rtseg_conv_desc *d, const void *x, const void *dy, float *dw, rtseg_stream_t stream) {
  RTSEG_REQUIRE(d && x && dy && dw, "NULL pointer");
  if (const char *why = wgrad_tc_why(d)) RTSEG_FAIL(RTSEG_ERR_UNSUPPORTED, "rtseg_conv_wgrad_tc: %s", why);
  RTSEG_REQUIRE(reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(dy) % 16 == 0 &&
                    reinterpret_cast<uintptr_t>(dw) % 16 == 0, "tensor pointers must be 16-byte aligned");
  int gh, gw;
  rtseg_out_hw(d, &gh, &gw);
  WgParams p{};
  pick_tile(gw, p.tw, p.th);
  p.tiles_w = (int)cdiv(gw, p.tw), p.tiles_h = (int)cdiv(gh, p.th);
  p.units = d->n * p.tiles_w * p.tiles_h;
  p.cin = d->cin, p.cout = d->cout, p.n = d->cout;
  p.g_c0 = d->y_coff;
  p.bs_a = d->cin % 64 == 0 ? 64 : 16, p.bs_b = d->cout % 64 == 0 ? 64 : 16;
  p.box_a = kTileM * p.bs_a * 2, p.box_b = kTileM * p.bs_b * 2;
  p.kstep_a = (16 * p.bs_a * 2) >> 4, p.sbo_a = (8 * p.bs_a * 2) >> 4, p.layout_a = p.bs_a == 64 ? 2 : 6;
  p.kstep_b = (16 * p.bs_b * 2) >> 4, p.sbo_b = (8 * p.bs_b * 2) >> 4, p.layout_b = p.bs_b == 64 ? 2 : 6;
  p.b_boxes = d->cout / p.bs_b;
  const int per = kTileM / p.bs_a;          // x boxes (accumulator row blocks) per MMA group
  p.a_region = per * p.box_a;               // an M = 128 MMA always walks `per` blocks from the first box
  p.stage_bytes = (p.a_region + p.b_boxes * p.box_b + 1023) / 1024 * 1024;
  p.stages = std::min<int>(kMaxStages, (int)(200 * 1024) / p.stage_bytes);
  RTSEG_REQUIRE(p.stages >= 2, "operand tiles do not fit in shared memory");

  // tensor maps: stride 1 = the plain NHWC view; stride 2 = one view per input parity (row, column)
  WgMaps maps;
  const int64_t xc = d->x_ctotal, gc = d->y_ctotal;
  if (d->stride == 1) {
    if (int rc = encode_nhwc_map(&maps.x[0], x, d->x_coff + d->cin, d->w, d->h, d->n, xc, xc * d->w, xc * d->w * d->h, p.bs_a, p.tw, p.th))
      return rc;
    for (int i = 1; i < 4; ++i) maps.x[i] = maps.x[0];
  } else {
    for (int ph = 0; ph < 2; ++ph)
      for (int pw = 0; pw < 2; ++pw) {
        const __nv_bfloat16 *base = static_cast<const __nv_bfloat16 *>(x) + ((int64_t)ph * d->w + pw) * xc;
        if (int rc = encode_nhwc_map(&maps.x[ph * 2 + pw], base, d->x_coff + d->cin, d->w / 2, d->h / 2, d->n, 2 * xc, 2 * xc * d->w,
                                     xc * d->w * d->h, p.bs_a, p.tw, p.th))
          return rc;
      }
  }
  if (int rc = encode_nhwc_map(&maps.g, dy, d->y_coff + d->cout, gw, gh, d->n, gc, gc * gw, gc * gw * gh, p.bs_b, p.tw, p.th))
    return rc;

  // every (tap, input-channel block) is one x box = one block of accumulator rows; `per` of them form an MMA group
  struct Box { int tap, ci, map, dx, dy; };
  Box boxes[9 * 16];
  int nboxes = 0;
  const int kblocks = d->cin / p.bs_a;
  RTSEG_REQUIRE(d->kh * d->kw * kblocks <= 9 * 16, "too many operand boxes");
  for (int t = 0; t < d->kh * d->kw; ++t) {
    const int oy = (t / d->kw) * d->dil - d->pad_h, ox = (t % d->kw) * d->dil - d->pad_w;   // input = stride * out + o
    int map = 0, dx = ox, dy_ = oy;
    if (d->stride == 2) {
      const int ph = ((oy % 2) + 2) % 2, pw = ((ox % 2) + 2) % 2;
      map = ph * 2 + pw, dy_ = (oy - ph) / 2, dx = (ox - pw) / 2;
    }
    for (int kb = 0; kb < kblocks; ++kb) boxes[nboxes++] = Box{t, kb * p.bs_a, map, dx, dy_};
  }
  const int groups_total = (nboxes + per - 1) / per;
  const int groups_per_launch = std::min(8, 512 / p.n);
  static bool attr_set[64] = {false};
  int dev = 0;
  RTSEG_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    RTSEG_CUDA(cudaFuncSetAttribute(conv_wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 201 * 1024 + 1024));
    attr_set[dev] = true;
  }
  const int smem = p.stages * p.stage_bytes + 1024;
  const unsigned grid = (unsigned)std::min(p.units, num_sms());
  for (int g0 = 0; g0 < groups_total; g0 += groups_per_launch) {   // more groups than tensor memory holds: several launches
    p.ngroups = std::min(groups_per_launch, groups_total - g0);
    for (int g = 0; g < 8; ++g) {
      p.a_n[g] = 0, p.a_lbo16[g] = 0;
      for (int b = 0; b < 8; ++b) p.tap_of_blk[g][b] = -1, p.ci_of_blk[g][b] = 0, p.a_map[g][b] = 0, p.a_c[g][b] = 0, p.a_dx[g][b] = 0, p.a_dy[g][b] = 0;
    }
    for (int g = 0; g < p.ngroups; ++g) {
      const int first = (g0 + g) * per, cnt = std::min(per, nboxes - first);
      p.a_n[g] = cnt;
      p.a_lbo16[g] = cnt > 1 ? p.box_a >> 4 : 0;
      for (int b = 0; b < cnt; ++b) {
        const Box &bx = boxes[first + b];
        p.a_map[g][b] = (short)bx.map, p.a_c[g][b] = (short)(d->x_coff + bx.ci), p.a_dx[g][b] = (short)bx.dx, p.a_dy[g][b] = (short)bx.dy;
        p.tap_of_blk[g][b] = (short)bx.tap, p.ci_of_blk[g][b] = (short)bx.ci;
      }
    }
    uint32_t cols = 32;
    while (cols < (uint32_t)(p.ngroups * p.n)) cols <<= 1;
    p.tmem_cols = (int)cols;
    conv_wgrad_tc_kernel<<<grid, kWgThreads, smem, as_stream(stream)>>>(maps, p, dw);
    RTSEG_LAUNCH_CHECK("conv_wgrad_tc_kernel");
  }
  return RTSEG_OK;
}
