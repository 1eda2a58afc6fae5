// dw_tma.cu — K1' for the big depthwise layers: TMA-staged halo tiles in shared memory.
//
// Same arithmetic as conv_dw.cu (DWConvBNAct, reference models/modules.py:46-59) for the case
// that dominates Fast-SCNN's depthwise traffic: 3x3, dilation 1, stride 1 or 2, bf16 NHWC.
// The register kernel cannot keep enough bytes in flight (every outstanding 16-byte load pins
// four registers until it is consumed); here the loads belong to the TMA engine instead:
//   * one persistent CTA per SM and per 64/32/16-channel block; a producer thread streams input
//     halo tiles [TH_in][TW_in][CB] through a 2-4 stage mbarrier ring with
//     cp.async.bulk.tensor.4d — out-of-image rows/columns arrive zero-filled (that is the padding);
//   * eight consumer warps read the tile with conflict-free LDS.128 (a warp touches four full
//     128-byte pixel rows per request), accumulate in fp32 with packed FFMA2, keep the nine weight
//     vectors, scale and shift of their channels in registers for the whole kernel, and store
//     16-byte vectors (full 128-byte lines per pixel).
// Input is read from HBM exactly once (+ halo), output written once.
#include <algorithm>
#include <atomic>
#include <mutex>

#include "common.cuh"
#include "ptx.cuh"

namespace rtseg {

constexpr int kDwConsumers = 256;
constexpr int kDwThreads = kDwConsumers + 32;
constexpr int kDwMaxStages = 4;

struct DwTmaParams {
  int n, h, w, ho, wo;
  int c;             // channels of this op
  int x_coff, y_ct, y_co;
  int tiles_w, tiles_h;
  int64_t num_tiles;  // spatial tiles per channel block = n * tiles_h * tiles_w
  int stages;
};

// S = stride, CB = channels per block (CB*2 bytes per pixel row in the tile), TH x TW = output tile
// Persistent-CTA tile walk: (image, tile row, tile column) of tile blockIdx.x + k * gridDim.x, advanced by the
// decomposed grid stride so that the tile loop has no division.
struct TileWalk {
  int twi, thi, img, sw, sh, sn, tiles_w, tiles_h;
  __device__ TileWalk(int tiles_w_, int tiles_h_) : tiles_w(tiles_w_), tiles_h(tiles_h_) {
    const int r0 = (int)blockIdx.x / tiles_w, rs = (int)gridDim.x / tiles_w;
    twi = (int)blockIdx.x % tiles_w, thi = r0 % tiles_h, img = r0 / tiles_h;
    sw = (int)gridDim.x % tiles_w, sh = rs % tiles_h, sn = rs / tiles_h;
  }
  __device__ __forceinline__ void next() {
    twi += sw;
    if (twi >= tiles_w) twi -= tiles_w, ++thi;
    thi += sh;
    if (thi >= tiles_h) thi -= tiles_h, ++img;
    img += sn;
  }
};

template <int S, int CB, int TH, int TW, int MODE>
__global__ void __launch_bounds__(kDwThreads, 1) dw_tma_kernel(const __grid_constant__ CUtensorMap xmap,
                                                              const DwTmaParams p,
                                                              const float *__restrict__ wgt, Epilogue ep,
                                                              __nv_bfloat16 *__restrict__ y) {
  constexpr int CVB = CB / 8;               // 16-byte channel vectors per pixel
  constexpr int PX = kDwConsumers * 8 / (TH * TW * CVB) * 0 + (TH * TW * CVB) / kDwConsumers;  // pixels per thread
  static_assert(TH * TW * CVB == kDwConsumers * PX && TW % PX == 0, "tile must split evenly over the consumers");
  constexpr int THI = (TH - 1) * S + 3, TWI = (TW - 1) * S + 3;  // input halo tile
  constexpr int NW = (PX - 1) * S + 3;                            // input columns feeding one thread's PX outputs
  constexpr uint32_t kRowBytes = CB * 2;
  constexpr uint32_t kStageBytes = (uint32_t)THI * TWI * kRowBytes;
  constexpr uint32_t kStagePitch = (kStageBytes + 127) / 128 * 128;

  extern __shared__ __align__(128) uint8_t smem_raw[];
  const uint32_t s_base = (smem_u32(smem_raw) + 127u) & ~127u;
  const uint32_t s_bars = s_base + (uint32_t)p.stages * kStagePitch;
  const uint32_t full_bar = s_bars, empty_bar = s_bars + 8 * kDwMaxStages;
  uint8_t *gen = smem_raw + (s_base - smem_u32(smem_raw));
  volatile uint32_t *dead = reinterpret_cast<uint32_t *>(gen + (empty_bar + 8 * kDwMaxStages - s_base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cblk = blockIdx.y;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(full_bar + 8 * s, 1);
      mbar_init(empty_bar + 8 * s, kDwConsumers / 32);
    }
    *dead = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (warp == kDwConsumers / 32) {
    // ===================================================== TMA producer ====
    if (lane == 0) {
      prefetch_tmap(&xmap);
      uint32_t stage = 0, phase = 0;
      TileWalk tw(p.tiles_w, p.tiles_h);
      for (int64_t tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, tw.next()) {
        const int twi = tw.twi, thi = tw.thi, img = tw.img;
        mbar_wait(empty_bar + 8 * stage, phase ^ 1, dead);
        mbar_expect_tx(full_bar + 8 * stage, kStageBytes);
        tma_load_4d(s_base + stage * kStagePitch, &xmap, full_bar + 8 * stage, p.x_coff + cblk * CB, twi * TW * S - 1,
                    thi * TH * S - 1, img);
        if (++stage == (uint32_t)p.stages) stage = 0, phase ^= 1;
      }
    }
    return;
  }

  // ======================================================= consumers ====
  const int cv = threadIdx.x % CVB;               // channel vector inside the block
  const int strip = threadIdx.x / CVB;            // which PX-pixel strip of the tile
  constexpr int kStripsPerRow = TW / PX;
  const int tr = strip / kStripsPerRow;           // output row inside the tile
  const int tc0 = (strip % kStripsPerRow) * PX;   // first output column inside the tile
  const int c0 = cblk * CB + cv * 8;              // first channel (of this op) this thread owns

  // weights, scale, shift (and slope) of this thread's 8 channels stay in registers for the whole kernel
  float2 wv[9][4];
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    const float4 a = __ldg(reinterpret_cast<const float4 *>(wgt + (int64_t)t * p.c + c0));
    const float4 b = __ldg(reinterpret_cast<const float4 *>(wgt + (int64_t)t * p.c + c0 + 4));
    wv[t][0] = make_float2(a.x, a.y), wv[t][1] = make_float2(a.z, a.w);
    wv[t][2] = make_float2(b.x, b.y), wv[t][3] = make_float2(b.z, b.w);
  }
  float sc[8], sh[8], sl[8];
  load_affine<8>(ep, c0, sc, sh, sl);
  const bool relu_only = ep.lo == 0.f && ep.hi > 3.0e38f && ep.res_mode == RTSEG_RES_NONE;   // uniform

  const uint32_t lane_off = (uint32_t)(tr * S * TWI + tc0 * S) * kRowBytes + (uint32_t)cv * 16;
  uint32_t stage = 0, phase = 0;
  TileWalk tw(p.tiles_w, p.tiles_h);
  for (int64_t tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, tw.next()) {
    const int twi = tw.twi, thi = tw.thi, img = tw.img;

    float2 acc[PX][4];
#pragma unroll
    for (int i = 0; i < PX; ++i)
#pragma unroll
      for (int v = 0; v < 4; ++v) acc[i][v] = make_float2(0.f, 0.f);

    mbar_wait(full_bar + 8 * stage, phase, dead);
    const uint32_t tile_base = s_base + stage * kStagePitch + lane_off;
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      // the row's NW input vectors are requested back to back (one exposed shared-memory latency per filter row; a load
      // consumed the moment it was issued cost one per vector)
      constexpr int kBatch = (NW + 1) / 2;   // half a row at a time: the whole row would not fit in the register budget
      float4 raw[kBatch];
#pragma unroll
      for (int j = 0; j < NW; ++j) {
        if (j % kBatch == 0) {
#pragma unroll
          for (int q = 0; q < kBatch; ++q)
            if (j + q < NW) raw[q] = lds128(tile_base + (uint32_t)(ky * TWI + j + q) * kRowBytes);
        }
        const float4 rawf = raw[j % kBatch];
        const uint32_t w4[4] = {__float_as_uint(rawf.x), __float_as_uint(rawf.y), __float_as_uint(rawf.z),
                                __float_as_uint(rawf.w)};
        float2 xv[4];
#pragma unroll
        for (int v = 0; v < 4; ++v)
          xv[v] = make_float2(__uint_as_float(w4[v] << 16), __uint_as_float(w4[v] & 0xffff0000u));
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          const int d = j - kx;
          if (d >= 0 && d % S == 0 && d / S < PX) {
#pragma unroll
            for (int v = 0; v < 4; ++v) acc[d / S][v] = __ffma2_rn(xv[v], wv[ky * 3 + kx][v], acc[d / S][v]);
          }
        }
      }
    }
    // the tile has been consumed: hand the stage back to the producer before the (long) store phase
    __syncwarp();
    if (lane == 0) mbar_arrive(empty_bar + 8 * stage);
    if (++stage == (uint32_t)p.stages) stage = 0, phase ^= 1;

    const int oy = thi * TH + tr;
    const int ox0 = twi * TW + tc0;
    if (oy < p.ho) {
      __nv_bfloat16 *yrow = y + (((int64_t)img * p.ho + oy) * p.wo + ox0) * p.y_ct + p.y_co + c0;
#pragma unroll
      for (int i = 0; i < PX; ++i) {
        if (ox0 + i < p.wo) {
          if (MODE == 0 && relu_only) {
            // plain ReLU (every depthwise layer of Fast-SCNN): the clamp rides in the bf16 conversion
            uint32_t pk[4];
#pragma unroll
            for (int v = 0; v < 4; ++v)
              pk[v] = pack_relu_bf16(fmaf(acc[i][v].x, sc[2 * v], sh[2 * v]), fmaf(acc[i][v].y, sc[2 * v + 1], sh[2 * v + 1]));
            *reinterpret_cast<uint4 *>(yrow + (int64_t)i * p.y_ct) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
          } else {
            float ov[8];
#pragma unroll
            for (int v = 0; v < 4; ++v) {
              ov[2 * v] = epi_elem<MODE, false>(acc[i][v].x, 0.f, sc[2 * v], sh[2 * v], sl[2 * v], ep);
              ov[2 * v + 1] = epi_elem<MODE, false>(acc[i][v].y, 0.f, sc[2 * v + 1], sh[2 * v + 1], sl[2 * v + 1], ep);
            }
            Vec16<__nv_bfloat16>::store(yrow + (int64_t)i * p.y_ct, ov);
          }
        }
      }
    }
  }
}

typedef CUresult (*EncodeTiledFnDw)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFnDw get_encode_fn_dw() {
  static EncodeTiledFnDw fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void *ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFnDw>(ptr);
  });
  return fn;
}

int dw_tma_timeout_flag() { return read_clear_timeout_flag(); }

static std::atomic<uint64_t> g_dw_tma_launches{0};

template <int S, int CB, int TH, int TW>
static int launch_dw_tma(const rtseg_conv_desc *d, const void *x, const float *w, const Epilogue &ep, void *y,
                         cudaStream_t st) {
  constexpr int THI = (TH - 1) * S + 3, TWI = (TW - 1) * S + 3;
  constexpr uint32_t kStagePitch = ((uint32_t)THI * TWI * CB * 2 + 127) / 128 * 128;
  DwTmaParams p;
  p.n = d->n, p.h = d->h, p.w = d->w, p.c = d->cin;
  rtseg_out_hw(d, &p.ho, &p.wo);
  p.x_coff = d->x_coff, p.y_ct = d->y_ctotal, p.y_co = d->y_coff;
  p.tiles_w = (int)cdiv(p.wo, TW), p.tiles_h = (int)cdiv(p.ho, TH);
  p.num_tiles = (int64_t)p.n * p.tiles_w * p.tiles_h;
  p.stages = (int)std::min<uint32_t>(kDwMaxStages, (200u * 1024u) / kStagePitch);
  if (p.stages < 2) RTSEG_FAIL(RTSEG_ERR_UNSUPPORTED, "dw_tma: tile does not fit in shared memory");
  const uint32_t smem = (uint32_t)p.stages * kStagePitch + 8 * 2 * kDwMaxStages + 16 + 128;
  static_assert(THI <= 256 && TWI <= 256 && CB * 2 % 16 == 0, "halo tile exceeds the TMA box limits");
  RTSEG_REQUIRE(smem <= 210 * 1024, "dw_tma: shared-memory plan exceeds the kernel's limit");
  RTSEG_REQUIRE(d->x_coff + d->cin <= d->x_ctotal && d->y_coff + d->cin <= d->y_ctotal, "dw_tma: channel slice outside its buffer");

  EncodeTiledFnDw enc = get_encode_fn_dw();
  if (!enc) RTSEG_FAIL(RTSEG_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  CUtensorMap map;
  const int64_t ct = d->x_ctotal;
  const cuuint64_t dims[4] = {(cuuint64_t)ct, (cuuint64_t)d->w, (cuuint64_t)d->h, (cuuint64_t)d->n};
  const cuuint64_t strides[3] = {(cuuint64_t)ct * 2, (cuuint64_t)ct * d->w * 2, (cuuint64_t)ct * d->w * d->h * 2};
  const cuuint32_t box[4] = {(cuuint32_t)CB, (cuuint32_t)TWI, (cuuint32_t)THI, 1};
  const cuuint32_t es[4] = {1, 1, 1, 1};
  const CUresult r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void *>(x), dims, strides, box, es,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) RTSEG_FAIL(RTSEG_ERR_CUDA, "cuTensorMapEncodeTiled(depthwise input) failed with CUresult %d", (int)r);

  const int cblocks = d->cin / CB;
  const int gx = (int)std::min<int64_t>(p.num_tiles, std::max(1, num_sms() / cblocks));
  dim3 grid((unsigned)gx, (unsigned)cblocks);
  auto yb = static_cast<__nv_bfloat16 *>(y);
#define RTSEG_DW_TMA(MODE_)                                                                                         \
  do {                                                                                                              \
    static bool attr_set[64] = {false};                                                                             \
    int dev = 0;                                                                                                    \
    RTSEG_CUDA(cudaGetDevice(&dev));                                                                                \
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {                                                                   \
      RTSEG_CUDA(cudaFuncSetAttribute(dw_tma_kernel<S, CB, TH, TW, MODE_>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                      210 * 1024));                                                                 \
      attr_set[dev] = true;                                                                                         \
    }                                                                                                               \
    dw_tma_kernel<S, CB, TH, TW, MODE_><<<grid, kDwThreads, smem, st>>>(map, p, w, ep, yb);                         \
  } while (0)
  if (ep.mode == 0)
    RTSEG_DW_TMA(0);
  else if (ep.mode == 1)
    RTSEG_DW_TMA(1);
  else
    RTSEG_DW_TMA(2);
#undef RTSEG_DW_TMA
  RTSEG_LAUNCH_CHECK("dw_tma_kernel");
  g_dw_tma_launches.fetch_add(1, std::memory_order_relaxed);
  return RTSEG_OK;
}

// Returns RTSEG_ERR_UNSUPPORTED (without setting work in motion) when the layer is not one this kernel
// handles; the caller (rtseg_conv_dw) then uses the register kernels.
int dw_tma_try(const rtseg_conv_desc *d, const void *x, const float *w, const Epilogue &ep, const void *res, void *y,
               cudaStream_t st) {
  const bool ok = d->dtype == RTSEG_BF16 && d->kh == 3 && d->kw == 3 && d->dil == 1 && d->pad_h == 1 && d->pad_w == 1 &&
                  (d->stride == 1 || d->stride == 2) && d->cout == d->cin && res == nullptr &&
                  d->res_mode == RTSEG_RES_NONE && d->cin % 16 == 0 && d->x_ctotal % 8 == 0 && d->x_coff % 8 == 0 &&
                  d->y_ctotal % 8 == 0 && d->y_coff % 8 == 0 && reinterpret_cast<uintptr_t>(x) % 16 == 0 &&
                  reinterpret_cast<uintptr_t>(y) % 16 == 0 && reinterpret_cast<uintptr_t>(w) % 16 == 0;
  if (!ok) return RTSEG_ERR_UNSUPPORTED;
  int ho, wo;
  rtseg_out_hw(d, &ho, &wo);
  if ((int64_t)d->n * ho * wo < 64 * 1024) return RTSEG_ERR_UNSUPPORTED;  // tiny maps: not worth a persistent launch
  const int cb = d->cin % 64 == 0 ? 64 : (d->cin % 32 == 0 ? 32 : 16);
  if (d->stride == 1) {
    if (cb == 64) return launch_dw_tma<1, 64, 8, 32>(d, x, w, ep, y, st);
    if (cb == 32) return launch_dw_tma<1, 32, 8, 32>(d, x, w, ep, y, st);
    return launch_dw_tma<1, 16, 8, 32>(d, x, w, ep, y, st);
  }
  if (cb == 64) return launch_dw_tma<2, 64, 4, 32>(d, x, w, ep, y, st);
  if (cb == 32) return launch_dw_tma<2, 32, 8, 32>(d, x, w, ep, y, st);
  return launch_dw_tma<2, 16, 8, 32>(d, x, w, ep, y, st);
}

}  // namespace rtseg

// Diagnostics: how many times the TMA-staged depthwise kernel has been launched by this process (tests assert that a
// layer above the size gate really took this kernel and not the register kernels of conv_dw.cu).
extern "C" uint64_t rtseg_dw_tma_launch_count(void) { return rtseg::g_dw_tma_launches.load(std::memory_order_relaxed); }
