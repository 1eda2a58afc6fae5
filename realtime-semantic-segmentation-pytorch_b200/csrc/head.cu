// head.cu — segmentation head and metric kernels.
//
//  * rtseg_upsample_bilinear  : F.interpolate(mode='bilinear') (reference
//    models/fastscnn.py:30,89, models/modules.py:153-156, models/enet.py:198-202)
//  * rtseg_upsample_argmax    : the same + preds.max(dim=1)[1]
//    (core/seg_trainer.py:172, models/ddrnet.py:55-58) + optionally the
//    torchmetrics confusion-matrix update (utils/metrics.py:4-7) in ONE pass;
//    the 19 x 1024 x 2048 full-resolution logits (159 MB fp32 per image) are
//    never written or re-read.
//  * rtseg_argmax_nhwc / _nchw, rtseg_confmat_update: the unfused pieces.
//  * layout plumbing: NCHW <-> NHWC transposes, AdaptiveAvgPool2d.
//
// Index arithmetic follows ATen's area_pixel_compute_source_index so that the
// sampled positions are the ones the reference computes:
//   align_corners:   src = dst * (in-1)/(out-1)
//   otherwise:       src = max(0, (dst+0.5) * in/out - 0.5)
// and  v = l0y*(l0x*v00 + l1x*v01) + l1y*(l0x*v10 + l1x*v11).
#include <algorithm>
#include <cfloat>

#include "common.cuh"

namespace rtseg {

// torch.max(dim) semantics: first maximal index, NaN beats everything.
__device__ __forceinline__ void argmax_step(float v, int c, float &best, int &idx) {
  if (v > best || (v != v && best == best)) {
    best = v;
    idx = c;
  }
}

// ---------------------------------------------------------------- bilinear ----
template <typename T, int V>
__global__ void __launch_bounds__(256) bilinear_kernel(int n, int hi, int wi, int c, int ho, int wo, int align,
                                                       float sy, float sx, const T *__restrict__ x, int x_ct,
                                                       int x_co, T *__restrict__ y, int y_ct, int y_co) {
  const int cv_n = c / V;
  const int64_t total = (int64_t)n * ho * wo * cv_n;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= total) return;
  const int c0 = (int)(tid % cv_n) * V;
  int64_t r = tid / cv_n;
  const int ox = (int)(r % wo);
  r /= wo;
  const int oy = (int)(r % ho);
  const int b = (int)(r / ho);
  int y0, y1, x0, x1;
  float ly0, ly1, lx0, lx1;
  src_index(oy, sy, hi, align, y0, y1, ly0, ly1);
  src_index(ox, sx, wi, align, x0, x1, lx0, lx1);
  const T *xb = x + (int64_t)b * hi * wi * x_ct + x_co + c0;
  const T *p00 = xb + ((int64_t)y0 * wi + x0) * x_ct, *p01 = xb + ((int64_t)y0 * wi + x1) * x_ct;
  const T *p10 = xb + ((int64_t)y1 * wi + x0) * x_ct, *p11 = xb + ((int64_t)y1 * wi + x1) * x_ct;
  T *yp = y + (((int64_t)b * ho + oy) * wo + ox) * y_ct + y_co + c0;
  if constexpr (V == 1) {
    const float v = ly0 * (lx0 * ld_f<T>(p00) + lx1 * ld_f<T>(p01)) + ly1 * (lx0 * ld_f<T>(p10) + lx1 * ld_f<T>(p11));
    st_f<T>(yp, v);
  } else {
    float a[V], bq[V], cq[V], dq[V], o[V];
    Vec16<T>::load(p00, a);
    Vec16<T>::load(p01, bq);
    Vec16<T>::load(p10, cq);
    Vec16<T>::load(p11, dq);
#pragma unroll
    for (int e = 0; e < V; ++e) o[e] = ly0 * (lx0 * a[e] + lx1 * bq[e]) + ly1 * (lx0 * cq[e] + lx1 * dq[e]);
    Vec16<T>::store(yp, o);
  }
}

// Upsampling variant: a thread owns one output column x V channels and walks R consecutive output rows.  The column's
// two source columns and weights are fixed; a source row is lerped horizontally once and reused by every output row
// that samples it (x4: ~1.5 source rows per 4 outputs instead of 8 row loads) — same expression tree as above, so the
// two kernels agree bit for bit.
template <typename T, int V, int R>
__global__ void __launch_bounds__(256) bilinear_rows_kernel(int n, int hi, int wi, int c, int ho, int wo, int align, float sy,
                                                            float sx, const T *__restrict__ x, int x_ct, int x_co,
                                                            T *__restrict__ y, int y_ct, int y_co) {
  const int cv_n = c / V;
  const int hq = (ho + R - 1) / R;
  const int64_t total = (int64_t)n * hq * wo * cv_n;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= total) return;
  const int c0 = (int)(tid % cv_n) * V;
  int64_t r = tid / cv_n;
  const int ox = (int)(r % wo);
  r /= wo;
  const int oyq = (int)(r % hq);
  const int b = (int)(r / hq);
  int x0, x1;
  float lx0, lx1;
  src_index(ox, sx, wi, align, x0, x1, lx0, lx1);
  const T *xb = x + (int64_t)b * hi * wi * x_ct + x_co + c0;
  int ya = -1, yb = -1;          // source rows held in ha / hb
  float ha[V], hb[V];
  auto hrow = [&](int yy, float (&h)[V]) {
    float a[V], bq[V];
    Vec16<T>::load(xb + ((int64_t)yy * wi + x0) * x_ct, a);
    Vec16<T>::load(xb + ((int64_t)yy * wi + x1) * x_ct, bq);
#pragma unroll
    for (int e = 0; e < V; ++e) h[e] = lx0 * a[e] + lx1 * bq[e];
  };
#pragma unroll
  for (int i = 0; i < R; ++i) {
    const int oy = oyq * R + i;
    if (oy >= ho) break;
    int y0, y1;
    float ly0, ly1;
    src_index(oy, sy, hi, align, y0, y1, ly0, ly1);
    if (y0 != ya) {
      if (y0 == yb) {
#pragma unroll
        for (int e = 0; e < V; ++e) ha[e] = hb[e];
      } else {
        hrow(y0, ha);
      }
      ya = y0;
    }
    if (y1 != yb) {
      if (y1 == ya) {
#pragma unroll
        for (int e = 0; e < V; ++e) hb[e] = ha[e];
      } else {
        hrow(y1, hb);
      }
      yb = y1;
    }
    float o[V];
#pragma unroll
    for (int e = 0; e < V; ++e) o[e] = ly0 * ha[e] + ly1 * hb[e];
    Vec16<T>::store(y + (((int64_t)b * ho + oy) * wo + ox) * y_ct + y_co + c0, o);
  }
}

// ------------------------------------------------- fused upsample + argmax ----
constexpr int kHeadThreads = 128;  // output columns per block
constexpr int kHeadRows = 32;      // output rows per thread (x8 heads: ~4 source-row pairs, each loaded once)
constexpr int kMaxBins = 32 * 32;

__device__ __forceinline__ void hist_add_aggregated(int *hist, int bin, bool valid) {
  // Lanes of a warp that hit the same bin (segmentation labels are spatially
  // coherent, so usually all 32) are merged into one shared-memory atomic.
  const unsigned active = __ballot_sync(0xffffffffu, valid);
  if (!valid) return;
  const unsigned peers = __match_any_sync(active, bin);
  const int leader = __ffs(peers) - 1;
  if ((int)(threadIdx.x & 31) == leader) atomicAdd(&hist[bin], __popc(peers));
}

// One thread owns one output column and walks kHeadRows output rows.  Per source-row pair it caches the
// horizontally interpolated row A and the difference D = B - A (channels >= c padded with -inf / 0),
// so that one output pixel costs, per class, one FFMA (A + ly1*D) and a branch-free compare/select —
// the x8 upsampled logits exist only in registers.  NaNs (which must win, torch.max semantics) are
// and infinities (inf - inf in D, inf + -inf in A + ly1*D would turn into NaNs torch does not produce) are detected
// when a source row is loaded; such threads evaluate torch's own expression ly0*(lx0*p00 + lx1*p01) + ly1*(...)
// straight from memory.  Padded channels hold -inf and the compare is strict, so the label never leaves [0, c).
template <typename T, int CMAX, bool EXACT = false, bool PLAIN = false>   // EXACT: c == CMAX, no padded classes to mask;
                                                                          // PLAIN: uint8 class map only (no int64 map, no confusion matrix)
__global__ void __launch_bounds__(kHeadThreads) upsample_argmax_kernel(
    int n, int hi, int wi, int c, int ct, int ho, int wo, int align, float sy, float sx, const T *__restrict__ x,
    uint8_t *__restrict__ lab8, int64_t *__restrict__ lab64, const int64_t *__restrict__ target, int ignore_index,
    unsigned long long *__restrict__ confmat, int vec) {
  __shared__ int hist[kMaxBins];
  __shared__ int4 s_rows[kHeadRows];   // per output row of this block: {y0, y1, bits(ly0), bits(ly1)} (the same for every column)
  const bool do_cm = !PLAIN && target != nullptr;
  if (do_cm) {
    for (int i = threadIdx.x; i < c * c; i += blockDim.x) hist[i] = 0;
  }
  if (threadIdx.x < kHeadRows) {
    int y0, y1;
    float ly0, ly1;
    src_index(min((int)(blockIdx.y * kHeadRows + threadIdx.x), ho - 1), sy, hi, align, y0, y1, ly0, ly1);
    s_rows[threadIdx.x] = make_int4(y0, y1, __float_as_int(ly0), __float_as_int(ly1));
  }
  __syncthreads();
  const int ox = blockIdx.x * kHeadThreads + threadIdx.x;
  const int oy_begin = blockIdx.y * kHeadRows;
  const int b = blockIdx.z;
  const bool col_ok = ox < wo;

  int x0 = 0, x1 = 0;
  float lx0 = 1.f, lx1 = 0.f;
  if (col_ok) src_index(ox, sx, wi, align, x0, x1, lx0, lx1);
  const T *xb = x + (int64_t)b * hi * wi * ct;  // ct = channels per pixel of the (possibly padded) logits buffer

  float rowA[CMAX], rowD[CMAX];  // A and (B - A)
  int ia = -1, ib = -1;
  int run_bin = 0, run_cnt = 0;  // confusion-matrix bin of the current run of rows and its length
  bool has_nan = false;

  // horizontally interpolated source row i, channels >= c padded so that they never win.  bf16 logits in a 16-byte
  // aligned buffer (`vec`) are read as 16-byte vectors: ceil(CMAX / 8) loads per corner pixel instead of CMAX.
  const int row_pitch = wi * ct, off0 = x0 * ct, off1 = x1 * ct;   // elements; one image fits 31 bits (checked by the launcher)
  auto load_row = [&](int i, float(&row)[CMAX]) {
    const T *p0 = xb + (i * row_pitch + off0);
    const T *p1 = xb + (i * row_pitch + off1);
    float bad = 0.f;   // max |v| with NaNs kept: one compare per row instead of one per class
    if (sizeof(T) == 2 && vec) {
      constexpr int NV = (CMAX + 7) / 8;
      uint32_t q0[NV * 4], q1[NV * 4];
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const uint4 a = __ldg(reinterpret_cast<const uint4 *>(p0) + j), bq = __ldg(reinterpret_cast<const uint4 *>(p1) + j);
        q0[4 * j] = a.x, q0[4 * j + 1] = a.y, q0[4 * j + 2] = a.z, q0[4 * j + 3] = a.w;
        q1[4 * j] = bq.x, q1[4 * j + 1] = bq.y, q1[4 * j + 2] = bq.z, q1[4 * j + 3] = bq.w;
      }
#pragma unroll
      for (int k = 0; k < CMAX; ++k) {
        const float a = __uint_as_float((k & 1) ? (q0[k >> 1] & 0xffff0000u) : (q0[k >> 1] << 16));
        const float bq = __uint_as_float((k & 1) ? (q1[k >> 1] & 0xffff0000u) : (q1[k >> 1] << 16));
        const float v = lx0 * a + lx1 * bq;
        bad = fmax_nan(bad, (EXACT || k < c) ? fabsf(v) : 0.f);
        row[k] = (EXACT || k < c) ? v : -INFINITY;
      }
    } else {
#pragma unroll
      for (int k = 0; k < CMAX; ++k) {
        float v = -INFINITY;
        if (k < c) {
          v = lx0 * ld_f<T>(p0 + k) + lx1 * ld_f<T>(p1 + k);
          bad = fmax_nan(bad, fabsf(v));
        }
        row[k] = v;
      }
    }
    has_nan |= !(bad <= FLT_MAX);   // NaN or +-inf somewhere in this row
  };

  int64_t opix = ((int64_t)b * ho + oy_begin) * wo + ox;
  for (int ry = 0; ry < kHeadRows; ++ry, opix += wo) {
    const int oy = oy_begin + ry;
    if (oy >= ho) break;  // uniform across the block
    const int4 sr = s_rows[ry];
    const int y0 = sr.x, y1 = sr.y;
    const float ly0 = __int_as_float(sr.z), ly1 = __int_as_float(sr.w);
    int label = 0;
    if (col_ok) {
      if (!(y0 == ia && y1 == ib)) {
        // new source-row pair: A <- row y0, D <- row y1 - A.  (Keeping B in registers to skip the reload of A when the
        // window moves down one row costs 19 registers: measured faster alone, slower with the confusion matrix.)
        float rowB[CMAX];
        if (y0 != ia) load_row(y0, rowA);  // (A + D would reproduce the old B only up to a rounding: reload)
        ia = y0;
        if (y1 == y0) {
#pragma unroll
          for (int k = 0; k < CMAX; ++k) rowD[k] = 0.f;
        } else {
          load_row(y1, rowB);
#pragma unroll
          for (int k = 0; k < CMAX; ++k) rowD[k] = (EXACT || k < c) ? rowB[k] - rowA[k] : 0.f;
        }
        ib = y1;
      }
      if (!has_nan) {
        float best = rowA[0] + ly1 * rowD[0];
#pragma unroll
        for (int k = 1; k < CMAX; ++k) {
          const float v = fmaf(ly1, rowD[k], rowA[k]);
          const bool g = v > best;  // strict: the first maximum keeps the label
          best = g ? v : best;
          label = g ? k : label;
        }
      } else {  // non-finite logits in this column: torch's expression and torch.max semantics (the first NaN wins)
        const T *r0 = xb + (int64_t)y0 * wi * ct, *r1 = xb + (int64_t)y1 * wi * ct;
        float best = 0.f;
        for (int k = 0; k < c; ++k) {
          const float a = lx0 * ld_f<T>(r0 + (int64_t)x0 * ct + k) + lx1 * ld_f<T>(r0 + (int64_t)x1 * ct + k);
          const float bb = lx0 * ld_f<T>(r1 + (int64_t)x0 * ct + k) + lx1 * ld_f<T>(r1 + (int64_t)x1 * ct + k);
          const float v = ly0 * a + ly1 * bb;
          if (k == 0) best = v; else argmax_step(v, k, best, label);
        }
      }
      if (PLAIN || lab8) lab8[opix] = (uint8_t)label;
      if (!PLAIN && lab64) lab64[opix] = (int64_t)label;
    }
    if (do_cm && col_ok) {
      // a thread walks one image column: consecutive rows mostly fall into the same (target, prediction) bin, so
      // the count runs in a register and reaches the shared histogram once per run, not once per pixel
      const int64_t t = __ldg(target + opix);
      if (t != (int64_t)ignore_index && t >= 0 && t < c) {
        const int bin = (int)t * c + label;
        if (bin == run_bin) {
          ++run_cnt;
        } else {
          if (run_cnt) atomicAdd(&hist[run_bin], run_cnt);
          run_bin = bin, run_cnt = 1;
        }
      }
    }
  }
  if (do_cm) {
    if (run_cnt) atomicAdd(&hist[run_bin], run_cnt);
    __syncthreads();
    for (int i = threadIdx.x; i < c * c; i += blockDim.x) {
      const int v = hist[i];
      if (v) atomicAdd(confmat + i, (unsigned long long)v);
    }
  }
}

// ------------------------------------------------ nearest resize of label maps ----
// F.interpolate(masks.unsqueeze(1).float(), size, mode='nearest') of the auxiliary-head path (reference
// core/seg_trainer.py:60-61): src = min(floor(dst * (in / out)), in - 1) with the ratio in fp32, like ATen's
// nearest_neighbor_compute_source_index.  Labels stay int64 (the float detour of the reference is exact for them).
__global__ void __launch_bounds__(256) resize_nearest_i64_kernel(int n, int hi, int wi, int ho, int wo, float sy, float sx,
                                                                 const int64_t *__restrict__ x, int64_t *__restrict__ y) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= (int64_t)n * ho * wo) return;
  const int ox = (int)(tid % wo);
  const int64_t r = tid / wo;
  const int oy = (int)(r % ho), b = (int)(r / ho);
  const int iy = min((int)floorf((float)oy * sy), hi - 1), ix = min((int)floorf((float)ox * sx), wi - 1);
  y[tid] = __ldg(x + ((int64_t)b * hi + iy) * wi + ix);
}

// ------------------------------------------------------------------ argmax ----
template <typename T>
__global__ void __launch_bounds__(256) argmax_nhwc_kernel(int64_t npix, int c, int c_total, const T *__restrict__ x,
                                                          uint8_t *__restrict__ lab8, int64_t *__restrict__ lab64) {
  const int64_t pix = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= npix) return;
  const T *p = x + pix * c_total;
  float best = ld_f<T>(p);
  int idx = 0;
  for (int k = 1; k < c; ++k) argmax_step(ld_f<T>(p + k), k, best, idx);
  if (lab8) lab8[pix] = (uint8_t)idx;
  if (lab64) lab64[pix] = idx;
}

template <typename T>
__global__ void __launch_bounds__(256) argmax_nchw_kernel(int n, int c, int64_t hw, const T *__restrict__ x,
                                                          uint8_t *__restrict__ lab8, int64_t *__restrict__ lab64) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= (int64_t)n * hw) return;
  const int64_t b = tid / hw, s = tid % hw;
  const T *p = x + b * c * hw + s;
  float best = ld_f<T>(p);
  int idx = 0;
  for (int k = 1; k < c; ++k) argmax_step(ld_f<T>(p + (int64_t)k * hw), k, best, idx);
  if (lab8) lab8[tid] = (uint8_t)idx;
  if (lab64) lab64[tid] = idx;
}

// ------------------------------------------------------------ confusion matrix ----
template <typename P>
__global__ void __launch_bounds__(256) confmat_kernel(int64_t npix, int c, int ignore_index, const P *__restrict__ pred,
                                                      const int64_t *__restrict__ target,
                                                      unsigned long long *__restrict__ confmat,
                                                      unsigned long long *__restrict__ bad) {
  __shared__ int hist[kMaxBins];
  __shared__ int nbad;
  for (int i = threadIdx.x; i < c * c; i += blockDim.x) hist[i] = 0;
  if (threadIdx.x == 0) nbad = 0;
  __syncthreads();
  // Each block walks a contiguous chunk; every warp iteration is full or fully masked.
  const int64_t per_block = (npix + gridDim.x - 1) / gridDim.x;
  const int64_t begin = (int64_t)blockIdx.x * per_block;
  const int64_t end = min(npix, begin + per_block);
  for (int64_t i0 = begin; i0 < end; i0 += blockDim.x) {
    const int64_t i = i0 + threadIdx.x;
    bool valid = false;
    int bin = 0;
    if (i < end) {
      const int64_t t = __ldg(target + i);
      const int64_t q = (int64_t)pred[i];
      if (t != (int64_t)ignore_index) {
        if (t >= 0 && t < c && q >= 0 && q < c) {
          valid = true;
          bin = (int)t * c + (int)q;
        } else {
          atomicAdd(&nbad, 1);
        }
      }
    }
    hist_add_aggregated(hist, bin, valid);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < c * c; i += blockDim.x) {
    const int v = hist[i];
    if (v) atomicAdd(confmat + i, (unsigned long long)v);
  }
  if (threadIdx.x == 0 && nbad && bad) atomicAdd(bad, (unsigned long long)nbad);
}

// ------------------------------------------------------------------- transposes ----
// dst[b][col][row] = src[b][row][col] on 32x32 tiles; rows/cols refer to src.
template <typename TS, typename TD>
__global__ void __launch_bounds__(256) transpose_kernel(int rows, int cols, const TS *__restrict__ src,
                                                        int64_t src_pitch, int64_t src_batch, TD *__restrict__ dst,
                                                        int64_t dst_pitch, int64_t dst_batch) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z;
  const int r0 = blockIdx.y * 32, c0 = blockIdx.x * 32;
  const int tx = threadIdx.x % 32, ty = threadIdx.x / 32;  // 32 x 8
  const TS *s = src + (int64_t)b * src_batch;
  TD *d = dst + (int64_t)b * dst_batch;
#pragma unroll
  for (int k = 0; k < 32; k += 8) {
    const int r = r0 + ty + k, c = c0 + tx;
    if (r < rows && c < cols) tile[ty + k][tx] = ld_f<TS>(s + (int64_t)r * src_pitch + c);
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 32; k += 8) {
    const int c = c0 + ty + k, r = r0 + tx;
    if (r < rows && c < cols) st_f<TD>(d + (int64_t)c * dst_pitch + r, tile[tx][ty + k]);
  }
}

template <typename TS, typename TD>
static int launch_transpose(int n, int rows, int cols, const void *src, int64_t src_pitch, int64_t src_batch,
                            void *dst, int64_t dst_pitch, int64_t dst_batch, cudaStream_t st) {
  // grid.y is limited to 65535: fold very tall matrices over several launches.
  const int max_rows = 65535 * 32;
  for (int r0 = 0; r0 < rows; r0 += max_rows) {
    const int rr = min(max_rows, rows - r0);
    dim3 grid((unsigned)cdiv(cols, 32), (unsigned)cdiv(rr, 32), (unsigned)n);
    transpose_kernel<TS, TD><<<grid, 256, 0, st>>>(rr, cols, static_cast<const TS *>(src) + (int64_t)r0 * src_pitch,
                                                   src_pitch, src_batch, static_cast<TD *>(dst) + r0, dst_pitch,
                                                   dst_batch);
    RTSEG_LAUNCH_CHECK("transpose_kernel");
  }
  return RTSEG_OK;
}

static int dispatch_transpose(int n, int rows, int cols, int sdt, const void *src, int64_t sp, int64_t sb, int ddt,
                              void *dst, int64_t dp, int64_t db, cudaStream_t st) {
  if (sdt == RTSEG_F32 && ddt == RTSEG_F32) return launch_transpose<float, float>(n, rows, cols, src, sp, sb, dst, dp, db, st);
  if (sdt == RTSEG_F32 && ddt == RTSEG_BF16)
    return launch_transpose<float, __nv_bfloat16>(n, rows, cols, src, sp, sb, dst, dp, db, st);
  if (sdt == RTSEG_BF16 && ddt == RTSEG_F32)
    return launch_transpose<__nv_bfloat16, float>(n, rows, cols, src, sp, sb, dst, dp, db, st);
  if (sdt == RTSEG_BF16 && ddt == RTSEG_BF16)
    return launch_transpose<__nv_bfloat16, __nv_bfloat16>(n, rows, cols, src, sp, sb, dst, dp, db, st);
  RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
}

// ------------------------------------------------------------- adaptive avg pool ----
template <typename T, typename TO>
__global__ void __launch_bounds__(256) adaptive_avgpool_kernel(int h, int w, int c, int p, const T *__restrict__ x,
                                                               int x_ct, int x_co, TO *__restrict__ y) {
  __shared__ float part[8][32];
  const int cell = blockIdx.x, b = blockIdx.y, c0 = blockIdx.z * 32;
  const int pi = cell / p, pj = cell % p;
  const int hs = (pi * h) / p, he = ((pi + 1) * h + p - 1) / p;
  const int wsx = (pj * w) / p, we = ((pj + 1) * w + p - 1) / p;
  const int lane = threadIdx.x % 32, ty = threadIdx.x / 32;
  const int ch = c0 + lane;
  const int ww = we - wsx, cnt = (he - hs) * ww;
  float s = 0.f;
  if (ch < c) {
    const T *xb = x + (int64_t)b * h * w * x_ct + x_co + ch;
    for (int i = ty; i < cnt; i += 8) {
      const int yy = hs + i / ww, xx = wsx + i % ww;
      s += ld_f<T>(xb + ((int64_t)yy * w + xx) * x_ct);
    }
  }
  part[ty][lane] = s;
  __syncthreads();
  if (ty == 0 && ch < c) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += part[k][lane];
    st_f<TO>(y + (((int64_t)b * p + pi) * p + pj) * c + ch, t / (float)cnt);
  }
}


// ---- pyramid pooling in one pass: AdaptiveAvgPool2d(1), (2), (4) and (6) of the same map ------------------------
// PyramidPoolingModule (reference models/modules.py:145-149) pools one feature map four times; each of those launches
// re-reads the map and the 1 x 1 one is a single long serial sum per block.  Here a block owns (image, 32 channels),
// reads its slice once — lanes = channels, a warp walks whole rows with eight loads in flight — and adds every pixel to
// its 4 x 4 cell (h, w multiples of 4: the 1-, 2- and 4-bin levels are exact unions of those cells) and to the up to
// four overlapping 6 x 6 bins.  Every warp sums into its own shared-memory copy (no atomics: the result does not depend
// on the order warps run in); the copies meet in a fixed order at the end.
template <typename T>
__global__ void __launch_bounds__(256) pyramid_pool_kernel(int h, int w, int c, const T *__restrict__ x, int x_ct, int x_co,
                                                           T *__restrict__ y1, T *__restrict__ y2, T *__restrict__ y4,
                                                           T *__restrict__ y6) {
  extern __shared__ float pp_smem[];   // [8 warps][16 + 36 bins][32 channels]
  const int b = blockIdx.x, c0 = blockIdx.y * 32;
  const int lane = threadIdx.x % 32, ty = threadIdx.x / 32;
  for (int i = threadIdx.x; i < 8 * 52 * 32; i += 256) pp_smem[i] = 0.f;
  __syncthreads();
  float *mine = pp_smem + ty * 52 * 32 + lane;      // bin k of this (warp, lane): mine[k * 32]
  const int ch = c0 + lane;
  if (ch < c) {
    const T *xb = x + (int64_t)b * h * w * x_ct + x_co + ch;
    const int ch4 = h / 4, cw4 = w / 4;
    int lo6[6], hi6[6];                           // adaptive 6-bin columns [lo, hi)
#pragma unroll
    for (int j = 0; j < 6; ++j) lo6[j] = (j * w) / 6, hi6[j] = ((j + 1) * w + 5) / 6;
    for (int yy = ty; yy < h; yy += 8) {
      const int cy = yy / ch4;
      int r6[2], nr = 0;                          // 6-bin rows that contain yy (adaptive bins overlap by one row)
      for (int i = 0; i < 6; ++i)
        if ((i * h) / 6 <= yy && yy < ((i + 1) * h + 5) / 6 && nr < 2) r6[nr++] = i;
      const T *xr = xb + (int64_t)yy * w * x_ct;
      int cx = 0, cx_end = cw4;                   // current 4 x 4 cell column and its end
      float acc = 0.f;
      // the row's six (overlapping) column bins are summed in registers, pixel order kept, and reach the shared
      // copies once per row: a shared-memory read-modify-write per pixel and bin was the whole kernel (0.10 ms)
      float a6[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      for (int x0 = 0; x0 < w; x0 += 8) {
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = (x0 + u < w) ? ld_f<T>(xr + (int64_t)(x0 + u) * x_ct) : 0.f;
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int xx = x0 + u;
          if (xx >= w) break;
          if (xx == cx_end) {
            mine[(cy * 4 + cx) * 32] += acc;
            acc = 0.f, ++cx, cx_end += cw4;
          }
          acc += v[u];
#pragma unroll
          for (int j = 0; j < 6; ++j)
            if (xx >= lo6[j] && xx < hi6[j]) a6[j] += v[u];
        }
      }
      mine[(cy * 4 + cx) * 32] += acc;
#pragma unroll
      for (int j = 0; j < 6; ++j)
        for (int k = 0; k < nr; ++k) mine[(16 + r6[k] * 6 + j) * 32] += a6[j];
    }
  }
  __syncthreads();
  auto total = [&](int bin) {   // the eight warps' copies, in a fixed order
    float t = 0.f;
#pragma unroll
    for (int wv = 0; wv < 8; ++wv) t += pp_smem[(wv * 52 + bin) * 32 + lane];
    return t;
  };
  if (ch < c) {
    const float n16 = 1.f / (float)((h / 4) * (w / 4));
    for (int i = ty; i < 16; i += 8) st_f<T>(y4 + ((int64_t)b * 16 + i) * c + ch, total(i) * n16);
    for (int i = ty; i < 4; i += 8) {
      const int qy = i / 2, qx = i % 2;
      const float t = total((2 * qy) * 4 + 2 * qx) + total((2 * qy) * 4 + 2 * qx + 1) + total((2 * qy + 1) * 4 + 2 * qx) +
                      total((2 * qy + 1) * 4 + 2 * qx + 1);
      st_f<T>(y2 + ((int64_t)b * 4 + i) * c + ch, t * n16 * 0.25f);
    }
    if (ty == 0) {
      float t = 0.f;
      for (int i = 0; i < 16; ++i) t += total(i);
      st_f<T>(y1 + (int64_t)b * c + ch, t * n16 * (1.f / 16.f));
    }
    for (int i = ty; i < 36; i += 8) {
      const int pi = i / 6, pj = i % 6;
      const int cnt = (((pi + 1) * h + 5) / 6 - (pi * h) / 6) * (((pj + 1) * w + 5) / 6 - (pj * w) / 6);
      st_f<T>(y6 + ((int64_t)b * 36 + i) * c + ch, total(16 + i) / (float)cnt);
    }
  }
}

// ------------------------------------------------- pyramid branches: 1x1 conv + bilinear resize + concat ----
// The four branches of PyramidPoolingModule behind the pooling (reference models/modules.py:145-156): conv1x1 of the
// 1x1 / 2x2 / 4x4 / 6x6 pooled maps (57 pixels in all) and the align_corners=True resize of each result to h x w, written
// into its channel slice of the concat buffer.  As eight launches (four tensor-core convolutions on 1-36 pixels, four
// resizes) this was latency only; here a block recomputes the 57 x hid conv outputs of its image into shared memory
// (bf16 inputs and bf16-rounded weights, fp32 accumulate, result rounded to bf16 like the stored tensor it replaces)
// and resizes its share of the output pixels, 16-byte channel vectors.
constexpr int kPpmPix = 1 + 4 + 16 + 36;
__global__ void __launch_bounds__(256) ppm_branches_kernel(int h, int w, int c, int hid, const __nv_bfloat16 *__restrict__ y1,
                                                           const __nv_bfloat16 *__restrict__ y2,
                                                           const __nv_bfloat16 *__restrict__ y4,
                                                           const __nv_bfloat16 *__restrict__ y6,
                                                           const float *__restrict__ wgt, const float *__restrict__ bias,
                                                           __nv_bfloat16 *__restrict__ cat, int y_ct, int y_co) {
  // shared memory: z [57][hid] fp32 | pooled inputs [57][c] bf16 | bf16-rounded weights [4 * hid][c + 8] (rows padded by
  // 16 bytes: a quarter warp's eight rows then cover 128 contiguous bank bytes).  Every output re-reading its weight row
  // from L2 made the first version slower than the eight launches it replaced.
  extern __shared__ __align__(16) unsigned char ppm_smem[];
  float *ppm_z = reinterpret_cast<float *>(ppm_smem);
  __nv_bfloat16 *sx = reinterpret_cast<__nv_bfloat16 *>(ppm_smem + (size_t)kPpmPix * hid * 4);
  __nv_bfloat16 *sw = sx + (size_t)kPpmPix * c;
  const int wp = c + 8;
  const int b = blockIdx.x;
  const int psz[4] = {1, 2, 4, 6}, pbase[4] = {0, 1, 5, 21};
  const __nv_bfloat16 *src[4] = {y1 + (int64_t)b * c, y2 + (int64_t)b * 4 * c, y4 + (int64_t)b * 16 * c, y6 + (int64_t)b * 36 * c};
  for (int br = 0; br < 4; ++br) {
    const int nv = psz[br] * psz[br] * c / 8;
    const uint4 *g = reinterpret_cast<const uint4 *>(src[br]);
    uint4 *d = reinterpret_cast<uint4 *>(sx + (size_t)pbase[br] * c);
    for (int i = threadIdx.x; i < nv; i += blockDim.x) d[i] = __ldg(g + i);
  }
  for (int i = threadIdx.x; i < 4 * hid * c / 4; i += blockDim.x) {
    const float4 f = __ldg(reinterpret_cast<const float4 *>(wgt) + i);
    const int row = (i * 4) / c, col = (i * 4) - row * c;
    const __nv_bfloat162 lo = __floats2bfloat162_rn(f.x, f.y), hi = __floats2bfloat162_rn(f.z, f.w);
    *reinterpret_cast<uint2 *>(sw + (size_t)row * wp + col) =
        make_uint2(*reinterpret_cast<const uint32_t *>(&lo), *reinterpret_cast<const uint32_t *>(&hi));
  }
  __syncthreads();
  for (int o = threadIdx.x; o < kPpmPix * hid; o += blockDim.x) {
    const int pix = o / hid, k = o - pix * hid;
    const int br = pix >= 21 ? 3 : (pix >= 5 ? 2 : (pix >= 1 ? 1 : 0));
    const uint4 *xp = reinterpret_cast<const uint4 *>(sx + (size_t)pix * c);
    const uint4 *wr = reinterpret_cast<const uint4 *>(sw + (size_t)(br * hid + k) * wp);
    float acc = 0.f;
    for (int i = 0; i < c / 8; ++i) {
      const uint4 q = xp[i], wq = wr[i];
      const uint32_t qs[4] = {q.x, q.y, q.z, q.w}, ws[4] = {wq.x, wq.y, wq.z, wq.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        acc = fmaf(__uint_as_float(qs[e] << 16), __uint_as_float(ws[e] << 16), acc);
        acc = fmaf(__uint_as_float(qs[e] & 0xffff0000u), __uint_as_float(ws[e] & 0xffff0000u), acc);
      }
    }
    if (bias) acc += bias[br * hid + k];
    ppm_z[o] = __bfloat162float(__float2bfloat16_rn(acc));
  }
  __syncthreads();
  const int vec_per_pix = 4 * hid / 8;                       // 16-byte vectors of one output pixel (all four branches)
  const int per_blk = (h * w + gridDim.y - 1) / gridDim.y;
  const int p0 = blockIdx.y * per_blk, p1 = min(h * w, p0 + per_blk);
  for (int t = p0 * vec_per_pix + threadIdx.x; t < p1 * vec_per_pix; t += blockDim.x) {
    const int pix = t / vec_per_pix, v = t - pix * vec_per_pix;
    const int oy = pix / w, ox = pix - oy * w;
    const int br = (v * 8) / hid, k0 = v * 8 - br * hid;
    const int ps = psz[br];
    int ya, yb, xa, xb;
    float ly0, ly1, lx0, lx1;
    src_index(oy, resize_scale(ps, h, 1), ps, 1, ya, yb, ly0, ly1);
    src_index(ox, resize_scale(ps, w, 1), ps, 1, xa, xb, lx0, lx1);
    const float *z = ppm_z + (int64_t)pbase[br] * hid + k0;
    const float *z00 = z + (ya * ps + xa) * hid, *z01 = z + (ya * ps + xb) * hid;
    const float *z10 = z + (yb * ps + xa) * hid, *z11 = z + (yb * ps + xb) * hid;
    uint32_t pk[4];
#pragma unroll
    for (int e = 0; e < 8; e += 2) {
      const float o0 = ly0 * (lx0 * z00[e] + lx1 * z01[e]) + ly1 * (lx0 * z10[e] + lx1 * z11[e]);
      const float o1 = ly0 * (lx0 * z00[e + 1] + lx1 * z01[e + 1]) + ly1 * (lx0 * z10[e + 1] + lx1 * z11[e + 1]);
      const __nv_bfloat162 hh = __floats2bfloat162_rn(o0, o1);
      pk[e >> 1] = *reinterpret_cast<const uint32_t *>(&hh);
    }
    *reinterpret_cast<uint4 *>(cat + ((int64_t)b * h * w + pix) * y_ct + y_co + v * 8) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
  }
}

// ------------------------------------------------- avg-pool 3x3 s2 p1, channel shuffle ----
// F.avg_pool2d(x, 3, 2, 1) of DABNet's input pyramid (reference models/dabnet.py:32-34): padding counts
// (count_include_pad=True, the default), i.e. always divide by 9.  Input strides are explicit so that the
// caller's NCHW image and our NHWC buffers both work; output is an NHWC channel slice.
template <typename TI, typename TO>
__global__ void __launch_bounds__(256) avgpool3x3s2_kernel(int n, int h, int w, int c, int ho, int wo, int64_t sn,
                                                           int64_t sc, int64_t sh, int64_t sw, const TI *__restrict__ x,
                                                           TO *__restrict__ y, int y_ct, int y_co, ImgNorm nz) {
  const int64_t total = (int64_t)n * ho * wo * c;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= total) return;
  const int ch = (int)(tid % c);
  int64_t r = tid / c;
  const int ox = (int)(r % wo);
  r /= wo;
  const int oy = (int)(r % ho);
  const int b = (int)(r / ho);
  float s = 0.f;
#pragma unroll
  for (int ky = 0; ky < 3; ++ky) {
    const int iy = oy * 2 - 1 + ky;
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      const int ix = ox * 2 - 1 + kx;
      if (iy >= 0 && iy < h && ix >= 0 && ix < w) s += ld_img<TI>(x + b * sn + ch * sc + iy * sh + ix * sw, nz, ch & 3);
    }
  }
  st_f<TO>(y + (((int64_t)b * ho + oy) * wo + ox) * y_ct + y_co + ch, s * (1.f / 9.f));
}

// channel_shuffle (reference models/modules.py:18-32): out[.., i*groups + g] = in[.., g*(c/groups) + i].
template <typename T>
__global__ void __launch_bounds__(256) channel_shuffle_kernel(int64_t npix, int c, int groups, const T *__restrict__ x,
                                                              int x_ct, int x_co, T *__restrict__ y, int y_ct, int y_co) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= npix * c) return;
  const int k = (int)(tid % c);
  const int64_t pix = tid / c;
  const int cpg = c / groups;
  const int g = k % groups, i = k / groups;
  y[pix * y_ct + y_co + k] = x[pix * x_ct + x_co + g * cpg + i];
}

// groups == 2, 16-byte output chunks: chunk k of a pixel interleaves 8 bytes of the first half with 8 bytes of the
// second half (out[2i] = in[i], out[2i+1] = in[c/2 + i]).  ES = element size in bytes (2 or 4).
template <int ES>
__global__ void __launch_bounds__(256) channel_shuffle2_kernel(int64_t npix, int chunks, int half_bytes,
                                                               const char *__restrict__ x, int64_t x_pitch,
                                                               char *__restrict__ y, int64_t y_pitch) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= npix * chunks) return;
  const int k = (int)(tid % chunks);
  const int64_t pix = tid / chunks;
  const char *xp = x + pix * x_pitch + k * 8;
  const uint2 a = __ldg(reinterpret_cast<const uint2 *>(xp));
  const uint2 b = __ldg(reinterpret_cast<const uint2 *>(xp + half_bytes));
  uint4 o;
  if constexpr (ES == 2) {
    o.x = __byte_perm(a.x, b.x, 0x5410), o.y = __byte_perm(a.x, b.x, 0x7632);
    o.z = __byte_perm(a.y, b.y, 0x5410), o.w = __byte_perm(a.y, b.y, 0x7632);
  } else {
    o = make_uint4(a.x, b.x, a.y, b.y);
  }
  *reinterpret_cast<uint4 *>(y + pix * y_pitch + k * 16) = o;
}

// ---- predict epilogue: class map -> colour image (+ alpha blend with the input picture) --------------------
// Reference core/seg_trainer.py:172-191: `colormap[preds.max(dim=1)[1]]` (a LUT gather into an int64 HxWx3 tensor),
// D2H, `astype(uint8)`, PIL `Image.blend(image, pred, alpha)`.  Here: four pixels per thread, LUT in shared
// memory, uint8 HWC out; the blend restates Pillow's ImagingBlend for 0 <= alpha <= 1,
// (UINT8)(in1 + alpha * (in2 - in1)) in single precision WITHOUT fused multiply-add, then truncation.
__global__ void __launch_bounds__(256) colorize_kernel(int64_t npix, const uint8_t *__restrict__ labels,
                                                       const uint8_t *__restrict__ lut, const uint8_t *__restrict__ image,
                                                       float alpha, uint8_t *__restrict__ out) {
  __shared__ uint8_t s_lut[768];
  for (int i = threadIdx.x; i < 768; i += blockDim.x) s_lut[i] = lut[i];
  __syncthreads();
  const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // group of four pixels
  const int64_t p0 = q * 4;
  if (p0 >= npix) return;
  uint8_t rgb[12];
  const bool full = p0 + 4 <= npix;
  uint32_t l4 = 0;
  if (full && (reinterpret_cast<uintptr_t>(labels) & 3) == 0) {
    l4 = __ldg(reinterpret_cast<const uint32_t *>(labels + p0));
  } else {
    for (int i = 0; i < 4 && p0 + i < npix; ++i) l4 |= (uint32_t)labels[p0 + i] << (8 * i);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int l = (l4 >> (8 * i)) & 255;
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) rgb[i * 3 + ch] = s_lut[l * 3 + ch];
  }
  if (image) {
#pragma unroll
    for (int i = 0; i < 12; ++i) {
      if (p0 * 3 + i < npix * 3) {
        const int in1 = image[p0 * 3 + i], in2 = rgb[i];
        const float t = __fadd_rn((float)in1, __fmul_rn(alpha, (float)(in2 - in1)));
        rgb[i] = (uint8_t)(int)t;
      }
    }
  }
  if (full && (reinterpret_cast<uintptr_t>(out) & 3) == 0) {
    uint32_t *o = reinterpret_cast<uint32_t *>(out + p0 * 3);
#pragma unroll
    for (int k = 0; k < 3; ++k)
      o[k] = (uint32_t)rgb[4 * k] | ((uint32_t)rgb[4 * k + 1] << 8) | ((uint32_t)rgb[4 * k + 2] << 16) | ((uint32_t)rgb[4 * k + 3] << 24);
  } else {
    for (int i = 0; i < 12 && p0 * 3 + i < npix * 3; ++i) out[p0 * 3 + i] = rgb[i];
  }
}

}  // namespace rtseg

using namespace rtseg;

extern "C" int rtseg_colorize_blend(int64_t npix, const uint8_t *labels, const uint8_t *lut, const uint8_t *image,
                                    float alpha, uint8_t *out, rtseg_stream_t stream) {
  RTSEG_REQUIRE(labels && lut && out, "NULL pointer");
  RTSEG_REQUIRE(npix > 0, "non-positive extent");
  RTSEG_REQUIRE(!image || (alpha >= 0.f && alpha <= 1.f), "blend alpha must be in [0, 1] (Pillow's interpolation branch)");
  const unsigned grid = (unsigned)cdiv(cdiv(npix, 4), 256);
  colorize_kernel<<<grid, 256, 0, as_stream(stream)>>>(npix, labels, lut, image, alpha, out);
  RTSEG_LAUNCH_CHECK("colorize_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_upsample_bilinear(int32_t n, int32_t hi, int32_t wi, int32_t c, int32_t ho, int32_t wo,
                                       int32_t align, int32_t dtype, const void *x, int32_t x_ct, int32_t x_co,
                                       void *y, int32_t y_ct, int32_t y_co, rtseg_stream_t stream) {
  RTSEG_REQUIRE(x && y, "NULL tensor pointer");
  RTSEG_REQUIRE(n > 0 && hi > 0 && wi > 0 && c > 0 && ho > 0 && wo > 0, "non-positive extent");
  RTSEG_REQUIRE(x_ct >= x_co + c && y_ct >= y_co + c && x_co >= 0 && y_co >= 0, "channel slice out of range");
  const float sy = resize_scale(hi, ho, align), sx = resize_scale(wi, wo, align);
  cudaStream_t st = as_stream(stream);
  const int V = dtype == RTSEG_BF16 ? 8 : 4;
  const bool vec = c % V == 0 && x_ct % V == 0 && x_co % V == 0 && y_ct % V == 0 && y_co % V == 0 &&
                   reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(y) % 16 == 0;
  const int64_t total = (int64_t)n * ho * wo * (vec ? c / V : c);
  const unsigned grid = (unsigned)cdiv(total, 256);
  static const int rows_knob = getenv("RTSEG_BILINEAR_ROWS") ? atoi(getenv("RTSEG_BILINEAR_ROWS")) : 1;
  const bool rows4 = rows_knob && vec && ho >= 2 * hi;   // upsampling: output rows share their source rows
  const unsigned grid_rows = (unsigned)cdiv((int64_t)n * cdiv(ho, 4) * wo * (c / V), 256);
  if (dtype == RTSEG_BF16) {
    auto xs = (const __nv_bfloat16 *)x;
    auto ys = (__nv_bfloat16 *)y;
    if (vec && rows4)
      bilinear_rows_kernel<__nv_bfloat16, 8, 4><<<grid_rows, 256, 0, st>>>(n, hi, wi, c, ho, wo, align, sy, sx, xs, x_ct, x_co, ys,
                                                                           y_ct, y_co);
    else if (vec)
      bilinear_kernel<__nv_bfloat16, 8><<<grid, 256, 0, st>>>(n, hi, wi, c, ho, wo, align, sy, sx, xs, x_ct, x_co, ys, y_ct, y_co);
    else
      bilinear_kernel<__nv_bfloat16, 1><<<grid, 256, 0, st>>>(n, hi, wi, c, ho, wo, align, sy, sx, xs, x_ct, x_co, ys, y_ct, y_co);
  } else if (dtype == RTSEG_F32) {
    auto xs = (const float *)x;
    auto ys = (float *)y;
    if (vec)
      bilinear_kernel<float, 4><<<grid, 256, 0, st>>>(n, hi, wi, c, ho, wo, align, sy, sx, xs, x_ct, x_co, ys, y_ct, y_co);
    else
      bilinear_kernel<float, 1><<<grid, 256, 0, st>>>(n, hi, wi, c, ho, wo, align, sy, sx, xs, x_ct, x_co, ys, y_ct, y_co);
  } else {
    RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
  }
  RTSEG_LAUNCH_CHECK("bilinear_kernel");
  return RTSEG_OK;
}

template <typename T>
static int launch_upsample_argmax(int n, int hi, int wi, int c, int ct, int ho, int wo, int align, const void *x,
                                  uint8_t *l8, int64_t *l64, const int64_t *target, int ignore, int64_t *confmat,
                                  cudaStream_t st) {
  const float sy = resize_scale(hi, ho, align), sx = resize_scale(wi, wo, align);
  dim3 grid((unsigned)cdiv(wo, kHeadThreads), (unsigned)cdiv(ho, kHeadRows), (unsigned)n);
  auto cm = reinterpret_cast<unsigned long long *>(confmat);
  const T *xs = static_cast<const T *>(x);
#define RTSEG_HEAD(CM)                                                                                              \
  if (c <= CM) {                                                                                                    \
    const int vec = sizeof(T) == 2 && ct % 8 == 0 && reinterpret_cast<uintptr_t>(x) % 16 == 0 && (CM + 7) / 8 * 8 <= ct; \
    upsample_argmax_kernel<T, CM><<<grid, kHeadThreads, 0, st>>>(n, hi, wi, c, ct, ho, wo, align, sy, sx, xs, l8, l64, \
                                                                 target, ignore, cm, vec);                         \
    RTSEG_LAUNCH_CHECK("upsample_argmax_kernel");                                                                   \
    return RTSEG_OK;                                                                                                \
  }
  if (c == 19) {  // Cityscapes: exact instantiation, no padded classes
    const int vec = sizeof(T) == 2 && ct % 8 == 0 && reinterpret_cast<uintptr_t>(x) % 16 == 0 && 24 <= ct;
    if (l8 && !l64 && !target)
      upsample_argmax_kernel<T, 19, true, true><<<grid, kHeadThreads, 0, st>>>(n, hi, wi, c, ct, ho, wo, align, sy, sx, xs, l8,
                                                                               l64, target, ignore, cm, vec);
    else
      upsample_argmax_kernel<T, 19, true><<<grid, kHeadThreads, 0, st>>>(n, hi, wi, c, ct, ho, wo, align, sy, sx, xs, l8, l64,
                                                                         target, ignore, cm, vec);
    RTSEG_LAUNCH_CHECK("upsample_argmax_kernel");
    return RTSEG_OK;
  }
  RTSEG_HEAD(8)
  RTSEG_HEAD(16)
  RTSEG_HEAD(24)
  RTSEG_HEAD(32)
#undef RTSEG_HEAD
  RTSEG_FAIL(RTSEG_ERR_UNSUPPORTED, "rtseg_upsample_argmax supports at most 32 classes (got %d)", c);
}

extern "C" int rtseg_upsample_argmax(int32_t n, int32_t hi, int32_t wi, int32_t c, int32_t x_ctotal, int32_t ho,
                                     int32_t wo, int32_t align, int32_t dtype, const void *x, uint8_t *labels_u8,
                                     int64_t *labels_i64, const int64_t *target, int32_t ignore_index,
                                     int64_t *confmat, rtseg_stream_t stream) {
  RTSEG_REQUIRE(x, "NULL logits pointer");
  RTSEG_REQUIRE(n > 0 && hi > 0 && wi > 0 && c > 0 && ho > 0 && wo > 0, "non-positive extent");
  RTSEG_REQUIRE(x_ctotal >= c, "x_ctotal (%d) smaller than the class count (%d)", x_ctotal, c);
  RTSEG_REQUIRE(n <= 65535 && cdiv(ho, kHeadRows) <= 65535, "batch or height too large for one launch");
  RTSEG_REQUIRE((int64_t)hi * wi * x_ctotal < INT32_MAX, "one image of logits must hold fewer than 2^31 elements");
  RTSEG_REQUIRE((target == nullptr) == (confmat == nullptr), "target and confmat must be given together");
  RTSEG_REQUIRE(labels_u8 || labels_i64 || confmat, "nothing to compute");
  cudaStream_t st = as_stream(stream);
  if (dtype == RTSEG_BF16)
    return launch_upsample_argmax<__nv_bfloat16>(n, hi, wi, c, x_ctotal, ho, wo, align, x, labels_u8, labels_i64, target,
                                                 ignore_index, confmat, st);
  if (dtype == RTSEG_F32)
    return launch_upsample_argmax<float>(n, hi, wi, c, x_ctotal, ho, wo, align, x, labels_u8, labels_i64, target, ignore_index,
                                         confmat, st);
  RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
}

extern "C" int rtseg_argmax_nhwc(int64_t npix, int32_t c, int32_t c_total, int32_t dtype, const void *x,
                                 uint8_t *labels_u8, int64_t *labels_i64, rtseg_stream_t stream) {
  RTSEG_REQUIRE(x && (labels_u8 || labels_i64), "NULL pointer");
  RTSEG_REQUIRE(npix > 0 && c > 0 && c_total >= c, "bad extent");
  RTSEG_REQUIRE(labels_u8 == nullptr || c <= 256, "uint8 labels need c <= 256");
  cudaStream_t st = as_stream(stream);
  const unsigned grid = (unsigned)cdiv(npix, 256);
  if (dtype == RTSEG_BF16)
    argmax_nhwc_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(npix, c, c_total, (const __nv_bfloat16 *)x, labels_u8, labels_i64);
  else if (dtype == RTSEG_F32)
    argmax_nhwc_kernel<float><<<grid, 256, 0, st>>>(npix, c, c_total, (const float *)x, labels_u8, labels_i64);
  else
    RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
  RTSEG_LAUNCH_CHECK("argmax_nhwc_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_argmax_nchw(int32_t n, int32_t c, int64_t hw, int32_t dtype, const void *x, uint8_t *labels_u8,
                                 int64_t *labels_i64, rtseg_stream_t stream) {
  RTSEG_REQUIRE(x && (labels_u8 || labels_i64), "NULL pointer");
  RTSEG_REQUIRE(n > 0 && c > 0 && hw > 0, "bad extent");
  RTSEG_REQUIRE(labels_u8 == nullptr || c <= 256, "uint8 labels need c <= 256");
  cudaStream_t st = as_stream(stream);
  const unsigned grid = (unsigned)cdiv((int64_t)n * hw, 256);
  if (dtype == RTSEG_BF16)
    argmax_nchw_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(n, c, hw, (const __nv_bfloat16 *)x, labels_u8, labels_i64);
  else if (dtype == RTSEG_F32)
    argmax_nchw_kernel<float><<<grid, 256, 0, st>>>(n, c, hw, (const float *)x, labels_u8, labels_i64);
  else
    RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
  RTSEG_LAUNCH_CHECK("argmax_nchw_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_confmat_update(int64_t npix, int32_t c, int32_t ignore_index, const void *pred,
                                    int32_t pred_is_i64, const int64_t *target, int64_t *confmat, int64_t *bad_count,
                                    rtseg_stream_t stream) {
  RTSEG_REQUIRE(pred && target && confmat, "NULL pointer");
  RTSEG_REQUIRE(npix >= 0 && c > 0 && c <= 32, "confusion matrix supports 1..32 classes (got %d)", c);
  if (npix == 0) return RTSEG_OK;
  cudaStream_t st = as_stream(stream);
  const unsigned grid = (unsigned)std::min<int64_t>(cdiv(npix, 256 * 8), (int64_t)num_sms() * 8);
  auto cm = reinterpret_cast<unsigned long long *>(confmat);
  auto bad = reinterpret_cast<unsigned long long *>(bad_count);
  if (pred_is_i64)
    confmat_kernel<int64_t><<<grid, 256, 0, st>>>(npix, c, ignore_index, (const int64_t *)pred, target, cm, bad);
  else
    confmat_kernel<uint8_t><<<grid, 256, 0, st>>>(npix, c, ignore_index, (const uint8_t *)pred, target, cm, bad);
  RTSEG_LAUNCH_CHECK("confmat_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_nchw_to_nhwc(int32_t n, int32_t c, int32_t h, int32_t w, int32_t src_dtype, const void *src,
                                  int32_t dst_dtype, void *dst, int32_t dst_ctotal, int32_t dst_coff,
                                  rtseg_stream_t stream) {
  RTSEG_REQUIRE(src && dst, "NULL pointer");
  RTSEG_REQUIRE(n > 0 && c > 0 && h > 0 && w > 0 && dst_ctotal >= dst_coff + c && dst_coff >= 0, "bad extent");
  const int64_t hw = (int64_t)h * w;
  RTSEG_REQUIRE(hw < (int64_t)INT32_MAX, "image too large");
  void *d = static_cast<char *>(dst) + (size_t)dst_coff * (dst_dtype == RTSEG_BF16 ? 2 : 4);
  // src per image: [c rows][hw cols]  ->  dst [hw][ctotal]
  return dispatch_transpose(n, c, (int)hw, src_dtype, src, hw, (int64_t)c * hw, dst_dtype, d, dst_ctotal,
                            hw * dst_ctotal, as_stream(stream));
}

extern "C" int rtseg_nhwc_to_nchw(int32_t n, int32_t c, int32_t h, int32_t w, int32_t src_dtype, const void *src,
                                  int32_t src_ctotal, int32_t src_coff, int32_t dst_dtype, void *dst,
                                  rtseg_stream_t stream) {
  RTSEG_REQUIRE(src && dst, "NULL pointer");
  RTSEG_REQUIRE(n > 0 && c > 0 && h > 0 && w > 0 && src_ctotal >= src_coff + c && src_coff >= 0, "bad extent");
  const int64_t hw = (int64_t)h * w;
  RTSEG_REQUIRE(hw < (int64_t)INT32_MAX, "image too large");
  const void *s = static_cast<const char *>(src) + (size_t)src_coff * (src_dtype == RTSEG_BF16 ? 2 : 4);
  // src per image: [hw rows][c cols] (pitch ctotal)  ->  dst [c][hw]
  return dispatch_transpose(n, (int)hw, c, src_dtype, s, src_ctotal, hw * src_ctotal, dst_dtype, dst, hw,
                            (int64_t)c * hw, as_stream(stream));
}

extern "C" int rtseg_adaptive_avgpool(int32_t n, int32_t h, int32_t w, int32_t c, int32_t p, int32_t dtype,
                                      const void *x, int32_t x_ct, int32_t x_co, int32_t y_dtype, void *y,
                                      rtseg_stream_t stream) {
  RTSEG_REQUIRE(x && y, "NULL pointer");
  RTSEG_REQUIRE(n > 0 && h > 0 && w > 0 && c > 0 && p > 0 && x_ct >= x_co + c && x_co >= 0, "bad extent");
  RTSEG_REQUIRE(n <= 65535 && cdiv(c, 32) <= 65535, "extent too large");
  dim3 grid((unsigned)(p * p), (unsigned)n, (unsigned)cdiv(c, 32));
  cudaStream_t st = as_stream(stream);
#define RTSEG_AAP(TI_, TO_) \
  adaptive_avgpool_kernel<TI_, TO_><<<grid, 256, 0, st>>>(h, w, c, p, (const TI_ *)x, x_ct, x_co, (TO_ *)y)
  if (dtype == RTSEG_BF16 && y_dtype == RTSEG_BF16) RTSEG_AAP(__nv_bfloat16, __nv_bfloat16);
  else if (dtype == RTSEG_BF16 && y_dtype == RTSEG_F32) RTSEG_AAP(__nv_bfloat16, float);
  else if (dtype == RTSEG_F32 && y_dtype == RTSEG_F32) RTSEG_AAP(float, float);
  else if (dtype == RTSEG_F32 && y_dtype == RTSEG_BF16) RTSEG_AAP(float, __nv_bfloat16);
  else RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
#undef RTSEG_AAP
  RTSEG_LAUNCH_CHECK("adaptive_avgpool_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_pyramid_avgpool(int32_t n, int32_t h, int32_t w, int32_t c, int32_t dtype, const void *x, int32_t x_ct,
                                     int32_t x_co, void *y1, void *y2, void *y4, void *y6, rtseg_stream_t stream) {
  RTSEG_REQUIRE(x && y1 && y2 && y4 && y6, "NULL pointer");
  RTSEG_REQUIRE(n > 0 && h >= 4 && w >= 4 && c > 0 && x_ct >= x_co + c && x_co >= 0, "bad extent");
  RTSEG_REQUIRE(h % 4 == 0 && w % 4 == 0, "rtseg_pyramid_avgpool: h and w must be multiples of 4");
  RTSEG_REQUIRE(cdiv(c, 32) <= 65535, "extent too large");
  dim3 grid((unsigned)n, (unsigned)cdiv(c, 32));
  cudaStream_t st = as_stream(stream);
  constexpr size_t kSmem = 8 * 52 * 32 * sizeof(float);   // 53,248 B: above the 48 KB default
  static bool attr_set[64] = {false};
  int dev = 0;
  RTSEG_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    RTSEG_CUDA(cudaFuncSetAttribute(pyramid_pool_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem));
    RTSEG_CUDA(cudaFuncSetAttribute(pyramid_pool_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem));
    attr_set[dev] = true;
  }
  if (dtype == RTSEG_BF16)
    pyramid_pool_kernel<__nv_bfloat16><<<grid, 256, kSmem, st>>>(h, w, c, (const __nv_bfloat16 *)x, x_ct, x_co, (__nv_bfloat16 *)y1,
                                                                 (__nv_bfloat16 *)y2, (__nv_bfloat16 *)y4, (__nv_bfloat16 *)y6);
  else if (dtype == RTSEG_F32)
    pyramid_pool_kernel<float><<<grid, 256, kSmem, st>>>(h, w, c, (const float *)x, x_ct, x_co, (float *)y1, (float *)y2,
                                                         (float *)y4, (float *)y6);
  else
    RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
  RTSEG_LAUNCH_CHECK("pyramid_pool_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_ppm_branches(int32_t n, int32_t h, int32_t w, int32_t c, int32_t hid, int32_t dtype, const void *y1,
                                  const void *y2, const void *y4, const void *y6, const float *weights, const float *bias,
                                  void *cat, int32_t y_ctotal, int32_t y_coff, rtseg_stream_t stream) {
  RTSEG_REQUIRE(y1 && y2 && y4 && y6 && weights && cat, "NULL pointer");
  RTSEG_REQUIRE(dtype == RTSEG_BF16, "rtseg_ppm_branches: bf16 activations only");
  RTSEG_REQUIRE(n > 0 && h > 1 && w > 1 && c > 0 && hid > 0 && n <= 65535, "bad extent");
  RTSEG_REQUIRE(c % 8 == 0 && hid % 8 == 0, "rtseg_ppm_branches: c and hid must be multiples of 8");
  RTSEG_REQUIRE(y_coff >= 0 && y_coff % 8 == 0 && y_ctotal % 8 == 0 && y_coff + 4 * hid <= y_ctotal, "channel slice outside its buffer or unaligned");
  RTSEG_REQUIRE(reinterpret_cast<uintptr_t>(cat) % 16 == 0 && reinterpret_cast<uintptr_t>(weights) % 16 == 0 &&
                    reinterpret_cast<uintptr_t>(y1) % 16 == 0 && reinterpret_cast<uintptr_t>(y2) % 16 == 0 &&
                    reinterpret_cast<uintptr_t>(y4) % 16 == 0 && reinterpret_cast<uintptr_t>(y6) % 16 == 0,
                "pointers must be 16-byte aligned");
  const size_t smem = (size_t)kPpmPix * hid * sizeof(float) + (size_t)kPpmPix * c * 2 + (size_t)4 * hid * (c + 8) * 2;
  RTSEG_REQUIRE(smem <= 200 * 1024, "rtseg_ppm_branches: c x hid too large for the shared-memory plan");
  static bool attr_set[64] = {false};
  int dev = 0;
  RTSEG_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    RTSEG_CUDA(cudaFuncSetAttribute(ppm_branches_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr_set[dev] = true;
  }
  // blocks per image: each recomputes the 57 x hid conv outputs, so few of them — but enough to fill the GPU
  const int split = std::max(1, std::min((int)cdiv((int64_t)h * w, 256), (int)cdiv(2 * num_sms(), n)));
  dim3 grid((unsigned)n, (unsigned)std::min(split, 16));
  ppm_branches_kernel<<<grid, 256, smem, as_stream(stream)>>>(h, w, c, hid, (const __nv_bfloat16 *)y1, (const __nv_bfloat16 *)y2,
                                                              (const __nv_bfloat16 *)y4, (const __nv_bfloat16 *)y6, weights, bias,
                                                              (__nv_bfloat16 *)cat, y_ctotal, y_coff);
  RTSEG_LAUNCH_CHECK("ppm_branches_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_upsample_argmax_host(int32_t n, int32_t hi, int32_t wi, int32_t c, int32_t ho, int32_t wo,
                                          int32_t align, int32_t dtype, const void *x_host, uint8_t *labels_host,
                                          void *dev_scratch, rtseg_stream_t stream) {
  RTSEG_REQUIRE(x_host && labels_host && dev_scratch, "NULL pointer");
  RTSEG_REQUIRE(n > 0 && hi > 0 && wi > 0 && c > 0 && ho > 0 && wo > 0, "non-positive extent");
  const size_t es = dtype == RTSEG_BF16 ? 2 : 4;
  size_t in_bytes = (size_t)n * hi * wi * c * es;
  const size_t out_bytes = (size_t)n * ho * wo;
  cudaStream_t st = as_stream(stream);
  // the uint8 class map sits directly behind the logits (a byte buffer needs no alignment): the documented scratch
  // size, logits bytes + n*ho*wo, is exactly what is touched (rtseg_upsample_argmax_host_scratch_bytes)
  uint8_t *dlab = static_cast<uint8_t *>(dev_scratch) + in_bytes;
  RTSEG_CUDA(cudaMemcpyAsync(dev_scratch, x_host, in_bytes, cudaMemcpyHostToDevice, st));
  if (int rc = rtseg_upsample_argmax(n, hi, wi, c, c, ho, wo, align, dtype, dev_scratch, dlab, nullptr, nullptr, 255,
                                     nullptr, stream))
    return rc;
  RTSEG_CUDA(cudaMemcpyAsync(labels_host, dlab, out_bytes, cudaMemcpyDeviceToHost, st));
  RTSEG_CUDA(cudaStreamSynchronize(st));
  return RTSEG_OK;
}

extern "C" int rtseg_resize_nearest_i64(int32_t n, int32_t hi, int32_t wi, int32_t ho, int32_t wo, const int64_t *x, int64_t *y,
                                        rtseg_stream_t stream) {
  RTSEG_REQUIRE(x && y, "NULL pointer");
  RTSEG_REQUIRE(n > 0 && hi > 0 && wi > 0 && ho > 0 && wo > 0, "non-positive extent");
  const int64_t total = (int64_t)n * ho * wo;
  resize_nearest_i64_kernel<<<(unsigned)cdiv(total, 256), 256, 0, as_stream(stream)>>>(
      n, hi, wi, ho, wo, (float)hi / (float)ho, (float)wi / (float)wo, x, y);
  RTSEG_LAUNCH_CHECK("resize_nearest_i64_kernel");
  return RTSEG_OK;
}

extern "C" size_t rtseg_upsample_argmax_host_scratch_bytes(int32_t n, int32_t hi, int32_t wi, int32_t c, int32_t ho,
                                                          int32_t wo, int32_t dtype) {
  if (n <= 0 || hi <= 0 || wi <= 0 || c <= 0 || ho <= 0 || wo <= 0) return 0;
  return (size_t)n * hi * wi * c * (dtype == RTSEG_BF16 ? 2 : 4) + (size_t)n * ho * wo;
}

static int avgpool3x3s2_impl(int32_t n, int32_t h, int32_t w, int32_t c, int32_t x_is_nchw, int32_t x_dtype,
                             const void *x, int32_t x_ctotal, int32_t x_coff, int32_t y_dtype, void *y,
                             int32_t y_ctotal, int32_t y_coff, const ImgNorm &nz, rtseg_stream_t stream) {
  RTSEG_REQUIRE(x && y, "NULL pointer");
  RTSEG_REQUIRE(n > 0 && h > 0 && w > 0 && c > 0 && y_ctotal >= y_coff + c && y_coff >= 0, "bad extent");
  const int ho = (h - 1) / 2 + 1, wo = (w - 1) / 2 + 1;
  int64_t sn, sc, sh, sw;
  if (x_is_nchw) {
    sn = (int64_t)c * h * w, sc = (int64_t)h * w, sh = w, sw = 1;
  } else {
    RTSEG_REQUIRE(x_ctotal >= x_coff + c && x_coff >= 0, "input channel slice out of range");
    sn = (int64_t)h * w * x_ctotal, sc = 1, sh = (int64_t)w * x_ctotal, sw = x_ctotal;
    x = static_cast<const char *>(x) + (size_t)x_coff * (x_dtype == RTSEG_BF16 ? 2 : (x_dtype == RTSEG_U8 ? 1 : 4));
  }
  const int64_t total = (int64_t)n * ho * wo * c;
  const unsigned grid = (unsigned)cdiv(total, 256);
  cudaStream_t st = as_stream(stream);
#define RTSEG_AVG(TI_, TO_)                                                                                     \
  avgpool3x3s2_kernel<TI_, TO_><<<grid, 256, 0, st>>>(n, h, w, c, ho, wo, sn, sc, sh, sw, (const TI_ *)x, (TO_ *)y, \
                                                      y_ctotal, y_coff, nz)
  if (x_dtype == RTSEG_U8 && y_dtype == RTSEG_F32) RTSEG_AVG(uint8_t, float);
  else if (x_dtype == RTSEG_U8 && y_dtype == RTSEG_BF16) RTSEG_AVG(uint8_t, __nv_bfloat16);
  else if (x_dtype == RTSEG_F32 && y_dtype == RTSEG_F32) RTSEG_AVG(float, float);
  else if (x_dtype == RTSEG_F32 && y_dtype == RTSEG_BF16) RTSEG_AVG(float, __nv_bfloat16);
  else if (x_dtype == RTSEG_BF16 && y_dtype == RTSEG_BF16) RTSEG_AVG(__nv_bfloat16, __nv_bfloat16);
  else if (x_dtype == RTSEG_BF16 && y_dtype == RTSEG_F32) RTSEG_AVG(__nv_bfloat16, float);
  else RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
#undef RTSEG_AVG
  RTSEG_LAUNCH_CHECK("avgpool3x3s2_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_avgpool3x3s2(int32_t n, int32_t h, int32_t w, int32_t c, int32_t x_is_nchw, int32_t x_dtype,
                                  const void *x, int32_t x_ctotal, int32_t x_coff, int32_t y_dtype, void *y,
                                  int32_t y_ctotal, int32_t y_coff, rtseg_stream_t stream) {
  RTSEG_REQUIRE(x_dtype == RTSEG_F32 || x_dtype == RTSEG_BF16, "unknown dtype");
  const ImgNorm nz{};
  return avgpool3x3s2_impl(n, h, w, c, x_is_nchw, x_dtype, x, x_ctotal, x_coff, y_dtype, y, y_ctotal, y_coff, nz, stream);
}

extern "C" int rtseg_avgpool3x3s2_u8(int32_t n, int32_t h, int32_t w, int32_t c, const uint8_t *x, const float *mean_host,
                                     const float *std_host, int32_t y_dtype, void *y, int32_t y_ctotal, int32_t y_coff,
                                     rtseg_stream_t stream) {
  RTSEG_REQUIRE(mean_host && std_host && c >= 1 && c <= 4, "uint8 images have 1 to 4 channels");
  ImgNorm nz{};
  RTSEG_REQUIRE(make_img_norm(c, mean_host, std_host, nz) == 0, "std must be positive");
  return avgpool3x3s2_impl(n, h, w, c, 0, RTSEG_U8, x, c, 0, y_dtype, y, y_ctotal, y_coff, nz, stream);
}

extern "C" int rtseg_channel_shuffle(int64_t npix, int32_t c, int32_t groups, int32_t dtype, const void *x,
                                     int32_t x_ctotal, int32_t x_coff, void *y, int32_t y_ctotal, int32_t y_coff,
                                     rtseg_stream_t stream) {
  RTSEG_REQUIRE(x && y && x != y, "NULL pointer or in-place shuffle");
  RTSEG_REQUIRE(npix > 0 && c > 0 && groups > 0 && c % groups == 0, "channels must be divisible by groups");
  RTSEG_REQUIRE(x_ctotal >= x_coff + c && y_ctotal >= y_coff + c && x_coff >= 0 && y_coff >= 0, "channel slice out of range");
  cudaStream_t st = as_stream(stream);
  RTSEG_REQUIRE(dtype == RTSEG_BF16 || dtype == RTSEG_F32, "unknown dtype");
  const int es = dtype == RTSEG_BF16 ? 2 : 4;
  const int V = 16 / es;
  if (groups == 2 && c % V == 0 && x_ctotal % V == 0 && x_coff % V == 0 && y_ctotal % V == 0 && y_coff % V == 0 &&
      (c / 2) % (V / 2) == 0 && reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(y) % 16 == 0) {
    const int chunks = c / V;
    const unsigned g2 = (unsigned)cdiv(npix * chunks, 256);
    const char *xb = static_cast<const char *>(x) + (size_t)x_coff * es;
    char *yb = static_cast<char *>(y) + (size_t)y_coff * es;
    if (es == 2)
      channel_shuffle2_kernel<2><<<g2, 256, 0, st>>>(npix, chunks, c / 2 * es, xb, (int64_t)x_ctotal * es, yb,
                                                     (int64_t)y_ctotal * es);
    else
      channel_shuffle2_kernel<4><<<g2, 256, 0, st>>>(npix, chunks, c / 2 * es, xb, (int64_t)x_ctotal * es, yb,
                                                     (int64_t)y_ctotal * es);
    RTSEG_LAUNCH_CHECK("channel_shuffle2_kernel");
    return RTSEG_OK;
  }
  const unsigned grid = (unsigned)cdiv(npix * c, 256);
  if (dtype == RTSEG_BF16)
    channel_shuffle_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(npix, c, groups, (const __nv_bfloat16 *)x, x_ctotal, x_coff,
                                                                (__nv_bfloat16 *)y, y_ctotal, y_coff);
  else if (dtype == RTSEG_F32)
    channel_shuffle_kernel<float><<<grid, 256, 0, st>>>(npix, c, groups, (const float *)x, x_ctotal, x_coff, (float *)y,
                                                        y_ctotal, y_coff);
  else
    RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
  RTSEG_LAUNCH_CHECK("channel_shuffle_kernel");
  return RTSEG_OK;
}
