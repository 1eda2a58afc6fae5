// Training kernels for the depthwise blocks and PReLU: data gradient and weight gradient of a
// depthwise convolution (reference models/modules.py:46-59 `DWConvBNAct` under autograd), and the slope gradient of
// nn.PReLU (reference models/modules.py:119, enet.py:165).  All three are HBM-bound per-channel passes over NHWC
// tensors; channels are read as 16-byte vectors where the slices allow it.
//
//   dx[n, iy, ix, c]   = sum_{ky,kx}  dy[n, (iy + ph - ky d) / s, (ix + pw - kx d) / s, c] * w[ky, kx, c]   (exact divisions only)
//   dw[ky, kx, c]      = sum_{n,oy,ox} dy[n, oy, ox, c] * x[n, oy s - ph + ky d, ox s - pw + kx d, c]
//   dslope[c]          = sum_{pixels}  dz * v * [v <= 0],   v = scale[c] * (y + res) + shift[c]
#include <algorithm>
#include <initializer_list>

#include "common.cuh"

namespace rtseg {

namespace {

struct DwGeom {
  int n, h, w;     // the forward input (= dx)
  int ho, wo;      // the forward output (= dy)
  int c;
  int kh, kw, stride, dil, pad_h, pad_w;
  int x_ct, x_co;  // x / dx slice
  int g_ct, g_co;  // dy slice
};

constexpr int kMaxTapsPerPass = 9;

template <typename T, int V>
__device__ __forceinline__ void load_vec(const T *p, float (&v)[V]) {
  if constexpr (V == 1) v[0] = ld_f<T>(p);
  else Vec16<T>::load(p, v);
}

template <typename T, int V>
__device__ __forceinline__ void store_vec(T *p, const float (&v)[V]) {
  if constexpr (V == 1) st_f<T>(p, v[0]);
  else Vec16<T>::store(p, v);
}

// One thread = one input pixel x V channels.
template <typename T, int V>
__global__ void __launch_bounds__(256) dw_bwd_data_kernel(DwGeom g, const T *__restrict__ dy, const float *__restrict__ w,
                                                          T *__restrict__ dx) {
  const int cvn = g.c / V;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (int64_t)g.n * g.h * g.w * cvn) return;
  const int c0 = (int)(idx % cvn) * V;
  int64_t p = idx / cvn;
  const int ix = (int)(p % g.w);
  p /= g.w;
  const int iy = (int)(p % g.h);
  const int img = (int)(p / g.h);
  float acc[V];
#pragma unroll
  for (int e = 0; e < V; ++e) acc[e] = 0.f;
  for (int ky = 0; ky < g.kh; ++ky) {
    const int ty = iy + g.pad_h - ky * g.dil;
    if (ty < 0 || ty % g.stride) continue;
    const int oy = ty / g.stride;
    if (oy >= g.ho) continue;
    for (int kx = 0; kx < g.kw; ++kx) {
      const int tx = ix + g.pad_w - kx * g.dil;
      if (tx < 0 || tx % g.stride) continue;
      const int ox = tx / g.stride;
      if (ox >= g.wo) continue;
      float d[V];
      load_vec<T, V>(dy + (((int64_t)img * g.ho + oy) * g.wo + ox) * g.g_ct + g.g_co + c0, d);
      const float *wp = w + (int64_t)(ky * g.kw + kx) * g.c + c0;
#pragma unroll
      for (int e = 0; e < V; ++e) acc[e] = fmaf(d[e], __ldg(wp + e), acc[e]);
    }
  }
  store_vec<T, V>(dx + (((int64_t)img * g.h + iy) * g.w + ix) * g.x_ct + g.x_co + c0, acc);
}

// A block owns `ppb` consecutive output pixels; thread (lane, cv) walks its lane's pixels for V channels and keeps
// up to kMaxTapsPerPass x V partial sums in registers (larger kernels take several passes over the block's pixels).
// Partial sums meet in shared memory tap by tap and leave as one fp32 atomic per (tap, channel) and block.
template <typename T, int V>
__global__ void __launch_bounds__(256) dw_wgrad_kernel(DwGeom g, int ppb, const T *__restrict__ x, const T *__restrict__ dy,
                                                       float *__restrict__ dw) {
  extern __shared__ float sm[];  // [lanes][cb * V]
  const int cvn = g.c / V;
  const int cb = cvn < (int)blockDim.x ? cvn : (int)blockDim.x;
  const int lanes = (int)blockDim.x / cb;
  const int tid = threadIdx.x, lane = tid / cb, cvl = tid % cb;
  const int width = cb * V;
  const int64_t npix = (int64_t)g.n * g.ho * g.wo;
  const int64_t p0 = (int64_t)blockIdx.x * ppb;
  const int64_t p1 = p0 + ppb < npix ? p0 + ppb : npix;
  const int ntaps = g.kh * g.kw;
  for (int cbase = 0; cbase < cvn; cbase += cb) {
    const int cv = cbase + cvl;
    for (int t0 = 0; t0 < ntaps; t0 += kMaxTapsPerPass) {
      const int nt = ntaps - t0 < kMaxTapsPerPass ? ntaps - t0 : kMaxTapsPerPass;
      float acc[kMaxTapsPerPass][V];
#pragma unroll
      for (int t = 0; t < kMaxTapsPerPass; ++t)
#pragma unroll
        for (int e = 0; e < V; ++e) acc[t][e] = 0.f;
      if (lane < lanes && cv < cvn) {
        for (int64_t p = p0 + lane; p < p1; p += lanes) {
          const int ox = (int)(p % g.wo);
          const int64_t q = p / g.wo;
          const int oy = (int)(q % g.ho);
          const int img = (int)(q / g.ho);
          float d[V];
          load_vec<T, V>(dy + p * g.g_ct + g.g_co + cv * V, d);
#pragma unroll
          for (int t = 0; t < kMaxTapsPerPass; ++t) {
            if (t >= nt) break;
            const int tap = t0 + t;
            const int iy = oy * g.stride - g.pad_h + (tap / g.kw) * g.dil;
            const int ix = ox * g.stride - g.pad_w + (tap % g.kw) * g.dil;
            if (iy < 0 || iy >= g.h || ix < 0 || ix >= g.w) continue;
            float v[V];
            load_vec<T, V>(x + (((int64_t)img * g.h + iy) * g.w + ix) * g.x_ct + g.x_co + cv * V, v);
#pragma unroll
            for (int e = 0; e < V; ++e) acc[t][e] = fmaf(d[e], v[e], acc[t][e]);
          }
        }
      }
#pragma unroll
      for (int t = 0; t < kMaxTapsPerPass; ++t) {
        if (t >= nt) break;
        if (lane < lanes) {
#pragma unroll
          for (int e = 0; e < V; ++e) sm[lane * width + cvl * V + e] = acc[t][e];
        }
        __syncthreads();
        for (int i = tid; i < width; i += blockDim.x) {
          const int ch = cbase * V + i;
          if (ch < g.c) {
            float s = 0.f;
            for (int l = 0; l < lanes; ++l) s += sm[l * width + i];
            atomicAdd(dw + (int64_t)(t0 + t) * g.c + ch, s);
          }
        }
        __syncthreads();
      }
    }
  }
}

// dslope[c] += sum over the block's pixels of dz * v * [v <= 0]
template <typename T, int V>
__global__ void __launch_bounds__(256) prelu_slope_grad_kernel(int64_t npix, int c, int ppb, const T *__restrict__ dz, int d_ct,
                                                               int d_co, const T *__restrict__ y, int y_ct, int y_co,
                                                               const T *__restrict__ res, int r_ct, int r_co,
                                                               const float *__restrict__ scale, const float *__restrict__ shift,
                                                               double *__restrict__ dslope) {
  extern __shared__ float sm[];  // [lanes][cb * V]
  const int cvn = c / V;
  const int cb = cvn < (int)blockDim.x ? cvn : (int)blockDim.x;
  const int lanes = (int)blockDim.x / cb;
  const int tid = threadIdx.x, lane = tid / cb, cvl = tid % cb;
  const int width = cb * V;
  const int64_t p0 = (int64_t)blockIdx.x * ppb;
  const int64_t p1 = p0 + ppb < npix ? p0 + ppb : npix;
  for (int cbase = 0; cbase < cvn; cbase += cb) {
    const int cv = cbase + cvl;
    float acc[V];
#pragma unroll
    for (int e = 0; e < V; ++e) acc[e] = 0.f;
    if (lane < lanes && cv < cvn) {
      const int c0 = cv * V;
      for (int64_t p = p0 + lane; p < p1; p += lanes) {
        float u[V], d[V];
        load_vec<T, V>(y + p * y_ct + y_co + c0, u);
        if (res) {
          float r[V];
          load_vec<T, V>(res + p * r_ct + r_co + c0, r);
#pragma unroll
          for (int e = 0; e < V; ++e) u[e] += r[e];
        }
        load_vec<T, V>(dz + p * d_ct + d_co + c0, d);
#pragma unroll
        for (int e = 0; e < V; ++e) {
          const float v = fmaf(u[e], scale ? __ldg(scale + c0 + e) : 1.f, shift ? __ldg(shift + c0 + e) : 0.f);
          if (v <= 0.f) acc[e] = fmaf(d[e], v, acc[e]);
        }
      }
    }
    if (lane < lanes) {
#pragma unroll
      for (int e = 0; e < V; ++e) sm[lane * width + cvl * V + e] = acc[e];
    }
    __syncthreads();
    for (int i = tid; i < width; i += blockDim.x) {
      const int ch = cbase * V + i;
      if (ch < c) {
        double s = 0.0;
        for (int l = 0; l < lanes; ++l) s += (double)sm[l * width + i];
        atomicAdd(dslope + ch, s);
      }
    }
    __syncthreads();
  }
}

// ---- adjoint of the bilinear resize: a gather per INPUT pixel over the output pixels that sampled it ----------
// The candidate range is a superset from the inverse map; membership and weights come from the forward's own
// src_index(), so the adjoint is exact (deterministic, no atomics).
__device__ __forceinline__ void dst_range(int i, float scale, int out, int align, int &lo, int &hi) {
  if (scale <= 0.f) {
    lo = 0, hi = out - 1;
    return;
  }
  const float off = align ? 0.f : 0.5f;
  const float a = ((float)(i - 1) + off) / scale - off, b = ((float)(i + 1) + off) / scale - off;
  lo = (int)floorf(a) - 2;
  hi = (int)ceilf(b) + 2;
  if (lo < 0) lo = 0;
  if (hi > out - 1) hi = out - 1;
}

template <typename T, int V>
__global__ void __launch_bounds__(256) bilinear_bwd_kernel(int n, int hi, int wi, int c, int ho, int wo, int align, float sy,
                                                           float sx, const T *__restrict__ dy, int g_ct, int g_co,
                                                           T *__restrict__ dx, int x_ct, int x_co) {
  const int cvn = c / V;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (int64_t)n * hi * wi * cvn) return;
  const int c0 = (int)(idx % cvn) * V;
  int64_t r = idx / cvn;
  const int ix = (int)(r % wi);
  r /= wi;
  const int iy = (int)(r % hi);
  const int b = (int)(r / hi);
  int oy_lo, oy_hi, ox_lo, ox_hi;
  dst_range(iy, sy, ho, align, oy_lo, oy_hi);
  dst_range(ix, sx, wo, align, ox_lo, ox_hi);
  float acc[V];
#pragma unroll
  for (int e = 0; e < V; ++e) acc[e] = 0.f;
  for (int oy = oy_lo; oy <= oy_hi; ++oy) {
    int y0, y1;
    float ly0, ly1;
    src_index(oy, sy, hi, align, y0, y1, ly0, ly1);
    const float wy = (y0 == iy ? ly0 : 0.f) + (y1 == iy ? ly1 : 0.f);
    if (y0 != iy && y1 != iy) continue;
    const T *row = dy + ((int64_t)b * ho + oy) * wo * g_ct + g_co + c0;
    for (int ox = ox_lo; ox <= ox_hi; ++ox) {
      int x0, x1;
      float lx0, lx1;
      src_index(ox, sx, wi, align, x0, x1, lx0, lx1);
      if (x0 != ix && x1 != ix) continue;
      const float wgt = wy * ((x0 == ix ? lx0 : 0.f) + (x1 == ix ? lx1 : 0.f));
      float d[V];
      load_vec<T, V>(row + (int64_t)ox * g_ct, d);
#pragma unroll
      for (int e = 0; e < V; ++e) acc[e] = fmaf(wgt, d[e], acc[e]);
    }
  }
  store_vec<T, V>(dx + (((int64_t)b * hi + iy) * wi + ix) * x_ct + x_co + c0, acc);
}

// ---- adjoint of AdaptiveAvgPool2d(p): every input pixel collects dy / count from the bins that contain it ----
template <typename T>
__global__ void __launch_bounds__(256) adaptive_avgpool_bwd_kernel(int n, int h, int w, int c, int p, const T *__restrict__ dy,
                                                                   T *__restrict__ dx, int x_ct, int x_co) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (int64_t)n * h * w * c) return;
  const int ch = (int)(idx % c);
  int64_t r = idx / c;
  const int ix = (int)(r % w);
  r /= w;
  const int iy = (int)(r % h);
  const int b = (int)(r / h);
  float acc = 0.f;
  for (int pi = 0; pi < p; ++pi) {
    const int hs = (pi * h) / p, he = ((pi + 1) * h + p - 1) / p;
    if (iy < hs || iy >= he) continue;
    for (int pj = 0; pj < p; ++pj) {
      const int ws = (pj * w) / p, we = ((pj + 1) * w + p - 1) / p;
      if (ix < ws || ix >= we) continue;
      acc += ld_f<T>(dy + (((int64_t)b * p + pi) * p + pj) * c + ch) / (float)((he - hs) * (we - ws));
    }
  }
  st_f<T>(dx + (((int64_t)b * h + iy) * w + ix) * x_ct + x_co + ch, acc);
}

// ---- adjoint of MaxUnpool2d(2, 2): every pooled position reads the gradient at the place its index points to ----
template <typename T>
__global__ void __launch_bounds__(256) gather2x2_kernel(int n, int hp, int wp, int c, int ho, int wo, const T *__restrict__ dy,
                                                        const uint8_t *__restrict__ idx, T *__restrict__ dx) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)n * hp * wp * c) return;
  const int ch = (int)(i % c);
  int64_t r = i / c;
  const int px = (int)(r % wp);
  r /= wp;
  const int py = (int)(r % hp);
  const int img = (int)(r / hp);
  const int pos = (int)idx[i];
  const int oy = 2 * py + (pos >> 1), ox = 2 * px + (pos & 1);
  const float v = (oy < ho && ox < wo) ? ld_f<T>(dy + (((int64_t)img * ho + oy) * wo + ox) * c + ch) : 0.f;
  st_f<T>(dx + i, v);
}

int fill_geom(const rtseg_conv_desc *d, DwGeom &g) {
  if (d->groups != d->cin || d->cout != d->cin) return -1;   // channel multiplier 1 only
  if (d->stride < 1 || d->dil < 1 || d->kh < 1 || d->kw < 1) return -1;
  g.n = d->n, g.h = d->h, g.w = d->w, g.c = d->cin;
  g.kh = d->kh, g.kw = d->kw, g.stride = d->stride, g.dil = d->dil, g.pad_h = d->pad_h, g.pad_w = d->pad_w;
  g.ho = (d->h + 2 * d->pad_h - d->dil * (d->kh - 1) - 1) / d->stride + 1;
  g.wo = (d->w + 2 * d->pad_w - d->dil * (d->kw - 1) - 1) / d->stride + 1;
  if (g.ho < 1 || g.wo < 1) return -1;
  g.x_ct = d->x_ctotal, g.x_co = d->x_coff, g.g_ct = d->y_ctotal, g.g_co = d->y_coff;
  if (g.x_co < 0 || g.g_co < 0 || g.x_ct < g.x_co + g.c || g.g_ct < g.g_co + g.c) return -1;
  return 0;
}

// 16-byte channel vectors when every slice allows it, else one channel per thread
bool vec_geom(int dtype, int c, std::initializer_list<int> pitches_and_offsets, std::initializer_list<const void *> ptrs) {
  const int V = dtype == RTSEG_BF16 ? 8 : 4;
  if (c % V) return false;
  for (int v : pitches_and_offsets)
    if (v % V) return false;
  for (const void *p : ptrs)
    if (reinterpret_cast<uintptr_t>(p) & 15) return false;
  return true;
}

void reduce_shape(int c, int V, int &lanes, size_t &smem) {
  const int cvn = c / V;
  const int cb = cvn < 256 ? cvn : 256;
  lanes = 256 / cb;
  smem = (size_t)lanes * cb * V * sizeof(float);
}

}  // namespace

}  // namespace rtseg

using namespace rtseg;

extern "C" int rtseg_conv_dw_bwd_data(const rtseg_conv_desc *d, const void *dy, const float *w, void *dx, rtseg_stream_t stream) {
  RTSEG_REQUIRE(d && dy && w && dx, "NULL pointer");
  DwGeom g;
  RTSEG_REQUIRE(fill_geom(d, g) == 0, "rtseg_conv_dw_bwd_data: depthwise convolution with channel multiplier 1 expected");
  const bool vec = vec_geom(d->dtype, g.c, {g.x_ct, g.x_co, g.g_ct, g.g_co}, {dy, dx});
  const int V = !vec ? 1 : d->dtype == RTSEG_BF16 ? 8 : 4;
  const int64_t threads = (int64_t)g.n * g.h * g.w * (g.c / V);
  RTSEG_REQUIRE(cdiv(threads, 256) <= INT32_MAX, "tensor too large");
  const unsigned grid = (unsigned)cdiv(threads, 256);
#define ARGS <<<grid, 256, 0, as_stream(stream)>>>(g, (const T *)dy, w, (T *)dx)
  if (d->dtype == RTSEG_BF16) {
    using T = __nv_bfloat16;
    if (vec) dw_bwd_data_kernel<T, 8> ARGS; else dw_bwd_data_kernel<T, 1> ARGS;
  } else if (d->dtype == RTSEG_F32) {
    using T = float;
    if (vec) dw_bwd_data_kernel<T, 4> ARGS; else dw_bwd_data_kernel<T, 1> ARGS;
  } else {
    RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
  }
#undef ARGS
  RTSEG_LAUNCH_CHECK("dw_bwd_data_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_conv_dw_wgrad(const rtseg_conv_desc *d, const void *x, const void *dy, float *dw, rtseg_stream_t stream) {
  RTSEG_REQUIRE(d && x && dy && dw, "NULL pointer");
  DwGeom g;
  RTSEG_REQUIRE(fill_geom(d, g) == 0, "rtseg_conv_dw_wgrad: depthwise convolution with channel multiplier 1 expected");
  const bool vec = vec_geom(d->dtype, g.c, {g.x_ct, g.x_co, g.g_ct, g.g_co}, {x, dy});
  const int V = !vec ? 1 : d->dtype == RTSEG_BF16 ? 8 : 4;
  int lanes;
  size_t smem;
  reduce_shape(g.c, V, lanes, smem);
  const int ppb = 64 * lanes;   // 64 output pixels x V channels per thread
  const int64_t blocks = cdiv((int64_t)g.n * g.ho * g.wo, ppb);
  RTSEG_REQUIRE(blocks <= INT32_MAX, "tensor too large");
  const unsigned grid = (unsigned)blocks;
#define ARGS <<<grid, 256, smem, as_stream(stream)>>>(g, ppb, (const T *)x, (const T *)dy, dw)
  if (d->dtype == RTSEG_BF16) {
    using T = __nv_bfloat16;
    if (vec) dw_wgrad_kernel<T, 8> ARGS; else dw_wgrad_kernel<T, 1> ARGS;
  } else if (d->dtype == RTSEG_F32) {
    using T = float;
    if (vec) dw_wgrad_kernel<T, 4> ARGS; else dw_wgrad_kernel<T, 1> ARGS;
  } else {
    RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
  }
#undef ARGS
  RTSEG_LAUNCH_CHECK("dw_wgrad_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_prelu_slope_grad(int64_t npix, int32_t c, int32_t dtype, const void *dz, int32_t d_ctotal, int32_t d_coff,
                                      const void *y, int32_t y_ctotal, int32_t y_coff, const void *res, int32_t r_ctotal,
                                      int32_t r_coff, const float *scale, const float *shift, double *dslope,
                                      rtseg_stream_t stream) {
  RTSEG_REQUIRE(dz && y && dslope, "NULL pointer");
  RTSEG_REQUIRE(npix > 0 && c > 0 && d_coff >= 0 && y_coff >= 0 && d_ctotal >= d_coff + c && y_ctotal >= y_coff + c, "bad extent");
  RTSEG_REQUIRE(!res || (r_coff >= 0 && r_ctotal >= r_coff + c), "bad residual slice");
  const bool vec = vec_geom(dtype, c, {d_ctotal, d_coff, y_ctotal, y_coff, res ? r_ctotal : 0, res ? r_coff : 0}, {dz, y, res});
  const int V = !vec ? 1 : dtype == RTSEG_BF16 ? 8 : 4;
  int lanes;
  size_t smem;
  reduce_shape(c, V, lanes, smem);
  const int ppb = 32 * lanes;
  const int64_t blocks = cdiv(npix, ppb);
  RTSEG_REQUIRE(blocks <= INT32_MAX, "tensor too large");
  const unsigned grid = (unsigned)blocks;
#define ARGS <<<grid, 256, smem, as_stream(stream)>>>(npix, c, ppb, (const T *)dz, d_ctotal, d_coff, (const T *)y, y_ctotal, y_coff, \
                                                      (const T *)res, r_ctotal, r_coff, scale, shift, dslope)
  if (dtype == RTSEG_BF16) {
    using T = __nv_bfloat16;
    if (vec) prelu_slope_grad_kernel<T, 8> ARGS; else prelu_slope_grad_kernel<T, 1> ARGS;
  } else if (dtype == RTSEG_F32) {
    using T = float;
    if (vec) prelu_slope_grad_kernel<T, 4> ARGS; else prelu_slope_grad_kernel<T, 1> ARGS;
  } else {
    RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
  }
#undef ARGS
  RTSEG_LAUNCH_CHECK("prelu_slope_grad_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_upsample_bilinear_bwd(int32_t n, int32_t hi, int32_t wi, int32_t c, int32_t ho, int32_t wo, int32_t align,
                                           int32_t dtype, const void *dy, int32_t g_ctotal, int32_t g_coff, void *dx,
                                           int32_t x_ctotal, int32_t x_coff, rtseg_stream_t stream) {
  RTSEG_REQUIRE(dy && dx, "NULL pointer");
  RTSEG_REQUIRE(n > 0 && hi > 0 && wi > 0 && c > 0 && ho > 0 && wo > 0, "non-positive extent");
  RTSEG_REQUIRE(g_coff >= 0 && x_coff >= 0 && g_ctotal >= g_coff + c && x_ctotal >= x_coff + c, "channel slice out of range");
  const float sy = resize_scale(hi, ho, align), sx = resize_scale(wi, wo, align);
  const bool vec = vec_geom(dtype, c, {g_ctotal, g_coff, x_ctotal, x_coff}, {dy, dx});
  const int V = !vec ? 1 : dtype == RTSEG_BF16 ? 8 : 4;
  const int64_t threads = (int64_t)n * hi * wi * (c / V);
  RTSEG_REQUIRE(cdiv(threads, 256) <= INT32_MAX, "tensor too large");
  const unsigned grid = (unsigned)cdiv(threads, 256);
#define ARGS <<<grid, 256, 0, as_stream(stream)>>>(n, hi, wi, c, ho, wo, align, sy, sx, (const T *)dy, g_ctotal, g_coff, (T *)dx, \
                                                   x_ctotal, x_coff)
  if (dtype == RTSEG_BF16) {
    using T = __nv_bfloat16;
    if (vec) bilinear_bwd_kernel<T, 8> ARGS; else bilinear_bwd_kernel<T, 1> ARGS;
  } else if (dtype == RTSEG_F32) {
    using T = float;
    if (vec) bilinear_bwd_kernel<T, 4> ARGS; else bilinear_bwd_kernel<T, 1> ARGS;
  } else {
    RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
  }
#undef ARGS
  RTSEG_LAUNCH_CHECK("bilinear_bwd_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_adaptive_avgpool_bwd(int32_t n, int32_t h, int32_t w, int32_t c, int32_t p, int32_t dtype, const void *dy,
                                          void *dx, int32_t x_ctotal, int32_t x_coff, rtseg_stream_t stream) {
  RTSEG_REQUIRE(dy && dx, "NULL pointer");
  RTSEG_REQUIRE(n > 0 && h > 0 && w > 0 && c > 0 && p > 0, "bad extent");
  RTSEG_REQUIRE(x_coff >= 0 && x_ctotal >= x_coff + c, "channel slice out of range");
  const int64_t threads = (int64_t)n * h * w * c;
  RTSEG_REQUIRE(cdiv(threads, 256) <= INT32_MAX, "tensor too large");
  const unsigned grid = (unsigned)cdiv(threads, 256);
  if (dtype == RTSEG_BF16)
    adaptive_avgpool_bwd_kernel<__nv_bfloat16><<<grid, 256, 0, as_stream(stream)>>>(n, h, w, c, p, (const __nv_bfloat16 *)dy,
                                                                                   (__nv_bfloat16 *)dx, x_ctotal, x_coff);
  else if (dtype == RTSEG_F32)
    adaptive_avgpool_bwd_kernel<float><<<grid, 256, 0, as_stream(stream)>>>(n, h, w, c, p, (const float *)dy, (float *)dx,
                                                                           x_ctotal, x_coff);
  else
    RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
  RTSEG_LAUNCH_CHECK("adaptive_avgpool_bwd_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_maxunpool2x2_bwd(int32_t n, int32_t hp, int32_t wp, int32_t c, int32_t ho, int32_t wo, int32_t dtype,
                                      const void *dy, const uint8_t *idx, void *dx, rtseg_stream_t stream) {
  RTSEG_REQUIRE(dy && idx && dx, "NULL pointer");
  RTSEG_REQUIRE(n > 0 && hp > 0 && wp > 0 && c > 0 && ho >= 2 * hp && wo >= 2 * wp, "bad extent");
  const int64_t threads = (int64_t)n * hp * wp * c;
  RTSEG_REQUIRE(cdiv(threads, 256) <= INT32_MAX, "tensor too large");
  const unsigned grid = (unsigned)cdiv(threads, 256);
  if (dtype == RTSEG_BF16)
    gather2x2_kernel<__nv_bfloat16><<<grid, 256, 0, as_stream(stream)>>>(n, hp, wp, c, ho, wo, (const __nv_bfloat16 *)dy, idx,
                                                                        (__nv_bfloat16 *)dx);
  else if (dtype == RTSEG_F32)
    gather2x2_kernel<float><<<grid, 256, 0, as_stream(stream)>>>(n, hp, wp, c, ho, wo, (const float *)dy, idx, (float *)dx);
  else
    RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
  RTSEG_LAUNCH_CHECK("gather2x2_kernel");
  return RTSEG_OK;
}
