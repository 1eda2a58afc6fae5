// train.cu — the training-mode pieces of the conv + BatchNorm + activation blocks:
// batch statistics, normalise + activation, their backward, weight gradients, max-pool backward.
//
// In training the reference runs Conv2d -> BatchNorm2d(batch statistics) -> activation as three autograd
// nodes (models/modules.py:73-85) and, in NonBt1DBlock, conv -> `+= x` -> BN -> act (models/erfnet.py:76-82).
// Here the convolution itself (forward and data gradient) is the same kernel as in inference — a data
// gradient is a convolution with transposed / flipped weights, the data gradient of a stride-2 3x3
// convolution is the 3x3 stride-2 transposed convolution and vice versa — and this file adds what inference
// does not need:
//   rtseg_bn_stats            per-channel sum and sum of squares of (y [+ res]) over all pixels (fp64 accumulators)
//   rtseg_affine_act          z = act(scale[c] * (y [+ res]) + shift[c])        (BN with batch statistics folded
//                             into scale / shift by the host from the sums above)
//   rtseg_bn_act_bwd_reduce   sums of g and g * xhat, g = dz * act'(.), xhat = (u - mean) * invstd
//   rtseg_bn_act_bwd_apply    du = gamma * invstd * (g - mean(g) - xhat * mean(g * xhat))
//   rtseg_conv_wgrad          dW[tap][ci][co] = sum over pixels of x[pixel + tap][ci] * dy[pixel][co]
//   rtseg_maxpool3x3s2_bwd    gradient of MaxPool2d(3, 2, 1) (InitialBlock's pool branch), gather form
// All tensors NHWC with channel slices, fp32 or bf16; reductions accumulate in fp32 per thread and in fp64
// across threads (atomics on the caller's zeroed buffer).  These are first, correct kernels on CUDA cores;
// the weight gradient in particular is a register-tiled split-K GEMM, not yet a tcgen05 one.
#include <mma.h>

#include <algorithm>
#include <initializer_list>

#include "common.cuh"

namespace rtseg {

namespace {

// d act / d v at pre-activation value v (the members that can be trained through here)
__device__ __forceinline__ float act_grad(float v, int act, float p0, float p1, float slope) {
  switch (act) {
    case RTSEG_ACT_NONE: return 1.f;
    case RTSEG_ACT_RELU: return v > 0.f ? 1.f : 0.f;
    case RTSEG_ACT_RELU6: return (v > 0.f && v < 6.f) ? 1.f : 0.f;
    case RTSEG_ACT_LEAKYRELU: return v > 0.f ? 1.f : p0;
    case RTSEG_ACT_PRELU: return v > 0.f ? 1.f : slope;
    case RTSEG_ACT_ELU: return v > 0.f ? 1.f : p0 * expf(v);
    case RTSEG_ACT_HARDTANH: return (v > p0 && v < p1) ? 1.f : 0.f;
    case RTSEG_ACT_SIGMOID: {
      const float s = 1.f / (1.f + expf(-v));
      return s * (1.f - s);
    }
    case RTSEG_ACT_TANH: {
      const float t = tanhf(v);
      return 1.f - t * t;
    }
    case RTSEG_ACT_SILU: {
      const float s = 1.f / (1.f + expf(-v));
      return s * (1.f + v * (1.f - s));
    }
    default: return 1.f;
  }
}

struct Slice {
  int ct, co;  // channel pitch and first channel of the slice
};

struct ActP {
  int act;
  float p0, p1;
  const float *slope;  // PReLU slopes or nullptr
};

// Per-channel reduction skeleton: a block owns `ppb` consecutive pixels; thread (lane, ch) with ch = tid % cb walks
// its lane's pixels, so that consecutive threads read consecutive channels of one pixel (coalesced).  F(pix, ch,
// &a, &b) adds one element's contribution to two running sums.
template <typename F>
__device__ __forceinline__ void reduce2_per_channel(int64_t npix, int c, int ppb, double *out0, double *out1, F f) {
  extern __shared__ float sm[];  // [2][lanes][cb]
  const int cb = c < (int)blockDim.x ? c : (int)blockDim.x;
  const int lanes = (int)blockDim.x / cb;
  const int tid = threadIdx.x;
  const int lane = tid / cb, chl = tid % cb;
  const int64_t p0 = (int64_t)blockIdx.x * ppb;
  const int64_t p1 = p0 + ppb < npix ? p0 + ppb : npix;
  for (int cbase = 0; cbase < c; cbase += cb) {
    const int ch = cbase + chl;
    float a = 0.f, b = 0.f;
    if (lane < lanes && ch < c)
      for (int64_t p = p0 + lane; p < p1; p += lanes) f(p, ch, a, b);
    if (lane < lanes) sm[lane * cb + chl] = a, sm[(lanes + lane) * cb + chl] = b;
    __syncthreads();
    if (tid < cb && cbase + tid < c) {
      double sa = 0.0, sb = 0.0;
      for (int l = 0; l < lanes; ++l) sa += (double)sm[l * cb + tid], sb += (double)sm[(lanes + l) * cb + tid];
      atomicAdd(out0 + cbase + tid, sa);
      if (out1) atomicAdd(out1 + cbase + tid, sb);
    }
    __syncthreads();
  }
}

template <typename T>
__global__ void __launch_bounds__(256) bn_stats_kernel(int64_t npix, int c, int ppb, const T *__restrict__ y, Slice ys,
                                                       const T *__restrict__ res, Slice rs, double *sum, double *sumsq) {
  reduce2_per_channel(npix, c, ppb, sum, sumsq, [&](int64_t p, int ch, float &a, float &b) {
    float u = ld_f<T>(y + p * ys.ct + ys.co + ch);
    if (res) u += ld_f<T>(res + p * rs.ct + rs.co + ch);
    a += u;
    b = fmaf(u, u, b);
  });
}

template <typename T>
__global__ void __launch_bounds__(256) affine_act_kernel(int64_t npix, int c, const T *__restrict__ y, Slice ys,
                                                         const T *__restrict__ res, Slice rs, const float *__restrict__ scale,
                                                         const float *__restrict__ shift, ActP ap, T *__restrict__ z, Slice zs) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= npix * c) return;
  const int ch = (int)(idx % c);
  const int64_t p = idx / c;
  float u = ld_f<T>(y + p * ys.ct + ys.co + ch);
  if (res) u += ld_f<T>(res + p * rs.ct + rs.co + ch);
  const float v = fmaf(u, scale ? __ldg(scale + ch) : 1.f, shift ? __ldg(shift + ch) : 0.f);
  st_f<T>(z + p * zs.ct + zs.co + ch, act_apply(v, ap.act, ap.p0, ap.p1, ap.slope ? __ldg(ap.slope + ch) : 0.f));
}

template <typename T>
__global__ void __launch_bounds__(256) bn_act_bwd_reduce_kernel(int64_t npix, int c, int ppb, const T *__restrict__ dz, Slice ds,
                                                                const T *__restrict__ y, Slice ys, const T *__restrict__ res, Slice rs,
                                                                const float *__restrict__ scale, const float *__restrict__ shift,
                                                                const float *__restrict__ mean, const float *__restrict__ invstd, ActP ap,
                                                                double *sum_g, double *sum_gx) {
  reduce2_per_channel(npix, c, ppb, sum_g, sum_gx, [&](int64_t p, int ch, float &a, float &b) {
    float u = ld_f<T>(y + p * ys.ct + ys.co + ch);
    if (res) u += ld_f<T>(res + p * rs.ct + rs.co + ch);
    const float v = fmaf(u, __ldg(scale + ch), __ldg(shift + ch));
    const float g = ld_f<T>(dz + p * ds.ct + ds.co + ch) * act_grad(v, ap.act, ap.p0, ap.p1, ap.slope ? __ldg(ap.slope + ch) : 0.f);
    a += g;
    b = fmaf(g, (u - __ldg(mean + ch)) * __ldg(invstd + ch), b);
  });
}

// du = k[c] * (g - mg[c] - xhat * mgx[c]);  k = gamma * invstd, mg = mean(g), mgx = mean(g * xhat)
template <typename T>
__global__ void __launch_bounds__(256) bn_act_bwd_apply_kernel(int64_t npix, int c, const T *__restrict__ dz, Slice ds,
                                                               const T *__restrict__ y, Slice ys, const T *__restrict__ res, Slice rs,
                                                               const float *__restrict__ scale, const float *__restrict__ shift,
                                                               const float *__restrict__ mean, const float *__restrict__ invstd,
                                                               const float *__restrict__ k, const float *__restrict__ mg,
                                                               const float *__restrict__ mgx, ActP ap, T *__restrict__ du, Slice us) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= npix * c) return;
  const int ch = (int)(idx % c);
  const int64_t p = idx / c;
  float u = ld_f<T>(y + p * ys.ct + ys.co + ch);
  if (res) u += ld_f<T>(res + p * rs.ct + rs.co + ch);
  const float v = fmaf(u, __ldg(scale + ch), __ldg(shift + ch));
  const float g = ld_f<T>(dz + p * ds.ct + ds.co + ch) * act_grad(v, ap.act, ap.p0, ap.p1, ap.slope ? __ldg(ap.slope + ch) : 0.f);
  const float xh = (u - __ldg(mean + ch)) * __ldg(invstd + ch);
  st_f<T>(du + p * us.ct + us.co + ch, __ldg(k + ch) * (g - __ldg(mg + ch) - xh * __ldg(mgx + ch)));
}

// ---- 16-byte-vector variants of the four BatchNorm kernels (c % V == 0, every slice 16-byte aligned) -------------
template <typename T>
__device__ __forceinline__ void load_u(const T *y, Slice ys, const T *res, Slice rs, int64_t p, int c0, float (&u)[Vec16<T>::N]) {
  Vec16<T>::load(y + p * ys.ct + ys.co + c0, u);
  if (res) {
    float r[Vec16<T>::N];
    Vec16<T>::load(res + p * rs.ct + rs.co + c0, r);
#pragma unroll
    for (int e = 0; e < Vec16<T>::N; ++e) u[e] += r[e];
  }
}

template <int V, typename F>
__device__ __forceinline__ void reduce2_vec(int64_t npix, int c, int ppb, double *out0, double *out1, F f) {
  extern __shared__ float sm[];  // [2][lanes][cb * V]
  const int cvn = c / V;
  const int cb = cvn < (int)blockDim.x ? cvn : (int)blockDim.x;
  const int lanes = (int)blockDim.x / cb;
  const int tid = threadIdx.x, lane = tid / cb, cvl = tid % cb;
  const int64_t p0 = (int64_t)blockIdx.x * ppb;
  const int64_t p1 = p0 + ppb < npix ? p0 + ppb : npix;
  const int width = cb * V;
  for (int cbase = 0; cbase < cvn; cbase += cb) {
    const int cv = cbase + cvl;
    float a[V], b[V];
#pragma unroll
    for (int e = 0; e < V; ++e) a[e] = b[e] = 0.f;
    if (lane < lanes && cv < cvn) {
#pragma unroll 4
      for (int64_t p = p0 + lane; p < p1; p += lanes) f(p, cv * V, a, b);   // unrolled: four pixels' loads in flight
    }
    if (lane < lanes) {
#pragma unroll
      for (int e = 0; e < V; ++e) sm[lane * width + cvl * V + e] = a[e], sm[(lanes + lane) * width + cvl * V + e] = b[e];
    }
    __syncthreads();
    for (int i = tid; i < width; i += blockDim.x) {
      const int ch = cbase * V + i;
      if (ch < c) {
        double sa = 0.0, sb = 0.0;
        for (int l = 0; l < lanes; ++l) sa += (double)sm[l * width + i], sb += (double)sm[(lanes + l) * width + i];
        atomicAdd(out0 + ch, sa);
        if (out1) atomicAdd(out1 + ch, sb);
      }
    }
    __syncthreads();
  }
}

template <typename T>
__global__ void __launch_bounds__(256) bn_stats_vec_kernel(int64_t npix, int c, int ppb, const T *__restrict__ y, Slice ys,
                                                           const T *__restrict__ res, Slice rs, double *sum, double *sumsq) {
  constexpr int V = Vec16<T>::N;
  reduce2_vec<V>(npix, c, ppb, sum, sumsq, [&](int64_t p, int c0, float (&a)[V], float (&b)[V]) {
    float u[V];
    load_u<T>(y, ys, res, rs, p, c0, u);
#pragma unroll
    for (int e = 0; e < V; ++e) a[e] += u[e], b[e] = fmaf(u[e], u[e], b[e]);
  });
}

template <typename T>
__global__ void __launch_bounds__(256) affine_act_vec_kernel(int64_t npix, int c, const T *__restrict__ y, Slice ys,
                                                             const T *__restrict__ res, Slice rs, const float *__restrict__ scale,
                                                             const float *__restrict__ shift, ActP ap, T *__restrict__ z, Slice zs) {
  constexpr int V = Vec16<T>::N;
  const int cvn = c / V;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= npix * cvn) return;
  const int c0 = (int)(idx % cvn) * V;
  const int64_t p = idx / cvn;
  float u[V], o[V];
  load_u<T>(y, ys, res, rs, p, c0, u);
#pragma unroll
  for (int e = 0; e < V; ++e) {
    const float v = fmaf(u[e], scale ? __ldg(scale + c0 + e) : 1.f, shift ? __ldg(shift + c0 + e) : 0.f);
    o[e] = act_apply(v, ap.act, ap.p0, ap.p1, ap.slope ? __ldg(ap.slope + c0 + e) : 0.f);
  }
  Vec16<T>::store(z + p * zs.ct + zs.co + c0, o);
}

template <typename T>
__global__ void __launch_bounds__(256) bn_act_bwd_reduce_vec_kernel(int64_t npix, int c, int ppb, const T *__restrict__ dz, Slice ds,
                                                                    const T *__restrict__ y, Slice ys, const T *__restrict__ res,
                                                                    Slice rs, const float *__restrict__ scale,
                                                                    const float *__restrict__ shift, const float *__restrict__ mean,
                                                                    const float *__restrict__ invstd, ActP ap, double *sum_g,
                                                                    double *sum_gx) {
  constexpr int V = Vec16<T>::N;
  reduce2_vec<V>(npix, c, ppb, sum_g, sum_gx, [&](int64_t p, int c0, float (&a)[V], float (&b)[V]) {
    float u[V], d[V];
    load_u<T>(y, ys, res, rs, p, c0, u);
    Vec16<T>::load(dz + p * ds.ct + ds.co + c0, d);
#pragma unroll
    for (int e = 0; e < V; ++e) {
      const float v = fmaf(u[e], __ldg(scale + c0 + e), __ldg(shift + c0 + e));
      const float g = d[e] * act_grad(v, ap.act, ap.p0, ap.p1, ap.slope ? __ldg(ap.slope + c0 + e) : 0.f);
      a[e] += g;
      b[e] = fmaf(g, (u[e] - __ldg(mean + c0 + e)) * __ldg(invstd + c0 + e), b[e]);
    }
  });
}

template <typename T>
__global__ void __launch_bounds__(256) bn_act_bwd_apply_vec_kernel(int64_t npix, int c, const T *__restrict__ dz, Slice ds,
                                                                   const T *__restrict__ y, Slice ys, const T *__restrict__ res, Slice rs,
                                                                   const float *__restrict__ scale, const float *__restrict__ shift,
                                                                   const float *__restrict__ mean, const float *__restrict__ invstd,
                                                                   const float *__restrict__ k, const float *__restrict__ mg,
                                                                   const float *__restrict__ mgx, ActP ap, T *__restrict__ du, Slice us) {
  constexpr int V = Vec16<T>::N;
  const int cvn = c / V;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= npix * cvn) return;
  const int c0 = (int)(idx % cvn) * V;
  const int64_t p = idx / cvn;
  float u[V], d[V], o[V];
  load_u<T>(y, ys, res, rs, p, c0, u);
  Vec16<T>::load(dz + p * ds.ct + ds.co + c0, d);
#pragma unroll
  for (int e = 0; e < V; ++e) {
    const int ch = c0 + e;
    const float v = fmaf(u[e], __ldg(scale + ch), __ldg(shift + ch));
    const float g = d[e] * act_grad(v, ap.act, ap.p0, ap.p1, ap.slope ? __ldg(ap.slope + ch) : 0.f);
    const float xh = (u[e] - __ldg(mean + ch)) * __ldg(invstd + ch);
    o[e] = __ldg(k + ch) * (g - __ldg(mg + ch) - xh * __ldg(mgx + ch));
  }
  Vec16<T>::store(du + p * us.ct + us.co + c0, o);
}

// can the slices be read / written as 16-byte channel vectors?
template <typename... P>
bool vec_ok(int dtype, int c, std::initializer_list<Slice> slices, P... ptrs) {
  const int V = dtype == RTSEG_BF16 ? 8 : 4;
  if (c % V) return false;
  for (const Slice &s : slices)
    if (s.ct % V || s.co % V) return false;
  const uintptr_t bits = (... | reinterpret_cast<uintptr_t>(ptrs));
  return (bits & 15) == 0;
}

int reduce_geometry_vec(int64_t npix, int c, int V, int &ppb, unsigned &grid, size_t &smem) {
  const int cvn = c / V;
  const int cb = cvn < 256 ? cvn : 256;
  const int lanes = 256 / cb;
  ppb = 16 * lanes;  // 16 pixels x V channels per thread
  const int64_t blocks = cdiv(npix, ppb);
  if (blocks > INT32_MAX) return -1;
  grid = (unsigned)blocks;
  smem = (size_t)2 * lanes * cb * V * sizeof(float);
  return 0;
}

// ---- weight gradient: split-K register-tiled GEMM, [64 ci] x [64 co] per block and tap, K = pixels ------------
struct WgradParams {
  int n, h, w;      // x
  int gh, gw;       // dy grid
  int stride, dil, pad_h, pad_w, kh, kw;
  int cin, cout;
  Slice xs, gs;
  int64_t chunk;    // base pixels per split
};

template <typename T>
__global__ void __launch_bounds__(256) conv_wgrad_kernel(WgradParams p, const T *__restrict__ x, const T *__restrict__ dy,
                                                         float *__restrict__ dw) {
  constexpr int TK = 16;
  __shared__ float As[TK][64 + 4];
  __shared__ float Bs[TK][64 + 4];
  const int tid = threadIdx.x;
  const int co_tiles = (p.cout + 63) / 64;
  const int ci0 = (blockIdx.x / co_tiles) * 64, co0 = (blockIdx.x % co_tiles) * 64;
  const int tap = blockIdx.y, ky = tap / p.kw, kx = tap % p.kw;
  const int64_t M = (int64_t)p.n * p.gh * p.gw;
  const int64_t m_begin = (int64_t)blockIdx.z * p.chunk;
  const int64_t m_end = m_begin + p.chunk < M ? m_begin + p.chunk : M;
  const int lk = tid / 16, lc = (tid % 16) * 4;  // loader: pixel lk of the chunk, channels lc..lc+3
  const int ty = tid / 16, tx = tid % 16;        // compute: ci rows ty*4.., co cols tx*4..
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int64_t m0 = m_begin; m0 < m_end; m0 += TK) {
    const int64_t m = m0 + lk;
    float av[4] = {0.f, 0.f, 0.f, 0.f}, bv[4] = {0.f, 0.f, 0.f, 0.f};
    if (m < m_end) {
      const int j = (int)(m % p.gw);
      const int64_t r = m / p.gw;
      const int i = (int)(r % p.gh), b = (int)(r / p.gh);
      const int hi = i * p.stride - p.pad_h + ky * p.dil, wi = j * p.stride - p.pad_w + kx * p.dil;
      if (hi >= 0 && hi < p.h && wi >= 0 && wi < p.w) {
        const T *xp = x + (((int64_t)b * p.h + hi) * p.w + wi) * p.xs.ct + p.xs.co + ci0 + lc;
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (ci0 + lc + e < p.cin) av[e] = ld_f<T>(xp + e);
        const T *gp = dy + m * p.gs.ct + p.gs.co + co0 + lc;
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (co0 + lc + e < p.cout) bv[e] = ld_f<T>(gp + e);
      }
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) As[lk][lc + e] = av[e], Bs[lk][lc + e] = bv[e];
    __syncthreads();
#pragma unroll
    for (int k = 0; k < TK; ++k) {
      float a[4], b[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) a[e] = As[k][ty * 4 + e], b[e] = Bs[k][tx * 4 + e];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
  float *dwt = dw + (int64_t)tap * p.cin * p.cout;
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int ci = ci0 + ty * 4 + i, co = co0 + tx * 4 + j;
      if (ci < p.cin && co < p.cout && acc[i][j] != 0.f) atomicAdd(dwt + (int64_t)ci * p.cout + co, acc[i][j]);
    }
}

// bf16 variant of the weight gradient on the (legacy, mma.sync) tensor cores: the pixel slab of x (at the tap's
// shifted positions) and of dy is staged in shared memory exactly as it lies in HBM, [pixel][channel]; read as a
// column-major A (M = ci, K = pixel) and a row-major B (K = pixel, N = co) it needs no transposition.  fp32
// accumulation, same split-K + atomics scheme as the CUDA-core kernel above.  (A tcgen05 version wants MN-major
// UMMA descriptors over the same tiles — next round.)
__global__ void __launch_bounds__(256) conv_wgrad_wmma_kernel(WgradParams p, const __nv_bfloat16 *__restrict__ x,
                                                              const __nv_bfloat16 *__restrict__ dy, float *__restrict__ dw) {
  using namespace nvcuda;
  constexpr int TK = 64, LD = 64 + 8;
  __shared__ __align__(32) unsigned char smem_raw[2 * TK * LD * 2];
  __nv_bfloat16(*As)[LD] = reinterpret_cast<__nv_bfloat16(*)[LD]>(smem_raw);
  __nv_bfloat16(*Bs)[LD] = reinterpret_cast<__nv_bfloat16(*)[LD]>(smem_raw + TK * LD * 2);
  const int tid = threadIdx.x, warp = tid >> 5;
  const int co_tiles = (p.cout + 63) / 64;
  const int ci0 = (blockIdx.x / co_tiles) * 64, co0 = (blockIdx.x % co_tiles) * 64;
  const int tap = blockIdx.y, ky = tap / p.kw, kx = tap % p.kw;
  const int64_t M = (int64_t)p.n * p.gh * p.gw;
  const int64_t m_begin = (int64_t)blockIdx.z * p.chunk;
  const int64_t m_end = m_begin + p.chunk < M ? m_begin + p.chunk : M;
  const int ci_sub = warp & 3, co_sub = (warp >> 2) * 2;
  const bool a_vec = p.xs.ct % 8 == 0 && p.xs.co % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0;
  const bool b_vec = p.gs.ct % 8 == 0 && p.gs.co % 8 == 0 && (reinterpret_cast<uintptr_t>(dy) & 15) == 0;
  wmma::fragment<wmma::accumulator, 16, 16, 16, float> acc[2];
  wmma::fill_fragment(acc[0], 0.f);
  wmma::fill_fragment(acc[1], 0.f);

  for (int64_t m0 = m_begin; m0 < m_end; m0 += TK) {
#pragma unroll
    for (int j = 0; j < 2; ++j) {  // 64 pixels x 8 sixteen-byte vectors per operand, two per thread
      const int px = (tid >> 3) + 32 * j, cv = (tid & 7) * 8;
      const int64_t m = m0 + px;
      uint4 av = make_uint4(0, 0, 0, 0), bv = av;
      if (m < m_end) {
        const int jj = (int)(m % p.gw);
        const int64_t r = m / p.gw;
        const int ii = (int)(r % p.gh), b = (int)(r / p.gh);
        const int hi = ii * p.stride - p.pad_h + ky * p.dil, wi = jj * p.stride - p.pad_w + kx * p.dil;
        if (hi >= 0 && hi < p.h && wi >= 0 && wi < p.w) {
          const __nv_bfloat16 *xp = x + (((int64_t)b * p.h + hi) * p.w + wi) * p.xs.ct + p.xs.co + ci0 + cv;
          const __nv_bfloat16 *gp = dy + m * p.gs.ct + p.gs.co + co0 + cv;
          if (a_vec && ci0 + cv + 8 <= p.cin) {
            av = __ldg(reinterpret_cast<const uint4 *>(xp));
          } else {
            __nv_bfloat16 t[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) t[e] = ci0 + cv + e < p.cin ? xp[e] : __float2bfloat16(0.f);
            av = *reinterpret_cast<const uint4 *>(t);
          }
          if (b_vec && co0 + cv + 8 <= p.cout) {
            bv = __ldg(reinterpret_cast<const uint4 *>(gp));
          } else {
            __nv_bfloat16 t[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) t[e] = co0 + cv + e < p.cout ? gp[e] : __float2bfloat16(0.f);
            bv = *reinterpret_cast<const uint4 *>(t);
          }
        }
      }
      *reinterpret_cast<uint4 *>(&As[px][cv]) = av;
      *reinterpret_cast<uint4 *>(&Bs[px][cv]) = bv;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < TK; kk += 16) {
      wmma::fragment<wmma::matrix_a, 16, 16, 16, __nv_bfloat16, wmma::col_major> a;
      wmma::load_matrix_sync(a, &As[kk][ci_sub * 16], LD);
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        wmma::fragment<wmma::matrix_b, 16, 16, 16, __nv_bfloat16, wmma::row_major> b;
        wmma::load_matrix_sync(b, &Bs[kk][(co_sub + j) * 16], LD);
        wmma::mma_sync(acc[j], a, b, acc[j]);
      }
    }
    __syncthreads();
  }
  float(*Cs)[64 + 4] = reinterpret_cast<float(*)[64 + 4]>(smem_raw);  // 64 x 68 floats = 17,408 B <= the operand tiles
#pragma unroll
  for (int j = 0; j < 2; ++j) wmma::store_matrix_sync(&Cs[ci_sub * 16][(co_sub + j) * 16], acc[j], 64 + 4, wmma::mem_row_major);
  __syncthreads();
  float *dwt = dw + (int64_t)tap * p.cin * p.cout;
  for (int i = tid; i < 64 * 64; i += 256) {
    const int ci = ci0 + i / 64, co = co0 + i % 64;
    const float v = Cs[i / 64][i % 64];
    if (ci < p.cin && co < p.cout && v != 0.f) atomicAdd(dwt + (int64_t)ci * p.cout + co, v);
  }
}

// ---- MaxPool2d(3, 2, 1) backward, gather form: an input pixel collects dy from every window whose FIRST maximum
// (row-major scan order, like ATen's max_pool2d_with_indices) it is.
template <typename T>
__global__ void __launch_bounds__(256) maxpool3x3s2_bwd_kernel(int n, int h, int w, int c, int ho, int wo, const T *__restrict__ x,
                                                               Slice xs, const T *__restrict__ dy, Slice gs, T *__restrict__ dx, Slice ds) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (int64_t)n * h * w * c) return;
  const int ch = (int)(idx % c);
  int64_t r = idx / c;
  const int ix = (int)(r % w);
  r /= w;
  const int iy = (int)(r % h), b = (int)(r / h);
  const T *xb = x + (int64_t)b * h * w * xs.ct + xs.co + ch;
  float g = 0.f;
  // windows (oy, ox) that contain (iy, ix): oy*2 - 1 <= iy <= oy*2 + 1
  for (int oy = (iy + 1) / 2 - ((iy + 1) % 2 == 0 ? 1 : 0); oy <= (iy + 1) / 2; ++oy) {
    if (oy < 0 || oy >= ho) continue;
    for (int ox = (ix + 1) / 2 - ((ix + 1) % 2 == 0 ? 1 : 0); ox <= (ix + 1) / 2; ++ox) {
      if (ox < 0 || ox >= wo) continue;
      float best = -INFINITY;
      int by = -1, bx = -1;
      for (int ky = 0; ky < 3; ++ky) {
        const int yy = oy * 2 - 1 + ky;
        if (yy < 0 || yy >= h) continue;
        for (int kx = 0; kx < 3; ++kx) {
          const int xx = ox * 2 - 1 + kx;
          if (xx < 0 || xx >= w) continue;
          const float v = ld_f<T>(xb + ((int64_t)yy * w + xx) * xs.ct);
          if (v > best || (v != v && best == best) || by < 0) best = v, by = yy, bx = xx;
        }
      }
      if (by == iy && bx == ix) g += ld_f<T>(dy + (((int64_t)b * ho + oy) * wo + ox) * gs.ct + gs.co + ch);
    }
  }
  st_f<T>(dx + (((int64_t)b * h + iy) * w + ix) * ds.ct + ds.co + ch, g);
}

// the same with 16-byte channel vectors (c % V == 0, aligned slices): 36 vector loads per V channels instead of per channel
template <typename T>
__global__ void __launch_bounds__(256) maxpool3x3s2_bwd_vec_kernel(int n, int h, int w, int c, int ho, int wo, const T *__restrict__ x,
                                                                   Slice xs, const T *__restrict__ dy, Slice gs, T *__restrict__ dx,
                                                                   Slice ds) {
  constexpr int V = Vec16<T>::N;
  const int cvn = c / V;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (int64_t)n * h * w * cvn) return;
  const int c0 = (int)(idx % cvn) * V;
  int64_t r = idx / cvn;
  const int ix = (int)(r % w);
  r /= w;
  const int iy = (int)(r % h), b = (int)(r / h);
  const T *xb = x + (int64_t)b * h * w * xs.ct + xs.co + c0;
  float g[V];
#pragma unroll
  for (int e = 0; e < V; ++e) g[e] = 0.f;
  for (int oy = (iy + 1) / 2 - ((iy + 1) % 2 == 0 ? 1 : 0); oy <= (iy + 1) / 2; ++oy) {
    if (oy < 0 || oy >= ho) continue;
    for (int ox = (ix + 1) / 2 - ((ix + 1) % 2 == 0 ? 1 : 0); ox <= (ix + 1) / 2; ++ox) {
      if (ox < 0 || ox >= wo) continue;
      float best[V];
      int bpos[V];  // ky * 3 + kx of the first maximum
#pragma unroll
      for (int e = 0; e < V; ++e) best[e] = -INFINITY, bpos[e] = -1;
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        const int yy = oy * 2 - 1 + t / 3, xx = ox * 2 - 1 + t % 3;
        if (yy < 0 || yy >= h || xx < 0 || xx >= w) continue;
        float v[V];
        Vec16<T>::load(xb + ((int64_t)yy * w + xx) * xs.ct, v);
#pragma unroll
        for (int e = 0; e < V; ++e)
          if (v[e] > best[e] || (v[e] != v[e] && best[e] == best[e]) || bpos[e] < 0) best[e] = v[e], bpos[e] = t;
      }
      const int mine = (iy - (oy * 2 - 1)) * 3 + (ix - (ox * 2 - 1));
      float d[V];
      Vec16<T>::load(dy + (((int64_t)b * ho + oy) * wo + ox) * gs.ct + gs.co + c0, d);
#pragma unroll
      for (int e = 0; e < V; ++e)
        if (bpos[e] == mine) g[e] += d[e];
    }
  }
  Vec16<T>::store(dx + (((int64_t)b * h + iy) * w + ix) * ds.ct + ds.co + c0, g);
}

// ---- the per-channel arithmetic between the reduction and the element-wise pass, one launch instead of ~20 tiny ones
// stats = [sum | sumsq] (fp64); out = [mean | invstd | scale | shift | k] (fp32), k = gamma * invstd; running statistics
// updated in place like nn.BatchNorm2d (momentum, unbiased variance).
__global__ void bn_finalize_kernel(int c, const double *__restrict__ stats, const double *__restrict__ count_dev, double count_host,
                                   const float *__restrict__ gamma, const float *__restrict__ beta, double eps, double momentum,
                                   float *__restrict__ running_mean, float *__restrict__ running_var, float *__restrict__ out) {
  const double count = count_dev ? *count_dev : count_host;
  for (int ch = blockIdx.x * blockDim.x + threadIdx.x; ch < c; ch += gridDim.x * blockDim.x) {
    const double mean = stats[ch] / count;
    double var = stats[c + ch] / count - mean * mean;
    var = var < 0.0 ? 0.0 : var;
    const double invstd = rsqrt(var + eps);
    const double g = gamma ? (double)gamma[ch] : 1.0, b = beta ? (double)beta[ch] : 0.0;
    out[ch] = (float)mean;
    out[c + ch] = (float)invstd;
    out[2 * c + ch] = (float)(g * invstd);
    out[3 * c + ch] = (float)(b - mean * g * invstd);
    out[4 * c + ch] = (float)(g * invstd);
    if (running_mean) {
      const double unbiased = var * (count / (count > 1.0 ? count - 1.0 : 1.0));
      running_mean[ch] = (float)((1.0 - momentum) * (double)running_mean[ch] + momentum * mean);
      running_var[ch] = (float)((1.0 - momentum) * (double)running_var[ch] + momentum * unbiased);
    }
  }
}

// local = this rank's [sum g | sum g*xhat], global = the same summed over ranks (== local without SyncBatchNorm);
// out = [mean g | mean g*xhat | dgamma | dbeta] (fp32): parameter gradients from the local sums (DDP averages them)
__global__ void bn_bwd_finalize_kernel(int c, const double *__restrict__ local, const double *__restrict__ global_,
                                       const double *__restrict__ count_dev, double count_host, float *__restrict__ out) {
  const double count = count_dev ? *count_dev : count_host;
  for (int ch = blockIdx.x * blockDim.x + threadIdx.x; ch < c; ch += gridDim.x * blockDim.x) {
    out[ch] = (float)(global_[ch] / count);
    out[c + ch] = (float)(global_[c + ch] / count);
    out[2 * c + ch] = (float)local[c + ch];
    out[3 * c + ch] = (float)local[ch];
  }
}

int reduce_geometry(int64_t npix, int c, int &ppb, unsigned &grid, size_t &smem) {
  const int cb = c < 256 ? c : 256;
  const int lanes = 256 / cb;
  ppb = 64 * lanes;  // 64 elements per thread
  const int64_t blocks = cdiv(npix, ppb);
  if (blocks > INT32_MAX) return -1;
  grid = (unsigned)blocks;
  smem = (size_t)2 * lanes * cb * sizeof(float);
  return 0;
}

}  // namespace

}  // namespace rtseg

using namespace rtseg;

#define RTSEG_BY_DTYPE(dtype, CALL)                       \
  do {                                                    \
    if ((dtype) == RTSEG_BF16) { using T = __nv_bfloat16; CALL; } \
    else if ((dtype) == RTSEG_F32) { using T = float; CALL; }      \
    else RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");  \
  } while (0)

extern "C" int rtseg_bn_stats(int64_t npix, int32_t c, int32_t dtype, const void *y, int32_t y_ctotal, int32_t y_coff,
                              const void *res, int32_t r_ctotal, int32_t r_coff, double *sum, double *sumsq,
                              rtseg_stream_t stream) {
  RTSEG_REQUIRE(y && sum, "NULL pointer");
  RTSEG_REQUIRE(npix > 0 && c > 0 && y_ctotal >= y_coff + c && y_coff >= 0, "bad extent");
  RTSEG_REQUIRE(!res || (r_ctotal >= r_coff + c && r_coff >= 0), "bad residual slice");
  int ppb;
  unsigned grid;
  size_t smem;
  RTSEG_REQUIRE(reduce_geometry(npix, c, ppb, grid, smem) == 0, "tensor too large");
  const Slice ys{y_ctotal, y_coff}, rs{r_ctotal, r_coff};
  if (vec_ok(dtype, c, {ys, rs}, y, res) && reduce_geometry_vec(npix, c, dtype == RTSEG_BF16 ? 8 : 4, ppb, grid, smem) == 0) {
    RTSEG_BY_DTYPE(dtype, (bn_stats_vec_kernel<T><<<grid, 256, smem, as_stream(stream)>>>(npix, c, ppb, (const T *)y, ys,
                                                                                          (const T *)res, rs, sum, sumsq)));
    RTSEG_LAUNCH_CHECK("bn_stats_vec_kernel");
    return RTSEG_OK;
  }
  RTSEG_BY_DTYPE(dtype, (bn_stats_kernel<T><<<grid, 256, smem, as_stream(stream)>>>(npix, c, ppb, (const T *)y, ys, (const T *)res, rs,
                                                                                    sum, sumsq)));
  RTSEG_LAUNCH_CHECK("bn_stats_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_affine_act(int64_t npix, int32_t c, int32_t dtype, const void *y, int32_t y_ctotal, int32_t y_coff,
                                const void *res, int32_t r_ctotal, int32_t r_coff, const float *scale, const float *shift,
                                int32_t act, float act_p0, float act_p1, const float *slope, void *z, int32_t z_ctotal,
                                int32_t z_coff, rtseg_stream_t stream) {
  RTSEG_REQUIRE(y && z, "NULL pointer");
  RTSEG_REQUIRE(npix > 0 && c > 0 && y_ctotal >= y_coff + c && z_ctotal >= z_coff + c && y_coff >= 0 && z_coff >= 0, "bad extent");
  RTSEG_REQUIRE(!res || (r_ctotal >= r_coff + c && r_coff >= 0), "bad residual slice");
  const Slice ys{y_ctotal, y_coff}, rs{r_ctotal, r_coff}, zs{z_ctotal, z_coff};
  const ActP ap{act, act_p0, act_p1, slope};
  if (vec_ok(dtype, c, {ys, rs, zs}, y, res, z)) {
    const unsigned gv = (unsigned)cdiv(npix * (c / (dtype == RTSEG_BF16 ? 8 : 4)), 256);
    RTSEG_BY_DTYPE(dtype, (affine_act_vec_kernel<T><<<gv, 256, 0, as_stream(stream)>>>(npix, c, (const T *)y, ys, (const T *)res, rs,
                                                                                       scale, shift, ap, (T *)z, zs)));
    RTSEG_LAUNCH_CHECK("affine_act_vec_kernel");
    return RTSEG_OK;
  }
  const unsigned grid = (unsigned)cdiv(npix * c, 256);
  RTSEG_BY_DTYPE(dtype, (affine_act_kernel<T><<<grid, 256, 0, as_stream(stream)>>>(npix, c, (const T *)y, ys, (const T *)res, rs, scale,
                                                                                   shift, ap, (T *)z, zs)));
  RTSEG_LAUNCH_CHECK("affine_act_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_bn_act_bwd_reduce(int64_t npix, int32_t c, int32_t dtype, const void *dz, int32_t d_ctotal, int32_t d_coff,
                                       const void *y, int32_t y_ctotal, int32_t y_coff, const void *res, int32_t r_ctotal,
                                       int32_t r_coff, const float *scale, const float *shift, const float *mean,
                                       const float *invstd, int32_t act, float act_p0, float act_p1, const float *slope,
                                       double *sum_g, double *sum_gx, rtseg_stream_t stream) {
  RTSEG_REQUIRE(dz && y && scale && shift && mean && invstd && sum_g && sum_gx, "NULL pointer");
  RTSEG_REQUIRE(npix > 0 && c > 0 && y_ctotal >= y_coff + c && d_ctotal >= d_coff + c, "bad extent");
  int ppb;
  unsigned grid;
  size_t smem;
  RTSEG_REQUIRE(reduce_geometry(npix, c, ppb, grid, smem) == 0, "tensor too large");
  const Slice ds{d_ctotal, d_coff}, ys{y_ctotal, y_coff}, rs{r_ctotal, r_coff};
  const ActP ap{act, act_p0, act_p1, slope};
  if (vec_ok(dtype, c, {ds, ys, rs}, dz, y, res) && reduce_geometry_vec(npix, c, dtype == RTSEG_BF16 ? 8 : 4, ppb, grid, smem) == 0) {
    RTSEG_BY_DTYPE(dtype, (bn_act_bwd_reduce_vec_kernel<T><<<grid, 256, smem, as_stream(stream)>>>(
                              npix, c, ppb, (const T *)dz, ds, (const T *)y, ys, (const T *)res, rs, scale, shift, mean, invstd, ap,
                              sum_g, sum_gx)));
    RTSEG_LAUNCH_CHECK("bn_act_bwd_reduce_vec_kernel");
    return RTSEG_OK;
  }
  RTSEG_BY_DTYPE(dtype, (bn_act_bwd_reduce_kernel<T><<<grid, 256, smem, as_stream(stream)>>>(
                            npix, c, ppb, (const T *)dz, ds, (const T *)y, ys, (const T *)res, rs, scale, shift, mean, invstd, ap,
                            sum_g, sum_gx)));
  RTSEG_LAUNCH_CHECK("bn_act_bwd_reduce_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_bn_act_bwd_apply(int64_t npix, int32_t c, int32_t dtype, const void *dz, int32_t d_ctotal, int32_t d_coff,
                                      const void *y, int32_t y_ctotal, int32_t y_coff, const void *res, int32_t r_ctotal,
                                      int32_t r_coff, const float *scale, const float *shift, const float *mean,
                                      const float *invstd, const float *k, const float *mean_g, const float *mean_gx,
                                      int32_t act, float act_p0, float act_p1, const float *slope, void *du, int32_t u_ctotal,
                                      int32_t u_coff, rtseg_stream_t stream) {
  RTSEG_REQUIRE(dz && y && scale && shift && mean && invstd && k && mean_g && mean_gx && du, "NULL pointer");
  RTSEG_REQUIRE(npix > 0 && c > 0 && y_ctotal >= y_coff + c && d_ctotal >= d_coff + c && u_ctotal >= u_coff + c, "bad extent");
  const Slice ds{d_ctotal, d_coff}, ys{y_ctotal, y_coff}, rs{r_ctotal, r_coff}, us{u_ctotal, u_coff};
  const ActP ap{act, act_p0, act_p1, slope};
  if (vec_ok(dtype, c, {ds, ys, rs, us}, dz, y, res, du)) {
    const unsigned gv = (unsigned)cdiv(npix * (c / (dtype == RTSEG_BF16 ? 8 : 4)), 256);
    RTSEG_BY_DTYPE(dtype, (bn_act_bwd_apply_vec_kernel<T><<<gv, 256, 0, as_stream(stream)>>>(
                              npix, c, (const T *)dz, ds, (const T *)y, ys, (const T *)res, rs, scale, shift, mean, invstd, k, mean_g,
                              mean_gx, ap, (T *)du, us)));
    RTSEG_LAUNCH_CHECK("bn_act_bwd_apply_vec_kernel");
    return RTSEG_OK;
  }
  const unsigned grid = (unsigned)cdiv(npix * c, 256);
  RTSEG_BY_DTYPE(dtype, (bn_act_bwd_apply_kernel<T><<<grid, 256, 0, as_stream(stream)>>>(
                            npix, c, (const T *)dz, ds, (const T *)y, ys, (const T *)res, rs, scale, shift, mean, invstd, k, mean_g,
                            mean_gx, ap, (T *)du, us)));
  RTSEG_LAUNCH_CHECK("bn_act_bwd_apply_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_conv_wgrad(const rtseg_conv_desc *d, const void *x, const void *dy, float *dw, rtseg_stream_t stream) {
  RTSEG_REQUIRE(d && x && dy && dw, "NULL pointer");
  RTSEG_REQUIRE(d->groups == 1, "rtseg_conv_wgrad: dense convolutions only");
  RTSEG_REQUIRE(d->n > 0 && d->h > 0 && d->w > 0 && d->cin > 0 && d->cout > 0 && d->kh > 0 && d->kw > 0, "non-positive extent");
  RTSEG_REQUIRE(d->x_ctotal >= d->x_coff + d->cin && d->y_ctotal >= d->y_coff + d->cout, "channel slice out of range");
  WgradParams p;
  p.n = d->n, p.h = d->h, p.w = d->w;
  rtseg_out_hw(d, &p.gh, &p.gw);
  p.stride = d->stride, p.dil = d->dil, p.pad_h = d->pad_h, p.pad_w = d->pad_w, p.kh = d->kh, p.kw = d->kw;
  p.cin = d->cin, p.cout = d->cout;
  p.xs = Slice{d->x_ctotal, d->x_coff}, p.gs = Slice{d->y_ctotal, d->y_coff};
  const int tiles = (int)(cdiv(p.cin, 64) * cdiv(p.cout, 64)), taps = p.kh * p.kw;
  const int64_t M = (int64_t)p.n * p.gh * p.gw;
  int splits = (int)std::max<int64_t>(1, std::min<int64_t>(cdiv(M, 256), (int64_t)num_sms() * 6 / std::max(1, tiles * taps)));
  splits = std::min(splits, 65535);
  p.chunk = cdiv(cdiv(M, splits), 64) * 64;
  splits = (int)cdiv(M, p.chunk);
  RTSEG_REQUIRE(taps <= 65535, "too many taps");
  dim3 grid((unsigned)tiles, (unsigned)taps, (unsigned)splits);
  static const int wmma_knob = getenv("RTSEG_WGRAD_WMMA") ? atoi(getenv("RTSEG_WGRAD_WMMA")) : 1;
  if (d->dtype == RTSEG_BF16 && wmma_knob) {
    conv_wgrad_wmma_kernel<<<grid, 256, 0, as_stream(stream)>>>(p, (const __nv_bfloat16 *)x, (const __nv_bfloat16 *)dy, dw);
    RTSEG_LAUNCH_CHECK("conv_wgrad_wmma_kernel");
    return RTSEG_OK;
  }
  RTSEG_BY_DTYPE(d->dtype, (conv_wgrad_kernel<T><<<grid, 256, 0, as_stream(stream)>>>(p, (const T *)x, (const T *)dy, dw)));
  RTSEG_LAUNCH_CHECK("conv_wgrad_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_maxpool3x3s2_bwd(int32_t n, int32_t h, int32_t w, int32_t c, int32_t dtype, const void *x, int32_t x_ctotal,
                                      int32_t x_coff, const void *dy, int32_t g_ctotal, int32_t g_coff, void *dx, int32_t d_ctotal,
                                      int32_t d_coff, rtseg_stream_t stream) {
  RTSEG_REQUIRE(x && dy && dx, "NULL pointer");
  RTSEG_REQUIRE(n > 0 && h > 0 && w > 0 && c > 0, "non-positive extent");
  RTSEG_REQUIRE(x_ctotal >= x_coff + c && g_ctotal >= g_coff + c && d_ctotal >= d_coff + c, "channel slice out of range");
  const int ho = (h - 1) / 2 + 1, wo = (w - 1) / 2 + 1;
  const Slice xs{x_ctotal, x_coff}, gs{g_ctotal, g_coff}, ds{d_ctotal, d_coff};
  if (vec_ok(dtype, c, {xs, gs, ds}, x, dy, dx)) {
    const unsigned gv = (unsigned)cdiv((int64_t)n * h * w * (c / (dtype == RTSEG_BF16 ? 8 : 4)), 256);
    RTSEG_BY_DTYPE(dtype, (maxpool3x3s2_bwd_vec_kernel<T><<<gv, 256, 0, as_stream(stream)>>>(n, h, w, c, ho, wo, (const T *)x, xs,
                                                                                            (const T *)dy, gs, (T *)dx, ds)));
    RTSEG_LAUNCH_CHECK("maxpool3x3s2_bwd_vec_kernel");
    return RTSEG_OK;
  }
  const unsigned grid = (unsigned)cdiv((int64_t)n * h * w * c, 256);
  RTSEG_BY_DTYPE(dtype, (maxpool3x3s2_bwd_kernel<T><<<grid, 256, 0, as_stream(stream)>>>(n, h, w, c, ho, wo, (const T *)x, xs,
                                                                                          (const T *)dy, gs, (T *)dx, ds)));
  RTSEG_LAUNCH_CHECK("maxpool3x3s2_bwd_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_bn_finalize(int32_t c, const double *stats, const double *count_dev, double count_host, const float *gamma,
                                 const float *beta, double eps, double momentum, float *running_mean, float *running_var,
                                 float *out, rtseg_stream_t stream) {
  RTSEG_REQUIRE(stats && out && c > 0, "NULL pointer or non-positive channel count");
  RTSEG_REQUIRE((running_mean == nullptr) == (running_var == nullptr), "running_mean and running_var go together");
  RTSEG_REQUIRE(count_dev || count_host > 0, "element count must be positive");
  bn_finalize_kernel<<<(unsigned)cdiv(c, 256), 256, 0, as_stream(stream)>>>(c, stats, count_dev, count_host, gamma, beta, eps, momentum,
                                                                         running_mean, running_var, out);
  RTSEG_LAUNCH_CHECK("bn_finalize_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_bn_bwd_finalize(int32_t c, const double *local_sums, const double *global_sums, const double *count_dev,
                                     double count_host, float *out, rtseg_stream_t stream) {
  RTSEG_REQUIRE(local_sums && global_sums && out && c > 0, "NULL pointer or non-positive channel count");
  RTSEG_REQUIRE(count_dev || count_host > 0, "element count must be positive");
  bn_bwd_finalize_kernel<<<(unsigned)cdiv(c, 256), 256, 0, as_stream(stream)>>>(c, local_sums, global_sums, count_dev, count_host, out);
  RTSEG_LAUNCH_CHECK("bn_bwd_finalize_kernel");
  return RTSEG_OK;
}
