// conv_dw.cu — K1': NHWC depthwise convolution + folded BatchNorm + activation.
//
// Replaces DWConvBNAct (reference models/modules.py:46-59): Conv2d(groups=cin)
// -> BatchNorm2d -> Activation, i.e. three ATen kernels and five tensor passes,
// with ONE pass: read x once, write y once.  The op is 1.5-4.5 flop/byte, so the
// only roofline that matters is HBM; the design rules are
//   * channels innermost (NHWC) so that a thread owns a 16-byte channel vector
//     (8 x bf16 / 4 x f32) and a warp reads/writes 512 contiguous bytes,
//   * each thread produces PX neighbouring output pixels so that every input
//     vector fetched from L1/L2 feeds up to KW outputs (dilation 1) and the
//     per-tap weight vectors are amortised over PX pixels,
//   * fp32 accumulation, one rounding at the store.
#include <cstdlib>

#include "common.cuh"

namespace rtseg {

int dw_tma_try(const rtseg_conv_desc *d, const void *x, const float *w, const Epilogue &ep, const void *res, void *y,
               cudaStream_t st);  // dw_tma.cu

struct DwParams {
  int n, h, w, c;  // c = cin = cout for the vector kernels
  int ho, wo;
  int kh, kw, stride, dil, pad_h, pad_w;
  int x_ct, x_co, y_ct, y_co, r_ct, r_co;
  int mult;  // channel multiplier (scalar kernel only)
  int cout;
};

// ---- dilation-1 kernel with a sliding register window -----------------------
template <typename T, int KH, int KW, int S, int PX, int MODE>
__global__ void __launch_bounds__(256) dw_win_kernel(DwParams p, const T *__restrict__ x,
                                                     const float *__restrict__ wgt, Epilogue ep,
                                                     const T *__restrict__ res, T *__restrict__ y) {
  constexpr int V = Vec16<T>::N;
  constexpr int NW = (PX - 1) * S + KW;  // input columns feeding PX outputs
  const int cvecs = p.c / V;
  const int wq_n = (p.wo + PX - 1) / PX;
  const int64_t total = (int64_t)p.n * p.ho * wq_n * cvecs;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= total) return;
  const int cv = (int)(tid % cvecs);
  int64_t r = tid / cvecs;
  const int wq = (int)(r % wq_n);
  r /= wq_n;
  const int ho = (int)(r % p.ho);
  const int n = (int)(r / p.ho);
  const int c0 = cv * V;

  // fp32 accumulation on channel PAIRS: sm_100's packed FFMA2 retires two fp32 FMAs per issue slot,
  // which matters because a stride-1 3x3 depthwise layer is issue-bound, not HBM-bound, on B200.
  constexpr int V2 = V / 2;
  float2 acc[PX][V2];
#pragma unroll
  for (int i = 0; i < PX; ++i)
#pragma unroll
    for (int v = 0; v < V2; ++v) acc[i][v] = make_float2(0.f, 0.f);

  const int wi0 = wq * PX * S - p.pad_w;
  const int hi0 = ho * S - p.pad_h;
  const T *xn = x + (int64_t)n * p.h * p.w * p.x_ct + p.x_co + c0;

#pragma unroll
  for (int ky = 0; ky < KH; ++ky) {
    const int hi = hi0 + ky;
    if (hi < 0 || hi >= p.h) continue;
    float2 wv[KW][V2];
#pragma unroll
    for (int kx = 0; kx < KW; ++kx) {
      const float *wp = wgt + (int64_t)(ky * KW + kx) * p.cout + c0;
#pragma unroll
      for (int v = 0; v < V; v += 4) {
        const float4 t = __ldg(reinterpret_cast<const float4 *>(wp + v));
        wv[kx][v / 2] = make_float2(t.x, t.y), wv[kx][v / 2 + 1] = make_float2(t.z, t.w);
      }
    }
    const T *xr = xn + (int64_t)hi * p.w * p.x_ct;
    // Issue the whole row of loads before touching any of them: NW independent 16-byte requests in
    // flight per thread instead of NW serial DRAM round trips (out-of-image columns read as zero).
    typename Vec16<T>::Raw raw[NW];
#pragma unroll
    for (int j = 0; j < NW; ++j) {
      const int wi = wi0 + j;
      raw[j] = (wi >= 0 && wi < p.w) ? Vec16<T>::ldg(xr + (int64_t)wi * p.x_ct) : Vec16<T>::zero();
    }
#pragma unroll
    for (int j = 0; j < NW; ++j) {
      float2 xv[V2];
      Vec16<T>::unpack2(raw[j], xv);
#pragma unroll
      for (int kx = 0; kx < KW; ++kx) {
        const int d = j - kx;
        if (d >= 0 && d % S == 0 && d / S < PX) {
#pragma unroll
          for (int v = 0; v < V2; ++v) acc[d / S][v] = __ffma2_rn(xv[v], wv[kx][v], acc[d / S][v]);
        }
      }
    }
  }

  const int64_t pix0 = ((int64_t)n * p.ho + ho) * p.wo + (int64_t)wq * PX;
  float sc[V], sh[V], sl[V];
  load_affine<V>(ep, c0, sc, sh, sl);
#pragma unroll
  for (int i = 0; i < PX; ++i) {
    if (wq * PX + i >= p.wo) break;
    float rv[V];
    if (ep.res_mode != RTSEG_RES_NONE) {
      Vec16<T>::load(res + (pix0 + i) * p.r_ct + p.r_co + c0, rv);
    } else {
#pragma unroll
      for (int v = 0; v < V; ++v) rv[v] = 0.f;
    }
    float ov[V];
#pragma unroll
    for (int v = 0; v < V2; ++v) {
      ov[2 * v] = epi_elem<MODE>(acc[i][v].x, rv[2 * v], sc[2 * v], sh[2 * v], sl[2 * v], ep);
      ov[2 * v + 1] = epi_elem<MODE>(acc[i][v].y, rv[2 * v + 1], sc[2 * v + 1], sh[2 * v + 1], sl[2 * v + 1], ep);
    }
    Vec16<T>::store(y + (pix0 + i) * p.y_ct + p.y_co + c0, ov);
  }
}

// ---- 3x1 / 1x3 with any dilation, stride 1 (DABNet's asymmetric depthwise pairs) -----------------
// All 3*PX tap loads and the PX residual loads are issued before the first FMA: one DRAM round trip per
// thread instead of three (the generic kernel below waits once per tap).
template <typename T, bool VERT, int PX, int MODE>
__global__ void __launch_bounds__(256) dw_tap3_kernel(DwParams p, const T *__restrict__ x,
                                                      const float *__restrict__ wgt, Epilogue ep,
                                                      const T *__restrict__ res, T *__restrict__ y) {
  constexpr int V = Vec16<T>::N;
  constexpr int V2 = V / 2;
  const int cvecs = p.c / V;
  const int wq_n = (p.wo + PX - 1) / PX;
  const int64_t total = (int64_t)p.n * p.ho * wq_n * cvecs;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= total) return;
  const int cv = (int)(tid % cvecs);
  int64_t r = tid / cvecs;
  const int wq = (int)(r % wq_n);
  r /= wq_n;
  const int ho = (int)(r % p.ho);
  const int n = (int)(r / p.ho);
  const int c0 = cv * V;
  const T *xn = x + (int64_t)n * p.h * p.w * p.x_ct + p.x_co + c0;
  const int64_t pix0 = ((int64_t)n * p.ho + ho) * p.wo + (int64_t)wq * PX;

  typename Vec16<T>::Raw raw[3][PX], rraw[PX];
#pragma unroll
  for (int t = 0; t < 3; ++t) {
    const int hi = VERT ? ho - p.pad_h + t * p.dil : ho - p.pad_h;
    const bool row_ok = hi >= 0 && hi < p.h;
    const T *xr = xn + (int64_t)hi * p.w * p.x_ct;
#pragma unroll
    for (int i = 0; i < PX; ++i) {
      const int wi = wq * PX + i - p.pad_w + (VERT ? 0 : t * p.dil);
      raw[t][i] = (row_ok && wi >= 0 && wi < p.w) ? Vec16<T>::ldg(xr + (int64_t)wi * p.x_ct) : Vec16<T>::zero();
    }
  }
  const bool has_res = ep.res_mode != RTSEG_RES_NONE;
#pragma unroll
  for (int i = 0; i < PX; ++i)
    rraw[i] = (has_res && wq * PX + i < p.wo) ? Vec16<T>::ldg(res + (pix0 + i) * p.r_ct + p.r_co + c0) : Vec16<T>::zero();

  float2 wv[3][V2];
#pragma unroll
  for (int t = 0; t < 3; ++t) {
    const float *wp = wgt + (int64_t)t * p.cout + c0;
#pragma unroll
    for (int v = 0; v < V; v += 4) {
      const float4 q = __ldg(reinterpret_cast<const float4 *>(wp + v));
      wv[t][v / 2] = make_float2(q.x, q.y), wv[t][v / 2 + 1] = make_float2(q.z, q.w);
    }
  }
  float sc[V], sh[V], sl[V];
  load_affine<V>(ep, c0, sc, sh, sl);
#pragma unroll
  for (int i = 0; i < PX; ++i) {
    if (wq * PX + i >= p.wo) break;
    float2 acc[V2];
#pragma unroll
    for (int v = 0; v < V2; ++v) acc[v] = make_float2(0.f, 0.f);
#pragma unroll
    for (int t = 0; t < 3; ++t) {
      float2 xv[V2];
      Vec16<T>::unpack2(raw[t][i], xv);
#pragma unroll
      for (int v = 0; v < V2; ++v) acc[v] = __ffma2_rn(xv[v], wv[t][v], acc[v]);
    }
    float rv[V], ov[V];
    Vec16<T>::unpack(rraw[i], rv);
#pragma unroll
    for (int v = 0; v < V2; ++v) {
      ov[2 * v] = epi_elem<MODE>(acc[v].x, rv[2 * v], sc[2 * v], sh[2 * v], sl[2 * v], ep);
      ov[2 * v + 1] = epi_elem<MODE>(acc[v].y, rv[2 * v + 1], sc[2 * v + 1], sh[2 * v + 1], sl[2 * v + 1], ep);
    }
    Vec16<T>::store(y + (pix0 + i) * p.y_ct + p.y_co + c0, ov);
  }
}

// ---- any kernel size / dilation / stride, vectorised over channels ------------
template <typename T, int PX, int MODE>
__global__ void __launch_bounds__(256) dw_vec_kernel(DwParams p, const T *__restrict__ x,
                                                     const float *__restrict__ wgt, Epilogue ep,
                                                     const T *__restrict__ res, T *__restrict__ y) {
  constexpr int V = Vec16<T>::N;
  const int cvecs = p.c / V;
  const int wq_n = (p.wo + PX - 1) / PX;
  const int64_t total = (int64_t)p.n * p.ho * wq_n * cvecs;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= total) return;
  const int cv = (int)(tid % cvecs);
  int64_t r = tid / cvecs;
  const int wq = (int)(r % wq_n);
  r /= wq_n;
  const int ho = (int)(r % p.ho);
  const int n = (int)(r / p.ho);
  const int c0 = cv * V;

  float acc[PX][V];
#pragma unroll
  for (int i = 0; i < PX; ++i)
#pragma unroll
    for (int v = 0; v < V; ++v) acc[i][v] = 0.f;

  const T *xn = x + (int64_t)n * p.h * p.w * p.x_ct + p.x_co + c0;
  for (int ky = 0; ky < p.kh; ++ky) {
    const int hi = ho * p.stride - p.pad_h + ky * p.dil;
    if (hi < 0 || hi >= p.h) continue;
    const T *xr = xn + (int64_t)hi * p.w * p.x_ct;
    for (int kx = 0; kx < p.kw; ++kx) {
      float wv[V];
      const float *wp = wgt + (int64_t)(ky * p.kw + kx) * p.cout + c0;
#pragma unroll
      for (int v = 0; v < V; v += 4) {
        const float4 t = __ldg(reinterpret_cast<const float4 *>(wp + v));
        wv[v] = t.x, wv[v + 1] = t.y, wv[v + 2] = t.z, wv[v + 3] = t.w;
      }
      typename Vec16<T>::Raw raw[PX];
#pragma unroll
      for (int i = 0; i < PX; ++i) {
        const int wi = (wq * PX + i) * p.stride - p.pad_w + kx * p.dil;
        raw[i] = (wi >= 0 && wi < p.w) ? Vec16<T>::ldg(xr + (int64_t)wi * p.x_ct) : Vec16<T>::zero();
      }
#pragma unroll
      for (int i = 0; i < PX; ++i) {
        float xv[V];
        Vec16<T>::unpack(raw[i], xv);
#pragma unroll
        for (int v = 0; v < V; ++v) acc[i][v] = fmaf(xv[v], wv[v], acc[i][v]);
      }
    }
  }

  const int64_t pix0 = ((int64_t)n * p.ho + ho) * p.wo + (int64_t)wq * PX;
  float sc[V], sh[V], sl[V];
  load_affine<V>(ep, c0, sc, sh, sl);
#pragma unroll
  for (int i = 0; i < PX; ++i) {
    if (wq * PX + i >= p.wo) break;
    float rv[V];
    if (ep.res_mode != RTSEG_RES_NONE) {
      Vec16<T>::load(res + (pix0 + i) * p.r_ct + p.r_co + c0, rv);
    } else {
#pragma unroll
      for (int v = 0; v < V; ++v) rv[v] = 0.f;
    }
    float ov[V];
#pragma unroll
    for (int v = 0; v < V; ++v) ov[v] = epi_elem<MODE>(acc[i][v], rv[v], sc[v], sh[v], sl[v], ep);
    Vec16<T>::store(y + (pix0 + i) * p.y_ct + p.y_co + c0, ov);
  }
}

// ---- channel multiplier m >= 2 (BiSeNetv2's x6 depthwise expansion, reference models/bisenetv2.py:139-149) ------
// One thread = one output pixel x one 16-byte vector of OUTPUT channels.  Output channel co reads input channel
// co / m, and the host only takes this kernel when every vector's V outputs come from at most two neighbouring
// input channels (m = 6, V = 8: always), so a tap costs two scalar input loads (L1 hits: neighbouring threads share
// them), two float4 weight loads and V FMAs.  The layer is bound by its output: 1/m of the bytes are reads.
template <typename T, int MODE>
__global__ void __launch_bounds__(256) dw_mult_kernel(DwParams p, const T *__restrict__ x, const float *__restrict__ wgt,
                                                      Epilogue ep, const T *__restrict__ res, T *__restrict__ y) {
  constexpr int V = Vec16<T>::N;
  const int cvecs = p.cout / V;
  const int64_t total = (int64_t)p.n * p.ho * p.wo * cvecs;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= total) return;
  const int c0 = (int)(tid % cvecs) * V;
  int64_t r = tid / cvecs;
  const int wo = (int)(r % p.wo);
  r /= p.wo;
  const int ho = (int)(r % p.ho);
  const int n = (int)(r / p.ho);
  const int ci0 = c0 / p.mult;
  const int phase = c0 - ci0 * p.mult;          // outputs with phase + e >= mult read input channel ci0 + 1
  const bool two = phase + V - 1 >= p.mult;
  float acc[V];
#pragma unroll
  for (int v = 0; v < V; ++v) acc[v] = 0.f;
  const T *xn = x + (int64_t)n * p.h * p.w * p.x_ct + p.x_co + ci0;
  for (int ky = 0; ky < p.kh; ++ky) {
    const int hi = ho * p.stride - p.pad_h + ky * p.dil;
    if (hi < 0 || hi >= p.h) continue;
    for (int kx = 0; kx < p.kw; ++kx) {
      const int wi = wo * p.stride - p.pad_w + kx * p.dil;
      if (wi < 0 || wi >= p.w) continue;
      const T *xp = xn + ((int64_t)hi * p.w + wi) * p.x_ct;
      const float x0 = ld_f<T>(xp), x1 = two ? ld_f<T>(xp + 1) : 0.f;
      const float *wp = wgt + (int64_t)(ky * p.kw + kx) * p.cout + c0;
#pragma unroll
      for (int v = 0; v < V; v += 4) {
        const float4 t = __ldg(reinterpret_cast<const float4 *>(wp + v));
        acc[v] = fmaf(phase + v >= p.mult ? x1 : x0, t.x, acc[v]);
        acc[v + 1] = fmaf(phase + v + 1 >= p.mult ? x1 : x0, t.y, acc[v + 1]);
        acc[v + 2] = fmaf(phase + v + 2 >= p.mult ? x1 : x0, t.z, acc[v + 2]);
        acc[v + 3] = fmaf(phase + v + 3 >= p.mult ? x1 : x0, t.w, acc[v + 3]);
      }
    }
  }
  const int64_t pix = ((int64_t)n * p.ho + ho) * p.wo + wo;
  float sc[V], sh[V], sl[V], rv[V], ov[V];
  load_affine<V>(ep, c0, sc, sh, sl);
  if (ep.res_mode != RTSEG_RES_NONE) {
    Vec16<T>::load(res + pix * p.r_ct + p.r_co + c0, rv);
  } else {
#pragma unroll
    for (int v = 0; v < V; ++v) rv[v] = 0.f;
  }
#pragma unroll
  for (int v = 0; v < V; ++v) ov[v] = epi_elem<MODE>(acc[v], rv[v], sc[v], sh[v], sl[v], ep);
  Vec16<T>::store(y + pix * p.y_ct + p.y_co + c0, ov);
}

// ---- channel multiplier, bf16, 3x3: one thread = ONE input channel and its M output channels, PX pixels of a row ----
// With the taps of its M filters (9 M floats, BatchNorm scale folded in) in registers for the whole walk, an output
// pixel costs 3 new input loads (a 3 x 3 window sliding along the row), 9 M FMAs and M / 2 four-byte stores — no
// per-tap weight loads and no input selects (dw_mult_kernel spends two thirds of its instructions on those).  Lanes
// are neighbouring input channels, so a warp stores 32 M contiguous bf16 per pixel.
template <int M, int S, int PX, int MODE>
__global__ void __launch_bounds__(128) dw_chan_kernel(DwParams p, const __nv_bfloat16 *__restrict__ x,
                                                      const float *__restrict__ wgt, Epilogue ep, __nv_bfloat16 *__restrict__ y) {
  const int wq_n = (p.wo + PX - 1) / PX;
  const int64_t total = (int64_t)p.n * p.ho * wq_n * p.c;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= total) return;
  const int ci = (int)(tid % p.c);
  int64_t r = tid / p.c;
  const int wq = (int)(r % wq_n);
  r /= wq_n;
  const int ho = (int)(r % p.ho);
  const int n = (int)(r / p.ho);
  const int co0 = ci * M;
  float w[9][M], sh[M], sl[M];
#pragma unroll
  for (int j = 0; j < M; ++j) {
    const float sc = ep.scale ? __ldg(ep.scale + co0 + j) : 1.f;
    sh[j] = ep.shift ? __ldg(ep.shift + co0 + j) : 0.f;
    sl[j] = (ep.slope && ep.act == RTSEG_ACT_PRELU) ? __ldg(ep.slope + co0 + j) : 0.f;
#pragma unroll
    for (int t = 0; t < 9; ++t) w[t][j] = __ldg(wgt + (int64_t)t * p.cout + co0 + j) * sc;
  }
  const __nv_bfloat16 *xn = x + (int64_t)n * p.h * p.w * p.x_ct + p.x_co + ci;
  const int hi0 = ho * S - p.pad_h;
  const __nv_bfloat16 *row[3];
  bool rok[3];
#pragma unroll
  for (int ky = 0; ky < 3; ++ky) {
    rok[ky] = hi0 + ky >= 0 && hi0 + ky < p.h;
    row[ky] = xn + (int64_t)(rok[ky] ? hi0 + ky : 0) * p.w * p.x_ct;
  }
  auto ld = [&](int ky, int wi) -> float {
    return (rok[ky] && wi >= 0 && wi < p.w) ? __bfloat162float(row[ky][(int64_t)wi * p.x_ct]) : 0.f;
  };
  const int wo0 = wq * PX;
  float win[3][3];   // win[ky][kx] for the current output pixel
  int wi = wo0 * S - p.pad_w;
#pragma unroll
  for (int ky = 0; ky < 3; ++ky) {
    win[ky][0] = ld(ky, wi), win[ky][1] = ld(ky, wi + 1);
  }
  __nv_bfloat16 *yp = y + (((int64_t)n * p.ho + ho) * p.wo + wo0) * p.y_ct + p.y_co + co0;
#pragma unroll
  for (int i = 0; i < PX; ++i) {
    if (wo0 + i >= p.wo) break;
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) win[ky][2] = ld(ky, wi + 2);
    float acc[M];
#pragma unroll
    for (int j = 0; j < M; ++j) acc[j] = 0.f;
#pragma unroll
    for (int ky = 0; ky < 3; ++ky)
#pragma unroll
      for (int kx = 0; kx < 3; ++kx)
#pragma unroll
        for (int j = 0; j < M; ++j) acc[j] = fmaf(win[ky][kx], w[ky * 3 + kx][j], acc[j]);
    uint32_t *out = reinterpret_cast<uint32_t *>(yp + (int64_t)i * p.y_ct);
#pragma unroll
    for (int j = 0; j < M; j += 2) {
      const float a = epi_elem<MODE, false>(acc[j], 0.f, 1.f, sh[j], sl[j], ep), b = epi_elem<MODE, false>(acc[j + 1], 0.f, 1.f, sh[j + 1], sl[j + 1], ep);
      const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
      out[j >> 1] = *reinterpret_cast<const uint32_t *>(&h);
    }
    // slide the window: stride 1 keeps two columns, stride 2 one
    if (S == 1) {
#pragma unroll
      for (int ky = 0; ky < 3; ++ky) win[ky][0] = win[ky][1], win[ky][1] = win[ky][2];
      wi += 1;
    } else {
      wi += 2;
#pragma unroll
      for (int ky = 0; ky < 3; ++ky) win[ky][0] = win[ky][2], win[ky][1] = ld(ky, wi + 1);
    }
  }
}

// ---- scalar fallback: channel multiplier, odd channel counts, unaligned slices ---
template <typename T>
__global__ void __launch_bounds__(256) dw_scalar_kernel(DwParams p, const T *__restrict__ x,
                                                        const float *__restrict__ wgt, Epilogue ep,
                                                        const T *__restrict__ res, T *__restrict__ y) {
  const int64_t total = (int64_t)p.n * p.ho * p.wo * p.cout;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= total) return;
  const int co = (int)(tid % p.cout);
  int64_t r = tid / p.cout;
  const int wo = (int)(r % p.wo);
  r /= p.wo;
  const int ho = (int)(r % p.ho);
  const int n = (int)(r / p.ho);
  const int ci = co / p.mult;
  float acc = 0.f;
  const T *xn = x + (int64_t)n * p.h * p.w * p.x_ct + p.x_co + ci;
  for (int ky = 0; ky < p.kh; ++ky) {
    const int hi = ho * p.stride - p.pad_h + ky * p.dil;
    if (hi < 0 || hi >= p.h) continue;
    for (int kx = 0; kx < p.kw; ++kx) {
      const int wi = wo * p.stride - p.pad_w + kx * p.dil;
      if (wi < 0 || wi >= p.w) continue;
      acc = fmaf(ld_f<T>(xn + ((int64_t)hi * p.w + wi) * p.x_ct), __ldg(wgt + (int64_t)(ky * p.kw + kx) * p.cout + co), acc);
    }
  }
  const int64_t pix = ((int64_t)n * p.ho + ho) * p.wo + wo;
  const float rv = ep.res_mode != RTSEG_RES_NONE ? ld_f<T>(res + pix * p.r_ct + p.r_co + co) : 0.f;
  st_f<T>(y + pix * p.y_ct + p.y_co + co, epilogue_apply(acc, rv, co, ep));
}

template <typename T>
static int launch_dw(const rtseg_conv_desc *d, const void *x_, const float *w, const Epilogue &ep,
                     const void *res_, void *y_, cudaStream_t st) {
  const T *x = static_cast<const T *>(x_);
  const T *res = static_cast<const T *>(res_);
  T *y = static_cast<T *>(y_);
  DwParams p;
  p.n = d->n, p.h = d->h, p.w = d->w, p.c = d->cin, p.cout = d->cout;
  rtseg_out_hw(d, &p.ho, &p.wo);
  p.kh = d->kh, p.kw = d->kw, p.stride = d->stride, p.dil = d->dil, p.pad_h = d->pad_h, p.pad_w = d->pad_w;
  p.x_ct = d->x_ctotal, p.x_co = d->x_coff, p.y_ct = d->y_ctotal, p.y_co = d->y_coff;
  p.r_ct = d->r_ctotal, p.r_co = d->r_coff;
  p.mult = d->cout / d->cin;

  constexpr int V = Vec16<T>::N;
  const bool aligned = (reinterpret_cast<uintptr_t>(x) % 16 == 0) && (reinterpret_cast<uintptr_t>(y) % 16 == 0) &&
                       (res == nullptr || reinterpret_cast<uintptr_t>(res) % 16 == 0) &&
                       (reinterpret_cast<uintptr_t>(w) % 16 == 0);
  const bool vec_ok = aligned && p.mult == 1 && d->cin % V == 0 && d->x_ctotal % V == 0 && d->x_coff % V == 0 &&
                      d->y_ctotal % V == 0 && d->y_coff % V == 0 &&
                      (d->res_mode == RTSEG_RES_NONE || (d->r_ctotal % V == 0 && d->r_coff % V == 0));
  const int threads = 256;
  // channel multiplier: vector kernel when every output vector reads at most two neighbouring input channels
  bool mult_ok = aligned && p.mult >= 2 && d->cout % V == 0 && d->y_ctotal % V == 0 && d->y_coff % V == 0 &&
                 (d->res_mode == RTSEG_RES_NONE || (d->r_ctotal % V == 0 && d->r_coff % V == 0));
  for (int c0 = 0; mult_ok && c0 < d->cout; c0 += V) mult_ok = (c0 % p.mult) + V - 1 < 2 * p.mult;
  // bf16 3x3 with multiplier 6 (BiSeNetv2's gather-expansion layers), no residual: one thread per input channel
  static const int chan_knob = getenv("RTSEG_DW_CHAN") ? atoi(getenv("RTSEG_DW_CHAN")) : 1;
  if (chan_knob && sizeof(T) == 2 && p.mult == 6 && d->kh == 3 && d->kw == 3 && d->dil == 1 && (d->stride == 1 || d->stride == 2) &&
      d->res_mode == RTSEG_RES_NONE && d->y_ctotal % 2 == 0 && d->y_coff % 2 == 0 && reinterpret_cast<uintptr_t>(y) % 4 == 0) {
    constexpr int PX = 8;
    const int64_t total = (int64_t)p.n * p.ho * cdiv(p.wo, PX) * p.c;
    const unsigned grid = (unsigned)cdiv(total, 128);
    auto xb = reinterpret_cast<const __nv_bfloat16 *>(x);
    auto yb = reinterpret_cast<__nv_bfloat16 *>(y);
#define RTSEG_DW_CHAN(S_)                                                                   \
    do {                                                                                      \
      if (ep.mode == 0) dw_chan_kernel<6, S_, PX, 0><<<grid, 128, 0, st>>>(p, xb, w, ep, yb);      \
      else if (ep.mode == 1) dw_chan_kernel<6, S_, PX, 1><<<grid, 128, 0, st>>>(p, xb, w, ep, yb); \
      else dw_chan_kernel<6, S_, PX, 2><<<grid, 128, 0, st>>>(p, xb, w, ep, yb);                   \
    } while (0)
    if (d->stride == 1) RTSEG_DW_CHAN(1); else RTSEG_DW_CHAN(2);
#undef RTSEG_DW_CHAN
    RTSEG_LAUNCH_CHECK("dw_chan_kernel");
    return RTSEG_OK;
  }
  if (mult_ok) {
    const int64_t total = (int64_t)p.n * p.ho * p.wo * (d->cout / V);
    const unsigned grid = (unsigned)cdiv(total, threads);
    if (ep.mode == 0) dw_mult_kernel<T, 0><<<grid, threads, 0, st>>>(p, x, w, ep, res, y);
    else if (ep.mode == 1) dw_mult_kernel<T, 1><<<grid, threads, 0, st>>>(p, x, w, ep, res, y);
    else dw_mult_kernel<T, 2><<<grid, threads, 0, st>>>(p, x, w, ep, res, y);
    RTSEG_LAUNCH_CHECK("dw_mult_kernel");
    return RTSEG_OK;
  }
  if (!vec_ok) {
    const int64_t total = (int64_t)p.n * p.ho * p.wo * p.cout;
    dw_scalar_kernel<T><<<(unsigned)cdiv(total, threads), threads, 0, st>>>(p, x, w, ep, res, y);
    RTSEG_LAUNCH_CHECK("dw_scalar_kernel");
    return RTSEG_OK;
  }
  const int cvecs = p.c / V;
#define RTSEG_DW_WIN(KH_, KW_, S_, PX_)                                                                     \
  do {                                                                                                      \
    const int64_t total = (int64_t)p.n * p.ho * cdiv(p.wo, PX_) * cvecs;                                    \
    const unsigned grid = (unsigned)cdiv(total, threads);                                                   \
    if (ep.mode == 0)                                                                                       \
      dw_win_kernel<T, KH_, KW_, S_, PX_, 0><<<grid, threads, 0, st>>>(p, x, w, ep, res, y);                \
    else if (ep.mode == 1)                                                                                  \
      dw_win_kernel<T, KH_, KW_, S_, PX_, 1><<<grid, threads, 0, st>>>(p, x, w, ep, res, y);                \
    else                                                                                                    \
      dw_win_kernel<T, KH_, KW_, S_, PX_, 2><<<grid, threads, 0, st>>>(p, x, w, ep, res, y);                \
    RTSEG_LAUNCH_CHECK("dw_win_kernel");                                                                    \
    return RTSEG_OK;                                                                                        \
  } while (0)
  if (d->dil == 1 && d->kh == 3 && d->kw == 3) {
    if (d->stride == 1) {
      static const int px_override = getenv("RTSEG_DW_PX") ? atoi(getenv("RTSEG_DW_PX")) : 0;  // tuning knob
      if (px_override == 2) RTSEG_DW_WIN(3, 3, 1, 2);
      if (px_override == 4) RTSEG_DW_WIN(3, 3, 1, 4);
      if (p.wo >= 64) RTSEG_DW_WIN(3, 3, 1, 8);
      RTSEG_DW_WIN(3, 3, 1, 4);
    }
    if (d->stride == 2) RTSEG_DW_WIN(3, 3, 2, 4);
  }
  static const int tap3_knob = getenv("RTSEG_DW_TAP3") ? atoi(getenv("RTSEG_DW_TAP3")) : 1;  // 0 off, 2 = also 1x3 d1
  if (d->dil == 1 && d->stride == 1 && d->kh == 1 && d->kw == 3 && tap3_knob != 2) RTSEG_DW_WIN(1, 3, 1, 8);
#undef RTSEG_DW_WIN
  if (tap3_knob && d->stride == 1 && ((d->kh == 3 && d->kw == 1) || (d->kh == 1 && d->kw == 3))) {
    constexpr int PX = 4;
    const int64_t total = (int64_t)p.n * p.ho * cdiv(p.wo, PX) * cvecs;
    const unsigned grid = (unsigned)cdiv(total, threads);
#define RTSEG_DW_TAP3(VERT_)                                                                   \
  do {                                                                                         \
    if (ep.mode == 0) dw_tap3_kernel<T, VERT_, PX, 0><<<grid, threads, 0, st>>>(p, x, w, ep, res, y);      \
    else if (ep.mode == 1) dw_tap3_kernel<T, VERT_, PX, 1><<<grid, threads, 0, st>>>(p, x, w, ep, res, y); \
    else dw_tap3_kernel<T, VERT_, PX, 2><<<grid, threads, 0, st>>>(p, x, w, ep, res, y);                   \
  } while (0)
    if (d->kh == 3) RTSEG_DW_TAP3(true); else RTSEG_DW_TAP3(false);
#undef RTSEG_DW_TAP3
    RTSEG_LAUNCH_CHECK("dw_tap3_kernel");
    return RTSEG_OK;
  }
  {
    constexpr int PX = 4;
    const int64_t total = (int64_t)p.n * p.ho * cdiv(p.wo, PX) * cvecs;
    const unsigned grid = (unsigned)cdiv(total, threads);
    if (ep.mode == 0)
      dw_vec_kernel<T, PX, 0><<<grid, threads, 0, st>>>(p, x, w, ep, res, y);
    else if (ep.mode == 1)
      dw_vec_kernel<T, PX, 1><<<grid, threads, 0, st>>>(p, x, w, ep, res, y);
    else
      dw_vec_kernel<T, PX, 2><<<grid, threads, 0, st>>>(p, x, w, ep, res, y);
    RTSEG_LAUNCH_CHECK("dw_vec_kernel");
  }
  return RTSEG_OK;
}

}  // namespace rtseg

extern "C" int rtseg_conv_dw(const rtseg_conv_desc *d, const void *x, const float *w, const float *scale,
                             const float *shift, const float *slope, const void *res, void *y,
                             rtseg_stream_t stream) {
  using namespace rtseg;
  if (int rc = validate_conv_desc(d, res)) return rc;
  RTSEG_REQUIRE(x && w && y, "NULL tensor pointer");
  RTSEG_REQUIRE(d->groups == d->cin, "rtseg_conv_dw needs groups == cin");
  RTSEG_REQUIRE(d->cout % d->cin == 0, "depthwise cout (%d) must be a multiple of cin (%d)", d->cout, d->cin);
  const Epilogue ep = make_epilogue(d, scale, shift, slope);
  // Large 3x3 bf16 layers: TMA-staged shared-memory tiles (dw_tma.cu); everything else: register kernels.
  static const bool use_tma = !(getenv("RTSEG_DW_TMA") && atoi(getenv("RTSEG_DW_TMA")) == 0);  // A/B knob
  if (use_tma) {
    const int rc = dw_tma_try(d, x, w, ep, res, y, as_stream(stream));
    if (rc != RTSEG_ERR_UNSUPPORTED) return rc;
  }
  if (d->dtype == RTSEG_BF16) return launch_dw<__nv_bfloat16>(d, x, w, ep, res, y, as_stream(stream));
  return launch_dw<float>(d, x, w, ep, res, y, as_stream(stream));
}
