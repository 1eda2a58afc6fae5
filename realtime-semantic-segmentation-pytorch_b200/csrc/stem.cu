// stem.cu — K4'+K8: the first layer of every net on the path.
//
// Replaces InitialBlock (reference models/enet.py:38-48; ERFNet's
// DownsamplerBlock alias, models/erfnet.py:12): cat([ConvBNAct(cin, cout-cin,
// 3, stride 2)(x), MaxPool2d(3, 2, 1)(x)], dim=1) — conv, BN, act, pool and
// cat are five launches in the reference and the 1024x2048 input is read
// twice.  Here one thread produces one output pixel: it reads the 3x3xcin
// input window once (straight from the caller's NCHW fp32 tensor, so no layout
// conversion pass exists), computes every conv channel from weights staged in
// shared memory (broadcast reads), takes the window max for the pool branch and
// writes cconv+cin contiguous NHWC channels.  With pool == 0 it is the plain
// stride-2 stem ConvBNAct(3, C, 3, 2) of Fast-SCNN / DABNet.
//
// Also here: the NHWC MaxPool2d(3,2,1) branch for the deeper InitialBlocks.
#include <cfloat>

#include "common.cuh"

namespace rtseg {

// (StemParams lives in common.cuh: stem_tc.cu shares it)
// Tensor-core variant (stem_tc.cu): returns RTSEG_OK when it launched, 1 when the shape is not its business.
int stem_tc_try(const StemParams &p, int x_dtype, const void *x, const float *w, const Epilogue &ep, void *y,
                cudaStream_t st);

template <typename TI, typename TO, int CIN, int CC>
__global__ void __launch_bounds__(256) stem_kernel(StemParams p, const TI *__restrict__ x,
                                                   const float *__restrict__ wgt, Epilogue ep,
                                                   TO *__restrict__ y) {
  // weights [3][3][CIN][cconv] padded to CC output channels
  __shared__ __align__(16) float ws[9 * CIN * CC];
  for (int i = threadIdx.x; i < 9 * CIN * CC; i += blockDim.x) {
    const int co = i % CC, t = i / CC;
    ws[i] = co < p.cconv ? __ldg(wgt + (int64_t)t * p.cconv + co) : 0.f;
  }
  __syncthreads();

  const int64_t total = (int64_t)p.n * p.ho * p.wo;
  const int64_t pix = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= total) return;
  const int wo = (int)(pix % p.wo);
  const int64_t r = pix / p.wo;
  const int ho = (int)(r % p.ho);
  const int n = (int)(r / p.ho);

  // channel pairs: sm_100's packed FFMA2 halves the FMA issue count of this FMA-bound kernel
  float2 acc2[CC / 2];
#pragma unroll
  for (int c = 0; c < CC / 2; ++c) acc2[c] = make_float2(0.f, 0.f);
  float mx[CIN];
#pragma unroll
  for (int c = 0; c < CIN; ++c) mx[c] = -FLT_MAX;

  const TI *xn = x + (int64_t)n * p.sn;
  // All 9*CIN window loads are issued up front (independent requests in flight), then consumed.
  float win[3][3][CIN];
  bool inb[3][3];
#pragma unroll
  for (int ky = 0; ky < 3; ++ky) {
    const int hi = ho * 2 - 1 + ky;
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      const int wi = wo * 2 - 1 + kx;
      inb[ky][kx] = hi >= 0 && hi < p.h && wi >= 0 && wi < p.w;
#pragma unroll
      for (int ci = 0; ci < CIN; ++ci)
        win[ky][kx][ci] = inb[ky][kx] ? ld_img<TI>(xn + ci * p.sc + hi * p.sh + wi * p.sw, p.nz, ci) : 0.f;
    }
  }
#pragma unroll
  for (int ky = 0; ky < 3; ++ky) {
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
#pragma unroll
      for (int ci = 0; ci < CIN; ++ci) {
        const float v = win[ky][kx][ci];
        if (inb[ky][kx]) mx[ci] = fmax_nan(mx[ci], v);
        const float *wp = &ws[((ky * 3 + kx) * CIN + ci) * CC];
        const float2 vv = make_float2(v, v);
#pragma unroll
        for (int c = 0; c < CC; c += 4) {
          const float4 w4 = *reinterpret_cast<const float4 *>(wp + c);
          acc2[c / 2] = __ffma2_rn(vv, make_float2(w4.x, w4.y), acc2[c / 2]);
          acc2[c / 2 + 1] = __ffma2_rn(vv, make_float2(w4.z, w4.w), acc2[c / 2 + 1]);
        }
      }
    }
  }

  float acc[CC];
#pragma unroll
  for (int c = 0; c < CC / 2; ++c) acc[2 * c] = acc2[c].x, acc[2 * c + 1] = acc2[c].y;
  TO *yp = y + pix * p.y_ct + p.y_co;
  const int ctot = p.cconv + (p.pool ? CIN : 0);
  // Assemble the output channel vector in registers, then store 16-byte chunks when possible.
  constexpr int CT = CC + 4;  // CC >= cconv, CIN <= 4
  float out[CT];
#pragma unroll
  for (int c = 0; c < CT; ++c) out[c] = 0.f;
  if (ep.mode == 0) {  // none / relu / relu6 / hardtanh: branch-free clamp
#pragma unroll
    for (int c = 0; c < CC; ++c)
      if (c < p.cconv)
        out[c] = fmin_nan(fmax_nan(fmaf(acc[c], ep.scale ? __ldg(ep.scale + c) : 1.f, ep.shift ? __ldg(ep.shift + c) : 0.f), ep.lo), ep.hi);
  } else if (ep.mode == 1) {  // prelu with a slope vector
#pragma unroll
    for (int c = 0; c < CC; ++c)
      if (c < p.cconv) {
        const float v = fmaf(acc[c], ep.scale ? __ldg(ep.scale + c) : 1.f, ep.shift ? __ldg(ep.shift + c) : 0.f);
        out[c] = fmaf(__ldg(ep.slope + c), fmin_nan(v, 0.f), fmax_nan(v, 0.f));
      }
  } else {
#pragma unroll
    for (int c = 0; c < CC; ++c)
      if (c < p.cconv) out[c] = epilogue_apply(acc[c], 0.f, c, ep);
  }
  if (p.pool) {  // static register indexing: place mx[] right behind the conv channels
#pragma unroll
    for (int c = 0; c < CT; ++c)
#pragma unroll
      for (int ci = 0; ci < CIN; ++ci)
        if (c == p.cconv + ci) out[c] = mx[ci];
  }
  constexpr int V = Vec16<TO>::N;
  const bool vec = (ctot % V == 0) && (p.y_ct % V == 0) && (p.y_co % V == 0) &&
                   (reinterpret_cast<uintptr_t>(y) % 16 == 0);
  if (vec) {
#pragma unroll
    for (int c = 0; c + V <= CT; c += V) {
      if (c < ctot) {
        float v[V];
#pragma unroll
        for (int e = 0; e < V; ++e) v[e] = out[c + e];
        Vec16<TO>::store(yp + c, v);
      }
    }
  } else {
#pragma unroll
    for (int c = 0; c < CT; ++c)
      if (c < ctot) st_f<TO>(yp + c, out[c]);
  }
}

struct PoolParams {
  int n, h, w, c, ho, wo;
  int x_ct, x_co, y_ct, y_co;
};

// RAGGED: 16-byte loads from an aligned input slice, element-wise stores (output slice not 16-byte aligned or a
// channel count that is not a multiple of V: DABNet's 35-channel pool written behind 29 conv channels).
template <typename T, int V, bool RAGGED = false>
__global__ void __launch_bounds__(256) maxpool3x3s2_kernel(PoolParams p, const T *__restrict__ x, T *__restrict__ y) {
  const int cv_n = RAGGED ? (p.c + V - 1) / V : p.c / V;
  const int64_t total = (int64_t)p.n * p.ho * p.wo * cv_n;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= total) return;
  const int c0 = (int)(tid % cv_n) * V;
  int64_t r = tid / cv_n;
  const int wo = (int)(r % p.wo);
  r /= p.wo;
  const int ho = (int)(r % p.ho);
  const int n = (int)(r / p.ho);
  float m[V];
#pragma unroll
  for (int e = 0; e < V; ++e) m[e] = -FLT_MAX;
  if constexpr (V == 1) {
    for (int ky = 0; ky < 3; ++ky) {
      const int hi = ho * 2 - 1 + ky;
      if (hi < 0 || hi >= p.h) continue;
      for (int kx = 0; kx < 3; ++kx) {
        const int wi = wo * 2 - 1 + kx;
        if (wi < 0 || wi >= p.w) continue;
        m[0] = fmax_nan(m[0], ld_f<T>(x + (((int64_t)n * p.h + hi) * p.w + wi) * p.x_ct + p.x_co + c0));
      }
    }
  } else {
    // nine independent 16-byte loads in flight, then the reduction (out-of-image taps replicate the centre)
    typename Vec16<T>::Raw raw[9];
    const T *centre = x + (((int64_t)n * p.h + ho * 2) * p.w + wo * 2) * p.x_ct + p.x_co + c0;
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      const int hi = ho * 2 - 1 + ky;
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const int wi = wo * 2 - 1 + kx;
        const bool ok = hi >= 0 && hi < p.h && wi >= 0 && wi < p.w;
        raw[ky * 3 + kx] = Vec16<T>::ldg(ok ? x + (((int64_t)n * p.h + hi) * p.w + wi) * p.x_ct + p.x_co + c0 : centre);
      }
    }
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      float v[V];
      Vec16<T>::unpack(raw[t], v);
#pragma unroll
      for (int e = 0; e < V; ++e) m[e] = fmax_nan(m[e], v[e]);
    }
  }
  T *yp = y + (((int64_t)n * p.ho + ho) * p.wo + wo) * p.y_ct + p.y_co + c0;
  if constexpr (V == 1) {
    st_f<T>(yp, m[0]);
  } else if constexpr (RAGGED) {
#pragma unroll
    for (int e = 0; e < V; ++e)
      if (c0 + e < p.c) st_f<T>(yp + e, m[e]);
  } else {
    Vec16<T>::store(yp, m);
  }
}

// ---- ENet's pooling pair -------------------------------------------------------------------------
// nn.MaxPool2d(2, 2, return_indices=True) and nn.MaxUnpool2d(2, 2) of the down- / up-sampling Bottlenecks (reference
// models/enet.py:130-131, 138-139, 168-178).  NHWC, channel slices in and out.  The index of a window's maximum is
// kept as its position inside the 2x2 window (0..3, row-major, one byte per element) — every window owns its four
// output pixels, so the unpool is a pure gather: no scatter, no memset of the (three quarters zero) output.
// Selection follows ATen: scan the window row-major, take a value if it is greater than the running maximum or NaN.
struct Pool2Params {
  int n, h, w, c, ho, wo;   // input h x w, pooled ho x wo = floor(h / 2) x floor(w / 2)
  int x_ct, x_co, y_ct, y_co;
};

template <typename T, int V>
__global__ void __launch_bounds__(256) maxpool2x2_idx_kernel(Pool2Params p, const T *__restrict__ x, T *__restrict__ y,
                                                             uint8_t *__restrict__ idx) {
  const int cv_n = p.c / V;
  const int64_t total = (int64_t)p.n * p.ho * p.wo * cv_n;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= total) return;
  const int c0 = (int)(tid % cv_n) * V;
  int64_t r = tid / cv_n;
  const int wo = (int)(r % p.wo);
  r /= p.wo;
  const int ho = (int)(r % p.ho), n = (int)(r / p.ho);
  float best[V];
  int pos[V];
  const T *base = x + (((int64_t)n * p.h + ho * 2) * p.w + wo * 2) * p.x_ct + p.x_co + c0;
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const T *q = base + ((int64_t)(t >> 1) * p.w + (t & 1)) * p.x_ct;
    float v[V];
    if constexpr (V == 1) v[0] = ld_f<T>(q);
    else Vec16<T>::unpack(Vec16<T>::ldg(q), v);
#pragma unroll
    for (int e = 0; e < V; ++e)
      if (t == 0 || v[e] > best[e] || v[e] != v[e]) best[e] = v[e], pos[e] = t;
  }
  const int64_t opix = ((int64_t)n * p.ho + ho) * p.wo + wo;
  T *yp = y + opix * p.y_ct + p.y_co + c0;
  if constexpr (V == 1) st_f<T>(yp, best[0]);
  else Vec16<T>::store(yp, best);
#pragma unroll
  for (int e = 0; e < V; ++e) idx[opix * p.c + c0 + e] = (uint8_t)pos[e];
}

// y[n][oy][ox][c] = x[n][oy/2][ox/2][c] if idx says so, else 0; output rows / columns beyond 2*hp (odd sizes) are zero
template <typename T, int V>
__global__ void __launch_bounds__(256) maxunpool2x2_kernel(Pool2Params p, const T *__restrict__ x,
                                                           const uint8_t *__restrict__ idx, T *__restrict__ y) {
  const int cv_n = p.c / V;
  const int64_t total = (int64_t)p.n * p.h * p.w * cv_n;   // here h x w is the OUTPUT (unpooled) size, ho x wo the pooled input
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= total) return;
  const int c0 = (int)(tid % cv_n) * V;
  int64_t r = tid / cv_n;
  const int ox = (int)(r % p.w);
  r /= p.w;
  const int oy = (int)(r % p.h), n = (int)(r / p.h);
  float out[V];
#pragma unroll
  for (int e = 0; e < V; ++e) out[e] = 0.f;
  const int hp = oy >> 1, wp = ox >> 1;
  if (hp < p.ho && wp < p.wo) {
    const int64_t ipix = ((int64_t)n * p.ho + hp) * p.wo + wp;
    const int me = (oy & 1) * 2 + (ox & 1);
    float v[V];
    if constexpr (V == 1) v[0] = ld_f<T>(x + ipix * p.x_ct + p.x_co + c0);
    else Vec16<T>::unpack(Vec16<T>::ldg(x + ipix * p.x_ct + p.x_co + c0), v);
#pragma unroll
    for (int e = 0; e < V; ++e)
      if ((int)idx[ipix * p.c + c0 + e] == me) out[e] = v[e];
  }
  T *yp = y + (((int64_t)n * p.h + oy) * p.w + ox) * p.y_ct + p.y_co + c0;
  if constexpr (V == 1) st_f<T>(yp, out[0]);
  else Vec16<T>::store(yp, out);
}

template <typename TI, typename TO>
static int launch_stem(const StemParams &p, const void *x, const float *w, const Epilogue &ep, void *y,
                       cudaStream_t st) {
  const int64_t total = (int64_t)p.n * p.ho * p.wo;
  const unsigned grid = (unsigned)cdiv(total, 256);
  const TI *xi = static_cast<const TI *>(x);
  TO *yo = static_cast<TO *>(y);
#define RTSEG_STEM(CIN_, CC_)                                                    \
  if (p.cin == CIN_ && p.cconv <= CC_) {                                         \
    stem_kernel<TI, TO, CIN_, CC_><<<grid, 256, 0, st>>>(p, xi, w, ep, yo);      \
    RTSEG_LAUNCH_CHECK("stem_kernel");                                           \
    return RTSEG_OK;                                                             \
  }
  RTSEG_STEM(3, 16)
  RTSEG_STEM(3, 32)
  RTSEG_STEM(1, 16)
  RTSEG_STEM(1, 32)
  RTSEG_STEM(4, 16)
  RTSEG_STEM(4, 32)
#undef RTSEG_STEM
  RTSEG_FAIL(RTSEG_ERR_UNSUPPORTED, "rtseg_stem: cin=%d cconv=%d has no kernel (cin in {1,3,4}, cconv <= 32)", p.cin,
             p.cconv);
}

}  // namespace rtseg

static int stem_impl(const rtseg_conv_desc *d, int32_t x_is_nchw, int32_t x_dtype, int32_t pool, const void *x,
                     const rtseg::ImgNorm &nz, const float *w, const float *scale, const float *shift, const float *slope,
                     void *y, rtseg_stream_t stream) {
  using namespace rtseg;
  RTSEG_REQUIRE(d && x && w && y, "NULL argument");
  RTSEG_REQUIRE(d->kh == 3 && d->kw == 3 && d->stride == 2 && d->dil == 1 && d->pad_h == 1 && d->pad_w == 1 &&
                    d->groups == 1,
                "rtseg_stem is the 3x3 stride-2 pad-1 stem only");
  RTSEG_REQUIRE(d->res_mode == RTSEG_RES_NONE, "stem has no residual");
  RTSEG_REQUIRE(d->n > 0 && d->h > 0 && d->w > 0 && d->cin > 0 && d->cout > 0, "non-positive extent");
  StemParams p;
  p.n = d->n, p.h = d->h, p.w = d->w, p.cin = d->cin, p.cconv = d->cout;
  p.ho = (d->h - 1) / 2 + 1, p.wo = (d->w - 1) / 2 + 1;
  p.pool = pool ? 1 : 0;
  p.nz = nz;
  p.y_ct = d->y_ctotal, p.y_co = d->y_coff;
  RTSEG_REQUIRE(p.y_ct >= p.y_co + p.cconv + (p.pool ? p.cin : 0), "output channel slice too small");
  if (x_is_nchw) {
    p.sn = (int64_t)d->cin * d->h * d->w, p.sc = (int64_t)d->h * d->w, p.sh = d->w, p.sw = 1;
  } else {
    p.sn = (int64_t)d->h * d->w * d->x_ctotal, p.sc = 1, p.sh = (int64_t)d->w * d->x_ctotal, p.sw = d->x_ctotal;
    x = static_cast<const char *>(x) + (size_t)d->x_coff * (x_dtype == RTSEG_BF16 ? 2 : (x_dtype == rtseg::RTSEG_U8 ? 1 : 4));
  }
  const Epilogue ep = make_epilogue(d, scale, shift, slope);
  cudaStream_t st = as_stream(stream);
  if (d->dtype == RTSEG_BF16) {  // bf16 activations: im2col + tcgen05 (bf16 operands, fp32 accumulation)
    const int rc = stem_tc_try(p, x_dtype, x, w, ep, y, st);
    if (rc <= 0) return rc;
  }
  if (x_dtype == RTSEG_F32 && d->dtype == RTSEG_F32) return launch_stem<float, float>(p, x, w, ep, y, st);
  if (x_dtype == RTSEG_F32 && d->dtype == RTSEG_BF16) return launch_stem<float, __nv_bfloat16>(p, x, w, ep, y, st);
  if (x_dtype == RTSEG_BF16 && d->dtype == RTSEG_BF16)
    return launch_stem<__nv_bfloat16, __nv_bfloat16>(p, x, w, ep, y, st);
  if (x_dtype == RTSEG_BF16 && d->dtype == RTSEG_F32) return launch_stem<__nv_bfloat16, float>(p, x, w, ep, y, st);
  if (x_dtype == RTSEG_U8 && d->dtype == RTSEG_F32) return launch_stem<uint8_t, float>(p, x, w, ep, y, st);
  if (x_dtype == RTSEG_U8 && d->dtype == RTSEG_BF16) return launch_stem<uint8_t, __nv_bfloat16>(p, x, w, ep, y, st);
  RTSEG_FAIL(RTSEG_ERR_INVALID, "bad dtype combination");
}

extern "C" int rtseg_stem(const rtseg_conv_desc *d, int32_t x_is_nchw, int32_t x_dtype, int32_t pool, const void *x,
                          const float *w, const float *scale, const float *shift, const float *slope, void *y,
                          rtseg_stream_t stream) {
  RTSEG_REQUIRE(x_dtype == RTSEG_F32 || x_dtype == RTSEG_BF16, "rtseg_stem takes fp32 or bf16 images (uint8: rtseg_stem_u8)");
  const rtseg::ImgNorm nz{};
  return stem_impl(d, x_is_nchw, x_dtype, pool, x, nz, w, scale, shift, slope, y, stream);
}

extern "C" int rtseg_stem_u8(const rtseg_conv_desc *d, int32_t pool, const uint8_t *x, const float *mean_host,
                             const float *std_host, const float *w, const float *scale, const float *shift,
                             const float *slope, void *y, rtseg_stream_t stream) {
  RTSEG_REQUIRE(d && mean_host && std_host, "NULL argument");
  RTSEG_REQUIRE(d->cin >= 1 && d->cin <= 4, "uint8 images have 1 to 4 channels");
  rtseg::ImgNorm nz{};
  RTSEG_REQUIRE(rtseg::make_img_norm(d->cin, mean_host, std_host, nz) == 0, "std must be positive");
  return stem_impl(d, 0, rtseg::RTSEG_U8, pool, x, nz, w, scale, shift, slope, y, stream);
}

extern "C" int rtseg_maxpool3x3s2(const rtseg_conv_desc *d, const void *x, void *y, rtseg_stream_t stream) {
  using namespace rtseg;
  RTSEG_REQUIRE(d && x && y, "NULL argument");
  RTSEG_REQUIRE(d->n > 0 && d->h > 0 && d->w > 0 && d->cin > 0, "non-positive extent");
  PoolParams p;
  p.n = d->n, p.h = d->h, p.w = d->w, p.c = d->cin;
  p.ho = (d->h - 1) / 2 + 1, p.wo = (d->w - 1) / 2 + 1;
  p.x_ct = d->x_ctotal, p.x_co = d->x_coff, p.y_ct = d->y_ctotal, p.y_co = d->y_coff;
  RTSEG_REQUIRE(p.x_ct >= p.x_co + p.c && p.y_ct >= p.y_co + p.c, "channel slice out of range");
  cudaStream_t st = as_stream(stream);
  const int V = d->dtype == RTSEG_BF16 ? 8 : 4;
  const bool vec = p.c % V == 0 && p.x_ct % V == 0 && p.x_co % V == 0 && p.y_ct % V == 0 && p.y_co % V == 0 &&
                   reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(y) % 16 == 0;
  const bool ragged = !vec && p.x_ct % V == 0 && p.x_co % V == 0 && reinterpret_cast<uintptr_t>(x) % 16 == 0 &&
                      p.x_ct >= p.x_co + (p.c + V - 1) / V * V;
  const int64_t total = (int64_t)p.n * p.ho * p.wo * (vec ? p.c / V : ragged ? (p.c + V - 1) / V : p.c);
  const unsigned grid = (unsigned)cdiv(total, 256);
  if (ragged) {
    if (d->dtype == RTSEG_BF16)
      maxpool3x3s2_kernel<__nv_bfloat16, 8, true><<<grid, 256, 0, st>>>(p, (const __nv_bfloat16 *)x, (__nv_bfloat16 *)y);
    else if (d->dtype == RTSEG_F32)
      maxpool3x3s2_kernel<float, 4, true><<<grid, 256, 0, st>>>(p, (const float *)x, (float *)y);
    else
      RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
    RTSEG_LAUNCH_CHECK("maxpool3x3s2_kernel");
    return RTSEG_OK;
  }
  if (d->dtype == RTSEG_BF16) {
    if (vec)
      maxpool3x3s2_kernel<__nv_bfloat16, 8><<<grid, 256, 0, st>>>(p, (const __nv_bfloat16 *)x, (__nv_bfloat16 *)y);
    else
      maxpool3x3s2_kernel<__nv_bfloat16, 1><<<grid, 256, 0, st>>>(p, (const __nv_bfloat16 *)x, (__nv_bfloat16 *)y);
  } else if (d->dtype == RTSEG_F32) {
    if (vec)
      maxpool3x3s2_kernel<float, 4><<<grid, 256, 0, st>>>(p, (const float *)x, (float *)y);
    else
      maxpool3x3s2_kernel<float, 1><<<grid, 256, 0, st>>>(p, (const float *)x, (float *)y);
  } else {
    RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
  }
  RTSEG_LAUNCH_CHECK("maxpool3x3s2_kernel");
  return RTSEG_OK;
}


static bool pool2_vec_ok(const rtseg::Pool2Params &p, int V, const void *x, const void *y) {
  return p.c % V == 0 && p.x_ct % V == 0 && p.x_co % V == 0 && p.y_ct % V == 0 && p.y_co % V == 0 &&
         reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(y) % 16 == 0;
}

extern "C" int rtseg_maxpool2x2_idx(int32_t n, int32_t h, int32_t w, int32_t c, int32_t dtype, const void *x, int32_t x_ctotal,
                                    int32_t x_coff, void *y, int32_t y_ctotal, int32_t y_coff, uint8_t *idx,
                                    rtseg_stream_t stream) {
  using namespace rtseg;
  RTSEG_REQUIRE(x && y && idx, "NULL argument");
  RTSEG_REQUIRE(n > 0 && h >= 2 && w >= 2 && c > 0, "extent too small for a 2x2 window");
  RTSEG_REQUIRE(x_coff >= 0 && y_coff >= 0 && x_ctotal >= x_coff + c && y_ctotal >= y_coff + c, "channel slice out of range");
  Pool2Params p{n, h, w, c, h / 2, w / 2, x_ctotal, x_coff, y_ctotal, y_coff};
  cudaStream_t st = as_stream(stream);
  const int V = dtype == RTSEG_BF16 ? 8 : 4;
  const bool vec = pool2_vec_ok(p, V, x, y);
  const int64_t total = (int64_t)n * p.ho * p.wo * (vec ? c / V : c);
  const unsigned grid = (unsigned)cdiv(total, 256);
  if (dtype == RTSEG_BF16) {
    if (vec) maxpool2x2_idx_kernel<__nv_bfloat16, 8><<<grid, 256, 0, st>>>(p, (const __nv_bfloat16 *)x, (__nv_bfloat16 *)y, idx);
    else maxpool2x2_idx_kernel<__nv_bfloat16, 1><<<grid, 256, 0, st>>>(p, (const __nv_bfloat16 *)x, (__nv_bfloat16 *)y, idx);
  } else if (dtype == RTSEG_F32) {
    if (vec) maxpool2x2_idx_kernel<float, 4><<<grid, 256, 0, st>>>(p, (const float *)x, (float *)y, idx);
    else maxpool2x2_idx_kernel<float, 1><<<grid, 256, 0, st>>>(p, (const float *)x, (float *)y, idx);
  } else {
    RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
  }
  RTSEG_LAUNCH_CHECK("maxpool2x2_idx_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_maxunpool2x2(int32_t n, int32_t hp, int32_t wp, int32_t c, int32_t ho, int32_t wo, int32_t dtype,
                                  const void *x, int32_t x_ctotal, int32_t x_coff, const uint8_t *idx, void *y,
                                  int32_t y_ctotal, int32_t y_coff, rtseg_stream_t stream) {
  using namespace rtseg;
  RTSEG_REQUIRE(x && y && idx, "NULL argument");
  RTSEG_REQUIRE(n > 0 && hp > 0 && wp > 0 && c > 0 && ho >= 2 * hp && wo >= 2 * wp && ho <= 2 * hp + 1 && wo <= 2 * wp + 1,
                "output size must be 2*h (+1) x 2*w (+1)");
  RTSEG_REQUIRE(x_coff >= 0 && y_coff >= 0 && x_ctotal >= x_coff + c && y_ctotal >= y_coff + c, "channel slice out of range");
  Pool2Params p{n, ho, wo, c, hp, wp, x_ctotal, x_coff, y_ctotal, y_coff};
  cudaStream_t st = as_stream(stream);
  const int V = dtype == RTSEG_BF16 ? 8 : 4;
  const bool vec = pool2_vec_ok(p, V, x, y);
  const int64_t total = (int64_t)n * ho * wo * (vec ? c / V : c);
  const unsigned grid = (unsigned)cdiv(total, 256);
  if (dtype == RTSEG_BF16) {
    if (vec) maxunpool2x2_kernel<__nv_bfloat16, 8><<<grid, 256, 0, st>>>(p, (const __nv_bfloat16 *)x, idx, (__nv_bfloat16 *)y);
    else maxunpool2x2_kernel<__nv_bfloat16, 1><<<grid, 256, 0, st>>>(p, (const __nv_bfloat16 *)x, idx, (__nv_bfloat16 *)y);
  } else if (dtype == RTSEG_F32) {
    if (vec) maxunpool2x2_kernel<float, 4><<<grid, 256, 0, st>>>(p, (const float *)x, idx, (float *)y);
    else maxunpool2x2_kernel<float, 1><<<grid, 256, 0, st>>>(p, (const float *)x, idx, (float *)y);
  } else {
    RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
  }
  RTSEG_LAUNCH_CHECK("maxunpool2x2_kernel");
  return RTSEG_OK;
}
