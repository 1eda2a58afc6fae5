// conv_dense.cu — CUDA-core implicit GEMM for dense convolutions and the 3x3
// stride-2 transposed convolution, + folded BatchNorm + activation + residual.
//
// This is the fp32 parity path (BASELINE: fp32 logits within 1e-4 of the
// reference) and the catch-all for shapes the tensor-core kernel refuses
// (odd channel counts such as 3/13/19/29/35, stride 2, 5x1 kernels ...).
// Replaces ConvBNAct / PWConvBNAct / conv1x1 / conv3x3 (reference
// models/modules.py:7-15,63-85) and DeConvBNAct (:89-108).
//
// One kernel serves both directions through a *tap table*: output-grid pixel
// (i, j) reads input (i*istride + dy[t], j*istride + dx[t]) for tap t and
// writes output (i*ostride + oy0, j*ostride + ox0).  A forward conv has
// istride = stride, ostride = 1 and kh*kw taps; a transposed conv is four
// launches ("sub-pixel phases") with istride = 1, ostride = 2 and 1/2/2/4 taps.
//
// Tiling: 64 pixels x 64 output channels per CTA, K-step 16, 256 threads, each
// thread a 4x4 register tile, fp32 accumulation.
#include "common.cuh"

namespace rtseg {

constexpr int kMaxTaps = 49;

struct TapTable {
  int ntaps;
  short dy[kMaxTaps], dx[kMaxTaps], widx[kMaxTaps];
};

struct DenseParams {
  int n, h, w;        // input
  int gh, gw;         // output grid enumerated by this launch
  int istride;        // input coord = grid coord * istride + tap offset
  int ostride, oy0, ox0;
  int oh, ow;         // full output size
  int cin, cout;
  int x_ct, x_co, y_ct, y_co, r_ct, r_co;
  int x_vec, y_vec, r_vec, w_vec;  // 4-element vector access allowed
};

constexpr int BM = 64, BN = 64, BK = 16;

template <typename T>
__device__ __forceinline__ void load4(const T *p, bool vec, int nvalid, float (&v)[4]);

template <>
__device__ __forceinline__ void load4<float>(const float *p, bool vec, int nvalid, float (&v)[4]) {
  if (vec && nvalid == 4) {
    const float4 t = __ldg(reinterpret_cast<const float4 *>(p));
    v[0] = t.x, v[1] = t.y, v[2] = t.z, v[3] = t.w;
  } else {
#pragma unroll
    for (int e = 0; e < 4; ++e) v[e] = e < nvalid ? __ldg(p + e) : 0.f;
  }
}
template <>
__device__ __forceinline__ void load4<__nv_bfloat16>(const __nv_bfloat16 *p, bool vec, int nvalid, float (&v)[4]) {
  if (vec && nvalid == 4) {
    const uint2 t = __ldg(reinterpret_cast<const uint2 *>(p));
    v[0] = __uint_as_float(t.x << 16), v[1] = __uint_as_float(t.x & 0xffff0000u);
    v[2] = __uint_as_float(t.y << 16), v[3] = __uint_as_float(t.y & 0xffff0000u);
  } else {
#pragma unroll
    for (int e = 0; e < 4; ++e) v[e] = e < nvalid ? __bfloat162float(p[e]) : 0.f;
  }
}

template <typename T>
__device__ __forceinline__ void store4(T *p, bool vec, int nvalid, const float (&v)[4]);
template <>
__device__ __forceinline__ void store4<float>(float *p, bool vec, int nvalid, const float (&v)[4]) {
  if (vec && nvalid == 4) {
    *reinterpret_cast<float4 *>(p) = make_float4(v[0], v[1], v[2], v[3]);
  } else {
#pragma unroll
    for (int e = 0; e < 4; ++e)
      if (e < nvalid) p[e] = v[e];
  }
}
template <>
__device__ __forceinline__ void store4<__nv_bfloat16>(__nv_bfloat16 *p, bool vec, int nvalid, const float (&v)[4]) {
  if (vec && nvalid == 4) {
    const __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
    uint2 t;
    t.x = *reinterpret_cast<const uint32_t *>(&a);
    t.y = *reinterpret_cast<const uint32_t *>(&b);
    *reinterpret_cast<uint2 *>(p) = t;
  } else {
#pragma unroll
    for (int e = 0; e < 4; ++e)
      if (e < nvalid) p[e] = __float2bfloat16_rn(v[e]);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) dense_kernel(DenseParams p, TapTable taps, const T *__restrict__ x,
                                                    const float *__restrict__ wgt, Epilogue ep,
                                                    const T *__restrict__ res, T *__restrict__ y) {
  __shared__ float As[BM][BK + 1];
  __shared__ __align__(16) float Bs[BK][BN];

  const int tid = threadIdx.x;
  const int tx = tid % 16, ty = tid / 16;
  const int64_t M = (int64_t)p.n * p.gh * p.gw;
  const int64_t m0 = (int64_t)blockIdx.x * BM;
  const int co0 = blockIdx.y * BN;

  // Loader roles.
  const int a_pix = tid / 4, a_cg = (tid % 4) * 4;  // A: 64 px x 16 ci
  const int b_k = tid / 16, b_cg = (tid % 16) * 4;  // B: 16 k x 64 co
  const int64_t am = m0 + a_pix;
  const bool a_valid = am < M;
  int a_n = 0, a_i = 0, a_j = 0;
  if (a_valid) {
    a_j = (int)(am % p.gw);
    const int64_t r = am / p.gw;
    a_i = (int)(r % p.gh);
    a_n = (int)(r / p.gh);
  }
  const T *xn = x + (int64_t)a_n * p.h * p.w * p.x_ct + p.x_co;

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int t = 0; t < taps.ntaps; ++t) {
    const int hi = a_i * p.istride + taps.dy[t];
    const int wi = a_j * p.istride + taps.dx[t];
    const bool in_ok = a_valid && hi >= 0 && hi < p.h && wi >= 0 && wi < p.w;
    const T *xp = xn + ((int64_t)hi * p.w + wi) * p.x_ct;
    const float *wt = wgt + (int64_t)taps.widx[t] * p.cin * p.cout;
    for (int ci0 = 0; ci0 < p.cin; ci0 += BK) {
      {  // A tile
        float v[4] = {0.f, 0.f, 0.f, 0.f};
        const int ci = ci0 + a_cg;
        const int nv = min(4, p.cin - ci);
        if (in_ok && nv > 0) load4<T>(xp + ci, p.x_vec, nv, v);
#pragma unroll
        for (int e = 0; e < 4; ++e) As[a_pix][a_cg + e] = v[e];
      }
      {  // B tile
        float v[4] = {0.f, 0.f, 0.f, 0.f};
        const int ci = ci0 + b_k;
        const int co = co0 + b_cg;
        const int nv = min(4, p.cout - co);
        if (ci < p.cin && nv > 0) load4<float>(wt + (int64_t)ci * p.cout + co, p.w_vec, nv, v);
        *reinterpret_cast<float4 *>(&Bs[b_k][b_cg]) = make_float4(v[0], v[1], v[2], v[3]);
      }
      __syncthreads();
#pragma unroll
      for (int k = 0; k < BK; ++k) {
        float a[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = As[ty * 4 + i][k];
        const float4 bv = *reinterpret_cast<const float4 *>(&Bs[k][tx * 4]);
        const float b[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
      }
      __syncthreads();
    }
  }

  const int co = co0 + tx * 4;
  const int nv = min(4, p.cout - co);
  if (nv <= 0) return;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t m = m0 + ty * 4 + i;
    if (m >= M) break;
    const int j = (int)(m % p.gw);
    const int64_t r = m / p.gw;
    const int gi = (int)(r % p.gh);
    const int n = (int)(r / p.gh);
    const int64_t opix = ((int64_t)n * p.oh + (gi * p.ostride + p.oy0)) * p.ow + (j * p.ostride + p.ox0);
    float rv[4] = {0.f, 0.f, 0.f, 0.f};
    if (ep.res_mode != RTSEG_RES_NONE) load4<T>(res + opix * p.r_ct + p.r_co + co, p.r_vec, nv, rv);
    float ov[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) ov[e] = e < nv ? epilogue_apply(acc[i][e], rv[e], co + e, ep) : 0.f;
    store4<T>(y + opix * p.y_ct + p.y_co + co, p.y_vec, nv, ov);
  }
}

template <typename T>
static int launch_dense(const DenseParams &p, const TapTable &taps, const void *x, const float *w,
                        const Epilogue &ep, const void *res, void *y, cudaStream_t st) {
  const int64_t M = (int64_t)p.n * p.gh * p.gw;
  dim3 grid((unsigned)cdiv(M, BM), (unsigned)cdiv(p.cout, BN));
  dense_kernel<T><<<grid, 256, 0, st>>>(p, taps, static_cast<const T *>(x), w, ep, static_cast<const T *>(res),
                                        static_cast<T *>(y));
  RTSEG_LAUNCH_CHECK("dense_kernel");
  return RTSEG_OK;
}

static void fill_vec_flags(DenseParams &p, const rtseg_conv_desc *d, const void *x, const float *w,
                           const void *res, const void *y) {
  const size_t es = d->dtype == RTSEG_BF16 ? 2 : 4;
  const size_t av = 4 * es;  // bytes of a 4-element vector
  p.x_vec = (d->x_ctotal % 4 == 0) && (d->x_coff % 4 == 0) && (reinterpret_cast<uintptr_t>(x) % av == 0);
  p.y_vec = (d->y_ctotal % 4 == 0) && (d->y_coff % 4 == 0) && (reinterpret_cast<uintptr_t>(y) % av == 0);
  p.r_vec = res && (d->r_ctotal % 4 == 0) && (d->r_coff % 4 == 0) && (reinterpret_cast<uintptr_t>(res) % av == 0);
  p.w_vec = (d->cout % 4 == 0) && (reinterpret_cast<uintptr_t>(w) % 16 == 0);
}

}  // namespace rtseg

extern "C" int rtseg_conv_dense(const rtseg_conv_desc *d, const void *x, const float *w, const float *scale,
                                const float *shift, const float *slope, const void *res, void *y,
                                rtseg_stream_t stream) {
  using namespace rtseg;
  if (int rc = validate_conv_desc(d, res)) return rc;
  RTSEG_REQUIRE(x && w && y, "NULL tensor pointer");
  RTSEG_REQUIRE(d->groups == 1, "rtseg_conv_dense needs groups == 1");
  RTSEG_REQUIRE(d->kh * d->kw <= kMaxTaps, "too many taps");
  DenseParams p;
  p.n = d->n, p.h = d->h, p.w = d->w;
  rtseg_out_hw(d, &p.gh, &p.gw);
  p.oh = p.gh, p.ow = p.gw;
  p.istride = d->stride, p.ostride = 1, p.oy0 = 0, p.ox0 = 0;
  p.cin = d->cin, p.cout = d->cout;
  p.x_ct = d->x_ctotal, p.x_co = d->x_coff, p.y_ct = d->y_ctotal, p.y_co = d->y_coff;
  p.r_ct = d->r_ctotal, p.r_co = d->r_coff;
  fill_vec_flags(p, d, x, w, res, y);
  TapTable taps;
  taps.ntaps = 0;
  for (int ky = 0; ky < d->kh; ++ky)
    for (int kx = 0; kx < d->kw; ++kx) {
      taps.dy[taps.ntaps] = (short)(ky * d->dil - d->pad_h);
      taps.dx[taps.ntaps] = (short)(kx * d->dil - d->pad_w);
      taps.widx[taps.ntaps] = (short)(ky * d->kw + kx);
      ++taps.ntaps;
    }
  const Epilogue ep = make_epilogue(d, scale, shift, slope);
  if (d->dtype == RTSEG_BF16) return launch_dense<__nv_bfloat16>(p, taps, x, w, ep, res, y, as_stream(stream));
  return launch_dense<float>(p, taps, x, w, ep, res, y, as_stream(stream));
}

// Transposed conv k=3, s=2, p=1, output_padding=1 (DeConvBNAct with scale_factor 2,
// reference models/modules.py:92-102): oy = 2*iy - 1 + ky.  Output row 2i takes
// ky = 1 from input row i; output row 2i+1 takes ky = 0 from row i+1 and ky = 2
// from row i.  Same along x.  Four launches, one per output parity.
extern "C" int rtseg_deconv3x3s2(const rtseg_conv_desc *d, const void *x, const float *w, const float *scale,
                                 const float *shift, const float *slope, void *y, uint8_t *labels,
                                 rtseg_stream_t stream) {
  using namespace rtseg;
  RTSEG_REQUIRE(d != nullptr, "conv descriptor is NULL");
  RTSEG_REQUIRE(d->kh == 3 && d->kw == 3 && d->stride == 2 && d->groups == 1 && d->dil == 1,
                "rtseg_deconv3x3s2 handles k=3, s=2, dense only");
  RTSEG_REQUIRE(d->res_mode == RTSEG_RES_NONE, "deconv has no residual input");
  RTSEG_REQUIRE(x && w && y, "NULL tensor pointer (the CUDA-core deconv needs y even when labels are requested)");
  RTSEG_REQUIRE(d->n > 0 && d->h > 0 && d->w > 0 && d->cin > 0 && d->cout > 0, "non-positive extent");
  RTSEG_REQUIRE(d->dtype == RTSEG_F32 || d->dtype == RTSEG_BF16, "unknown dtype");
  RTSEG_REQUIRE(d->x_ctotal >= d->x_coff + d->cin && d->y_ctotal >= d->y_coff + d->cout, "channel slice out of range");
  DenseParams p;
  p.n = d->n, p.h = d->h, p.w = d->w;
  p.gh = d->h, p.gw = d->w;
  p.oh = 2 * d->h, p.ow = 2 * d->w;
  p.istride = 1, p.ostride = 2;
  p.cin = d->cin, p.cout = d->cout;
  p.x_ct = d->x_ctotal, p.x_co = d->x_coff, p.y_ct = d->y_ctotal, p.y_co = d->y_coff;
  p.r_ct = 0, p.r_co = 0;
  fill_vec_flags(p, d, x, w, nullptr, y);
  const Epilogue ep = make_epilogue(d, scale, shift, slope);
  for (int py = 0; py < 2; ++py)
    for (int px = 0; px < 2; ++px) {
      TapTable taps;
      taps.ntaps = 0;
      // (ky, dy) pairs valid for this parity
      const int kys[2] = {py == 0 ? 1 : 0, 2}, dys[2] = {py == 0 ? 0 : 1, 0};
      const int kxs[2] = {px == 0 ? 1 : 0, 2}, dxs[2] = {px == 0 ? 0 : 1, 0};
      const int ny = py == 0 ? 1 : 2, nx = px == 0 ? 1 : 2;
      for (int a = 0; a < ny; ++a)
        for (int b = 0; b < nx; ++b) {
          taps.dy[taps.ntaps] = (short)dys[a];
          taps.dx[taps.ntaps] = (short)dxs[b];
          taps.widx[taps.ntaps] = (short)(kys[a] * 3 + kxs[b]);
          ++taps.ntaps;
        }
      p.oy0 = py, p.ox0 = px;
      int rc = d->dtype == RTSEG_BF16
                   ? launch_dense<__nv_bfloat16>(p, taps, x, w, ep, nullptr, y, as_stream(stream))
                   : launch_dense<float>(p, taps, x, w, ep, nullptr, y, as_stream(stream));
      if (rc) return rc;
    }
  if (labels) {
    return rtseg_argmax_nhwc((int64_t)d->n * p.oh * p.ow, d->cout, d->y_ctotal, d->dtype,
                             static_cast<const char *>(y) + (size_t)d->y_coff * (d->dtype == RTSEG_BF16 ? 2 : 4),
                             labels, nullptr, stream);
  }
  return RTSEG_OK;
}
