// common.cuh — shared host/device helpers for librtseg (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/rtseg.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "librtseg is written for sm_100a (B200) only"
#endif

namespace rtseg {

// ---- error plumbing -------------------------------------------------------
void set_error(const char *fmt, ...);
void count_launch(int n = 1);

#define RTSEG_FAIL(code, ...)      \
  do {                             \
    rtseg::set_error(__VA_ARGS__); \
    return (code);                 \
  } while (0)

#define RTSEG_REQUIRE(cond, ...)                            \
  do {                                                      \
    if (!(cond)) RTSEG_FAIL(RTSEG_ERR_INVALID, __VA_ARGS__); \
  } while (0)

#define RTSEG_CUDA(call)                                                                \
  do {                                                                                  \
    cudaError_t e__ = (call);                                                           \
    if (e__ != cudaSuccess)                                                             \
      RTSEG_FAIL(RTSEG_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), \
                 __FILE__, __LINE__);                                                   \
  } while (0)

// Call right after a <<<>>> launch.
#define RTSEG_LAUNCH_CHECK(name)                                                       \
  do {                                                                                 \
    cudaError_t e__ = cudaGetLastError();                                              \
    if (e__ != cudaSuccess)                                                            \
      RTSEG_FAIL(RTSEG_ERR_CUDA, "launch of %s failed: %s", name, cudaGetErrorString(e__)); \
    rtseg::count_launch();                                                             \
  } while (0)

static inline cudaStream_t as_stream(rtseg_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

static inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }

int num_sms();

// ---- epilogue parameters passed by value to kernels -------------------------
struct Epilogue {
  const float *scale;  // [cout] or nullptr (=1)
  const float *shift;  // [cout] or nullptr (=0)
  const float *slope;  // [cout] PReLU slopes or nullptr
  int act, act2, res_mode;
  float p0, p1;
  // Fast forms chosen on the host (epilogue_mode): 0 = clamp(v, lo, hi) covers none / relu / relu6 /
  // hardtanh; 1 = max(v,0) + slope[c]*min(v,0) covers prelu; 2 = general switch.
  int mode;
  float lo, hi;
  float lo2, hi2;  // modes 0/1 with a POST residual: clamp(v + res, lo2, hi2) restates act2 in {none, relu, relu6}
};

static inline int epilogue_mode(int act, int act2, int res_mode, const float *slope, float act_p0, float act_p1,
                                float &lo, float &hi, float &lo2, float &hi2) {
  lo = lo2 = -3.402823466e+38f, hi = hi2 = 3.402823466e+38f;
  if (res_mode == RTSEG_RES_POST) {
    switch (act2) {
      case RTSEG_ACT_NONE: break;
      case RTSEG_ACT_RELU: lo2 = 0.f; break;
      case RTSEG_ACT_RELU6: lo2 = 0.f, hi2 = 6.f; break;
      default: return 2;
    }
  }
  switch (act) {
    case RTSEG_ACT_NONE: return 0;
    case RTSEG_ACT_RELU: lo = 0.f; return 0;
    case RTSEG_ACT_RELU6: lo = 0.f, hi = 6.f; return 0;
    case RTSEG_ACT_HARDTANH: lo = act_p0, hi = act_p1; return 0;
    case RTSEG_ACT_PRELU: return slope ? 1 : 2;
    default: return 2;
  }
}

static inline Epilogue make_epilogue(const rtseg_conv_desc *d, const float *scale, const float *shift,
                                     const float *slope) {
  Epilogue e;
  e.scale = scale;
  e.shift = shift;
  e.slope = slope;
  e.act = d->act;
  e.act2 = d->act2;
  e.res_mode = d->res_mode;
  e.p0 = d->act_p0;
  e.p1 = d->act_p1;
  e.mode = epilogue_mode(d->act, d->act2, d->res_mode, slope, d->act_p0, d->act_p1, e.lo, e.hi, e.lo2, e.hi2);
  return e;
}

int validate_conv_desc(const rtseg_conv_desc *d, const void *res);

// Geometry of the stride-2 3x3 image stem (stem.cu, stem_tc.cu).
// Dataset normalisation of a uint8 image, exactly as albumentations.Normalize computes it in fp32 (reference
// datasets/cityscapes.py:122-123): v = (float(x) - mean*255) * (1 / (std*255)).  Ignored for float / bf16 inputs.
struct ImgNorm {
  float m255[4];
  float rden[4];
};

// Internal dtype code of uint8 image inputs (the public rtseg_dtype covers activations only).
constexpr int RTSEG_U8 = 2;

// mean / std -> the two fp32 constants per channel, rounded where albumentations rounds them: float32 arrays
// multiplied by max_pixel_value = 255 in float32, then np.reciprocal in float32.
static inline int make_img_norm(int c, const float *mean_host, const float *std_host, ImgNorm &nz) {
  for (int i = 0; i < 4; ++i) nz.m255[i] = 0.f, nz.rden[i] = 1.f;
  for (int i = 0; i < c; ++i) {
    const float s255 = std_host[i] * 255.0f;
    if (!(s255 > 0.f)) return -1;
    nz.m255[i] = mean_host[i] * 255.0f;
    nz.rden[i] = 1.0f / s255;
  }
  return 0;
}

struct StemParams {
  int n, h, w, cin, cconv, ho, wo;
  int64_t sn, sc, sh, sw;  // input strides in elements
  int y_ct, y_co;
  int pool;
  ImgNorm nz;              // uint8 inputs only
  int u8_words = 0;        // uint8 HWC, 3 channels, 4-byte aligned rows: windows are read as aligned 32-bit words
};

// ---- device side ---------------------------------------------------------------
#ifdef __CUDACC__

// NaN-propagating max / min: torch's relu / relu6 / hardtanh / max_pool2d hand a NaN through, fmaxf / fminf drop it.
__device__ __forceinline__ float fmax_nan(float a, float b) {
  float d;
  asm("max.NaN.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
  return d;
}
__device__ __forceinline__ float fmin_nan(float a, float b) {
  float d;
  asm("min.NaN.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
  return d;
}

__device__ __forceinline__ float act_apply(float v, int act, float p0, float p1, float slope) {
  switch (act) {
    case RTSEG_ACT_NONE: return v;
    case RTSEG_ACT_RELU: return fmax_nan(v, 0.f);
    case RTSEG_ACT_RELU6: return fmin_nan(fmax_nan(v, 0.f), 6.f);
    case RTSEG_ACT_LEAKYRELU: return v > 0.f ? v : v * p0;
    case RTSEG_ACT_PRELU: return v > 0.f ? v : v * slope;
    case RTSEG_ACT_ELU: return v > 0.f ? v : p0 * expm1f(v);
    case RTSEG_ACT_CELU: return fmax_nan(v, 0.f) + fmin_nan(0.f, p0 * expm1f(v / p0));
    case RTSEG_ACT_SELU: {
      const float alpha = 1.6732632423543772848170429916717f;
      const float lam = 1.0507009873554804934193349852946f;
      return lam * (v > 0.f ? v : alpha * expm1f(v));
    }
    case RTSEG_ACT_HARDSWISH: return v * fmin_nan(fmax_nan(v + 3.f, 0.f), 6.f) * (1.f / 6.f);
    case RTSEG_ACT_HARDTANH: return fmin_nan(fmax_nan(v, p0), p1);
    case RTSEG_ACT_GELU: return 0.5f * v * (1.f + erff(v * 0.70710678118654752440f));
    case RTSEG_ACT_SILU: return v / (1.f + expf(-v));
    case RTSEG_ACT_SIGMOID: return 1.f / (1.f + expf(-v));
    case RTSEG_ACT_TANH: return tanhf(v);
    default: return v;
  }
}

// v = acc (+res PRE) -> affine -> act (+res POST) -> act2
__device__ __forceinline__ float epilogue_apply(float acc, float res, int c, const Epilogue &e) {
  float v = acc;
  if (e.res_mode == RTSEG_RES_PRE) v += res;
  const float sc = e.scale ? __ldg(e.scale + c) : 1.f;
  const float sh = e.shift ? __ldg(e.shift + c) : 0.f;
  v = fmaf(v, sc, sh);
  const float sl = (e.act == RTSEG_ACT_PRELU && e.slope) ? __ldg(e.slope + c) : 0.f;
  v = act_apply(v, e.act, e.p0, e.p1, sl);
  if (e.res_mode == RTSEG_RES_POST) {
    v += res;
    v = act_apply(v, e.act2, e.p0, e.p1, sl);
  }
  return v;
}

// One output element with the per-channel affine already in registers.  MODE as in Epilogue::mode.
// RES = false: the caller has no residual input (the add / select per element would be wasted issue slots).
template <int MODE, bool RES = true>
__device__ __forceinline__ float epi_elem(float acc, float rv, float sc, float sh, float sl, const Epilogue &e) {
  float v = (RES && e.res_mode == RTSEG_RES_PRE) ? acc + rv : acc;
  v = fmaf(v, sc, sh);
  if constexpr (MODE == 0) {
    v = fmin_nan(fmax_nan(v, e.lo), e.hi);
    if (RES && e.res_mode == RTSEG_RES_POST) v = fmin_nan(fmax_nan(v + rv, e.lo2), e.hi2);
  } else if constexpr (MODE == 1) {
    v = fmaf(sl, fmin_nan(v, 0.f), fmax_nan(v, 0.f));
    if (RES && e.res_mode == RTSEG_RES_POST) v = fmin_nan(fmax_nan(v + rv, e.lo2), e.hi2);
  } else {
    v = act_apply(v, e.act, e.p0, e.p1, sl);
    if (RES && e.res_mode == RTSEG_RES_POST) v = act_apply(v + rv, e.act2, e.p0, e.p1, sl);
  }
  return v;
}

// Per-channel epilogue vectors of V consecutive channels (scale = 1, shift = 0, slope = 0 when absent).
template <int V>
__device__ __forceinline__ void load_affine(const Epilogue &e, int c0, float (&sc)[V], float (&sh)[V], float (&sl)[V]) {
#pragma unroll
  for (int v = 0; v < V; ++v) {
    sc[v] = e.scale ? __ldg(e.scale + c0 + v) : 1.f;
    sh[v] = e.shift ? __ldg(e.shift + c0 + v) : 0.f;
    sl[v] = (e.slope && e.act == RTSEG_ACT_PRELU) ? __ldg(e.slope + c0 + v) : 0.f;
  }
}

// Scalar typed load/store.
template <typename T>
__device__ __forceinline__ float ld_f(const T *p);
template <>
__device__ __forceinline__ float ld_f<float>(const float *p) {
  return __ldg(p);
}
template <>
__device__ __forceinline__ float ld_f<__nv_bfloat16>(const __nv_bfloat16 *p) {
  return __bfloat162float(*p);
}
// Image element -> float: a uint8 sample is normalised on the way in (two fp32 operations, never contracted into an
// FMA, so the fp32 path reproduces albumentations' arithmetic bit for bit); float / bf16 samples are taken as they are.
template <typename T>
__device__ __forceinline__ float ld_img(const T *p, const ImgNorm &nz, int ch) {
  if constexpr (sizeof(T) == 1) {
    return __fmul_rn(__fsub_rn((float)__ldg(reinterpret_cast<const unsigned char *>(p)), nz.m255[ch]), nz.rden[ch]);
  } else {
    return ld_f<T>(p);
  }
}
template <typename T>
__device__ __forceinline__ void st_f(T *p, float v);
template <>
__device__ __forceinline__ void st_f<float>(float *p, float v) {
  *p = v;
}
template <>
__device__ __forceinline__ void st_f<__nv_bfloat16>(__nv_bfloat16 *p, float v) {
  *p = __float2bfloat16_rn(v);
}

// 16-byte vectors: 4 x f32 or 8 x bf16.
template <typename T>
struct Vec16;
template <>
struct Vec16<float> {
  static constexpr int N = 4;
  using Raw = float4;
  __device__ static __forceinline__ Raw zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }
  __device__ static __forceinline__ Raw ldg(const float *p) { return __ldg(reinterpret_cast<const float4 *>(p)); }
  __device__ static __forceinline__ void unpack(const Raw &t, float (&v)[4]) {
    v[0] = t.x, v[1] = t.y, v[2] = t.z, v[3] = t.w;
  }
  // channel pairs for the packed fp32 FMA (FFMA2) of sm_100
  __device__ static __forceinline__ void unpack2(const Raw &t, float2 (&v)[2]) {
    v[0] = make_float2(t.x, t.y), v[1] = make_float2(t.z, t.w);
  }
  __device__ static __forceinline__ void load(const float *p, float (&v)[4]) {
    const float4 t = __ldg(reinterpret_cast<const float4 *>(p));
    v[0] = t.x, v[1] = t.y, v[2] = t.z, v[3] = t.w;
  }
  __device__ static __forceinline__ void store(float *p, const float (&v)[4]) {
    *reinterpret_cast<float4 *>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
};
template <>
struct Vec16<__nv_bfloat16> {
  static constexpr int N = 8;
  using Raw = uint4;
  __device__ static __forceinline__ Raw zero() { return make_uint4(0, 0, 0, 0); }
  __device__ static __forceinline__ Raw ldg(const __nv_bfloat16 *p) { return __ldg(reinterpret_cast<const uint4 *>(p)); }
  __device__ static __forceinline__ void unpack(const Raw &t, float (&v)[8]) {
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[2 * i] = __uint_as_float(w[i] << 16);
      v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  __device__ static __forceinline__ void unpack2(const Raw &t, float2 (&v)[4]) {
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = make_float2(__uint_as_float(w[i] << 16), __uint_as_float(w[i] & 0xffff0000u));
  }
  __device__ static __forceinline__ void load(const __nv_bfloat16 *p, float (&v)[8]) {
    const uint4 t = __ldg(reinterpret_cast<const uint4 *>(p));
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[2 * i] = __uint_as_float(w[i] << 16);
      v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  __device__ static __forceinline__ void store(__nv_bfloat16 *p, const float (&v)[8]) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
      w[i] = *reinterpret_cast<const uint32_t *>(&h);
    }
    *reinterpret_cast<uint4 *>(p) = make_uint4(w[0], w[1], w[2], w[3]);
  }
};

// ---- bilinear resize geometry (F.interpolate(mode='bilinear')), shared by the forward and adjoint kernels ----
__host__ __device__ __forceinline__ float resize_scale(int in, int out, int align) {
  if (align) return out > 1 ? (float)(in - 1) / (float)(out - 1) : 0.f;
  return (float)in / (float)out;
}

__device__ __forceinline__ void src_index(int dst, float scale, int in, int align, int &i0, int &i1, float &l0,
                                          float &l1) {
  float s;
  if (align) {
    s = scale * (float)dst;
  } else {
    s = scale * ((float)dst + 0.5f) - 0.5f;
    if (s < 0.f) s = 0.f;
  }
  i0 = (int)s;
  if (i0 > in - 1) i0 = in - 1;
  i1 = i0 + (i0 < in - 1 ? 1 : 0);
  l1 = s - (float)i0;
  l0 = 1.f - l1;
}

#endif  // __CUDACC__

}  // namespace rtseg
