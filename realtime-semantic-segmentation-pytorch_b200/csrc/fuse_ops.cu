// Fused glue of the two-branch nets: everything BiSeNetv2's aggregation layer and DDRNet's
// bilateral fusion / DAPPM do BETWEEN convolutions, as one pass each instead of 2-4 ATen kernels.
//
//   rtseg_combine_up : y = act( a (+|*) g(resize(b)) )   resize = F.interpolate(bilinear, align_corners=True) of a
//                      lower-resolution (or 1x1: a broadcast) operand, g = identity or sigmoid.  Restates
//                      reference models/bisenetv2.py:190-201 (`x_d_high * x_s_high`, `x_d_low * x_s_low`,
//                      `x_high + interpolate(x_low)`, `res + x` of the context embedding :166-170) and
//                      models/ddrnet.py:165 (`high_conv2(x_high) + x_low`), :238-243 (BilateralFusion),
//                      :277-291 (DAPPM `y1 + y2` ...).  The resized operand never exists in memory.
//   rtseg_avgpool2d  : nn.AvgPool2d(k, s, p) with count_include_pad=True (the default): DAPPM's 5/2, 9/4, 17/8
//                      pools (models/ddrnet.py:266-275) and BiSeNetv2's 3/2/1 pool (:182).
//
// Both are HBM-bound elementwise passes: one thread = one output pixel x one 16-byte channel vector.
#include <initializer_list>

#include "common.cuh"

namespace rtseg {

namespace {

template <typename T, int V>
__device__ __forceinline__ void ldv(const T *p, float (&v)[V]) {
  if constexpr (V == 1) v[0] = ld_f<T>(p);
  else Vec16<T>::load(p, v);
}

template <typename T, int V>
__device__ __forceinline__ void stv(T *p, const float (&v)[V]) {
  if constexpr (V == 1) st_f<T>(p, v[0]);
  else Vec16<T>::store(p, v);
}

struct CombineParams {
  int n, h, w, c;
  int hb, wb;
  int a_ct, a_co, b_ct, b_co, y_ct, y_co;
  int op, b_sigmoid, act;
  float p0, p1, sy, sx;
};

template <typename T, int V>
__global__ void __launch_bounds__(256) combine_up_kernel(CombineParams q, const T *__restrict__ a, const T *__restrict__ b,
                                                         T *__restrict__ y) {
  const int cvn = q.c / V;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (int64_t)q.n * q.h * q.w * cvn) return;
  const int c0 = (int)(idx % cvn) * V;
  int64_t r = idx / cvn;
  const int ox = (int)(r % q.w);
  r /= q.w;
  const int oy = (int)(r % q.h);
  const int img = (int)(r / q.h);
  const int64_t pix = ((int64_t)img * q.h + oy) * q.w + ox;
  float av[V], bv[V];
  ldv<T, V>(a + pix * q.a_ct + q.a_co + c0, av);
  const T *bb = b + (int64_t)img * q.hb * q.wb * q.b_ct + q.b_co + c0;
  if (q.hb == q.h && q.wb == q.w) {
    ldv<T, V>(bb + ((int64_t)oy * q.wb + ox) * q.b_ct, bv);
  } else {
    int y0, y1, x0, x1;
    float ly0, ly1, lx0, lx1;
    src_index(oy, q.sy, q.hb, 1, y0, y1, ly0, ly1);
    src_index(ox, q.sx, q.wb, 1, x0, x1, lx0, lx1);
    float p00[V], p01[V], p10[V], p11[V];
    ldv<T, V>(bb + ((int64_t)y0 * q.wb + x0) * q.b_ct, p00);
    ldv<T, V>(bb + ((int64_t)y0 * q.wb + x1) * q.b_ct, p01);
    ldv<T, V>(bb + ((int64_t)y1 * q.wb + x0) * q.b_ct, p10);
    ldv<T, V>(bb + ((int64_t)y1 * q.wb + x1) * q.b_ct, p11);
#pragma unroll
    for (int e = 0; e < V; ++e) bv[e] = ly0 * (lx0 * p00[e] + lx1 * p01[e]) + ly1 * (lx0 * p10[e] + lx1 * p11[e]);
  }
  float o[V];
#pragma unroll
  for (int e = 0; e < V; ++e) {
    float g = bv[e];
    if (q.b_sigmoid) g = 1.f / (1.f + expf(-g));
    const float v = q.op ? av[e] * g : av[e] + g;
    o[e] = act_apply(v, q.act, q.p0, q.p1, 0.f);
  }
  stv<T, V>(y + pix * q.y_ct + q.y_co + c0, o);
}

struct PoolParams {
  int n, h, w, c, ho, wo, k, s, p;
  int x_ct, x_co, y_ct, y_co;
};

template <typename T, int V>
__global__ void __launch_bounds__(256) avgpool2d_kernel(PoolParams q, const T *__restrict__ x, T *__restrict__ y) {
  const int cvn = q.c / V;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (int64_t)q.n * q.ho * q.wo * cvn) return;
  const int c0 = (int)(idx % cvn) * V;
  int64_t r = idx / cvn;
  const int ox = (int)(r % q.wo);
  r /= q.wo;
  const int oy = (int)(r % q.ho);
  const int img = (int)(r / q.ho);
  const int y_lo = max(oy * q.s - q.p, 0), y_hi = min(oy * q.s - q.p + q.k, q.h);
  const int x_lo = max(ox * q.s - q.p, 0), x_hi = min(ox * q.s - q.p + q.k, q.w);
  float acc[V];
#pragma unroll
  for (int e = 0; e < V; ++e) acc[e] = 0.f;
  const T *xb = x + (int64_t)img * q.h * q.w * q.x_ct + q.x_co + c0;
  for (int iy = y_lo; iy < y_hi; ++iy)
    for (int ix = x_lo; ix < x_hi; ++ix) {
      float v[V];
      ldv<T, V>(xb + ((int64_t)iy * q.w + ix) * q.x_ct, v);
#pragma unroll
      for (int e = 0; e < V; ++e) acc[e] += v[e];
    }
  // count_include_pad=True: the divisor is the window clipped to the PADDED extent (torch's pool_size), which is
  // k * k except where the window runs past the padded border (ceil_mode only) — never here (floor mode)
  const float inv = 1.f / (float)(q.k * q.k);
#pragma unroll
  for (int e = 0; e < V; ++e) acc[e] *= inv;
  stv<T, V>(y + (((int64_t)img * q.ho + oy) * q.wo + ox) * q.y_ct + q.y_co + c0, acc);
}

bool vec_slices(int dtype, int c, std::initializer_list<int> pitches_and_offsets, std::initializer_list<const void *> ptrs) {
  const int V = dtype == RTSEG_BF16 ? 8 : 4;
  if (c % V) return false;
  for (int v : pitches_and_offsets)
    if (v % V) return false;
  for (const void *p : ptrs)
    if (reinterpret_cast<uintptr_t>(p) & 15) return false;
  return true;
}

}  // namespace

}  // namespace rtseg

using namespace rtseg;

extern "C" int rtseg_combine_up(int32_t n, int32_t h, int32_t w, int32_t c, int32_t dtype, const void *a, int32_t a_ctotal,
                                int32_t a_coff, const void *b, int32_t hb, int32_t wb, int32_t b_ctotal, int32_t b_coff,
                                int32_t op, int32_t b_sigmoid, int32_t act, float act_p0, float act_p1, void *y,
                                int32_t y_ctotal, int32_t y_coff, rtseg_stream_t stream) {
  RTSEG_REQUIRE(a && b && y, "NULL pointer");
  RTSEG_REQUIRE(n > 0 && h > 0 && w > 0 && c > 0 && hb > 0 && wb > 0, "non-positive extent");
  RTSEG_REQUIRE(a_coff >= 0 && b_coff >= 0 && y_coff >= 0 && a_ctotal >= a_coff + c && b_ctotal >= b_coff + c &&
                    y_ctotal >= y_coff + c,
                "channel slice out of range");
  RTSEG_REQUIRE(op == 0 || op == 1, "op must be 0 (add) or 1 (multiply)");
  RTSEG_REQUIRE(act != RTSEG_ACT_PRELU, "rtseg_combine_up: no slope vector here");
  CombineParams q{n, h, w, c, hb, wb, a_ctotal, a_coff, b_ctotal, b_coff, y_ctotal, y_coff, op, b_sigmoid, act, act_p0, act_p1,
                  resize_scale(hb, h, 1), resize_scale(wb, w, 1)};
  const bool vec = vec_slices(dtype, c, {a_ctotal, a_coff, b_ctotal, b_coff, y_ctotal, y_coff}, {a, b, y});
  const int V = !vec ? 1 : dtype == RTSEG_BF16 ? 8 : 4;
  const int64_t threads = (int64_t)n * h * w * (c / V);
  RTSEG_REQUIRE(cdiv(threads, 256) <= INT32_MAX, "tensor too large");
  const unsigned grid = (unsigned)cdiv(threads, 256);
#define ARGS <<<grid, 256, 0, as_stream(stream)>>>(q, (const T *)a, (const T *)b, (T *)y)
  if (dtype == RTSEG_BF16) {
    using T = __nv_bfloat16;
    if (vec) combine_up_kernel<T, 8> ARGS; else combine_up_kernel<T, 1> ARGS;
  } else if (dtype == RTSEG_F32) {
    using T = float;
    if (vec) combine_up_kernel<T, 4> ARGS; else combine_up_kernel<T, 1> ARGS;
  } else {
    RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
  }
#undef ARGS
  RTSEG_LAUNCH_CHECK("combine_up_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_avgpool2d(int32_t n, int32_t h, int32_t w, int32_t c, int32_t k, int32_t s, int32_t p, int32_t dtype,
                               const void *x, int32_t x_ctotal, int32_t x_coff, void *y, int32_t y_ctotal, int32_t y_coff,
                               rtseg_stream_t stream) {
  RTSEG_REQUIRE(x && y, "NULL pointer");
  RTSEG_REQUIRE(n > 0 && h > 0 && w > 0 && c > 0 && k > 0 && s > 0 && p >= 0 && 2 * p <= k, "bad pooling geometry");
  RTSEG_REQUIRE(h + 2 * p >= k && w + 2 * p >= k, "window larger than the padded input");
  RTSEG_REQUIRE(x_coff >= 0 && y_coff >= 0 && x_ctotal >= x_coff + c && y_ctotal >= y_coff + c, "channel slice out of range");
  PoolParams q{n, h, w, c, (h + 2 * p - k) / s + 1, (w + 2 * p - k) / s + 1, k, s, p, x_ctotal, x_coff, y_ctotal, y_coff};
  const bool vec = vec_slices(dtype, c, {x_ctotal, x_coff, y_ctotal, y_coff}, {x, y});
  const int V = !vec ? 1 : dtype == RTSEG_BF16 ? 8 : 4;
  const int64_t threads = (int64_t)n * q.ho * q.wo * (c / V);
  RTSEG_REQUIRE(cdiv(threads, 256) <= INT32_MAX, "tensor too large");
  const unsigned grid = (unsigned)cdiv(threads, 256);
#define ARGS <<<grid, 256, 0, as_stream(stream)>>>(q, (const T *)x, (T *)y)
  if (dtype == RTSEG_BF16) {
    using T = __nv_bfloat16;
    if (vec) avgpool2d_kernel<T, 8> ARGS; else avgpool2d_kernel<T, 1> ARGS;
  } else if (dtype == RTSEG_F32) {
    using T = float;
    if (vec) avgpool2d_kernel<T, 4> ARGS; else avgpool2d_kernel<T, 1> ARGS;
  } else {
    RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
  }
#undef ARGS
  RTSEG_LAUNCH_CHECK("avgpool2d_kernel");
  return RTSEG_OK;
}
