// ohem.cu — K12': OHEM cross entropy, forward and backward.
//
// Replaces OhemCELoss.forward (reference core/loss.py:13-20):
//     n_min     = labels[labels != ignore].numel() // 16
//     loss      = CrossEntropyLoss(ignore_index, reduction='none')(logits, labels).view(-1)
//     loss_hard = loss[loss > thresh]
//     if loss_hard.numel() < n_min: loss_hard, _ = loss.topk(n_min)
//     return mean(loss_hard)
// The reference materialises log_softmax (637 MB at 8x19x1024x1024), does two
// boolean-mask gathers (two host syncs), branches on the host and runs a
// radix-select top-k.  Here:
//   pass 1  one read of the logits: per-pixel CE (fp32), #valid, #hard,
//           sum(hard) and a 2048-bin histogram of the order-preserving loss key
//           (shared-memory atomics, flushed once per block);
//   select  a one-block kernel decides the branch ON THE DEVICE;
//   top-k   (only if needed) two more 11/10-bit histogram refinements over the
//           4-byte-per-pixel loss buffer find the exact k-th largest loss, a last
//           pass sums everything above it:  mean = (sum_gt + r*kth) / k.
// No host synchronisation, no sort, no gather.  Ignored pixels have loss 0 and
// take part in the top-k exactly as in the reference (core/loss.py:18).
//
// Backward: d/dlogits = (softmax - onehot) * weight(pixel) * grad_out / n_sel,
// weight = [loss > thresh] or, in the top-k branch, 1 above the k-th value and
// r/count_eq on exact ties with it (torch.topk picks an arbitrary subset of
// the ties; the symmetric choice has the same loss value and is deterministic).
#include <algorithm>
#include <cfloat>
#include <mutex>

#include "common.cuh"

namespace rtseg {

constexpr int kBins1 = 2048, kBins2 = 2048, kBins3 = 1024;

struct OhemWs {
  unsigned long long n_valid, n_hard, k, count_gt, count_eq, need_eq;
  double sum_hard, sum_gt;
  unsigned int branch, bin1, bin2, bin3, kth_key, pad;
  unsigned long long rank1, rank2;  // elements still to take inside the selected bin
  float kth, thresh;
  unsigned int hist1[kBins1], hist2[kBins2], hist3[kBins3];
};

__device__ __forceinline__ unsigned int loss_key(float v) {
  const unsigned int b = __float_as_uint(v);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float key_to_loss(unsigned int k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

template <typename T>
__device__ __forceinline__ float logit_at(const T *base, int64_t cstride, int c) {
  return ld_f<T>(base + (int64_t)c * cstride);
}

// pass 1 ------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) ohem_ce_kernel(int64_t npix, int c, int64_t hw, int layout, int pitch,
                                                      const T *__restrict__ logits, const int64_t *__restrict__ labels,
                                                      int ignore_index, float thresh, float *__restrict__ loss_px,
                                                      OhemWs *__restrict__ ws) {
  __shared__ unsigned int hist[kBins1];
  __shared__ unsigned long long s_valid, s_hard;
  __shared__ double s_sum;
  for (int i = threadIdx.x; i < kBins1; i += blockDim.x) hist[i] = 0;
  if (threadIdx.x == 0) s_valid = 0, s_hard = 0, s_sum = 0.0;
  __syncthreads();

  unsigned int my_valid = 0, my_hard = 0;
  float my_sum = 0.f;
  for (int64_t pix = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; pix < npix;
       pix += (int64_t)gridDim.x * blockDim.x) {
    const int64_t lab = __ldg(labels + pix);
    float loss = 0.f;
    if (lab != (int64_t)ignore_index) {
      ++my_valid;
      const T *base;
      int64_t cs;
      if (layout == 0) {  // NCHW
        const int64_t b = pix / hw, s = pix % hw;
        base = logits + b * c * hw + s;
        cs = hw;
      } else {  // NHWC, pixels `pitch` elements apart (pitch >= c: a class map inside a wider, aligned buffer)
        base = logits + pix * pitch;
        cs = 1;
      }
      float m = -FLT_MAX;
      for (int k = 0; k < c; ++k) m = fmaxf(m, logit_at<T>(base, cs, k));
      float sum = 0.f;
      for (int k = 0; k < c; ++k) sum += expf(logit_at<T>(base, cs, k) - m);
      const float xl = (lab >= 0 && lab < c) ? logit_at<T>(base, cs, (int)lab) : 0.f;
      // -log_softmax[label] = -(x_l - m - log(sum))
      loss = -((xl - m) - logf(sum));
      if (loss > thresh) {
        ++my_hard;
        my_sum += loss;
      }
    }
    loss_px[pix] = loss;
    atomicAdd(&hist[loss_key(loss) >> 21], 1u);
  }
  // block reduction of the three scalars
  for (int o = 16; o > 0; o >>= 1) {
    my_valid += __shfl_xor_sync(0xffffffffu, my_valid, o);
    my_hard += __shfl_xor_sync(0xffffffffu, my_hard, o);
    my_sum += __shfl_xor_sync(0xffffffffu, my_sum, o);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(&s_valid, (unsigned long long)my_valid);
    atomicAdd(&s_hard, (unsigned long long)my_hard);
    atomicAdd(&s_sum, (double)my_sum);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kBins1; i += blockDim.x)
    if (hist[i]) atomicAdd(&ws->hist1[i], hist[i]);
  if (threadIdx.x == 0) {
    atomicAdd(&ws->n_valid, s_valid);
    atomicAdd(&ws->n_hard, s_hard);
    atomicAdd(&ws->sum_hard, s_sum);
  }
}

// Walk a histogram from the top bin down until `need` elements are covered.
__device__ void pick_bin(const unsigned int *hist, int nbins, unsigned long long need, unsigned int &bin,
                         unsigned long long &remaining) {
  unsigned long long above = 0;
  int b = nbins - 1;
  for (; b > 0; --b) {
    if (above + hist[b] >= need) break;
    above += hist[b];
  }
  bin = (unsigned int)b;
  remaining = need - above;  // >= 1 elements still to take from bin b
}

__global__ void ohem_select1_kernel(OhemWs *ws, float thresh, float *out) {
  if (threadIdx.x != 0) return;
  const unsigned long long k = ws->n_valid / 16;
  ws->k = k;
  ws->thresh = thresh;
  if (ws->n_hard >= k) {
    ws->branch = 0;
    const float loss = (float)(ws->sum_hard / (double)ws->n_hard);  // 0/0 -> NaN like torch.mean of empty
    out[0] = loss;
    out[1] = thresh;
    out[2] = (float)ws->n_hard;
    out[3] = 0.f;
    return;
  }
  ws->branch = 1;
  pick_bin(ws->hist1, kBins1, k, ws->bin1, ws->rank1);
}

__global__ void __launch_bounds__(256) ohem_refine_kernel(int64_t npix, const float *__restrict__ loss_px,
                                                          OhemWs *__restrict__ ws, int level) {
  if (ws->branch == 0) return;
  __shared__ unsigned int hist[kBins2];
  const int nb = level == 2 ? kBins2 : kBins3;
  for (int i = threadIdx.x; i < nb; i += blockDim.x) hist[i] = 0;
  __syncthreads();
  const unsigned int prefix = level == 2 ? ws->bin1 : ((ws->bin1 << 11) | ws->bin2);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (int64_t)gridDim.x * blockDim.x) {
    const unsigned int key = loss_key(loss_px[i]);
    if (level == 2) {
      if ((key >> 21) == prefix) atomicAdd(&hist[(key >> 10) & 0x7ffu], 1u);
    } else {
      if ((key >> 10) == prefix) atomicAdd(&hist[key & 0x3ffu], 1u);
    }
  }
  __syncthreads();
  unsigned int *g = level == 2 ? ws->hist2 : ws->hist3;
  for (int i = threadIdx.x; i < nb; i += blockDim.x)
    if (hist[i]) atomicAdd(&g[i], hist[i]);
}

__global__ void ohem_select23_kernel(OhemWs *ws, int level) {
  if (threadIdx.x != 0 || ws->branch == 0) return;
  if (level == 2) {
    pick_bin(ws->hist2, kBins2, ws->rank1, ws->bin2, ws->rank2);
  } else {
    unsigned long long need_eq;
    pick_bin(ws->hist3, kBins3, ws->rank2, ws->bin3, need_eq);
    ws->need_eq = need_eq;
    ws->count_eq = ws->hist3[ws->bin3];
    ws->count_gt = ws->k - need_eq;
    ws->kth_key = (ws->bin1 << 21) | (ws->bin2 << 10) | ws->bin3;
    ws->kth = key_to_loss(ws->kth_key);
  }
}

__global__ void __launch_bounds__(256) ohem_sum_gt_kernel(int64_t npix, const float *__restrict__ loss_px,
                                                          OhemWs *__restrict__ ws) {
  if (ws->branch == 0) return;
  const unsigned int kth_key = ws->kth_key;
  float s = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = loss_px[i];
    if (loss_key(v) > kth_key) s += v;
  }
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  __shared__ double acc;
  if (threadIdx.x == 0) acc = 0.0;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) atomicAdd(&acc, (double)s);
  __syncthreads();
  if (threadIdx.x == 0) atomicAdd(&ws->sum_gt, acc);
}

__global__ void ohem_finalize_kernel(OhemWs *ws, float *out) {
  if (threadIdx.x != 0 || ws->branch == 0) return;
  const double total = ws->sum_gt + (double)ws->need_eq * (double)ws->kth;
  out[0] = (float)(total / (double)ws->k);
  out[1] = ws->kth;
  out[2] = (float)ws->k;
  out[3] = 1.f;
}

// backward ------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) ohem_bwd_kernel(int64_t npix, int c, int64_t hw, int layout, int pitch,
                                                       const T *__restrict__ logits, const int64_t *__restrict__ labels,
                                                       int ignore_index, const float *__restrict__ loss_px,
                                                       const OhemWs *__restrict__ ws, const float *__restrict__ grad_out,
                                                       T *__restrict__ grad) {
  const int64_t pix = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= npix) return;
  const T *base;
  T *gbase;
  int64_t cs;
  if (layout == 0) {
    const int64_t b = pix / hw, s = pix % hw;
    base = logits + b * c * hw + s;
    gbase = grad + b * c * hw + s;
    cs = hw;
  } else {
    base = logits + pix * pitch;
    gbase = grad + pix * pitch;
    cs = 1;
    for (int k = c; k < pitch; ++k) st_f<T>(gbase + k, 0.f);  // the padding channels carry no gradient
  }
  const int64_t lab = __ldg(labels + pix);
  float wgt = 0.f;
  if (lab != (int64_t)ignore_index) {
    const float loss = loss_px[pix];
    if (ws->branch == 0) {
      wgt = loss > ws->thresh ? 1.f / (float)ws->n_hard : 0.f;
    } else {
      const unsigned int key = loss_key(loss);
      if (key > ws->kth_key)
        wgt = 1.f / (float)ws->k;
      else if (key == ws->kth_key)
        wgt = (float)((double)ws->need_eq / (double)ws->count_eq / (double)ws->k);
    }
  }
  wgt *= __ldg(grad_out);
  if (wgt == 0.f) {
    for (int k = 0; k < c; ++k) st_f<T>(gbase + (int64_t)k * cs, 0.f);
    return;
  }
  float m = -FLT_MAX;
  for (int k = 0; k < c; ++k) m = fmaxf(m, logit_at<T>(base, cs, k));
  float sum = 0.f;
  for (int k = 0; k < c; ++k) sum += expf(logit_at<T>(base, cs, k) - m);
  const float inv = 1.f / sum;
  for (int k = 0; k < c; ++k) {
    const float pk = expf(logit_at<T>(base, cs, k) - m) * inv;
    st_f<T>(gbase + (int64_t)k * cs, (pk - (k == (int)lab ? 1.f : 0.f)) * wgt);
  }
}

// bf16 NHWC class maps whose pixels are 16-byte aligned (pitch % 8 == 0, e.g. 19 classes in a pitch-32 buffer): the
// whole pixel is read and written as 16-byte vectors, same arithmetic (and order) as ohem_bwd_kernel.
template <int CP>  // CP = pitch, a multiple of 8, <= 32
__global__ void __launch_bounds__(256) ohem_bwd_nhwc_vec_kernel(int64_t npix, int c, const __nv_bfloat16 *__restrict__ logits,
                                                                const int64_t *__restrict__ labels, int ignore_index,
                                                                const float *__restrict__ loss_px, const OhemWs *__restrict__ ws,
                                                                const float *__restrict__ grad_out, __nv_bfloat16 *__restrict__ grad) {
  const int64_t pix = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= npix) return;
  const int64_t lab = __ldg(labels + pix);
  float wgt = 0.f;
  if (lab != (int64_t)ignore_index) {
    const float loss = loss_px[pix];
    if (ws->branch == 0) {
      wgt = loss > ws->thresh ? 1.f / (float)ws->n_hard : 0.f;
    } else {
      const unsigned int key = loss_key(loss);
      if (key > ws->kth_key)
        wgt = 1.f / (float)ws->k;
      else if (key == ws->kth_key)
        wgt = (float)((double)ws->need_eq / (double)ws->count_eq / (double)ws->k);
    }
  }
  wgt *= __ldg(grad_out);
  uint4 *gp = reinterpret_cast<uint4 *>(grad + pix * CP);
  if (wgt == 0.f) {
#pragma unroll
    for (int q = 0; q < CP / 8; ++q) gp[q] = make_uint4(0, 0, 0, 0);
    return;
  }
  float x[CP];
  const uint4 *lp = reinterpret_cast<const uint4 *>(logits + pix * CP);
#pragma unroll
  for (int q = 0; q < CP / 8; ++q) {
    const uint4 t = __ldg(lp + q);
    const uint32_t w4[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) x[q * 8 + 2 * i] = __uint_as_float(w4[i] << 16), x[q * 8 + 2 * i + 1] = __uint_as_float(w4[i] & 0xffff0000u);
  }
  float m = -FLT_MAX;
#pragma unroll
  for (int k = 0; k < CP; ++k)
    if (k < c) m = fmaxf(m, x[k]);
  float sum = 0.f;
#pragma unroll
  for (int k = 0; k < CP; ++k)
    if (k < c) sum += expf(x[k] - m);
  const float inv = 1.f / sum;
  float g[CP];
#pragma unroll
  for (int k = 0; k < CP; ++k) g[k] = k < c ? (expf(x[k] - m) * inv - (k == (int)lab ? 1.f : 0.f)) * wgt : 0.f;
#pragma unroll
  for (int q = 0; q < CP / 8; ++q) {
    uint32_t o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __nv_bfloat162 h = __floats2bfloat162_rn(g[q * 8 + 2 * i], g[q * 8 + 2 * i + 1]);
      o[i] = *reinterpret_cast<const uint32_t *>(&h);
    }
    gp[q] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

}  // namespace rtseg

using namespace rtseg;

extern "C" size_t rtseg_ohem_workspace_bytes(void) { return (sizeof(OhemWs) + 255) / 256 * 256; }

extern "C" int rtseg_ohem_fwd(int32_t n, int32_t c, int32_t h, int32_t w, int32_t layout, int32_t dtype,
                              const void *logits, const int64_t *labels, int32_t ignore_index, float thresh,
                              float *loss_px, void *workspace, float *out, rtseg_stream_t stream) {
  RTSEG_REQUIRE(logits && labels && loss_px && workspace && out, "NULL pointer");
  RTSEG_REQUIRE(n > 0 && c > 0 && h > 0 && w > 0, "non-positive extent");
  const int pitch = (layout >> 8) ? (layout >> 8) : c;   // NHWC pixel pitch in elements, encoded as layout = 1 | pitch << 8
  layout &= 0xff;
  RTSEG_REQUIRE(layout == 0 || layout == 1, "layout must be 0 (NCHW) or 1 (NHWC, optionally | pitch << 8)");
  RTSEG_REQUIRE(pitch >= c && (layout == 1 || pitch == c), "pixel pitch must be >= c and needs the NHWC layout");
  RTSEG_REQUIRE(reinterpret_cast<uintptr_t>(workspace) % 8 == 0, "workspace must be 8-byte aligned");
  cudaStream_t st = as_stream(stream);
  OhemWs *ws = static_cast<OhemWs *>(workspace);
  const int64_t hw = (int64_t)h * w, npix = (int64_t)n * hw;
  RTSEG_CUDA(cudaMemsetAsync(ws, 0, sizeof(OhemWs), st));
  const unsigned grid = (unsigned)std::min<int64_t>(cdiv(npix, 256), (int64_t)num_sms() * 16);
  if (dtype == RTSEG_BF16)
    ohem_ce_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(npix, c, hw, layout, pitch, (const __nv_bfloat16 *)logits, labels,
                                                        ignore_index, thresh, loss_px, ws);
  else if (dtype == RTSEG_F32)
    ohem_ce_kernel<float><<<grid, 256, 0, st>>>(npix, c, hw, layout, pitch, (const float *)logits, labels, ignore_index,
                                                thresh, loss_px, ws);
  else
    RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
  RTSEG_LAUNCH_CHECK("ohem_ce_kernel");
  ohem_select1_kernel<<<1, 32, 0, st>>>(ws, thresh, out);
  RTSEG_LAUNCH_CHECK("ohem_select1_kernel");
  const unsigned rgrid = (unsigned)std::min<int64_t>(cdiv(npix, 256 * 4), (int64_t)num_sms() * 8);
  ohem_refine_kernel<<<rgrid, 256, 0, st>>>(npix, loss_px, ws, 2);
  RTSEG_LAUNCH_CHECK("ohem_refine_kernel");
  ohem_select23_kernel<<<1, 32, 0, st>>>(ws, 2);
  RTSEG_LAUNCH_CHECK("ohem_select23_kernel");
  ohem_refine_kernel<<<rgrid, 256, 0, st>>>(npix, loss_px, ws, 3);
  RTSEG_LAUNCH_CHECK("ohem_refine_kernel");
  ohem_select23_kernel<<<1, 32, 0, st>>>(ws, 3);
  RTSEG_LAUNCH_CHECK("ohem_select23_kernel");
  ohem_sum_gt_kernel<<<rgrid, 256, 0, st>>>(npix, loss_px, ws);
  RTSEG_LAUNCH_CHECK("ohem_sum_gt_kernel");
  ohem_finalize_kernel<<<1, 32, 0, st>>>(ws, out);
  RTSEG_LAUNCH_CHECK("ohem_finalize_kernel");
  return RTSEG_OK;
}

extern "C" int rtseg_ohem_bwd(int32_t n, int32_t c, int32_t h, int32_t w, int32_t layout, int32_t dtype,
                              const void *logits, const int64_t *labels, int32_t ignore_index, const float *loss_px,
                              const void *workspace, const float *grad_out, void *grad_logits,
                              rtseg_stream_t stream) {
  RTSEG_REQUIRE(logits && labels && loss_px && workspace && grad_out && grad_logits, "NULL pointer");
  RTSEG_REQUIRE(n > 0 && c > 0 && h > 0 && w > 0, "non-positive extent");
  const int pitch = (layout >> 8) ? (layout >> 8) : c;   // NHWC pixel pitch in elements, encoded as layout = 1 | pitch << 8
  layout &= 0xff;
  RTSEG_REQUIRE(layout == 0 || layout == 1, "layout must be 0 (NCHW) or 1 (NHWC, optionally | pitch << 8)");
  RTSEG_REQUIRE(pitch >= c && (layout == 1 || pitch == c), "pixel pitch must be >= c and needs the NHWC layout");
  cudaStream_t st = as_stream(stream);
  const OhemWs *ws = static_cast<const OhemWs *>(workspace);
  const int64_t hw = (int64_t)h * w, npix = (int64_t)n * hw;
  const unsigned grid = (unsigned)cdiv(npix, 256);
  if (dtype == RTSEG_BF16 && layout == 1 && (pitch == 24 || pitch == 32) && reinterpret_cast<uintptr_t>(logits) % 16 == 0 &&
      reinterpret_cast<uintptr_t>(grad_logits) % 16 == 0) {
    if (pitch == 32)
      ohem_bwd_nhwc_vec_kernel<32><<<grid, 256, 0, st>>>(npix, c, (const __nv_bfloat16 *)logits, labels, ignore_index, loss_px, ws,
                                                         grad_out, (__nv_bfloat16 *)grad_logits);
    else
      ohem_bwd_nhwc_vec_kernel<24><<<grid, 256, 0, st>>>(npix, c, (const __nv_bfloat16 *)logits, labels, ignore_index, loss_px, ws,
                                                         grad_out, (__nv_bfloat16 *)grad_logits);
    RTSEG_LAUNCH_CHECK("ohem_bwd_nhwc_vec_kernel");
    return RTSEG_OK;
  }
  if (dtype == RTSEG_BF16)
    ohem_bwd_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(npix, c, hw, layout, pitch, (const __nv_bfloat16 *)logits, labels,
                                                         ignore_index, loss_px, ws, grad_out, (__nv_bfloat16 *)grad_logits);
  else if (dtype == RTSEG_F32)
    ohem_bwd_kernel<float><<<grid, 256, 0, st>>>(npix, c, hw, layout, pitch, (const float *)logits, labels, ignore_index,
                                                 loss_px, ws, grad_out, (float *)grad_logits);
  else
    RTSEG_FAIL(RTSEG_ERR_INVALID, "unknown dtype");
  RTSEG_LAUNCH_CHECK("ohem_bwd_kernel");
  return RTSEG_OK;
}

// ---------------------------------------------------------------------------------------------
// OHEM forward for callers that cannot allocate device memory.
//
// rtseg_ohem_fwd needs two scratch buffers from its caller (the per-pixel loss and the selection
// workspace).  Some callers have no allocator to get them from: code running inside a CUDA-graph
// capture or a real-time loop where cudaMalloc is forbidden, a host that owns its logits/labels
// through another runtime (a Go or Java service reaching this library over cgo/JNI with nothing
// but the two tensors it was handed), or a plugin sandbox that may launch kernels but not grow the
// heap.  For them the scratch lives in the library image itself: one file-scope device array per
// GPU, sized for kOhemStaticMaxPixels pixels, present from module load until process exit.
//
// `g_ohem_static_scratch` below is a *device* symbol.  In host code its name is only a handle for
// the runtime — not an address a kernel can dereference — so the device address is looked up once
// per device with cudaGetSymbolAddress (with that device current) and cached.
namespace rtseg {

constexpr int64_t kOhemStaticMaxPixels = 1 << 21;  // 2,097,152 = one 1024 x 2048 image (or 8 x 512 x 512 ...)
constexpr size_t kOhemStaticWsBytes = (sizeof(OhemWs) + 255) / 256 * 256;
constexpr size_t kOhemStaticBytes = kOhemStaticWsBytes + (size_t)kOhemStaticMaxPixels * sizeof(float);

static __device__ __align__(256) unsigned char g_ohem_static_scratch[kOhemStaticBytes];

// Device address of the scratch array on the CURRENT device (resolved on first use, then cached).
static int ohem_static_scratch_ptr(unsigned char **out) {
  static std::mutex mu;
  static unsigned char *cached[64] = {nullptr};
  int dev = 0;
  RTSEG_CUDA(cudaGetDevice(&dev));
  RTSEG_REQUIRE(dev >= 0 && dev < 64, "device index %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (cached[dev] == nullptr) {
    void *p = nullptr;
    RTSEG_CUDA(cudaGetSymbolAddress(&p, g_ohem_static_scratch));
    RTSEG_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a NULL scratch address");
    cached[dev] = static_cast<unsigned char *>(p);
  }
  *out = cached[dev];
  return RTSEG_OK;
}

}  // namespace rtseg

extern "C" int64_t rtseg_ohem_fwd_static_max_pixels(void) { return rtseg::kOhemStaticMaxPixels; }

extern "C" int rtseg_ohem_fwd_static(int32_t n, int32_t c, int32_t h, int32_t w, int32_t layout, int32_t dtype,
                                     const void *logits, const int64_t *labels, int32_t ignore_index, float thresh,
                                     float *out, rtseg_stream_t stream) {
  RTSEG_REQUIRE(n > 0 && c > 0 && h > 0 && w > 0, "non-positive extent");
  const int64_t npix = (int64_t)n * h * w;
  if (npix > kOhemStaticMaxPixels)
    RTSEG_FAIL(RTSEG_ERR_UNSUPPORTED,
               "rtseg_ohem_fwd_static: %lld pixels exceed the built-in scratch (%lld pixels); use rtseg_ohem_fwd with "
               "caller-provided buffers",
               (long long)npix, (long long)kOhemStaticMaxPixels);
  unsigned char *scratch = nullptr;
  if (int rc = ohem_static_scratch_ptr(&scratch)) return rc;
  // Layout of the array: [selection workspace | per-pixel loss].  rtseg_ohem_fwd zeroes the workspace
  // itself (cudaMemsetAsync on `stream`) and overwrites every loss it later reads, so nothing stale
  // from the previous call survives into this one.
  float *loss_px = reinterpret_cast<float *>(scratch + kOhemStaticWsBytes);
  return rtseg_ohem_fwd(n, c, h, w, layout, dtype, logits, labels, ignore_index, thresh, loss_px, scratch, out, stream);
}
