// optim.cu — the per-iteration parameter update of the training loop as ONE launch.
//
// The reference's step (core/seg_trainer.py:104-108) runs torch.optim.SGD (utils/optimizer.py:8-12: momentum, weight
// decay) — a handful of ATen kernels per parameter tensor, ~150 tensors for ERFNet — and then ModelEmaV2.update
// (utils/model_ema.py:28-40): one `decay * e + (1 - decay) * m` + copy_ per entry of the state dict (447 entries incl.
// BatchNorm statistics and the int64 num_batches_tracked counters).  Both are pure streaming arithmetic over ~8 MB.
// Here a chunk table built once (device memory, owned by the caller) describes every tensor; each block owns one
// chunk of at most 4096 elements and does, with 16-byte accesses where alignment allows:
//     g' = g + wd * p;   buf = mu * buf + g';   p -= lr * buf            (SGD, dampening 0, no Nesterov)
//     e  = decay * e + (1 - decay) * p_new    (or e = p_new when EMA is off: the reference still mirrors the weights)
// lr / momentum / weight decay / decay are read from a 4-float device buffer, so a CUDA graph that captured the step
// picks up the scheduler's new learning rate without being re-captured.
#include "common.cuh"

namespace rtseg {

namespace {

constexpr int kOptThreads = 256;
constexpr int kOptChunk = 4096;

__global__ void __launch_bounds__(kOptThreads) sgd_ema_kernel(const rtseg_opt_chunk *__restrict__ chunks,
                                                               const float *__restrict__ hyper, int use_ema) {
  const rtseg_opt_chunk ck = chunks[blockIdx.x];
  const float lr = hyper[0], mu = hyper[1], wd = hyper[2], decay = hyper[3];
  if (ck.flags & RTSEG_OPT_INT64) {   // num_batches_tracked: no gradient; the mirror is int64 like the source
    const long long *src = reinterpret_cast<const long long *>(ck.p);
    long long *dst = reinterpret_cast<long long *>(ck.e);
    if (!dst) return;
    for (int i = threadIdx.x; i < ck.n; i += kOptThreads) {
      // torch: decay * e (float32) + (1 - decay) * m (float32), copy_ into the int64 tensor truncates toward zero
      dst[i] = use_ema ? (long long)(decay * (float)dst[i] + (1.f - decay) * (float)src[i]) : src[i];
    }
    return;
  }
  float *p = ck.p;
  const float *g = ck.g;
  float *m = ck.m;
  float *e = reinterpret_cast<float *>(ck.e);
  const bool vec = (ck.n % 4 == 0) && (reinterpret_cast<uintptr_t>(p) % 16 == 0) &&
                   (!g || reinterpret_cast<uintptr_t>(g) % 16 == 0) && (!m || reinterpret_cast<uintptr_t>(m) % 16 == 0) &&
                   (!e || reinterpret_cast<uintptr_t>(e) % 16 == 0);
  auto update = [&](float pv, float gv, float &mv, float ev, float &pn, float &en) {
    pn = pv;
    if (g) {
      const float gg = fmaf(wd, pv, gv);
      mv = fmaf(mu, mv, gg);
      pn = fmaf(-lr, mv, pv);
    }
    en = use_ema ? decay * ev + (1.f - decay) * pn : pn;
  };
  if (vec) {
    for (int i = threadIdx.x * 4; i < ck.n; i += kOptThreads * 4) {
      float4 pv = *reinterpret_cast<const float4 *>(p + i);
      float4 gv = g ? *reinterpret_cast<const float4 *>(g + i) : make_float4(0, 0, 0, 0);
      float4 mv = (g && m) ? *reinterpret_cast<const float4 *>(m + i) : make_float4(0, 0, 0, 0);
      float4 ev = (e && use_ema) ? *reinterpret_cast<const float4 *>(e + i) : make_float4(0, 0, 0, 0);
      float4 pn, en;
      update(pv.x, gv.x, mv.x, ev.x, pn.x, en.x);
      update(pv.y, gv.y, mv.y, ev.y, pn.y, en.y);
      update(pv.z, gv.z, mv.z, ev.z, pn.z, en.z);
      update(pv.w, gv.w, mv.w, ev.w, pn.w, en.w);
      if (g) {
        *reinterpret_cast<float4 *>(p + i) = pn;
        if (m) *reinterpret_cast<float4 *>(m + i) = mv;
      }
      if (e) *reinterpret_cast<float4 *>(e + i) = en;
    }
  } else {
    for (int i = threadIdx.x; i < ck.n; i += kOptThreads) {
      float mv = (g && m) ? m[i] : 0.f, pn, en;
      update(p[i], g ? g[i] : 0.f, mv, (e && use_ema) ? e[i] : 0.f, pn, en);
      if (g) {
        p[i] = pn;
        if (m) m[i] = mv;
      }
      if (e) e[i] = en;
    }
  }
}

}  // namespace

}  // namespace rtseg

using namespace rtseg;

extern "C" int32_t rtseg_opt_chunk_elems(void) { return kOptChunk; }

extern "C" int rtseg_sgd_ema_step(const rtseg_opt_chunk *chunks_dev, int32_t nchunks, const float *hyper_dev, int32_t use_ema,
                                  rtseg_stream_t stream) {
  RTSEG_REQUIRE(chunks_dev && hyper_dev, "NULL argument");
  RTSEG_REQUIRE(nchunks >= 0, "negative chunk count");
  if (nchunks == 0) return RTSEG_OK;
  sgd_ema_kernel<<<(unsigned)nchunks, kOptThreads, 0, as_stream(stream)>>>(chunks_dev, hyper_dev, use_ema ? 1 : 0);
  RTSEG_LAUNCH_CHECK("sgd_ema_kernel");
  return RTSEG_OK;
}
