// runtime.cu — library-level entry points: version, errors, launch accounting,
// descriptor validation shared by every conv kernel.
#include <atomic>
#include <cstdarg>
#include <cstring>

#include "common.cuh"

namespace rtseg {

static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

int validate_conv_desc(const rtseg_conv_desc *d, const void *res) {
  RTSEG_REQUIRE(d != nullptr, "conv descriptor is NULL");
  RTSEG_REQUIRE(d->n > 0 && d->h > 0 && d->w > 0 && d->cin > 0 && d->cout > 0,
                "non-positive tensor extent (n=%d h=%d w=%d cin=%d cout=%d)", d->n, d->h, d->w, d->cin,
                d->cout);
  RTSEG_REQUIRE(d->kh >= 1 && d->kw >= 1 && d->kh <= 7 && d->kw <= 7, "kernel size %dx%d out of range",
                d->kh, d->kw);
  RTSEG_REQUIRE(d->stride >= 1 && d->dil >= 1 && d->pad_h >= 0 && d->pad_w >= 0,
                "bad stride/dilation/padding");
  RTSEG_REQUIRE(d->dtype == RTSEG_F32 || d->dtype == RTSEG_BF16, "unknown dtype %d", d->dtype);
  RTSEG_REQUIRE(d->act >= RTSEG_ACT_NONE && d->act <= RTSEG_ACT_TANH, "unknown activation %d", d->act);
  RTSEG_REQUIRE(d->act2 >= RTSEG_ACT_NONE && d->act2 <= RTSEG_ACT_TANH, "unknown activation2 %d", d->act2);
  RTSEG_REQUIRE(d->groups == 1 || d->groups == d->cin, "groups must be 1 or cin (got %d)", d->groups);
  RTSEG_REQUIRE(d->x_ctotal >= d->x_coff + d->cin && d->x_coff >= 0, "input channel slice out of range");
  RTSEG_REQUIRE(d->y_ctotal >= d->y_coff + d->cout && d->y_coff >= 0, "output channel slice out of range");
  RTSEG_REQUIRE(d->res_mode >= RTSEG_RES_NONE && d->res_mode <= RTSEG_RES_POST, "bad res_mode");
  if (d->res_mode != RTSEG_RES_NONE) {
    RTSEG_REQUIRE(res != nullptr, "res_mode set but residual pointer is NULL");
    RTSEG_REQUIRE(d->r_ctotal >= d->r_coff + d->cout && d->r_coff >= 0, "residual channel slice out of range");
  }
  int32_t ho, wo;
  if (rtseg_out_hw(d, &ho, &wo) != 0) return RTSEG_ERR_INVALID;
  return RTSEG_OK;
}

}  // namespace rtseg

extern "C" {

int rtseg_version(void) { return RTSEG_VERSION; }

const char *rtseg_last_error(void) { return rtseg::g_err; }

uint64_t rtseg_launch_count(void) { return rtseg::g_launches.load(std::memory_order_relaxed); }

int rtseg_device_check(void) {
  int dev = 0;
  RTSEG_CUDA(cudaGetDevice(&dev));
  int major = 0;
  RTSEG_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  if (major != 10) RTSEG_FAIL(RTSEG_ERR_ARCH, "device %d has compute capability %d.x; librtseg needs sm_100", dev, major);
  return RTSEG_OK;
}

int rtseg_out_hw(const rtseg_conv_desc *d, int32_t *ho, int32_t *wo) {
  RTSEG_REQUIRE(d && ho && wo, "NULL argument");
  const int eh = d->dil * (d->kh - 1) + 1, ew = d->dil * (d->kw - 1) + 1;
  const int nh = d->h + 2 * d->pad_h - eh, nw = d->w + 2 * d->pad_w - ew;
  RTSEG_REQUIRE(nh >= 0 && nw >= 0, "kernel extent larger than padded input");
  *ho = nh / d->stride + 1;
  *wo = nw / d->stride + 1;
  return RTSEG_OK;
}

}  // extern "C"

// The library links the CUDA runtime statically; the host (PyTorch) has its own
// runtime instance.  Both follow the thread's current driver context, but the
// Python binding also calls this whenever torch's current device changes.
extern "C" int rtseg_set_device(int device) {
  RTSEG_CUDA(cudaSetDevice(device));
  return RTSEG_OK;
}
