/*
 * rtseg.h — C ABI of librtseg.so, the B200 (sm_100a) kernels behind the
 * real-time-segmentation hot path.
 *
 * The reference (acai66/realtime-semantic-segmentation-pytorch) has no FFI
 * layer at all: its hot path is `nn.Module.forward` of the blocks in
 * models/modules.py, dispatched by PyTorch to ATen/cuDNN.  Every entry point
 * below therefore cites the reference *module or function* whose arithmetic
 * it replaces (file:line relative to the reference tree); the Python host in
 * realtime-semantic-segmentation-pytorch_b200/ binds these symbols with
 * ctypes and exposes them under the reference's own class names.
 *
 * Conventions
 *  - plain C types only; every pointer is a DEVICE pointer unless its name
 *    ends in `_host`; the caller owns all memory (inputs, outputs, scratch).
 *  - activations are dense NHWC ("channels_last"); a tensor may be a channel
 *    slice [c_off, c_off+c) of a wider NHWC buffer with `c_total` channels
 *    per pixel (this is how concat / split are fused away).
 *  - all work is enqueued on the `stream` argument (a cudaStream_t); nothing
 *    synchronises, nothing uses the legacy default stream.
 *  - return value: 0 on success, negative rtseg_status otherwise; the text is
 *    available from rtseg_last_error() (thread local).  There is no CPU
 *    fallback anywhere: an unsupported shape is an error.
 */
#ifndef RTSEG_H_
#define RTSEG_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTSEG_VERSION 100

typedef void *rtseg_stream_t; /* cudaStream_t */

typedef enum rtseg_status {
  RTSEG_OK = 0,
  RTSEG_ERR_INVALID = -1,     /* bad argument / inconsistent descriptor      */
  RTSEG_ERR_UNSUPPORTED = -2, /* legal request this library has no kernel for */
  RTSEG_ERR_CUDA = -3,        /* a CUDA runtime / driver call failed          */
  RTSEG_ERR_ARCH = -4         /* device is not sm_100                         */
} rtseg_status;

typedef enum rtseg_dtype { RTSEG_F32 = 0, RTSEG_BF16 = 1 } rtseg_dtype;

/* Activation hub of models/modules.py:111-131 (elementwise members only;
 * 'glu' and 'softmax' are not pointwise and are rejected by the host). */
typedef enum rtseg_act {
  RTSEG_ACT_NONE = 0,
  RTSEG_ACT_RELU = 1,
  RTSEG_ACT_RELU6 = 2,
  RTSEG_ACT_LEAKYRELU = 3, /* p0 = negative_slope                      */
  RTSEG_ACT_PRELU = 4,     /* per-channel slope vector (`slope`)       */
  RTSEG_ACT_ELU = 5,       /* p0 = alpha                               */
  RTSEG_ACT_CELU = 6,      /* p0 = alpha                               */
  RTSEG_ACT_SELU = 7,
  RTSEG_ACT_HARDSWISH = 8,
  RTSEG_ACT_HARDTANH = 9,  /* p0 = min_val, p1 = max_val               */
  RTSEG_ACT_GELU = 10,     /* exact erf form                           */
  RTSEG_ACT_SILU = 11,
  RTSEG_ACT_SIGMOID = 12,
  RTSEG_ACT_TANH = 13
} rtseg_act;

/* Where the residual tensor enters the epilogue
 *   v = acc (+ res if PRE);  v = v*scale[c] + shift[c];  v = act(v);
 *   (+ res if POST);  v = act2(v)
 * PRE  = NonBt1DBlock, models/erfnet.py:77-82 (add, then BN, then act);
 *        FeatureFusionModule, models/fastscnn.py:86-93.
 * POST = InvertedResidual models/fastscnn.py:117-121, DABModule
 *        models/dabnet.py:93-96, SSnbtUnit models/lednet.py:80-87 (act2). */
typedef enum rtseg_res_mode { RTSEG_RES_NONE = 0, RTSEG_RES_PRE = 1, RTSEG_RES_POST = 2 } rtseg_res_mode;

/* One fused "conv -> folded BatchNorm -> activation (-> residual)" layer. */
typedef struct rtseg_conv_desc {
  int32_t n, h, w;          /* input batch and spatial size                      */
  int32_t cin, cout;        /* logical channels of this op                       */
  int32_t kh, kw;           /* kernel size (3,1) / (1,3) / (3,3) / (1,1) / ...   */
  int32_t stride;           /* applied to both dims, like nn.Conv2d(int)         */
  int32_t dil;              /* applied to both dims                              */
  int32_t pad_h, pad_w;
  int32_t groups;           /* 1 = dense, cin = depthwise (cout = m*cin)         */
  int32_t dtype;            /* rtseg_dtype of x, res and y                       */
  int32_t act;              /* rtseg_act                                         */
  float act_p0, act_p1;
  int32_t res_mode;         /* rtseg_res_mode                                    */
  int32_t act2;             /* second activation after a POST residual           */
  int32_t x_ctotal, x_coff; /* x is channels [x_coff, x_coff+cin) of x_ctotal   */
  int32_t y_ctotal, y_coff; /* y is channels [y_coff, y_coff+cout) of y_ctotal  */
  int32_t r_ctotal, r_coff; /* residual slice (same spatial size as y)          */
} rtseg_conv_desc;

/* ---- library ---------------------------------------------------------- */
int rtseg_version(void);
const char *rtseg_last_error(void);
/* 0 if the current device is a compute-capability 10.x part, else RTSEG_ERR_ARCH. */
int rtseg_device_check(void);
/* Make `device` current for this library's (statically linked) CUDA runtime on
 * the calling thread; the Python host calls it when torch's device changes. */
int rtseg_set_device(int device);
/* Number of kernel launches this library has enqueued from the calling
 * process since load (monotonic; bench.py reports the delta as gpu_launches). */
uint64_t rtseg_launch_count(void);
int rtseg_out_hw(const rtseg_conv_desc *d, int32_t *ho, int32_t *wo);

/* ---- K1': depthwise conv + BN + act ------------------------------------
 * Replaces DWConvBNAct, models/modules.py:46-59 (nn.Conv2d(groups=in) :55,
 * nn.BatchNorm2d :57, Activation :58).  Any kh,kw in {1,3,5}, stride 1/2,
 * dilation >= 1, channel multiplier m = cout/cin >= 1.
 *   w      fp32 [kh][kw][cout]
 *   scale, shift  fp32 [cout]  (BN folded: scale = g/sqrt(var+eps),
 *                               shift = b - mean*scale + conv_bias*scale)
 *   slope  fp32 [cout] or NULL (PReLU only)                               */
int rtseg_conv_dw(const rtseg_conv_desc *d, const void *x, const float *w, const float *scale,
                  const float *shift, const float *slope, const void *res, void *y,
                  rtseg_stream_t stream);

/* ---- K3'/K4' CUDA-core path: dense conv + BN + act ----------------------
 * Replaces ConvBNAct, models/modules.py:73-85, PWConvBNAct :63-69, conv1x1 /
 * conv3x3 :7-15 for every shape (fp32 parity path, odd channel counts).
 *   w fp32 [kh][kw][cin][cout]                                              */
int rtseg_conv_dense(const rtseg_conv_desc *d, const void *x, const float *w, const float *scale,
                     const float *shift, const float *slope, const void *res, void *y,
                     rtseg_stream_t stream);

/* ---- K2'/K3'/K4' tensor-core path (tcgen05 + TMEM + TMA), bf16 only ------
 * Same arithmetic as rtseg_conv_dense with bf16 operands and fp32
 * accumulation.  Constraints (else RTSEG_ERR_UNSUPPORTED, see
 * rtseg_conv_tc_supported): groups == 1, dtype bf16, stride 1 or 2 (even H, W),
 * kh*kw <= 9, channel slices 16-byte aligned (x_ctotal, x_coff multiples of
 * 8; cin itself is free: DABNet's 35/131/259-channel concatenations,
 * models/dabnet.py:42-58, run here), and the weights of one output-channel
 * tile must fit in shared memory.
 *   w  bf16 [kh*kw][cout_pad][cin_pad]  (K-major per tap; cout_pad / cin_pad
 *      = cout / cin rounded up to 16, padding zero).  The kernel never reads
 *      input channels past cin: the TMA unit zero-fills the last K block.   */
int rtseg_conv_tc_supported(const rtseg_conv_desc *d);
int rtseg_conv_tc(const rtseg_conv_desc *d, const void *x, const void *w_bf16, const float *scale,
                  const float *shift, const float *slope, const void *res, void *y,
                  rtseg_stream_t stream);
/* ---- two convolutions of a NonBt1D block in one launch (tcgen05, bf16) -----
 * Replaces ConvBNAct(c, c, (3,1)) -> ConvBNAct(c, c, (1,3)) — models/erfnet.py:65-66 — and the block tail
 * ConvBNAct(c, c, (3,1)) -> Conv2d(c, c, (1,3)) -> `x += residual` -> BatchNorm2d -> Activation —
 * models/erfnet.py:67-70, 77-81 — for undilated blocks with c in {16, 64}: the intermediate tensor stays in
 * shared memory (written by the first epilogue in the layout the second MMA reads), so the pair moves
 * (read x, write y) instead of (read x, write t, read t, write y) through HBM.
 *   d       describes the FIRST conv (kh = 3, kw = 1, pad (1, 0), stride 1, dil 1, cin == cout) and carries the
 *           second stage's activation (act, act_p0, act_p1), res_mode (NONE or PRE) and the x / y / r slices
 *   w1, w2  bf16 [3][c][c] (tap, cout, cin); scale / shift: folded BatchNorm of each stage (NULL = identity)
 *   act1    activation after the first conv: none / relu / relu6
 *   w_scaled  bit 0 / bit 1: w1 / w2 already carry their BatchNorm scale (w' = bf16(scale[co] * w)); the kernel then
 *           adds the shift through one extra MMA per tile and that stage's epilogue does no arithmetic (scale1 /
 *           scale2 are then only used for a PRE residual: v = acc + scale2 * res)                            */
int rtseg_conv_pair_tc_supported(const rtseg_conv_desc *d, int32_t act1);
int rtseg_conv_pair_tc(const rtseg_conv_desc *d, const void *x, const void *w1_bf16, const float *scale1,
                       const float *shift1, int32_t act1, const void *w2_bf16, const float *scale2,
                       const float *shift2, int32_t w_scaled, const void *res, void *y, rtseg_stream_t stream);
/* Every mbarrier wait in the tensor-core kernels carries a ~2 s watchdog: a
 * lost TMA or a wrong descriptor ends the kernel with garbage output and sets
 * a device flag instead of hanging the GPU.  This call synchronises `stream`,
 * returns the flag (1 = a kernel gave up since the last call) and clears it. */
int rtseg_tc_timeout_flag(rtseg_stream_t stream);
/* Diagnostics (no reference counterpart; the reference has no kernels to plan):
 * the execution plan rtseg_conv_tc would choose for `d`, so that a test can
 * assert which regime it exercised.  info = {K block, N tile, N tiles,
 * sub-tiles per unit, ring stages, staging buffers, halo-sharing mode, unit
 * tiles, grid.x (persistent CTAs), shared-memory bytes, TMEM columns, residual
 * boxes per unit}.  units / grid.x = unit tiles every CTA walks. */
int rtseg_conv_tc_plan_info(const rtseg_conv_desc *d, int32_t info[12]);
/* Diagnostics (no reference counterpart): launches of the TMA-staged depthwise
 * kernel (dw_tma_kernel) by this process.  rtseg_conv_dw picks that kernel for
 * bf16 3x3 dilation-1 layers with at least 65,536 output pixels. */
uint64_t rtseg_dw_tma_launch_count(void);

/* ---- K4'+K8: network stem ------------------------------------------------
 * Replaces InitialBlock, models/enet.py:38-48 (conv 3x3 s2 + BN + act on
 * channels [0,cconv) || MaxPool2d(3,2,1) of the input on channels
 * [cconv, cconv+cin)) and the plain stride-2 stems ConvBNAct(3, C, 3, 2)
 * (models/fastscnn.py:38, models/dabnet.py:19) when pool == 0.
 * The input is the caller's NCHW tensor (x_is_nchw=1, x_dtype f32/bf16) or an
 * NHWC one; cin <= 4.  Output is NHWC `dtype`, y_ctotal channels per pixel.
 *   w fp32 [3][3][cin][cconv]                                              */
int rtseg_stem(const rtseg_conv_desc *d, int32_t x_is_nchw, int32_t x_dtype, int32_t pool,
               const void *x, const float *w, const float *scale, const float *shift,
               const float *slope, void *y, rtseg_stream_t stream);

/* The same stem fed by the decoded image itself: x is uint8 [n][h][w][cin] (HWC, cin <= 4).  Replaces the dataset's
 * AT.Normalize(mean, std) + ToTensorV2 (datasets/cityscapes.py:122-123) and images.to(device, float32)
 * (core/seg_trainer.py:42) in front of InitialBlock / the stride-2 stems: every sample is normalised while the stem
 * builds its window, v = (float(x) - mean*255) * (1 / (std*255)) — the two fp32 operations albumentations performs —
 * so a step ships 3 bytes per pixel to the device instead of 12.  mean_host / std_host: cin floats on the HOST. */
int rtseg_stem_u8(const rtseg_conv_desc *d, int32_t pool, const uint8_t *x, const float *mean_host,
                  const float *std_host, const float *w, const float *scale, const float *shift,
                  const float *slope, void *y, rtseg_stream_t stream);

/* MaxPool2d(3,2,1) branch of InitialBlock (models/enet.py:43) for NHWC inputs
 * with many channels; writes a channel slice of y. Uses d->n,h,w,cin,
 * x_ctotal/x_coff, y_ctotal/y_coff, dtype. */
int rtseg_maxpool3x3s2(const rtseg_conv_desc *d, const void *x, void *y, rtseg_stream_t stream);

/* ---- f1: ENet's pooling pair ------------------------------------------------
 * nn.MaxPool2d(2, 2, return_indices=True) of the downsampling Bottleneck (models/enet.py:130-131, 168-170) and
 * nn.MaxUnpool2d(2, 2) of the upsampling one (:138-139, 176-178).  NHWC channel slices.  `idx` is this library's
 * own index format — one byte per pooled element [n][h/2][w/2][c], the position 0..3 of the maximum inside its 2x2
 * window (row-major); torch's flat index is (2*hp + idx/2) * w + 2*wp + idx%2.  Selection = ATen's: first maximum
 * in row-major order, a NaN always taken.  The unpool writes every output element (zeros included; output size
 * 2*hp (+1) x 2*wp (+1) for odd originals), so its output needs no clearing. */
int rtseg_maxpool2x2_idx(int32_t n, int32_t h, int32_t w, int32_t c, int32_t dtype, const void *x, int32_t x_ctotal,
                         int32_t x_coff, void *y, int32_t y_ctotal, int32_t y_coff, uint8_t *idx, rtseg_stream_t stream);
int rtseg_maxunpool2x2(int32_t n, int32_t hp, int32_t wp, int32_t c, int32_t ho, int32_t wo, int32_t dtype, const void *x,
                       int32_t x_ctotal, int32_t x_coff, const uint8_t *idx, void *y, int32_t y_ctotal, int32_t y_coff,
                       rtseg_stream_t stream);

/* ---- K5': transposed conv 3x3 s2 p1 op1 + BN + act -----------------------
 * Replaces DeConvBNAct, models/modules.py:89-108 (scale_factor 2).
 * d->h,w are the INPUT size; output is 2h x 2w.  d->kh=kw=3, stride=2.
 *   w fp32 [3][3][cin][cout]  with w[ky][kx][ci][co] = torch weight[ci][co][ky][kx]
 * If labels != NULL the per-pixel first-max channel index (torch.max(dim=1)
 * semantics, models/ddrnet.py:55-58, core/seg_trainer.py:172) is written as
 * uint8 [n][2h][2w].  The CUDA-core version always needs y (it takes the
 * argmax in a second kernel); the tensor-core version fuses the argmax into
 * its epilogue and accepts y == NULL, so the logits are never materialised. */
int rtseg_deconv3x3s2(const rtseg_conv_desc *d, const void *x, const float *w, const float *scale,
                      const float *shift, const float *slope, void *y, uint8_t *labels,
                      rtseg_stream_t stream);
/* tcgen05 variant: bf16, cin % 16 == 0; w bf16 [9][cout_pad][cin], cout_pad = cout rounded up to 16. */
int rtseg_deconv3x3s2_tc_supported(const rtseg_conv_desc *d);
int rtseg_deconv3x3s2_tc(const rtseg_conv_desc *d, const void *x, const void *w_bf16,
                         const float *scale, const float *shift, const float *slope, void *y,
                         uint8_t *labels, rtseg_stream_t stream);

/* ---- K10'/K11: bilinear resize, fused head -------------------------------
 * Replaces F.interpolate(mode='bilinear') at models/fastscnn.py:30,89,
 * models/modules.py:153-156, models/dabnet.py:60, models/enet.py:198-202
 * (align_corners = 0 there).  NHWC in, NHWC out (channel slices allowed via
 * x_ctotal/x_coff, y_ctotal/y_coff in elements).                           */
int rtseg_upsample_bilinear(int32_t n, int32_t hi, int32_t wi, int32_t c, int32_t ho, int32_t wo,
                            int32_t align_corners, int32_t dtype, const void *x, int32_t x_ctotal,
                            int32_t x_coff, void *y, int32_t y_ctotal, int32_t y_coff,
                            rtseg_stream_t stream);
/* Fused bilinear upsample + first-max argmax over channels (+ optional
 * confusion-matrix update when target != NULL): the full-resolution logits
 * never exist.  x holds x_ctotal >= c channels per pixel (the classifier may write a padded,
 * 16-byte-aligned buffer); only the first c are classes.  labels_u8 and/or labels_i64 may be NULL.  confmat is int64
 * [c][c], rows = target (utils/metrics.py:4-7 semantics).                   */
int rtseg_upsample_argmax(int32_t n, int32_t hi, int32_t wi, int32_t c, int32_t x_ctotal, int32_t ho,
                          int32_t wo, int32_t align_corners, int32_t dtype, const void *x, uint8_t *labels_u8,
                          int64_t *labels_i64, const int64_t *target, int32_t ignore_index,
                          int64_t *confmat, rtseg_stream_t stream);
/* argmax over channels of an NHWC tensor (preds.max(dim=1)[1], core/seg_trainer.py:172). */
int rtseg_argmax_nhwc(int64_t npix, int32_t c, int32_t c_total, int32_t dtype, const void *x,
                      uint8_t *labels_u8, int64_t *labels_i64, rtseg_stream_t stream);
/* argmax over dim 1 of an NCHW tensor (what the reference's models return). */
int rtseg_argmax_nchw(int32_t n, int32_t c, int64_t hw, int32_t dtype, const void *x,
                      uint8_t *labels_u8, int64_t *labels_i64, rtseg_stream_t stream);

/* Label maps for the auxiliary heads: F.interpolate(masks.unsqueeze(1).float(), aux_size, mode='nearest') then
 * .long(), core/seg_trainer.py:58-62 — int64 [n][hi][wi] -> [n][ho][wo], src = min(floor(dst * in/out), in - 1). */
int rtseg_resize_nearest_i64(int32_t n, int32_t hi, int32_t wi, int32_t ho, int32_t wo, const int64_t *x, int64_t *y,
                             rtseg_stream_t stream);

/* ---- K13': confusion matrix ----------------------------------------------
 * Replaces torchmetrics JaccardIndex.update as called at utils/metrics.py:4-7
 * / core/seg_trainer.py:131: drop target == ignore_index, confmat[t][p] += 1.
 * pred is uint8 (pred_is_i64 = 0) or int64 class indices.  Pixels whose
 * target or pred is outside [0,c) and not ignore_index set *bad_count (may be
 * NULL) instead of being counted.                                           */
int rtseg_confmat_update(int64_t npix, int32_t c, int32_t ignore_index, const void *pred,
                         int32_t pred_is_i64, const int64_t *target, int64_t *confmat,
                         int64_t *bad_count, rtseg_stream_t stream);

/* ---- K12': OHEM cross entropy ---------------------------------------------
 * Replaces OhemCELoss.forward, core/loss.py:13-20.  logits are NCHW
 * (layout = 0, what the reference trainer passes) or NHWC (layout = 1; with
 * layout = 1 | pitch << 8 the pixels are `pitch` >= c elements apart: the
 * classes are the first c channels of a wider, aligned buffer, and the
 * gradient of rtseg_ohem_bwd has the same geometry with zeroed padding),
 * f32 or bf16; labels int64 [n][h][w].
 *   thresh      = -log(config.ohem_thrs) as fp32 (core/loss.py:9)
 *   loss_px     fp32 [n*h*w] scratch: per-pixel CE (0 where ignored)
 *   workspace   rtseg_ohem_workspace_bytes() bytes of scratch, any content
 *   out         fp32 [4]: {loss, kth_value (top-k branch threshold or thresh),
 *                          n_selected, branch (0 = loss>thresh, 1 = top-k)}
 * The branch is decided on the device; no host synchronisation.            */
size_t rtseg_ohem_workspace_bytes(void);
int rtseg_ohem_fwd(int32_t n, int32_t c, int32_t h, int32_t w, int32_t layout, int32_t dtype,
                   const void *logits, const int64_t *labels, int32_t ignore_index, float thresh,
                   float *loss_px, void *workspace, float *out, rtseg_stream_t stream);
/* d loss / d logits = (softmax - onehot) * [pixel selected] * grad_out / n_selected,
 * written in the logits' layout and dtype.  `loss_px` and `workspace` are the
 * buffers rtseg_ohem_fwd filled for the same logits; grad_out is a device scalar. */
int rtseg_ohem_bwd(int32_t n, int32_t c, int32_t h, int32_t w, int32_t layout, int32_t dtype,
                   const void *logits, const int64_t *labels, int32_t ignore_index,
                   const float *loss_px, const void *workspace, const float *grad_out,
                   void *grad_logits, rtseg_stream_t stream);

/* OHEM forward without caller-provided scratch — NO COUNTERPART IN THE REFERENCE (which lets
 * PyTorch allocate whatever core/loss.py:14-20 needs).  It exists for callers that cannot
 * allocate device memory: a host embedding this library from another runtime (cgo / JNI / a
 * plugin sandbox) that owns only the logits and labels it was handed, or a loop that must not
 * call cudaMalloc (real-time inference, code between cudaStreamBeginCapture/EndCapture).
 * The per-pixel losses and the selection workspace live in a device array inside the library
 * (about 8.4 MB per GPU that loads it, held until the process exits), so:
 *   - n*h*w must not exceed rtseg_ohem_fwd_static_max_pixels() (2,097,152 — one 1024x2048
 *     image); larger inputs are refused with RTSEG_ERR_UNSUPPORTED and a message;
 *   - the call allocates nothing (no cudaMalloc / cudaFree / library allocator) and frees nothing;
 *   - the first call on each device looks the array's address up (cudaGetSymbolAddress) and
 *     caches it — make that first call outside any stream capture;
 *   - every call on a device shares the one array, which still holds the previous call's
 *     contents: the library re-zeroes the part that must start at zero, but two calls whose GPU
 *     work overlaps (different streams, or threads) race on it — serialise them.
 * Arguments and `out` are those of rtseg_ohem_fwd; there is no matching backward.              */
int64_t rtseg_ohem_fwd_static_max_pixels(void);
int rtseg_ohem_fwd_static(int32_t n, int32_t c, int32_t h, int32_t w, int32_t layout, int32_t dtype,
                          const void *logits, const int64_t *labels, int32_t ignore_index,
                          float thresh, float *out, rtseg_stream_t stream);

/* ---- layout plumbing --------------------------------------------------------
 * NCHW <-> NHWC with dtype conversion (the reference's tensors are NCHW fp32,
 * tools/test_speed.py:29; ours are NHWC bf16/fp32).                         */
int rtseg_nchw_to_nhwc(int32_t n, int32_t c, int32_t h, int32_t w, int32_t src_dtype,
                       const void *src, int32_t dst_dtype, void *dst, int32_t dst_ctotal,
                       int32_t dst_coff, rtseg_stream_t stream);
int rtseg_nhwc_to_nchw(int32_t n, int32_t c, int32_t h, int32_t w, int32_t src_dtype,
                       const void *src, int32_t src_ctotal, int32_t src_coff, int32_t dst_dtype,
                       void *dst, rtseg_stream_t stream);
/* AdaptiveAvgPool2d(p) of PyramidPoolingModule, models/modules.py:145-149, and of LEDNet's attention
 * pyramid, models/lednet.py:104.  NHWC in (channel slice allowed), dense NHWC [n][p][p][c] out in
 * y_dtype (an fp32 output lets a caller pool a large map in two exact stages). */
int rtseg_adaptive_avgpool(int32_t n, int32_t h, int32_t w, int32_t c, int32_t p, int32_t dtype,
                           const void *x, int32_t x_ctotal, int32_t x_coff, int32_t y_dtype, void *y,
                           rtseg_stream_t stream);

/* ---- depthwise 3x3 + pointwise 1x1 in one launch (SURVEY §7 step 6 / VERDICT r1 #8) ------------------
 * `DSConvBNAct` (models/modules.py:36-42), the depthwise + projection tail of `InvertedResidual`
 * (models/fastscnn.py:113-121) and FeatureFusionModule's low-resolution branch (:80-92): the depthwise output
 * (for InvertedResidual the 6x-expanded tensor) never reaches HBM — CUDA cores write it as the bf16 A tile of the
 * pointwise convolution's tcgen05.mma.  dw: the depthwise op (3x3, padding 1, multiplier 1, x slice, act none / relu /
 * relu6); pw: the 1x1 op on the depthwise op's output grid (y slice, residual slice and mode, act / act2).  Two shapes:
 * stride 1 with channels a multiple of 64 (<= 512) and 64 / 128 outputs (residual PRE / POST allowed); stride 2 on an
 * even-sized input with channels and outputs multiples of 16 (<= 64), no residual (Fast-SCNN's learning-to-downsample
 * pairs, models/fastscnn.py:38-40).  w_dw: fp32 [3][3][c]; w_pw_bf16: [cout][c] K-major.  bf16 NHWC only. */
int rtseg_dwpw_tc_supported(const rtseg_conv_desc *dw, const rtseg_conv_desc *pw);
int rtseg_dwpw_tc(const rtseg_conv_desc *dw, const void *x, const float *w_dw, const float *scale_dw, const float *shift_dw,
                  const rtseg_conv_desc *pw, const void *w_pw_bf16, const float *scale_pw, const float *shift_pw,
                  const void *res, void *y, rtseg_stream_t stream);

/* ---- f1: fused glue of the two-branch nets (config 4) ------------------------------------------
 * y = act( a (+|*) g(resize(b)) ): `b` ([n][hb][wb], channel slice allowed) is resized to a's (h, w) with
 * F.interpolate(mode='bilinear', align_corners=True) on the fly (hb = wb = 1: a broadcast), g = sigmoid when
 * b_sigmoid, op 0 = add / 1 = multiply, act a pointwise rtseg_act (no PReLU).  One pass instead of the reference's
 * interpolate / sigmoid / mul / add / relu kernels: models/bisenetv2.py:166-170, :190-201; models/ddrnet.py:165,
 * :238-243, :277-291.  y may alias a. */
int rtseg_combine_up(int32_t n, int32_t h, int32_t w, int32_t c, int32_t dtype, const void *a, int32_t a_ctotal,
                     int32_t a_coff, const void *b, int32_t hb, int32_t wb, int32_t b_ctotal, int32_t b_coff, int32_t op,
                     int32_t b_sigmoid, int32_t act, float act_p0, float act_p1, void *y, int32_t y_ctotal, int32_t y_coff,
                     rtseg_stream_t stream);
/* nn.AvgPool2d(k, s, p), count_include_pad=True, floor mode (models/ddrnet.py:266-275, models/bisenetv2.py:182). */
int rtseg_avgpool2d(int32_t n, int32_t h, int32_t w, int32_t c, int32_t k, int32_t s, int32_t p, int32_t dtype, const void *x,
                    int32_t x_ctotal, int32_t x_coff, void *y, int32_t y_ctotal, int32_t y_coff, rtseg_stream_t stream);

/* The four poolings of PyramidPoolingModule (models/modules.py:145-149: AdaptiveAvgPool2d(1), (2), (4), (6)) in one pass
 * over the map (h, w multiples of 4); y1 / y2 / y4 / y6: dense NHWC [n][p][p][c] in `dtype`. */
int rtseg_pyramid_avgpool(int32_t n, int32_t h, int32_t w, int32_t c, int32_t dtype, const void *x, int32_t x_ctotal,
                          int32_t x_coff, void *y1, void *y2, void *y4, void *y6, rtseg_stream_t stream);
/* The four branches of PyramidPoolingModule behind the pooling, reference
 * models/modules.py:145-156: conv1x1 (no BatchNorm, optional bias) of the 1x1 /
 * 2x2 / 4x4 / 6x6 pooled maps y1..y6 (bf16 NHWC, c channels, as written by
 * rtseg_pyramid_avgpool) and F.interpolate(..., (h, w), bilinear,
 * align_corners=True) of each result, written to channels [y_coff + b*hid,
 * +hid) of the NHWC concat buffer `cat` (torch.cat, modules.py:156).
 * weights: fp32 [4][hid][c] (rounded to bf16 in the kernel, like the tensor-core
 * convolutions it replaces), bias: fp32 [4][hid] or NULL. */
int rtseg_ppm_branches(int32_t n, int32_t h, int32_t w, int32_t c, int32_t hid, int32_t dtype, const void *y1,
                       const void *y2, const void *y4, const void *y6, const float *weights, const float *bias,
                       void *cat, int32_t y_ctotal, int32_t y_coff, rtseg_stream_t stream);

/* ---- training-mode pieces ---------------------------------------------
 * In training the reference's ConvBNAct is Conv2d -> BatchNorm2d(batch statistics) -> activation
 * as three autograd nodes (models/modules.py:73-85); NonBt1DBlock adds `+= x` before the last
 * BatchNorm (models/erfnet.py:76-82).  The convolution itself — forward and data gradient — is
 * rtseg_conv_* above (a data gradient is a convolution with transposed, flipped weights; the data
 * gradient of a 3x3 stride-2 convolution is rtseg_deconv3x3s2 and vice versa).  The entry points
 * below are the rest.  `res` (optional) is added to `y` before everything else (u = y + res).
 * All reductions land in caller-zeroed fp64 buffers of `c` elements.                           */
/* sum[c] += sum over pixels of u, sumsq[c] += sum of u^2 (sumsq may be NULL: bias gradients). */
int rtseg_bn_stats(int64_t npix, int32_t c, int32_t dtype, const void *y, int32_t y_ctotal, int32_t y_coff,
                   const void *res, int32_t r_ctotal, int32_t r_coff, double *sum, double *sumsq,
                   rtseg_stream_t stream);
/* z = act(scale[c] * u + shift[c]): BatchNorm with the batch statistics folded into scale/shift. */
int rtseg_affine_act(int64_t npix, int32_t c, int32_t dtype, const void *y, int32_t y_ctotal, int32_t y_coff,
                     const void *res, int32_t r_ctotal, int32_t r_coff, const float *scale, const float *shift,
                     int32_t act, float act_p0, float act_p1, const float *slope, void *z, int32_t z_ctotal,
                     int32_t z_coff, rtseg_stream_t stream);
/* g = dz * act'(scale*u + shift), xhat = (u - mean) * invstd: sum_g[c] += sum g, sum_gx[c] += sum g*xhat. */
int rtseg_bn_act_bwd_reduce(int64_t npix, int32_t c, int32_t dtype, const void *dz, int32_t d_ctotal, int32_t d_coff,
                            const void *y, int32_t y_ctotal, int32_t y_coff, const void *res, int32_t r_ctotal,
                            int32_t r_coff, const float *scale, const float *shift, const float *mean,
                            const float *invstd, int32_t act, float act_p0, float act_p1, const float *slope,
                            double *sum_g, double *sum_gx, rtseg_stream_t stream);
/* du = k[c] * (g - mean_g[c] - xhat * mean_gx[c]), k = gamma * invstd: the gradient wrt u (= wrt y and wrt res). */
int rtseg_bn_act_bwd_apply(int64_t npix, int32_t c, int32_t dtype, const void *dz, int32_t d_ctotal, int32_t d_coff,
                           const void *y, int32_t y_ctotal, int32_t y_coff, const void *res, int32_t r_ctotal,
                           int32_t r_coff, const float *scale, const float *shift, const float *mean,
                           const float *invstd, const float *k, const float *mean_g, const float *mean_gx,
                           int32_t act, float act_p0, float act_p1, const float *slope, void *du, int32_t u_ctotal,
                           int32_t u_coff, rtseg_stream_t stream);
/* The per-channel arithmetic between those passes, one launch each (nn.BatchNorm2d, torch/nn/modules/batchnorm.py
 * semantics as called by models/modules.py:81): stats = [sum | sumsq] from rtseg_bn_stats; the element count comes
 * from count_dev (a device double, e.g. all-reduced under SyncBatchNorm) or, when NULL, count_host.
 * out (5*c floats) = [mean | invstd | scale | shift | gamma*invstd]; running statistics (optional) are updated in
 * place with `momentum` and the unbiased variance. */
int rtseg_bn_finalize(int32_t c, const double *stats, const double *count_dev, double count_host, const float *gamma,
                      const float *beta, double eps, double momentum, float *running_mean, float *running_var,
                      float *out, rtseg_stream_t stream);
/* local_sums = this rank's [sum g | sum g*xhat] from rtseg_bn_act_bwd_reduce, global_sums = the same summed over
 * ranks (the same pointer without SyncBatchNorm); out (4*c floats) = [mean g | mean g*xhat | dgamma | dbeta]. */
int rtseg_bn_bwd_finalize(int32_t c, const double *local_sums, const double *global_sums, const double *count_dev,
                          double count_host, float *out, rtseg_stream_t stream);
/* Weight gradient of the dense convolution described by d (x: input slice, dy: gradient of the output slice
 * d->y_ctotal / d->y_coff): dw[kh][kw][cin][cout] (fp32, caller-zeroed) += sum over pixels. */
int rtseg_conv_wgrad(const rtseg_conv_desc *d, const void *x, const void *dy, float *dw, rtseg_stream_t stream);
/* The same gradient on tcgen05 for ERFNet's asymmetric convolutions (bf16, stride 1, 3x1 / 1x3, cin and cout in
 * {64, 128}): the pixel index is the GEMM's K, x and dy tiles are MN-major UMMA operands exactly as TMA lands
 * them, the accumulators stay in tensor memory for the whole kernel.  Same arguments and dw layout. */
int rtseg_conv_wgrad_tc_supported(const rtseg_conv_desc *d);
int rtseg_conv_wgrad_tc(const rtseg_conv_desc *d, const void *x, const void *dy, float *dw, rtseg_stream_t stream);
/* Row T, depthwise blocks (reference models/modules.py:46-59 `DWConvBNAct` under autograd; channel multiplier 1):
 * d describes the FORWARD convolution (n, h, w = its input; x_ctotal / x_coff = the slice of x and dx, y_ctotal /
 * y_coff = the slice of dy).  w and dw are fp32 [kh][kw][c]; dw is caller-zeroed and accumulated into. */
int rtseg_conv_dw_bwd_data(const rtseg_conv_desc *d, const void *dy, const float *w, void *dx, rtseg_stream_t stream);
int rtseg_conv_dw_wgrad(const rtseg_conv_desc *d, const void *x, const void *dy, float *dw, rtseg_stream_t stream);
/* Slope gradient of nn.PReLU (reference models/modules.py:119; enet.py:165 `self.act(x + x_right)`):
 * dslope[c] (fp64, caller-zeroed) += sum over pixels of dz * v * [v <= 0], v = scale[c] * (y + res) + shift[c];
 * scale / shift / res may be NULL (v = y). */
int rtseg_prelu_slope_grad(int64_t npix, int32_t c, int32_t dtype, const void *dz, int32_t d_ctotal, int32_t d_coff,
                           const void *y, int32_t y_ctotal, int32_t y_coff, const void *res, int32_t r_ctotal,
                           int32_t r_coff, const float *scale, const float *shift, double *dslope, rtseg_stream_t stream);
/* Adjoints of the resize / pooling steps inside Fast-SCNN's pyramid pooling, feature fusion and output head
 * (reference models/modules.py:145-156, fastscnn.py:30, 88) for the training step: dx of rtseg_upsample_bilinear
 * (a gather per input pixel with the forward's own index arithmetic: deterministic) and of rtseg_adaptive_avgpool
 * (dy: dense NHWC [n][p][p][c] in `dtype`). */
int rtseg_upsample_bilinear_bwd(int32_t n, int32_t hi, int32_t wi, int32_t c, int32_t ho, int32_t wo, int32_t align_corners,
                                int32_t dtype, const void *dy, int32_t g_ctotal, int32_t g_coff, void *dx, int32_t x_ctotal,
                                int32_t x_coff, rtseg_stream_t stream);
int rtseg_adaptive_avgpool_bwd(int32_t n, int32_t h, int32_t w, int32_t c, int32_t p, int32_t dtype, const void *dy, void *dx,
                               int32_t x_ctotal, int32_t x_coff, rtseg_stream_t stream);
/* Adjoint of rtseg_maxunpool2x2 (ENet's upsampling bottleneck, models/enet.py:176-178): dx [n][hp][wp][c] dense NHWC =
 * dy [n][ho][wo][c] at the position each one-byte index points to.  (The adjoint of rtseg_maxpool2x2_idx is
 * rtseg_maxunpool2x2 itself, applied to the gradient.) */
int rtseg_maxunpool2x2_bwd(int32_t n, int32_t hp, int32_t wp, int32_t c, int32_t ho, int32_t wo, int32_t dtype, const void *dy,
                           const uint8_t *idx, void *dx, rtseg_stream_t stream);
/* ---- f3: optimizer step + weight mirror in one launch ----------------------
 * Replaces torch.optim.SGD.step() as the reference configures it (utils/optimizer.py:8-12: lr, momentum, weight
 * decay; dampening 0, no Nesterov) followed by ModelEmaV2.update (utils/model_ema.py:28-40:
 * e = decay * e + (1 - decay) * m for EVERY state-dict entry, or e = m when EMA is off), called once per iteration
 * at core/seg_trainer.py:104-108.  The caller describes all tensors once in a device-resident chunk table (one
 * entry per <= rtseg_opt_chunk_elems() consecutive elements; build it with the pointers of the parameters, their
 * gradients, momentum buffers and mirror tensors) and then launches one kernel per iteration.
 *   p   parameter or buffer (fp32; int64 with RTSEG_OPT_INT64: num_batches_tracked)
 *   g   gradient or NULL (buffers: mirrored only)        m   momentum buffer (zero before the first step) or NULL
 *   e   mirror (same dtype as p) or NULL
 *   hyper_dev  4 floats on the device: lr, momentum, weight_decay, decay — rewritten by the host between steps
 *              (a captured CUDA graph sees the scheduler's new learning rate)                                   */
#define RTSEG_OPT_INT64 1
typedef struct rtseg_opt_chunk {
  float *p;
  const float *g;
  float *m;
  void *e;
  int32_t n;
  int32_t flags;
} rtseg_opt_chunk;
int32_t rtseg_opt_chunk_elems(void);
int rtseg_sgd_ema_step(const rtseg_opt_chunk *chunks_dev, int32_t nchunks, const float *hyper_dev, int32_t use_ema,
                       rtseg_stream_t stream);

/* Gradient of MaxPool2d(3, 2, 1) (InitialBlock's pool branch, models/enet.py:44): dx = dy routed to the first
 * maximum of each window. */
int rtseg_maxpool3x3s2_bwd(int32_t n, int32_t h, int32_t w, int32_t c, int32_t dtype, const void *x, int32_t x_ctotal,
                           int32_t x_coff, const void *dy, int32_t g_ctotal, int32_t g_coff, void *dx, int32_t d_ctotal,
                           int32_t d_coff, rtseg_stream_t stream);

/* F.avg_pool2d(x, 3, 2, 1) of DABNet's input pyramid, models/dabnet.py:32-34 (padding counted: /9).
 * Input: the caller's NCHW tensor (x_is_nchw = 1) or an NHWC channel slice; output: NHWC channel slice. */
int rtseg_avgpool3x3s2(int32_t n, int32_t h, int32_t w, int32_t c, int32_t x_is_nchw, int32_t x_dtype,
                       const void *x, int32_t x_ctotal, int32_t x_coff, int32_t y_dtype, void *y,
                       int32_t y_ctotal, int32_t y_coff, rtseg_stream_t stream);
/* The first level of that pyramid from the uint8 HWC image (normalised like rtseg_stem_u8, datasets/cityscapes.py:122-123). */
int rtseg_avgpool3x3s2_u8(int32_t n, int32_t h, int32_t w, int32_t c, const uint8_t *x, const float *mean_host,
                          const float *std_host, int32_t y_dtype, void *y, int32_t y_ctotal, int32_t y_coff,
                          rtseg_stream_t stream);
/* channel_shuffle, models/modules.py:18-32 (used by SSnbtUnit, models/lednet.py:87): NHWC slices,
 * out[.., i*groups + g] = in[.., g*(c/groups) + i]; x and y must not alias.                      */
int rtseg_channel_shuffle(int64_t npix, int32_t c, int32_t groups, int32_t dtype, const void *x,
                          int32_t x_ctotal, int32_t x_coff, void *y, int32_t y_ctotal, int32_t y_coff,
                          rtseg_stream_t stream);

/* Predict epilogue, core/seg_trainer.py:172-191 + utils/utils.py:59-78: out[p] = lut[labels[p]] (uint8 RGB, HWC;
 * lut = 256 x 3 bytes, the colormap padded to 256 rows) and, when `image` (uint8 HWC, same size) is given,
 * Pillow's Image.blend(image, pred, alpha) = (uint8)(image + alpha * (pred - image)), 0 <= alpha <= 1. */
int rtseg_colorize_blend(int64_t npix, const uint8_t *labels, const uint8_t *lut, const uint8_t *image,
                         float alpha, uint8_t *out, rtseg_stream_t stream);

/* ---- host-buffer convenience (pinned or pageable HOST pointers) ------------
 * One call = H2D copy of the low-resolution logits, the fused head kernel,
 * D2H copy of the uint8 class map, all on `stream`, then a stream
 * synchronise.  `dev_scratch` must hold logits bytes + n*ho*wo bytes
 * (rtseg_upsample_argmax_host_scratch_bytes returns exactly that).          */
size_t rtseg_upsample_argmax_host_scratch_bytes(int32_t n, int32_t hi, int32_t wi, int32_t c, int32_t ho, int32_t wo,
                                                int32_t dtype);
int rtseg_upsample_argmax_host(int32_t n, int32_t hi, int32_t wi, int32_t c, int32_t ho, int32_t wo,
                               int32_t align_corners, int32_t dtype, const void *x_host,
                               uint8_t *labels_host, void *dev_scratch, rtseg_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* RTSEG_H_ */
