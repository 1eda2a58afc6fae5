"""Tensor-level host API over the C ABI (include/rtseg.h).

PyTorch is used here for device memory, streams and views only; every FLOP
goes through librtseg.so.  Tensors are NCHW-*logical* torch tensors stored
channels_last (dense NHWC), which is what the kernels read and write; a
channel slice of a wider NHWC buffer is passed as (buffer, c_offset, channels).
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import ACT_CODES, BF16, F32, RES_NONE, RES_POST, RES_PRE, ConvDesc

_DT = {torch.float32: F32, torch.bfloat16: BF16}


def _dtype_code(t: torch.dtype) -> int:
    try:
        return _DT[t]
    except KeyError:
        raise TypeError(f"rtseg kernels support float32 and bfloat16 tensors, got {t}") from None


def _require_cuda(x: torch.Tensor, what: str) -> None:
    if not x.is_cuda:
        raise RuntimeError(
            f"{what}: input is on {x.device}; the rtseg path has no CPU implementation "
            "(the CPU restatement lives in oracle/ and is test infrastructure only)")
    _lib.sync_device(x.device.index)


def _no_tape(x: torch.Tensor, what: str) -> None:
    """Raw pointer launches are invisible to autograd: refuse inputs that are on the tape instead of silently
    cutting the graph (the training path wraps its kernels in autograd Functions, train.py)."""
    if torch.is_grad_enabled() and x.requires_grad:
        raise NotImplementedError(f"{what} has no backward kernel: it cannot run on a tensor that requires grad "
                                  "(call under torch.no_grad() / model.eval(), or use the modules train.py covers)")


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def is_nhwc(x: torch.Tensor) -> bool:
    return x.dim() == 4 and x.is_contiguous(memory_format=torch.channels_last)


def empty_nhwc(n: int, c: int, h: int, w: int, dtype, device) -> torch.Tensor:
    return torch.empty((n, c, h, w), dtype=dtype, device=device, memory_format=torch.channels_last)


def to_nhwc(x: torch.Tensor, dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """Dense-NHWC copy of an NCHW tensor (our transpose kernel), or x itself if already NHWC."""
    _require_cuda(x, "to_nhwc")
    if x.dtype == torch.uint8:   # a decoded image outside the stem kernels (training, unusual stems)
        return normalize_u8(x, dtype or torch.float32)
    dtype = dtype or x.dtype
    if is_nhwc(x) and x.dtype == dtype:
        return x
    if torch.is_grad_enabled() and x.requires_grad:
        # on the autograd tape (training, a leaf in NCHW): torch's own differentiable layout / dtype change —
        # the raw transpose kernel below would silently cut the graph
        return x.to(dtype).contiguous(memory_format=torch.channels_last)
    n, c, h, w = x.shape
    if is_nhwc(x):  # dtype change only: treat as [n*h*w, c] -> same layout
        src = x.permute(0, 2, 3, 1).reshape(1, 1, n * h * w, c)  # view, "NCHW" with c'=1
        out = empty_nhwc(n, c, h, w, dtype, x.device)
        # a [rows=1][cols] transpose is a dtype-converting copy
        _lib.check(_lib.load().rtseg_nchw_to_nhwc(1, 1, 1, n * h * w * c, _dtype_code(x.dtype), src.data_ptr(),
                                                  _dtype_code(dtype), out.data_ptr(), 1, 0, _stream()), "convert")
        return out
    if not x.is_contiguous():
        x = x.contiguous()
    out = empty_nhwc(n, c, h, w, dtype, x.device)
    _lib.check(_lib.load().rtseg_nchw_to_nhwc(n, c, h, w, _dtype_code(x.dtype), x.data_ptr(), _dtype_code(dtype),
                                              out.data_ptr(), c, 0, _stream()), "rtseg_nchw_to_nhwc")
    return out


def to_nchw(x: torch.Tensor, dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """NCHW-contiguous copy of an NHWC tensor (what the reference's callers index into)."""
    _require_cuda(x, "to_nchw")
    dtype = dtype or x.dtype
    n, c, h, w = x.shape
    if not is_nhwc(x):
        return x.contiguous().to(dtype)
    out = torch.empty((n, c, h, w), dtype=dtype, device=x.device)
    _lib.check(_lib.load().rtseg_nhwc_to_nchw(n, c, h, w, _dtype_code(x.dtype), x.data_ptr(), c, 0,
                                              _dtype_code(dtype), out.data_ptr(), _stream()), "rtseg_nhwc_to_nchw")
    return out


# --------------------------------------------------------------------- folding ----
@dataclass
class Folded:
    """Device-resident, kernel-ready parameters of one conv(+BN)(+act) layer."""

    kind: str  # 'dense' | 'dw' | 'deconv'
    cin: int
    cout: int
    kh: int
    kw: int
    stride: int
    dil: int
    pad_h: int
    pad_w: int
    w_f32: torch.Tensor            # dense/deconv [kh][kw][cin][cout]; dw [kh][kw][cout]
    w_tc: Optional[torch.Tensor]   # bf16 [kh*kw][cout_pad][cin] or None
    scale: Optional[torch.Tensor]
    shift: Optional[torch.Tensor]
    slope: Optional[torch.Tensor]
    act: int = 0
    p0: float = 0.0
    p1: float = 0.0
    w_tc_scaled: Optional[torch.Tensor] = None   # bf16(scale[co] * w): the pair kernel's BatchNorm-in-the-MMA form

    @property
    def groups(self) -> int:
        return self.cin if self.kind == "dw" else 1


def act_spec(module) -> Tuple[int, float, float, Optional[torch.Tensor], Optional[torch.nn.Module]]:
    """(code, p0, p1, prelu_weight, post_module) for an nn activation module.

    `post_module` is returned for the two non-pointwise hub members (GLU, Softmax),
    which cannot live in a conv epilogue; the kernel then runs with act = none.
    Mirrors the name table of reference models/modules.py:114-123.
    """
    nn = torch.nn
    m = module
    if m is None or isinstance(m, nn.Identity):
        return ACT_CODES["none"], 0.0, 0.0, None, None
    if isinstance(m, nn.ReLU6):
        return ACT_CODES["relu6"], 0.0, 0.0, None, None
    if isinstance(m, nn.ReLU):
        return ACT_CODES["relu"], 0.0, 0.0, None, None
    if isinstance(m, nn.LeakyReLU):
        return ACT_CODES["leakyrelu"], float(m.negative_slope), 0.0, None, None
    if isinstance(m, nn.PReLU):
        return ACT_CODES["prelu"], 0.0, 0.0, m.weight, None
    if isinstance(m, nn.ELU):
        return ACT_CODES["elu"], float(m.alpha), 0.0, None, None
    if isinstance(m, nn.CELU):
        return ACT_CODES["celu"], float(m.alpha), 0.0, None, None
    if isinstance(m, nn.SELU):
        return ACT_CODES["selu"], 0.0, 0.0, None, None
    if isinstance(m, nn.Hardswish):
        return ACT_CODES["hardswish"], 0.0, 0.0, None, None
    if isinstance(m, nn.Hardtanh):
        return ACT_CODES["hardtanh"], float(m.min_val), float(m.max_val), None, None
    if isinstance(m, nn.GELU):
        if getattr(m, "approximate", "none") != "none":
            raise NotImplementedError("GELU(approximate='tanh') has no fused epilogue")
        return ACT_CODES["gelu"], 0.0, 0.0, None, None
    if isinstance(m, nn.SiLU):
        return ACT_CODES["silu"], 0.0, 0.0, None, None
    if isinstance(m, nn.Sigmoid):
        return ACT_CODES["sigmoid"], 0.0, 0.0, None, None
    if isinstance(m, nn.Tanh):
        return ACT_CODES["tanh"], 0.0, 0.0, None, None
    if isinstance(m, (nn.GLU, nn.Softmax)):
        return ACT_CODES["none"], 0.0, 0.0, None, m
    raise NotImplementedError(f"no fused epilogue for activation {type(m).__name__}")


def fold_affine(cout: int, bias: Optional[torch.Tensor], bn: Optional[torch.nn.BatchNorm2d], host: bool = False):
    """Eval-mode BatchNorm (+ conv bias) as per-channel fp32 (scale, shift), folded in fp64."""
    # host=True (inference fold caches): a few hundred bytes per layer, folded on the CPU so that no device kernels are
    # spent on parameter preparation.  host=False (training, inside a captured step): stays on the device, no sync.
    f64 = (lambda t: t.detach().to("cpu", torch.float64)) if host else (lambda t: t.detach().double())
    if bn is None:
        if bias is None:
            return None, None
        return None, f64(bias)
    if bn.running_mean is None or bn.running_var is None:
        raise NotImplementedError("BatchNorm2d(track_running_stats=False) cannot be folded")
    mean, var = f64(bn.running_mean), f64(bn.running_var)
    g = f64(bn.weight) if bn.weight is not None else torch.ones_like(mean)
    b = f64(bn.bias) if bn.bias is not None else torch.zeros_like(g)
    scale = g / torch.sqrt(var + bn.eps)
    shift = b - mean * scale
    if bias is not None:
        shift = shift + f64(bias) * scale
    return scale, shift


def _f32(t: Optional[torch.Tensor], device) -> Optional[torch.Tensor]:
    return None if t is None else t.to(device=device, dtype=torch.float32).contiguous()


def fold_conv(conv: torch.nn.Module, bn=None, act_module=None, device=None,
              in_slice: Optional[Tuple[int, int]] = None, tc_small_cin: bool = False, host: bool = False
              ) -> Tuple[Folded, Optional[torch.nn.Module]]:
    """Pack nn.Conv2d / nn.ConvTranspose2d (+BN, +activation) for the kernels.

    in_slice=(lo, hi) folds only input channels [lo, hi) of a dense conv (a caller that splits K across two
    launches); tc_small_cin also builds tensor-core weights when that leaves <= 4 input channels."""
    nn = torch.nn
    device = device or conv.weight.device
    code, p0, p1, prelu_w, post = act_spec(act_module)
    # host=True: packed on the CPU and uploaded once per layout (inference caches); else packed on the device
    w = conv.weight.detach().to(device="cpu" if host else device, dtype=torch.float32)
    kh, kw = conv.kernel_size
    sh, sw = conv.stride
    dh, dw_ = conv.dilation
    ph, pw = conv.padding if not isinstance(conv.padding, str) else (None, None)
    if ph is None:
        raise NotImplementedError("string padding modes are not supported")
    if sh != sw or dh != dw_:
        raise NotImplementedError("anisotropic stride/dilation is not supported")
    if getattr(conv, "padding_mode", "zeros") != "zeros":
        raise NotImplementedError("only zero padding is supported")
    if getattr(conv, "transposed", False):   # nn.ConvTranspose2d (or train._Spec around a bare weight tensor)
        cin, cout = conv.in_channels, conv.out_channels
        if not (kh == 3 and kw == 3 and sh == 2 and ph == 1 and pw == 1 and tuple(conv.output_padding) == (1, 1)
                and conv.groups == 1 and dh == 1):
            raise NotImplementedError("only ConvTranspose2d(k=3, s=2, p=1, output_padding=1) has a kernel "
                                      "(DeConvBNAct with scale_factor=2, reference models/modules.py:92-102)")
        kind = "deconv"
        w_f32 = w.permute(2, 3, 0, 1).contiguous()          # [kh][kw][cin][cout]
        w_kmaj = w.permute(2, 3, 1, 0)                        # [kh][kw][cout][cin]
    else:
        cin, cout = conv.in_channels, conv.out_channels
        if in_slice is not None:
            if conv.groups != 1:
                raise NotImplementedError("in_slice is for dense convolutions")
            w = w[:, in_slice[0]:in_slice[1]]
            cin = w.shape[1]
        if conv.groups == 1:
            kind = "dense"
            w_f32 = w.permute(2, 3, 1, 0).contiguous()       # [kh][kw][cin][cout]
            w_kmaj = w.permute(2, 3, 0, 1)                    # [kh][kw][cout][cin]
        elif conv.groups == cin and cout % cin == 0:
            kind = "dw"
            w_f32 = w[:, 0].permute(1, 2, 0).contiguous()    # [kh][kw][cout]
            w_kmaj = None
        else:
            raise NotImplementedError(f"grouped convolution groups={conv.groups} has no kernel (only 1 or cin)")
    w_tc = None
    if w_kmaj is not None and (cin % 16 == 0 or (kind == "dense" and (cin > 4 or tc_small_cin))):
        # K (and N) zero-padded to multiples of 16; the conv kernel's input tensor map stops at the last real channel
        cpad, kpad = (cout + 15) // 16 * 16, (cin + 15) // 16 * 16
        w_tc = torch.zeros((kh * kw, cpad, kpad), dtype=torch.bfloat16, device=w.device)
        w_tc[:, :cout, :cin] = w_kmaj.reshape(kh * kw, cout, cin).to(torch.bfloat16)
        w_tc = w_tc.to(device)
    scale, shift = fold_affine(cout, conv.bias, bn, host)
    w_tc_scaled = None
    if w_tc is not None and scale is not None and kind == "dense" and host and cin == cout and cin in (16, 64) \
            and (kh, kw) in ((3, 1), (1, 3)):
        # pair kernel (pair_tc.cu): the BatchNorm scale rides in the weights, rounded to bf16 once after scaling
        ws = torch.zeros((kh * kw, cpad, kpad), dtype=torch.bfloat16)
        ws[:, :cout, :cin] = (w_kmaj.reshape(kh * kw, cout, cin).double() * scale.view(1, cout, 1)).to(torch.bfloat16)
        w_tc_scaled = ws.to(device)
    slope = None
    if prelu_w is not None:
        slope = prelu_w.detach().to(w.device, torch.float32).reshape(-1)
        slope = slope.expand(cout) if slope.numel() == 1 else slope
        if slope.numel() != cout:
            raise ValueError("PReLU num_parameters must be 1 or the number of output channels")
    f = Folded(kind, cin, cout, kh, kw, sh, dh, ph, pw, w_f32.to(device), w_tc, _f32(scale, device), _f32(shift, device),
               _f32(slope, device), code, p0, p1, w_tc_scaled)
    return f, post


# ------------------------------------------------------------------- conv calls ----
def _desc(f: Folded, n, h, w, dtype_code, x_ct, x_co, y_ct, y_co, res_mode=RES_NONE, r_ct=0, r_co=0, act2=0,
          act: Optional[int] = None) -> ConvDesc:
    return ConvDesc(n, h, w, f.cin, f.cout, f.kh, f.kw, f.stride, f.dil, f.pad_h, f.pad_w, f.groups, dtype_code,
                    f.act if act is None else act, f.p0, f.p1, res_mode, act2, x_ct, x_co, y_ct, y_co, r_ct, r_co)


def out_hw(f: Folded, h: int, w: int) -> Tuple[int, int]:
    if f.kind == "deconv":
        return 2 * h, 2 * w
    ho = (h + 2 * f.pad_h - f.dil * (f.kh - 1) - 1) // f.stride + 1
    wo = (w + 2 * f.pad_w - f.dil * (f.kw - 1) - 1) // f.stride + 1
    return ho, wo


USE_TENSOR_CORES = True  # module-level switch used by tests to force the CUDA-core path


def conv(f: Folded, x: torch.Tensor, *, res: Optional[torch.Tensor] = None, res_mode: int = RES_NONE,
         out: Optional[torch.Tensor] = None, x_coff: int = 0, y_coff: int = 0, r_coff: int = 0, act2: int = 0,
         labels: Optional[torch.Tensor] = None, want_logits: bool = True, allow_tc: Optional[bool] = None):
    """One fused conv/deconv (+BN +act +residual) layer.

    x, res, out are NHWC buffers; the op reads channels [x_coff, x_coff+cin) of x and
    writes channels [y_coff, y_coff+cout) of out (allocated dense if None).
    For kind == 'deconv', `labels` (uint8 [n, 2h, 2w]) requests the fused argmax.
    """
    _require_cuda(x, "conv")
    if not is_nhwc(x):
        raise ValueError("conv expects a dense NHWC (channels_last) tensor; use ops.to_nhwc")
    lib = _lib.load()
    n, x_ct, h, w = x.shape
    dt = _dtype_code(x.dtype)
    ho, wo = out_hw(f, h, w)
    allow_tc = USE_TENSOR_CORES if allow_tc is None else allow_tc
    if out is None and want_logits:
        out = empty_nhwc(n, f.cout, ho, wo, x.dtype, x.device)
    y_ct = out.shape[1] if out is not None else f.cout
    if out is not None:
        if not is_nhwc(out) or out.dtype != x.dtype or tuple(out.shape[2:]) != (ho, wo) or out.shape[0] != n:
            raise ValueError("bad output buffer")
    r_ct = 0
    if res_mode != RES_NONE:
        if res is None or not is_nhwc(res) or res.dtype != x.dtype or tuple(res.shape[2:]) != (ho, wo):
            raise ValueError("bad residual tensor")
        r_ct = res.shape[1]
    d = _desc(f, n, h, w, dt, x_ct, x_coff, y_ct, y_coff, res_mode, r_ct, r_coff, act2)
    st = _stream()
    args = (_ptr(f.scale), _ptr(f.shift), _ptr(f.slope))
    if f.kind == "dw":
        _lib.check(lib.rtseg_conv_dw(C.byref(d), x.data_ptr(), f.w_f32.data_ptr(), *args, _ptr(res), out.data_ptr(), st),
                   "rtseg_conv_dw")
    elif f.kind == "dense":
        if allow_tc and dt == BF16 and f.w_tc is not None and lib.rtseg_conv_tc_supported(C.byref(d)):
            _lib.check(lib.rtseg_conv_tc(C.byref(d), x.data_ptr(), f.w_tc.data_ptr(), *args, _ptr(res), out.data_ptr(), st),
                       "rtseg_conv_tc")
        else:
            _lib.check(lib.rtseg_conv_dense(C.byref(d), x.data_ptr(), f.w_f32.data_ptr(), *args, _ptr(res),
                                            out.data_ptr(), st), "rtseg_conv_dense")
    elif f.kind == "deconv":
        if allow_tc and dt == BF16 and f.w_tc is not None and lib.rtseg_deconv3x3s2_tc_supported(C.byref(d)):
            _lib.check(lib.rtseg_deconv3x3s2_tc(C.byref(d), x.data_ptr(), f.w_tc.data_ptr(), *args, _ptr(out),
                                                _ptr(labels), st), "rtseg_deconv3x3s2_tc")
        else:
            if out is None:  # the CUDA-core deconv takes its argmax from materialised logits
                out = empty_nhwc(n, f.cout, ho, wo, x.dtype, x.device)
                d.y_ctotal = f.cout
            _lib.check(lib.rtseg_deconv3x3s2(C.byref(d), x.data_ptr(), f.w_f32.data_ptr(), *args, out.data_ptr(),
                                             _ptr(labels), st), "rtseg_deconv3x3s2")
    else:
        raise ValueError(f.kind)
    return out


# Normalisation applied to uint8 images inside the stem kernels: the constants of reference
# datasets/cityscapes.py:122-123 (AT.Normalize).  set_image_norm() changes them for other datasets.
IMAGE_MEAN = (0.485, 0.456, 0.406)
IMAGE_STD = (0.229, 0.224, 0.225)


def set_image_norm(mean, std) -> None:
    global IMAGE_MEAN, IMAGE_STD
    mean, std = tuple(float(v) for v in mean), tuple(float(v) for v in std)
    if len(mean) != len(std) or not 1 <= len(mean) <= 4 or min(std) <= 0:
        raise ValueError("mean / std: 1 to 4 channels, std > 0")
    IMAGE_MEAN, IMAGE_STD = mean, std


def _norm_args(cin: int):
    if cin > len(IMAGE_MEAN):
        raise ValueError(f"a {cin}-channel uint8 image needs {cin} mean / std values (ops.set_image_norm)")
    arr = C.c_float * cin
    return arr(*IMAGE_MEAN[:cin]), arr(*IMAGE_STD[:cin])


def as_image_u8(x: torch.Tensor) -> torch.Tensor:
    """A decoded uint8 image batch as the NCHW-logical, HWC-stored tensor the stems read.

    Accepts [N, H, W, C] (what a decoder produces) or an [N, C, H, W] tensor that is already channels_last."""
    if x.dtype != torch.uint8 or x.dim() != 4:
        raise TypeError("expected a 4-d uint8 image batch")
    if x.shape[1] <= 4 and is_nhwc(x):
        return x
    if x.shape[3] <= 4 and x.is_contiguous():
        return x.permute(0, 3, 1, 2)
    raise ValueError("uint8 images must be stored HWC: [N, H, W, C] contiguous or [N, C, H, W] channels_last, C <= 4")


def image_view(x: torch.Tensor) -> torch.Tensor:
    """Entry point of every net / stem: a uint8 batch becomes its NCHW-logical HWC-stored view, anything else passes."""
    return as_image_u8(x) if x.dtype == torch.uint8 else x


def normalize_u8(x: torch.Tensor, dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """The normalised float image the reference's data pipeline would hand to the model (NHWC storage): only for
    callers that need the tensor itself; the stems normalise while they read and never materialise it."""
    _require_cuda(x, "normalize_u8")
    x = as_image_u8(x)
    n, c, h, w = x.shape
    mean = torch.tensor(IMAGE_MEAN[:c], dtype=torch.float32, device=x.device) * 255.0
    rden = torch.reciprocal(torch.tensor(IMAGE_STD[:c], dtype=torch.float32, device=x.device) * 255.0)
    return ((x.float() - mean.view(1, c, 1, 1)) * rden.view(1, c, 1, 1)).to(dtype).contiguous(memory_format=torch.channels_last)


USE_PAIR_FUSION = True   # module-level switch (tests / A-B runs): fused 3x1 -> 1x3 pairs of NonBt1D blocks
PAIR_FOLD_BN = True      # ... with the BatchNorm scale in the weights and the shift added by a bias MMA


def conv_pair_supported(fa: Folded, fb: Folded, x: torch.Tensor, res: Optional[torch.Tensor] = None) -> bool:
    """Can `conv(fb, conv(fa, x)) (+ res before fb's BatchNorm)` run as one rtseg_conv_pair_tc launch?"""
    if not (USE_PAIR_FUSION and USE_TENSOR_CORES and x.is_cuda and x.dtype == torch.bfloat16 and is_nhwc(x)):
        return False
    if fa.kind != "dense" or fb.kind != "dense" or fa.w_tc is None or fb.w_tc is None or fa.slope is not None or fb.slope is not None:
        return False
    if (fa.kh, fa.kw, fa.pad_h, fa.pad_w, fa.stride, fa.dil) != (3, 1, 1, 0, 1, 1) or \
            (fb.kh, fb.kw, fb.pad_h, fb.pad_w, fb.stride, fb.dil) != (1, 3, 0, 1, 1, 1):
        return False
    if not (fa.cin == fa.cout == fb.cin == fb.cout == x.shape[1]):
        return False
    n, ct, h, w = x.shape
    d = _desc(fa, n, h, w, BF16, ct, 0, fb.cout, 0, RES_PRE if res is not None else RES_NONE, fb.cout if res is not None else 0, 0,
              act=fb.act)
    d.act_p0, d.act_p1 = fb.p0, fb.p1
    return bool(_lib.load().rtseg_conv_pair_tc_supported(C.byref(d), fa.act))


def conv_pair(fa: Folded, fb: Folded, x: torch.Tensor, res: Optional[torch.Tensor] = None,
              out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """act_b(bn_b(conv1x3_b(act_a(bn_a(conv3x1_a(x)))) [+ res])) in one launch (reference models/erfnet.py:65-71, 77-81)."""
    _require_cuda(x, "conv_pair")
    n, ct, h, w = x.shape
    if out is None:
        out = empty_nhwc(n, fb.cout, h, w, x.dtype, x.device)
    r_ct = 0
    if res is not None:
        if not is_nhwc(res) or res.dtype != x.dtype or tuple(res.shape) != (n, res.shape[1], h, w):
            raise ValueError("bad residual tensor")
        r_ct = res.shape[1]
    d = _desc(fa, n, h, w, BF16, ct, 0, out.shape[1], 0, RES_PRE if res is not None else RES_NONE, r_ct, 0, act=fb.act)
    d.act_p0, d.act_p1 = fb.p0, fb.p1
    wa, wb, flags = fa.w_tc, fb.w_tc, 0
    if PAIR_FOLD_BN and fa.w_tc_scaled is not None:
        wa, flags = fa.w_tc_scaled, flags | 1
    if PAIR_FOLD_BN and fb.w_tc_scaled is not None:
        wb, flags = fb.w_tc_scaled, flags | 2
    _lib.check(_lib.load().rtseg_conv_pair_tc(C.byref(d), x.data_ptr(), wa.data_ptr(), _ptr(fa.scale), _ptr(fa.shift),
                                              fa.act, wb.data_ptr(), _ptr(fb.scale), _ptr(fb.shift), flags, _ptr(res),
                                              out.data_ptr(), _stream()), "rtseg_conv_pair_tc")
    return out


USE_DWPW_FUSION = True   # module-level switch (tests / A-B runs): depthwise 3x3 + pointwise 1x1 in one launch


def _dwpw_descs(fdw: Folded, fpw: Folded, x: torch.Tensor, x_coff: int, y_ct: int, y_coff: int, res_mode: int, r_ct: int,
                r_coff: int, act2: int):
    n, x_ct, h, w = x.shape
    ho, wo = out_hw(fdw, h, w)
    ddw = _desc(fdw, n, h, w, BF16, x_ct, x_coff, fdw.cout, 0)
    dpw = _desc(fpw, n, ho, wo, BF16, fdw.cout, 0, y_ct, y_coff, res_mode, r_ct, r_coff, act2)
    return ddw, dpw


def dwpw_supported(fdw: Folded, fpw: Folded, x: torch.Tensor, res_mode: int = RES_NONE, res: Optional[torch.Tensor] = None,
                   out: Optional[torch.Tensor] = None, x_coff: int = 0, y_coff: int = 0, r_coff: int = 0, act2: int = 0) -> bool:
    """Can `conv(fpw, conv(fdw, x))` (depthwise 3x3 -> pointwise 1x1, + residual) run as one rtseg_dwpw_tc launch?"""
    if not (USE_DWPW_FUSION and USE_TENSOR_CORES and x.is_cuda and x.dtype == torch.bfloat16 and is_nhwc(x)):
        return False
    if fdw.kind != "dw" or fpw.kind != "dense" or fpw.w_tc is None or fdw.slope is not None or fpw.slope is not None:
        return False
    if fpw.w_tc.shape[1] != fpw.cout or fpw.w_tc.shape[2] != fpw.cin:     # padded weights: not this kernel's layout
        return False
    y_ct = out.shape[1] if out is not None else fpw.cout
    r_ct = res.shape[1] if res is not None else 0
    ddw, dpw = _dwpw_descs(fdw, fpw, x, x_coff, y_ct, y_coff, res_mode, r_ct, r_coff, act2)
    return bool(_lib.load().rtseg_dwpw_tc_supported(C.byref(ddw), C.byref(dpw)))


def dwpw(fdw: Folded, fpw: Folded, x: torch.Tensor, res: Optional[torch.Tensor] = None, res_mode: int = RES_NONE,
         out: Optional[torch.Tensor] = None, x_coff: int = 0, y_coff: int = 0, r_coff: int = 0, act2: int = 0) -> torch.Tensor:
    """DWConvBNAct(3x3) -> PWConvBNAct / conv1x1 (+ BN, + residual) in one launch (reference models/modules.py:36-42,
    models/fastscnn.py:80-92, 113-121): the depthwise output stays in shared memory as the pointwise MMA's A tile."""
    _require_cuda(x, "dwpw")
    _no_tape(x, "dwpw")
    n = x.shape[0]
    h, w = out_hw(fdw, x.shape[2], x.shape[3])      # the pointwise op's grid (stride-2 depthwise: half the input)
    if out is None:
        out = empty_nhwc(n, fpw.cout, h, w, x.dtype, x.device)
    elif not is_nhwc(out) or out.dtype != x.dtype or tuple(out.shape[2:]) != (h, w) or out.shape[0] != n:
        raise ValueError("bad output buffer")
    if res_mode != RES_NONE and (res is None or not is_nhwc(res) or res.dtype != x.dtype or tuple(res.shape[2:]) != (h, w)):
        raise ValueError("bad residual tensor")
    ddw, dpw = _dwpw_descs(fdw, fpw, x, x_coff, out.shape[1], y_coff, res_mode, res.shape[1] if res is not None else 0,
                           r_coff, act2)
    _lib.check(_lib.load().rtseg_dwpw_tc(C.byref(ddw), x.data_ptr(), fdw.w_f32.data_ptr(), _ptr(fdw.scale), _ptr(fdw.shift),
                                         C.byref(dpw), fpw.w_tc.data_ptr(), _ptr(fpw.scale), _ptr(fpw.shift), _ptr(res),
                                         out.data_ptr(), _stream()), "rtseg_dwpw_tc")
    return out


def stem(f: Folded, x: torch.Tensor, pool: bool, out_dtype: torch.dtype, out: Optional[torch.Tensor] = None,
         y_coff: int = 0) -> torch.Tensor:
    """InitialBlock / stride-2 stem straight from the caller's NCHW (or NHWC) tensor, or from the uint8 HWC image
    (normalisation of reference datasets/cityscapes.py:122-123 applied while the window is read)."""
    _require_cuda(x, "stem")
    if x.dtype == torch.uint8:
        x = as_image_u8(x)
    n, cin, h, w = x.shape
    if x.dtype == torch.uint8:
        ho, wo = (h - 1) // 2 + 1, (w - 1) // 2 + 1
        if out is None:
            out = empty_nhwc(n, f.cout + (cin if pool else 0), ho, wo, out_dtype, x.device)
        d = _desc(f, n, h, w, _dtype_code(out_dtype), cin, 0, out.shape[1], y_coff)
        mean, std = _norm_args(cin)
        _lib.check(_lib.load().rtseg_stem_u8(C.byref(d), 1 if pool else 0, x.data_ptr(), mean, std, f.w_f32.data_ptr(),
                                             _ptr(f.scale), _ptr(f.shift), _ptr(f.slope), out.data_ptr(), _stream()),
                   "rtseg_stem_u8")
        return out
    if x.is_contiguous():
        nchw = 1
    elif is_nhwc(x):
        nchw = 0
    else:
        x, nchw = x.contiguous(), 1
    ho, wo = (h - 1) // 2 + 1, (w - 1) // 2 + 1
    ctot = f.cout + (cin if pool else 0)
    if out is None:
        out = empty_nhwc(n, ctot, ho, wo, out_dtype, x.device)
    d = _desc(f, n, h, w, _dtype_code(out_dtype), cin, 0, out.shape[1], y_coff)
    _lib.check(_lib.load().rtseg_stem(C.byref(d), nchw, _dtype_code(x.dtype), 1 if pool else 0, x.data_ptr(),
                                      f.w_f32.data_ptr(), _ptr(f.scale), _ptr(f.shift), _ptr(f.slope), out.data_ptr(),
                                      _stream()), "rtseg_stem")
    return out


def maxpool3x3s2(x: torch.Tensor, out: torch.Tensor, y_coff: int, x_coff: int = 0, c: Optional[int] = None) -> torch.Tensor:
    _require_cuda(x, "maxpool3x3s2")
    n, x_ct, h, w = x.shape
    c = c if c is not None else x_ct
    d = ConvDesc(n, h, w, c, c, 3, 3, 2, 1, 1, 1, 1, _dtype_code(x.dtype), 0, 0.0, 0.0, 0, 0, x_ct, x_coff,
                 out.shape[1], y_coff, 0, 0)
    _lib.check(_lib.load().rtseg_maxpool3x3s2(C.byref(d), x.data_ptr(), out.data_ptr(), _stream()), "rtseg_maxpool3x3s2")
    return out


def maxpool2x2_with_indices(x: torch.Tensor, out: Optional[torch.Tensor] = None, y_coff: int = 0, x_coff: int = 0,
                            c: Optional[int] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """nn.MaxPool2d(2, 2, return_indices=True) (reference models/enet.py:130-131) on an NHWC tensor.
    Returns (pooled NHWC tensor, uint8 [N, H/2, W/2, C] window positions — see include/rtseg.h)."""
    _require_cuda(x, "maxpool2x2_with_indices")
    _no_tape(x, "maxpool2x2_with_indices")
    x = to_nhwc(x)
    n, x_ct, h, w = x.shape
    c = c if c is not None else x_ct - x_coff
    ho, wo = h // 2, w // 2
    if out is None:
        out = empty_nhwc(n, c, ho, wo, x.dtype, x.device)
    idx = torch.empty((n, ho, wo, c), dtype=torch.uint8, device=x.device)
    _lib.check(_lib.load().rtseg_maxpool2x2_idx(n, h, w, c, _dtype_code(x.dtype), x.data_ptr(), x_ct, x_coff, out.data_ptr(),
                                                out.shape[1], y_coff, idx.data_ptr(), _stream()), "rtseg_maxpool2x2_idx")
    return out, idx


def maxunpool2x2(x: torch.Tensor, idx: torch.Tensor, size: Optional[Tuple[int, int]] = None) -> torch.Tensor:
    """nn.MaxUnpool2d(2, 2)(x, indices) (reference models/enet.py:138-139, 176-178) with the indices of
    maxpool2x2_with_indices; `size` = (H, W) of the tensor that was pooled (default 2h x 2w)."""
    _require_cuda(x, "maxunpool2x2")
    _no_tape(x, "maxunpool2x2")
    x = to_nhwc(x)
    n, c, hp, wp = x.shape
    if idx.dtype != torch.uint8 or tuple(idx.shape) != (n, hp, wp, c) or not idx.is_contiguous():
        raise ValueError("indices must be the uint8 [N, h, w, C] tensor maxpool2x2_with_indices returned")
    ho, wo = (2 * hp, 2 * wp) if size is None else (int(size[0]), int(size[1]))
    out = empty_nhwc(n, c, ho, wo, x.dtype, x.device)
    _lib.check(_lib.load().rtseg_maxunpool2x2(n, hp, wp, c, ho, wo, _dtype_code(x.dtype), x.data_ptr(), c, 0, idx.data_ptr(),
                                              out.data_ptr(), c, 0, _stream()), "rtseg_maxunpool2x2")
    return out


# -------------------------------------------------------------------- head ops ----
def upsample_bilinear(x: torch.Tensor, size: Tuple[int, int], align_corners: bool, out: Optional[torch.Tensor] = None,
                      y_coff: int = 0, x_coff: int = 0, c: Optional[int] = None) -> torch.Tensor:
    _require_cuda(x, "upsample_bilinear")
    _no_tape(x, "upsample_bilinear")
    if not is_nhwc(x):
        x = to_nhwc(x)
    n, x_ct, hi, wi = x.shape
    c = c if c is not None else x_ct
    ho, wo = int(size[0]), int(size[1])
    if out is None:
        out = empty_nhwc(n, c, ho, wo, x.dtype, x.device)
    _lib.check(_lib.load().rtseg_upsample_bilinear(n, hi, wi, c, ho, wo, int(bool(align_corners)), _dtype_code(x.dtype),
                                                   x.data_ptr(), x_ct, x_coff, out.data_ptr(), out.shape[1], y_coff,
                                                   _stream()), "rtseg_upsample_bilinear")
    return out


def upsample_argmax(logits: torch.Tensor, size: Tuple[int, int], align_corners: bool = True, *,
                    out_dtype: torch.dtype = torch.uint8, target: Optional[torch.Tensor] = None,
                    confmat: Optional[torch.Tensor] = None, ignore_index: int = 255,
                    want_labels: bool = True, num_classes: Optional[int] = None) -> Optional[torch.Tensor]:
    """Fused F.interpolate(bilinear) + max(dim=1)[1] (+ confusion-matrix update).

    Replaces reference models/fastscnn.py:30 followed by core/seg_trainer.py:172 /
    utils/metrics.py.  `logits` is the low-resolution [N, C, h, w] tensor.
    """
    _require_cuda(logits, "upsample_argmax")
    _no_tape(logits, "upsample_argmax")
    x = to_nhwc(logits)
    n, x_ct, hi, wi = x.shape
    c = num_classes if num_classes is not None else x_ct  # a padded logits buffer carries extra channels
    if not 0 < c <= x_ct:
        raise ValueError("num_classes must be in (0, channels of logits]")
    ho, wo = int(size[0]), int(size[1])
    l8 = l64 = None
    if want_labels:
        if out_dtype == torch.uint8:
            l8 = torch.empty((n, ho, wo), dtype=torch.uint8, device=x.device)
        elif out_dtype == torch.int64:
            l64 = torch.empty((n, ho, wo), dtype=torch.int64, device=x.device)
        else:
            raise TypeError("labels are uint8 or int64")
    if (target is None) != (confmat is None):
        raise ValueError("target and confmat go together")
    if target is not None:
        if target.dtype != torch.int64 or tuple(target.shape) != (n, ho, wo) or not target.is_contiguous():
            raise ValueError("target must be a contiguous int64 [N, H, W] tensor")
        if confmat.dtype != torch.int64 or tuple(confmat.shape) != (c, c) or not confmat.is_contiguous():
            raise ValueError("confmat must be a contiguous int64 [C, C] tensor")
    _lib.check(_lib.load().rtseg_upsample_argmax(n, hi, wi, c, x_ct, ho, wo, int(bool(align_corners)), _dtype_code(x.dtype),
                                                 x.data_ptr(), _ptr(l8), _ptr(l64), _ptr(target), int(ignore_index),
                                                 _ptr(confmat), _stream()), "rtseg_upsample_argmax")
    return l8 if l8 is not None else l64


def resize_labels_nearest(masks: torch.Tensor, size: Tuple[int, int]) -> torch.Tensor:
    """int64 label maps [N, H, W] -> [N, h, w] with 'nearest' sampling: what reference core/seg_trainer.py:58-62 gets
    from F.interpolate(masks.unsqueeze(1).float(), size, mode='nearest').squeeze(1).long()."""
    _require_cuda(masks, "resize_labels_nearest")
    if masks.dim() != 3:
        raise ValueError("masks must be [N, H, W]")
    m = masks.long().contiguous()
    n, hi, wi = m.shape
    out = torch.empty((n, int(size[0]), int(size[1])), dtype=torch.int64, device=m.device)
    _lib.check(_lib.load().rtseg_resize_nearest_i64(n, hi, wi, int(size[0]), int(size[1]), m.data_ptr(), out.data_ptr(),
                                                    _stream()), "rtseg_resize_nearest_i64")
    return out


def argmax_channels(x: torch.Tensor, out_dtype: torch.dtype = torch.int64) -> torch.Tensor:
    """preds.max(dim=1)[1] (reference core/seg_trainer.py:172) for NCHW or NHWC storage."""
    _require_cuda(x, "argmax_channels")
    n, c, h, w = x.shape
    out = torch.empty((n, h, w), dtype=out_dtype, device=x.device)
    l8, l64 = (out, None) if out_dtype == torch.uint8 else (None, out)
    if out_dtype not in (torch.uint8, torch.int64):
        raise TypeError("labels are uint8 or int64")
    lib = _lib.load()
    if is_nhwc(x) and not (x.is_contiguous() and c > 1 and h * w > 1):
        _lib.check(lib.rtseg_argmax_nhwc(n * h * w, c, c, _dtype_code(x.dtype), x.data_ptr(), _ptr(l8), _ptr(l64),
                                         _stream()), "rtseg_argmax_nhwc")
    else:
        x = x.contiguous()
        _lib.check(lib.rtseg_argmax_nchw(n, c, h * w, _dtype_code(x.dtype), x.data_ptr(), _ptr(l8), _ptr(l64),
                                         _stream()), "rtseg_argmax_nchw")
    return out


def confmat_update(confmat: torch.Tensor, pred: torch.Tensor, target: torch.Tensor, ignore_index: int = 255,
                   bad_count: Optional[torch.Tensor] = None) -> torch.Tensor:
    _require_cuda(pred, "confmat_update")
    c = confmat.shape[0]
    if confmat.dtype != torch.int64 or not confmat.is_contiguous() or confmat.shape != (c, c):
        raise ValueError("confmat must be a contiguous int64 [C, C] tensor")
    if pred.dtype not in (torch.uint8, torch.int64) or target.dtype != torch.int64:
        raise TypeError("pred must be uint8/int64 and target int64")
    if pred.numel() != target.numel():
        raise ValueError("pred and target differ in size")
    if pred.numel() == 0:
        return confmat
    pred, target = pred.contiguous(), target.contiguous()
    _lib.check(_lib.load().rtseg_confmat_update(pred.numel(), c, int(ignore_index), pred.data_ptr(),
                                                int(pred.dtype == torch.int64), target.data_ptr(), confmat.data_ptr(),
                                                _ptr(bad_count), _stream()), "rtseg_confmat_update")
    return confmat


def adaptive_avgpool(x: torch.Tensor, p: int, x_coff: int = 0, c: Optional[int] = None,
                     out_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """nn.AdaptiveAvgPool2d(p) on an NHWC tensor (channel slice allowed).

    Large cells (LEDNet pools a 128x256 map to 1x1) run in two stages with an fp32 intermediate: first to a
    grid of equal cells that the target cells tile exactly, then to p x p — the mean of equal-sized cell means
    is the cell mean, and every block of the first stage reads a short run of pixels."""
    _require_cuda(x, "adaptive_avgpool")
    _no_tape(x, "adaptive_avgpool")
    n, x_ct, h, w = x.shape
    c = c if c is not None else x_ct
    out_dtype = out_dtype or x.dtype
    lib = _lib.load()
    if (h // p) * (w // p) >= 4096:
        for q in (16, 8, 4):
            if h % (p * q) == 0 and w % (p * q) == 0:
                mid = empty_nhwc(n, c, p * q, p * q, torch.float32, x.device)
                _lib.check(lib.rtseg_adaptive_avgpool(n, h, w, c, p * q, _dtype_code(x.dtype), x.data_ptr(), x_ct, x_coff,
                                                      _dtype_code(mid.dtype), mid.data_ptr(), _stream()),
                           "rtseg_adaptive_avgpool")
                x, x_ct, x_coff, h, w = mid, c, 0, p * q, p * q
                break
    out = empty_nhwc(n, c, p, p, out_dtype, x.device)
    _lib.check(lib.rtseg_adaptive_avgpool(n, h, w, c, p, _dtype_code(x.dtype), x.data_ptr(), x_ct, x_coff,
                                          _dtype_code(out.dtype), out.data_ptr(), _stream()), "rtseg_adaptive_avgpool")
    return out


def pyramid_avgpool(x: torch.Tensor, x_coff: int = 0, c: Optional[int] = None):
    """AdaptiveAvgPool2d(1), (2), (4), (6) of one NHWC map (channel slice allowed) in ONE pass — the four poolings of
    PyramidPoolingModule (reference models/modules.py:145-149).  h and w must be multiples of 4."""
    _require_cuda(x, "pyramid_avgpool")
    _no_tape(x, "pyramid_avgpool")
    n, x_ct, h, w = x.shape
    c = c if c is not None else x_ct - x_coff
    outs = [empty_nhwc(n, c, p, p, x.dtype, x.device) for p in (1, 2, 4, 6)]
    _lib.check(_lib.load().rtseg_pyramid_avgpool(n, h, w, c, _dtype_code(x.dtype), x.data_ptr(), x_ct, x_coff,
                                                 *[o.data_ptr() for o in outs], _stream()), "rtseg_pyramid_avgpool")
    return outs


USE_PPM_FUSION = os.environ.get("RTSEG_PPM_FUSION", "1") != "0"   # module-level switch (tests / A-B runs): the pyramid's four conv1x1 + resize branches in one launch


def ppm_branches(pooled, weights: torch.Tensor, bias: Optional[torch.Tensor], cat: torch.Tensor, y_coff: int) -> torch.Tensor:
    """conv1x1 + bilinear resize (align_corners=True) of the four pooled maps of PyramidPoolingModule into channels
    [y_coff, y_coff + 4 * hid) of the NHWC concat buffer `cat` (reference models/modules.py:145-156) — one launch.

    pooled: the four tensors of pyramid_avgpool(); weights: fp32 [4, hid, c]; bias: fp32 [4, hid] or None."""
    _require_cuda(cat, "ppm_branches")
    _no_tape(cat, "ppm_branches")
    n, y_ct, h, w = cat.shape
    c = pooled[0].shape[1]
    hid = weights.shape[1]
    if cat.dtype != torch.bfloat16 or any(t.dtype != torch.bfloat16 for t in pooled):
        raise TypeError("ppm_branches: bf16 activations only")
    if weights.dtype != torch.float32 or not weights.is_contiguous() or tuple(weights.shape) != (4, hid, c):
        raise ValueError("ppm_branches: weights must be contiguous fp32 [4, hid, c]")
    if [tuple(t.shape) for t in pooled] != [(n, c, p, p) for p in (1, 2, 4, 6)]:
        raise ValueError("ppm_branches: pooled maps must be [n, c, p, p] for p = 1, 2, 4, 6")
    _lib.check(_lib.load().rtseg_ppm_branches(n, h, w, c, hid, _dtype_code(cat.dtype), *[t.data_ptr() for t in pooled],
                                              weights.data_ptr(), bias.data_ptr() if bias is not None else None,
                                              cat.data_ptr(), y_ct, y_coff, _stream()), "rtseg_ppm_branches")
    return cat


CITYSCAPES_COLORMAP = [(128, 64, 128), (244, 35, 232), (70, 70, 70), (102, 102, 156), (190, 153, 153), (153, 153, 153),
                       (250, 170, 30), (220, 220, 0), (107, 142, 35), (152, 251, 152), (70, 130, 180), (220, 20, 60),
                       (255, 0, 0), (0, 0, 142), (0, 0, 70), (0, 60, 100), (0, 80, 100), (0, 0, 230), (119, 11, 32)]
"""The 'cityscapes' table of reference utils/utils.py:59-66 (data, the only colormap the reference defines)."""


def colorize(labels: torch.Tensor, colormap=None, image: Optional[torch.Tensor] = None, alpha: float = 0.0) -> torch.Tensor:
    """uint8 [N, H, W, 3] colour picture of a uint8 class map, optionally alpha-blended with `image` (uint8 NHWC).

    Replaces `colormap[preds.max(dim=1)[1]]` + `Image.blend(image, pred, blend_alpha)` of reference
    core/seg_trainer.py:172-191 with one kernel on the device class map (no int64 map, no HxWx3 int64 gather)."""
    _require_cuda(labels, "colorize")
    if labels.dtype != torch.uint8 or not labels.is_contiguous():
        raise TypeError("labels must be a contiguous uint8 class map")
    colormap = CITYSCAPES_COLORMAP if colormap is None else colormap
    lut = torch.zeros((256, 3), dtype=torch.uint8)
    lut[:len(colormap)] = torch.as_tensor(colormap, dtype=torch.uint8)
    lut = lut.to(labels.device)
    out = torch.empty((*labels.shape, 3), dtype=torch.uint8, device=labels.device)
    if image is not None:
        if image.dtype != torch.uint8 or tuple(image.shape) != tuple(out.shape) or not image.is_contiguous() or not image.is_cuda:
            raise ValueError("image must be a contiguous uint8 CUDA tensor shaped like the output (N, H, W, 3)")
    _lib.check(_lib.load().rtseg_colorize_blend(labels.numel(), labels.data_ptr(), lut.data_ptr(), _ptr(image), float(alpha),
                                                out.data_ptr(), _stream()), "rtseg_colorize_blend")
    return out


def tc_timeout_flag() -> int:
    return int(_lib.load().rtseg_tc_timeout_flag(_stream()))


def avgpool3x3s2(x: torch.Tensor, out_dtype: Optional[torch.dtype] = None, out: Optional[torch.Tensor] = None,
                 y_coff: int = 0, x_coff: int = 0, c: Optional[int] = None) -> torch.Tensor:
    """F.avg_pool2d(x, 3, 2, 1) (reference models/dabnet.py:32-34) from an NCHW or NHWC tensor into an NHWC slice."""
    _require_cuda(x, "avgpool3x3s2")
    _no_tape(x, "avgpool3x3s2")
    if x.dtype == torch.uint8:
        x = as_image_u8(x)
    n, x_ct, h, w = x.shape
    c = c if c is not None else x_ct
    if x.dtype == torch.uint8:   # the decoded image: normalised while it is read
        if x_coff or c != x_ct or out is None:
            raise ValueError("uint8 images are pooled whole, into a caller-provided buffer")
        mean, std = _norm_args(c)
        _lib.check(_lib.load().rtseg_avgpool3x3s2_u8(n, h, w, c, x.data_ptr(), mean, std, _dtype_code(out.dtype),
                                                     out.data_ptr(), out.shape[1], y_coff, _stream()), "rtseg_avgpool3x3s2_u8")
        return out
    if x.is_contiguous() and x_coff == 0 and c == x_ct:
        nchw = 1
    elif is_nhwc(x):
        nchw = 0
    else:
        x, nchw = x.contiguous(), 1
        if x_coff or c != x_ct:
            raise ValueError("channel slices need an NHWC tensor")
    out_dtype = out_dtype or (out.dtype if out is not None else x.dtype)
    if out is None:
        out = empty_nhwc(n, c, (h - 1) // 2 + 1, (w - 1) // 2 + 1, out_dtype, x.device)
    _lib.check(_lib.load().rtseg_avgpool3x3s2(n, h, w, c, nchw, _dtype_code(x.dtype), x.data_ptr(), x_ct, x_coff,
                                              _dtype_code(out.dtype), out.data_ptr(), out.shape[1], y_coff, _stream()),
               "rtseg_avgpool3x3s2")
    return out


def combine_up(a: torch.Tensor, b: torch.Tensor, op: str = "add", b_sigmoid: bool = False, act_module=None,
               out: Optional[torch.Tensor] = None, y_coff: int = 0, a_coff: int = 0, b_coff: int = 0,
               c: Optional[int] = None) -> torch.Tensor:
    """act(a (+|*) g(resize(b))) in one pass: `b` is bilinearly resized (align_corners=True) to a's size on the fly
    ([n, c, 1, 1]: a broadcast), g = sigmoid when b_sigmoid.  The glue between the convolutions of BiSeNetv2's
    aggregation layer and DDRNet's bilateral fusion / DAPPM (reference models/bisenetv2.py:166-201,
    models/ddrnet.py:165, 238-243, 277-291)."""
    _require_cuda(a, "combine_up")
    _no_tape(a, "combine_up")
    _no_tape(b, "combine_up")
    if not is_nhwc(a) or not is_nhwc(b) or a.dtype != b.dtype or a.shape[0] != b.shape[0]:
        raise ValueError("combine_up expects two NHWC tensors of one dtype and batch size")
    n, a_ct, h, w = a.shape
    c = c if c is not None else a_ct - a_coff
    code, p0, p1, prelu_w, post = act_spec(act_module)
    if prelu_w is not None or post is not None:
        raise NotImplementedError("combine_up: pointwise activations without parameters only")
    if out is None:
        out = empty_nhwc(n, c, h, w, a.dtype, a.device)
    elif not is_nhwc(out) or out.dtype != a.dtype or tuple(out.shape[2:]) != (h, w) or out.shape[0] != n:
        raise ValueError("bad output buffer")
    _lib.check(_lib.load().rtseg_combine_up(n, h, w, c, _dtype_code(a.dtype), a.data_ptr(), a_ct, a_coff, b.data_ptr(),
                                            b.shape[2], b.shape[3], b.shape[1], b_coff, {"add": 0, "mul": 1}[op],
                                            int(bool(b_sigmoid)), code, p0, p1, out.data_ptr(), out.shape[1], y_coff,
                                            _stream()), "rtseg_combine_up")
    return out


def avgpool2d(x: torch.Tensor, k: int, s: int, p: int, x_coff: int = 0, c: Optional[int] = None) -> torch.Tensor:
    """nn.AvgPool2d(k, s, p) (count_include_pad=True) on an NHWC tensor — reference models/ddrnet.py:266-275."""
    _require_cuda(x, "avgpool2d")
    _no_tape(x, "avgpool2d")
    if not is_nhwc(x):
        raise ValueError("avgpool2d expects an NHWC tensor")
    n, x_ct, h, w = x.shape
    c = c if c is not None else x_ct - x_coff
    out = empty_nhwc(n, c, (h + 2 * p - k) // s + 1, (w + 2 * p - k) // s + 1, x.dtype, x.device)
    _lib.check(_lib.load().rtseg_avgpool2d(n, h, w, c, k, s, p, _dtype_code(x.dtype), x.data_ptr(), x_ct, x_coff,
                                           out.data_ptr(), c, 0, _stream()), "rtseg_avgpool2d")
    return out


def copy_into_slice(x: torch.Tensor, out: torch.Tensor, y_coff: int = 0, x_coff: int = 0, c: Optional[int] = None) -> torch.Tensor:
    """out[:, y_coff : y_coff + c] = x[:, x_coff : x_coff + c] for NHWC buffers: the permutation kernel with one group
    (the identity permutation) — how a concat's pass-through operand reaches its slice without an ATen copy."""
    _require_cuda(x, "copy_into_slice")
    n, x_ct, h, w = x.shape
    c = c if c is not None else x_ct - x_coff
    if not (is_nhwc(x) and is_nhwc(out)) or x.dtype != out.dtype or tuple(out.shape[2:]) != (h, w) or out.shape[0] != n:
        raise ValueError("copy_into_slice needs NHWC tensors of one dtype and one spatial size")
    _lib.check(_lib.load().rtseg_channel_shuffle(n * h * w, c, 1, _dtype_code(x.dtype), x.data_ptr(), x_ct, x_coff,
                                                 out.data_ptr(), out.shape[1], y_coff, _stream()), "rtseg_channel_shuffle")
    return out


def channel_shuffle(x: torch.Tensor, groups: int = 2) -> torch.Tensor:
    """channel_shuffle (reference models/modules.py:18-32) as one NHWC permutation kernel."""
    _require_cuda(x, "channel_shuffle")
    _no_tape(x, "channel_shuffle")
    x = to_nhwc(x)
    n, c, h, w = x.shape
    out = torch.empty_like(x, memory_format=torch.channels_last)
    _lib.check(_lib.load().rtseg_channel_shuffle(n * h * w, c, groups, _dtype_code(x.dtype), x.data_ptr(), c, 0,
                                                 out.data_ptr(), c, 0, _stream()), "rtseg_channel_shuffle")
    return out
