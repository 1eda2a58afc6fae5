"""Block-level drop-in: `convert(model)` and the fused-head shim.

`inject.install` makes the reference's unchanged model files build on the fused LEAF modules (ConvBNAct, ...).  What
stays on ATen in those files is the glue their own `forward`s run between the leaves: the bare `nn.Conv2d` at the
end of `NonBt1DBlock` (reference models/erfnet.py:69-70), `x += residual` and the trailing BatchNorm (:77-81),
`torch.cat` + `nn.MaxPool2d` in `InitialBlock` (models/enet.py:43-46), the full-resolution `F.interpolate` before
the argmax (models/fastscnn.py:30).  Two mechanisms close that gap without touching the reference:

* `convert(model)` walks a built model — the reference's classes or the injected ones — and re-classes every block
  it knows to the fused implementation of nets.py / modules.py, in the manner of
  `nn.SyncBatchNorm.convert_sync_batchnorm`.  Parameters, buffers, child modules and state-dict keys are the same
  objects before and after (the fused classes mirror the reference's attribute layout one to one), only `forward`
  changes.  A block is recognised by class name AND child structure, so a different class that merely shares a name
  in one of the other model files is left alone.
* `fused_head()` is a `TorchFunctionMode`: inside it, a bilinear `F.interpolate` of CUDA logits returns a deferred
  tensor; `max(dim=1)` / `argmax(dim=1)` on it (reference core/seg_trainer.py:172, models/ddrnet.py:55-58) runs the
  fused upsample+argmax kernel and the full-resolution logits are never written; any other use materialises them
  with the bilinear kernel, so results are unchanged for callers that do want the logits.
"""
from __future__ import annotations

from typing import Callable, Dict, Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.overrides import TorchFunctionMode

from . import modules as M
from . import nets as N
from . import nets_dual as D
from . import ops


# ------------------------------------------------------------------------------------------ structure checks ----
def _is_cba(m) -> bool:
    """Sequential(Conv2d, BatchNorm2d, Activation-with-.activation): ConvBNAct / DWConvBNAct / PWConvBNAct."""
    return (isinstance(m, nn.Sequential) and len(m) == 3 and isinstance(m[0], nn.Conv2d)
            and isinstance(m[1], nn.modules.batchnorm._BatchNorm) and hasattr(m[2], "activation"))


def _seq(m, n) -> bool:
    return isinstance(m, nn.Sequential) and len(m) == n


def _chk_nonbt1d(m) -> bool:
    return (_seq(getattr(m, "conv", None), 4) and all(_is_cba(m.conv[i]) for i in range(3)) and isinstance(m.conv[3], nn.Conv2d)
            and _seq(getattr(m, "bn_act", None), 2) and isinstance(m.bn_act[0], nn.modules.batchnorm._BatchNorm)
            and hasattr(m.bn_act[1], "activation"))


def _chk_initial(m) -> bool:
    return _is_cba(getattr(m, "conv", None)) and isinstance(getattr(m, "pool", None), nn.MaxPool2d) and m.conv[0].stride == (2, 2)


def _chk_inverted(m) -> bool:
    return (_seq(getattr(m, "conv", None), 3) and all(_is_cba(m.conv[i]) for i in range(3)) and hasattr(m, "use_res_connect")
            and m.conv[0][0].kernel_size == (1, 1) and m.conv[1][0].groups == m.conv[1][0].in_channels
            and m.conv[2][0].kernel_size == (1, 1))


def _chk_ffm(m) -> bool:
    return (isinstance(getattr(m, "higher_res_conv", None), nn.Conv2d) and _seq(getattr(m, "lower_res_conv", None), 2)
            and _is_cba(m.lower_res_conv[0]) and isinstance(m.lower_res_conv[1], nn.Conv2d)
            and _seq(getattr(m, "non_linear", None), 2))


def _chk_ltd(m) -> bool:
    return _seq(m, 3) and _is_cba(m[0]) and all(_seq(m[i], 2) and _is_cba(m[i][0]) and _is_cba(m[i][1]) for i in (1, 2))


def _chk_dab(m) -> bool:
    return (_is_cba(getattr(m, "init_conv", None)) and _is_cba(getattr(m, "last_conv", None))
            and all(_seq(getattr(m, b, None), 2) and _is_cba(getattr(m, b)[0]) and _is_cba(getattr(m, b)[1])
                    for b in ("left_branch", "right_branch")))


def _chk_ssnbt(m) -> bool:
    def branch(b):
        return (_seq(b, 6) and isinstance(b[0], nn.Conv2d) and hasattr(b[1], "activation") and _is_cba(b[2])
                and isinstance(b[3], nn.Conv2d) and hasattr(b[4], "activation") and _is_cba(b[5]))
    return hasattr(m, "split_channels") and branch(getattr(m, "left_branch", None)) and branch(getattr(m, "right_branch", None)) \
        and hasattr(getattr(m, "act", None), "activation")


def _chk_apn(m) -> bool:
    names = ("left_conv1_1", "left_conv1_2", "left_conv2_1", "left_conv2_2", "mid_branch")
    return all(_is_cba(getattr(m, k, None)) for k in names) and _seq(getattr(m, "left_conv3", None), 2) \
        and _seq(getattr(m, "right_branch", None), 2) and isinstance(m.right_branch[0], nn.AdaptiveAvgPool2d)


def _chk_ppm(m) -> bool:
    return all(_seq(getattr(m, f"stage{i}", None), 2) and isinstance(getattr(m, f"stage{i}")[1], nn.Conv2d) for i in range(1, 5)) \
        and _is_cba(getattr(m, "conv", None))


def _chk_deconv(m) -> bool:
    u = getattr(m, "up_conv", None)
    return _seq(u, 3) and isinstance(u[0], nn.ConvTranspose2d) and hasattr(u[2], "activation")


def _chk_ds(m) -> bool:
    return _seq(m, 2) and _is_cba(m[0]) and _is_cba(m[1]) and m[0][0].groups == m[0][0].in_channels and m[1][0].kernel_size == (1, 1)


def _chk_act(m) -> bool:
    return isinstance(getattr(m, "activation", None), nn.Module)


def _has(m, *names) -> bool:
    return all(isinstance(getattr(m, k, None), nn.Module) for k in names)


# whole networks: give them predict() / set_act_dtype() / the concat-free execution plan
def _net_erfnet(m) -> Optional[int]:
    if _has(m, "layer1", "layer2", "layer3_7", "layer8", "layer9_16", "layer17", "layer18_19", "layer20", "layer21_22", "layer23") \
            and _chk_deconv(m.layer23):
        return m.layer23.up_conv[0].out_channels
    return None


def _net_fastscnn(m) -> Optional[int]:
    if _has(m, "learning_to_downsample", "global_feature_extractor", "feature_fusion", "classifier") and _seq(m.classifier, 3):
        return m.classifier[2][0].out_channels
    return None


def _net_dabnet(m) -> Optional[int]:
    if _has(m, "layer1", "layer2", "layer3", "layer4", "layer5_7", "layer8", "layer9_10", "layer11_12", "layer13_14", "layer15") \
            and isinstance(m.layer15, nn.Conv2d):
        return m.layer15.out_channels
    return None


def _net_lednet(m) -> Optional[int]:
    if _has(m, "encoder", "apn") and _chk_apn(m.apn):
        return m.apn.mid_branch[0].out_channels
    return None


# ---- round 2: ENet (models/enet.py), BiSeNetv2 (models/bisenetv2.py), DDRNet (models/ddrnet.py) ----
def _chk_seghead(m) -> bool:
    return _seq(m, 2) and _is_cba(m[0]) and isinstance(m[1], nn.Conv2d)


def _chk_enet_upsample(m) -> bool:
    u = getattr(m, "up_conv", None)
    if isinstance(u, nn.ConvTranspose2d):
        return u.kernel_size == (3, 3) and u.stride == (2, 2) and u.padding == (1, 1) and u.output_padding == (1, 1) and u.bias is None
    return _seq(u, 2) and _is_cba(u[0]) and isinstance(u[1], nn.Upsample)


def _chk_enet_bottleneck(m) -> bool:
    if getattr(m, "conv_type", None) not in ("regular", "downsampling", "upsampling", "dilate", "asymmetric"):
        return False
    last = getattr(m, "right_last_conv", None)
    return (isinstance(getattr(m, "right_init_conv", None), nn.Sequential) and _is_cba(m.right_init_conv[0]) and _seq(last, 2)
            and isinstance(last[0], nn.Conv2d) and isinstance(last[1], nn.Dropout) and hasattr(getattr(m, "act", None), "activation"))


def _chk_ge(m) -> bool:
    lb = getattr(m, "left_branch", None)
    return (isinstance(lb, nn.Sequential) and len(lb) in (3, 4) and all(_is_cba(x) for x in lb) and hasattr(m, "stride")
            and hasattr(getattr(m, "act", None), "activation") and (m.stride != 2 or _seq(getattr(m, "right_branch", None), 2)))


def _chk_stemblock(m) -> bool:
    return (_is_cba(getattr(m, "conv_init", None)) and _seq(getattr(m, "left_branch", None), 2) and _is_cba(m.left_branch[0])
            and isinstance(getattr(m, "right_branch", None), nn.MaxPool2d) and _is_cba(getattr(m, "conv_last", None)))


def _chk_ceb(m) -> bool:
    return (_seq(getattr(m, "pool", None), 2) and isinstance(m.pool[0], nn.AdaptiveAvgPool2d)
            and isinstance(m.pool[1], nn.modules.batchnorm._BatchNorm) and _is_cba(getattr(m, "conv_mid", None))
            and isinstance(getattr(m, "conv_last", None), nn.Conv2d))


def _chk_bga(m) -> bool:
    return (_seq(getattr(m, "detail_high", None), 2) and _is_cba(m.detail_high[0]) and isinstance(m.detail_high[1], nn.Conv2d)
            and _seq(getattr(m, "detail_low", None), 2) and isinstance(m.detail_low[1], nn.AvgPool2d)
            and _seq(getattr(m, "semantic_high", None), 3) and isinstance(m.semantic_high[2], nn.Sigmoid)
            and _seq(getattr(m, "semantic_low", None), 3) and isinstance(m.semantic_low[1], nn.Conv2d)
            and _is_cba(getattr(m, "conv_last", None)))


def _chk_detail(m) -> bool:
    return _seq(m, 8) and all(_is_cba(x) for x in m)


def _chk_semantic(m) -> bool:
    return _has(m, "stage1to2", "stage3", "stage4", "stage5_1to4", "stage5_5") and hasattr(m, "use_aux")


def _chk_rb(m) -> bool:
    return (_is_cba(getattr(m, "conv1", None)) and _is_cba(getattr(m, "conv2", None)) and not hasattr(m, "conv3")
            and hasattr(m, "downsample") and isinstance(getattr(m, "act", None), nn.Module))


def _chk_rbb(m) -> bool:
    return all(_is_cba(getattr(m, k, None)) for k in ("conv1", "conv2", "conv3")) and hasattr(m, "downsample") \
        and hasattr(getattr(m, "act", None), "activation")


def _chk_bfusion(m) -> bool:
    return _is_cba(getattr(m, "conv_low", None)) and _is_cba(getattr(m, "conv_high", None)) and hasattr(getattr(m, "act", None), "activation")


def _chk_dappm(m) -> bool:
    return (all(_is_cba(getattr(m, f"conv{i}", None)) for i in range(6)) and _is_cba(getattr(m, "conv_last", None))
            and all(_seq(getattr(m, f"pool{i}", None), 2) and isinstance(getattr(m, f"pool{i}")[1], nn.Conv2d) for i in range(2, 6)))


def _chk_stage4(m) -> bool:
    return _has(m, "low_conv1", "high_conv1", "bilateral_fusion1") and hasattr(m, "extra_conv")


def _chk_stage5(m) -> bool:
    return _has(m, "low_conv1", "high_conv1", "bilateral_fusion", "low_conv2", "high_conv2", "dappm")


def _net_enet(m) -> Optional[int]:
    if _has(m, "initial", "bottleneck1", "bottleneck2", "bottleneck3", "bottleneck4", "bottleneck5", "fullconv") \
            and _chk_enet_upsample(m.fullconv):
        u = m.fullconv.up_conv
        return u.out_channels if isinstance(u, nn.ConvTranspose2d) else u[0][0].out_channels
    return None


def _net_bisenetv2(m) -> Optional[int]:
    if _has(m, "detail_branch", "semantic_branch", "bga_layer", "seg_head") and _chk_seghead(m.seg_head):
        return m.seg_head[1].out_channels
    return None


def _net_ddrnet(m) -> Optional[int]:
    if _has(m, "conv1", "conv2", "conv3", "conv4", "conv5", "seg_head") and _chk_seghead(m.seg_head) and _is_cba(m.conv1):
        return m.seg_head[1].out_channels
    return None


# class name -> (fused class, structure check)
_BLOCKS: Dict[str, Tuple[type, Callable]] = {
    "ConvBNAct": (M.ConvBNAct, _is_cba),
    "DWConvBNAct": (M.DWConvBNAct, _is_cba),
    "PWConvBNAct": (M.PWConvBNAct, _is_cba),
    "DSConvBNAct": (M.DSConvBNAct, _chk_ds),
    "DeConvBNAct": (M.DeConvBNAct, _chk_deconv),
    "Activation": (M.Activation, _chk_act),
    "PyramidPoolingModule": (M.PyramidPoolingModule, _chk_ppm),
    "InitialBlock": (N.InitialBlock, _chk_initial),
    "NonBt1DBlock": (N.NonBt1DBlock, _chk_nonbt1d),
    "InvertedResidual": (N.InvertedResidual, _chk_inverted),
    "LearningToDownsample": (N.LearningToDownsample, _chk_ltd),
    "FeatureFusionModule": (N.FeatureFusionModule, _chk_ffm),
    "DABModule": (N.DABModule, _chk_dab),
    "SSnbtUnit": (N.SSnbtUnit, _chk_ssnbt),
    "AttentionPyramidNetwork": (N.AttentionPyramidNetwork, _chk_apn),
    "SegHead": (M.SegHead, _chk_seghead),
    "Upsample": (N.Upsample, _chk_enet_upsample),
    "Bottleneck": (N.Bottleneck, _chk_enet_bottleneck),
    "GatherExpansionLayer": (D.GatherExpansionLayer, _chk_ge),
    "StemBlock": (D.StemBlock, _chk_stemblock),
    "ContextEmbeddingBlock": (D.ContextEmbeddingBlock, _chk_ceb),
    "BilateralGuidedAggregationLayer": (D.BilateralGuidedAggregationLayer, _chk_bga),
    "DetailBranch": (D.DetailBranch, _chk_detail),
    "SemanticBranch": (D.SemanticBranch, _chk_semantic),
    "RB": (D.RB, _chk_rb),
    "RBB": (D.RBB, _chk_rbb),
    "BilateralFusion": (D.BilateralFusion, _chk_bfusion),
    "DAPPM": (D.DAPPM, _chk_dappm),
    "Stage4": (D.Stage4, _chk_stage4),
    "Stage5": (D.Stage5, _chk_stage5),
}
_NETS: Dict[str, Tuple[type, Callable]] = {
    "ERFNet": (N.ERFNet, _net_erfnet), "FastSCNN": (N.FastSCNN, _net_fastscnn),
    "DABNet": (N.DABNet, _net_dabnet), "LEDNet": (N.LEDNet, _net_lednet),
    "ENet": (N.ENet, _net_enet), "BiSeNetv2": (D.BiSeNetv2, _net_bisenetv2), "DDRNet": (D.DDRNet, _net_ddrnet),
}


_BARE_CONV_OWNERS = (N.NonBt1DBlock, N.FeatureFusionModule, M.PyramidPoolingModule, N.SSnbtUnit, N.DABNet, M.SegHead, N.Bottleneck,
                     D.ContextEmbeddingBlock, D.BilateralGuidedAggregationLayer, D.DAPPM)


def convert(model: nn.Module, verbose: bool = False) -> nn.Module:
    """Re-class the blocks of a built model to their fused implementations, in place; returns the model.

    Works on a model built from the reference's own classes or from the injected ones.  Idempotent.  Blocks the
    function does not know keep their `forward` (their fused leaves still run through librtseg)."""
    swapped: Dict[str, int] = {}

    def note(name):
        swapped[name] = swapped.get(name, 0) + 1

    owners = []   # fused blocks whose forward hands keyword arguments to their bare convolutions
    for mod in list(model.modules()):
        cls = type(mod)
        name = cls.__name__
        if name in _NETS:
            fused, probe = _NETS[name]
            if cls is not fused:
                num_class = probe(mod)
                if num_class is not None:
                    mod.__class__ = fused
                    mod.__dict__.setdefault("num_class", int(num_class))
                    if fused is N.DABNet:
                        mod.__dict__.setdefault("act_dtype", None)
                    note(name)
        elif name in _BLOCKS:
            fused, ok = _BLOCKS[name]
            if cls is not fused and ok(mod):
                mod.__class__ = fused
                note(name)
        if isinstance(mod, _BARE_CONV_OWNERS):
            owners.append(mod)

    def reclass_bare_convs(mod):
        for child in mod.children():
            if type(child) is nn.Conv2d:
                if child.groups in (1, child.in_channels) and child.padding_mode == "zeros" and not isinstance(child.padding, str):
                    child.__class__ = M.Conv2d    # same storage; forward(x, res=..., bn=..., act=...) on the rtseg kernels
                    note("Conv2d")
            elif not _is_cba(child) and not _chk_deconv(child):
                reclass_bare_convs(child)
    for mod in owners:
        reclass_bare_convs(mod)
    if verbose:
        print("rtseg_b200.convert:", ", ".join(f"{v} x {k}" for k, v in sorted(swapped.items())) or "nothing to swap")
    model.__dict__["_rtseg_converted"] = swapped
    return model


# --------------------------------------------------------------------------------------------- fused head ----
class DeferredLogits(torch.Tensor):
    """What `F.interpolate(logits, size, mode='bilinear')` returns inside `fused_head()`: the shape, dtype and device of
    the full-resolution logits, backed by the LOW-resolution tensor.  `max(dim=1)` / `argmax(dim=1)` run the fused
    upsample+argmax kernel; every other operation first materialises the logits with the bilinear kernel."""

    @staticmethod
    def __new__(cls, low: torch.Tensor, size, align_corners: bool):
        n, c = low.shape[:2]
        t = torch.Tensor._make_wrapper_subclass(cls, (n, c, int(size[0]), int(size[1])), dtype=low.dtype, device=low.device)
        t._low, t._align, t._full = low, bool(align_corners), None
        return t

    def materialise(self) -> torch.Tensor:
        if self._full is None:
            self._full = ops.upsample_bilinear(self._low, tuple(self.shape[2:]), self._align)
        return self._full

    def labels(self, out_dtype=torch.int64) -> torch.Tensor:
        return ops.upsample_argmax(self._low, tuple(self.shape[2:]), self._align, out_dtype=out_dtype)

    def __repr__(self):
        return f"DeferredLogits(shape={tuple(self.shape)}, low={tuple(self._low.shape)}, align_corners={self._align})"

    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        kwargs = kwargs or {}
        name = getattr(func, "__name__", "")
        self = next((a for a in args if isinstance(a, DeferredLogits)), None)
        if self is not None and args and args[0] is self:
            if name in ("size", "dim", "__get__") or func in (torch.Tensor.size, torch.Tensor.dim):
                with torch._C.DisableTorchFunctionSubclass():
                    return func(*args, **kwargs)
            dim = kwargs.get("dim", args[1] if len(args) > 1 and isinstance(args[1], int) else None)
            keepdim = bool(kwargs.get("keepdim", args[2] if len(args) > 2 and isinstance(args[2], bool) else False))
            if dim in (1, -3) and len(args) <= 3 and set(kwargs) <= {"dim", "keepdim"}:
                if func in (torch.max, torch.Tensor.max):
                    return _LazyMax(self, keepdim)
                if func in (torch.argmax, torch.Tensor.argmax):
                    lab = self.labels()
                    return lab.unsqueeze(1) if keepdim else lab

        def unwrap(a):
            if isinstance(a, DeferredLogits):
                return a.materialise()
            if isinstance(a, (list, tuple)):
                return type(a)(unwrap(v) for v in a)
            return a
        with torch._C.DisableTorchFunctionSubclass():
            return func(*[unwrap(a) for a in args], **{k: unwrap(v) for k, v in kwargs.items()})


    @classmethod
    def __torch_dispatch__(cls, func, types, args=(), kwargs=None):
        # Reached only by calls that bypass __torch_function__ (C++-side callers): same rule, materialise and run.
        def unwrap(a):
            if isinstance(a, DeferredLogits):
                return a.materialise()
            if isinstance(a, (list, tuple)):
                return type(a)(unwrap(v) for v in a)
            return a
        return func(*[unwrap(a) for a in args], **{k: unwrap(v) for k, v in (kwargs or {}).items()})


class _LazyMax:
    """Result of `deferred.max(dim=1)`: `[1]` / `.indices` is the fused class map (int64, like torch.max);
    `[0]` / `.values` materialises the logits."""

    def __init__(self, d: DeferredLogits, keepdim: bool = False):
        self._d, self._idx, self._val, self._keep = d, None, None, keepdim

    @property
    def indices(self):
        if self._idx is None:
            lab = self._d.labels()
            self._idx = lab.unsqueeze(1) if self._keep else lab
        return self._idx

    @property
    def values(self):
        if self._val is None:
            self._val = self._d.materialise().max(dim=1, keepdim=self._keep)[0]
        return self._val

    def __getitem__(self, i):
        if i in (0, -2):
            return self.values
        if i in (1, -1):
            return self.indices
        raise IndexError("max(dim) returns (values, indices)")

    def __iter__(self):
        return iter((self.values, self.indices))

    def __len__(self):
        return 2


class fused_head(TorchFunctionMode):
    """with rtseg_b200.fused_head(): preds = model(images); labels = preds.max(dim=1)[1]

    Defers every bilinear `F.interpolate` whose input is a CUDA tensor with at most `max_channels` channels and whose
    output is larger than its input (the segmentation heads of reference models/fastscnn.py:30, dabnet.py:60,
    lednet.py:26, bisenetv2.py:35, ddrnet.py:53)."""

    def __init__(self, max_channels: int = 256):
        super().__init__()
        self.max_channels = max_channels

    def __torch_function__(self, func, types, args=(), kwargs=None):
        kwargs = kwargs or {}
        if func is F.interpolate:
            x = args[0] if args else kwargs.get("input")
            size = args[1] if len(args) > 1 else kwargs.get("size")
            mode = args[3] if len(args) > 3 else kwargs.get("mode", "nearest")
            align = args[4] if len(args) > 4 else kwargs.get("align_corners")
            if (isinstance(x, torch.Tensor) and not isinstance(x, DeferredLogits) and x.is_cuda and x.dim() == 4
                    and mode == "bilinear" and size is not None and kwargs.get("scale_factor") is None
                    and x.shape[1] <= self.max_channels and x.dtype in (torch.float32, torch.bfloat16)
                    and not (torch.is_grad_enabled() and x.requires_grad)):
                size = (int(size), int(size)) if isinstance(size, int) else tuple(int(s) for s in size)
                if size[0] * size[1] > x.shape[2] * x.shape[3]:
                    return DeferredLogits(x, size, bool(align))
        return func(*args, **kwargs)


def predict_labels(model: nn.Module, images: torch.Tensor, out_dtype=torch.int64) -> torch.Tensor:
    """`model(images).max(dim=1)[1]` (reference core/seg_trainer.py:170-172) with the head fused: the model's own
    predict() when convert() gave it one, else its forward under fused_head()."""
    with torch.no_grad():
        if hasattr(model, "predict"):
            lab = model.predict(images)
            return lab if out_dtype == torch.uint8 else lab.to(out_dtype)
        with fused_head():
            preds = model(images)
            if isinstance(preds, DeferredLogits):
                return preds.labels(out_dtype)
            return ops.argmax_channels(preds, out_dtype)
