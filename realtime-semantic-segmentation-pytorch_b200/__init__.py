"""rtseg-b200: the B200-native hot path of acai66/realtime-semantic-segmentation-pytorch.

Import name: `rtseg_b200` (this directory carries the repository's mandated,
hyphenated name, so a two-line shim package at the repo root points here).

    csrc/        hand-written sm_100a CUDA kernels + the C ABI (include/rtseg.h)
    _lib.py      ctypes binding of librtseg.so (fails loudly if absent)
    ops.py       tensor-level calls into the C ABI
    modules.py   drop-in ConvBNAct / DWConvBNAct / PWConvBNAct / ... (reference models/modules.py)
    nets.py      ERFNet, Fast-SCNN, DABNet, LEDNet, ENet built from fused blocks, reference-compatible state dicts
    nets_dual.py BiSeNetv2, DDRNet (config 4): fused gather-expansion / residual blocks, one-pass resize+combine glue
    train.py     training-mode forward / backward of the blocks as autograd nodes on rtseg kernels
    loss.py      OhemCELoss / get_loss_fn (reference core/loss.py)
    metrics.py   confusion-matrix mIoU / get_seg_metrics (reference utils/metrics.py)
    parallel.py  set_device / parallel_model / de_parallel (reference utils/parallel.py)
    engine.py    CUDA-graph inference engine with pinned-host I/O (what bench.py times)
    inject.py    installs modules.py under the reference's `models.modules` name
    optim.py     FusedSGD / FusedModelEma: optimizer step + EMA mirror in one launch (utils/optimizer.py, utils/model_ema.py)
    convert.py   convert(model): re-classes a built model's blocks to the fused ones; fused_head(): deferred final interpolate
"""
from . import _lib, ops, optim  # noqa: F401
from .convert import DeferredLogits, convert, fused_head, predict_labels  # noqa: F401
from .engine import InferenceEngine  # noqa: F401
from .loss import OhemCELoss, get_loss_fn, loss_with_aux_heads  # noqa: F401
from .metrics import SegMetrics, get_seg_metrics, iou_from_confmat  # noqa: F401
from .nets import BiSeNetv2, DDRNet, ERFNet, FastSCNN, get_model  # noqa: F401
from .ops import upsample_argmax  # noqa: F401

__version__ = "0.1.0"
