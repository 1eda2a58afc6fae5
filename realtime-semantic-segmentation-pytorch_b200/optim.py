"""Optimizer step + weight mirror as one kernel launch.

    opt = rtseg_b200.optim.FusedSGD(model.parameters(), lr=0.08, momentum=0.9, weight_decay=1e-4)
    ema = rtseg_b200.optim.FusedModelEma(model, use_ema=True, total_itrs=N)        # optional
    ...
    loss.backward()
    opt.step(ema=ema, cur_itrs=it)     # SGD update of every parameter + EMA / mirror of every state-dict entry: ONE launch

`FusedSGD` is a `torch.optim.Optimizer` (schedulers, `zero_grad`, `state_dict` work as usual; the momentum buffers live
in `state[p]['momentum_buffer']` like torch.optim.SGD's) restating the reference's configuration (utils/optimizer.py:8-12:
SGD(lr, momentum, weight_decay), dampening 0, no Nesterov).  `FusedModelEma` restates ModelEmaV2
(utils/model_ema.py:14-40): `.ema` is a deep copy of the de-parallelised model in eval mode, and `update` applies
`decay * e + (1 - decay) * m` (decay = clamp(cur_itrs / total_itrs, 0, 1)) to EVERY state-dict entry, or copies when
`use_ema` is off.  Without an optimizer in the loop, `ema.update(model, cur_itrs)` launches the same kernel alone.
"""
from __future__ import annotations

import ctypes as C
from copy import deepcopy
from typing import List, Optional

import torch

from . import _lib
from .parallel import de_parallel


class _Chunk(C.Structure):
    """struct rtseg_opt_chunk (include/rtseg.h)."""
    _fields_ = [("p", C.c_void_p), ("g", C.c_void_p), ("m", C.c_void_p), ("e", C.c_void_p), ("n", C.c_int32), ("flags", C.c_int32)]


RTSEG_OPT_INT64 = 1


def _table(entries, device) -> torch.Tensor:
    """entries: (p, g, m, e) tensors (g / m / e may be None).  Returns the device-resident chunk table."""
    step = int(_lib.load().rtseg_opt_chunk_elems())
    rows: List[_Chunk] = []
    for p, g, m, e in entries:
        if p.numel() == 0:
            continue
        for t in (p, g, m, e):
            if t is not None and (not t.is_cuda or not t.is_contiguous()):
                raise ValueError("fused optimizer: tensors must be contiguous CUDA tensors")
        if p.dtype == torch.int64:
            if e is None:
                continue
            rows.append(_Chunk(p.data_ptr(), None, None, e.data_ptr(), p.numel(), RTSEG_OPT_INT64))
            continue
        if p.dtype != torch.float32 or any(t is not None and t.dtype != torch.float32 for t in (g, m, e)):
            raise TypeError("fused optimizer: fp32 master parameters, gradients, momentum buffers and mirrors")
        es = 4
        for off in range(0, p.numel(), step):
            n = min(step, p.numel() - off)
            ptr = lambda t: None if t is None else t.data_ptr() + off * es
            rows.append(_Chunk(ptr(p), ptr(g), ptr(m), ptr(e), n, 0))
    arr = (_Chunk * max(1, len(rows)))(*rows)
    host = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).clone()
    return host.to(device), len(rows)


class FusedModelEma(torch.nn.Module):
    """ModelEmaV2 of the reference (utils/model_ema.py:14-40) with a one-launch update."""

    def __init__(self, model: torch.nn.Module, use_ema: bool = True, total_itrs: int = 1):
        super().__init__()
        self.ema = deepcopy(de_parallel(model))
        self.ema.eval()
        self.use_ema, self.total_itrs = bool(use_ema), int(total_itrs)
        self._plan = None

    def decay(self, cur_itrs: int) -> float:
        return min(max(cur_itrs / self.total_itrs, 0), 1)

    def pairs(self, model):
        """(model value, mirror) for every state-dict entry, in state-dict order like the reference's zip."""
        return list(zip(de_parallel(model).state_dict().values(), self.ema.state_dict().values()))

    @torch.no_grad()
    def update(self, model, cur_itrs: int):
        pairs = self.pairs(model)
        key = tuple((m.data_ptr(), e.data_ptr()) for m, e in pairs)
        if self._plan is None or self._plan[0] != key:
            dev = pairs[0][0].device
            table, n = _table([(m, None, None, e) for m, e in pairs], dev)
            self._plan = (key, table, n, torch.zeros(4, dtype=torch.float32, device=dev))
        _, table, n, hyper = self._plan
        hyper[3] = self.decay(cur_itrs)
        _lib.check(_lib.load().rtseg_sgd_ema_step(table.data_ptr(), n, hyper.data_ptr(), int(self.use_ema),
                                                  torch.cuda.current_stream().cuda_stream), "rtseg_sgd_ema_step")


class FusedSGD(torch.optim.Optimizer):
    """SGD with momentum and weight decay (reference utils/optimizer.py:8-12) as one kernel launch per step."""

    def __init__(self, params, lr: float, momentum: float = 0.0, weight_decay: float = 0.0):
        if lr < 0 or momentum < 0 or weight_decay < 0:
            raise ValueError("lr, momentum and weight_decay must be non-negative")
        super().__init__(params, dict(lr=lr, momentum=momentum, weight_decay=weight_decay))
        self._plan = None

    def _entries(self, group, ema_of):
        out = []
        for p in group["params"]:
            if p.grad is None:
                continue
            st = self.state[p]
            if "momentum_buffer" not in st:
                st["momentum_buffer"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            out.append((p.data, p.grad, st["momentum_buffer"] if group["momentum"] else None, ema_of.get(p.data_ptr())))
        return out

    @torch.no_grad()
    def step(self, closure=None, ema: Optional[FusedModelEma] = None, model=None, cur_itrs: int = 0):
        """One launch per parameter group.  With `ema` (and the `model` it mirrors) the EMA / mirror update of every
        state-dict entry rides in the same launch — parameters read their just-updated values, BatchNorm statistics
        and counters are mirrored as they are.  The chunk table is rebuilt whenever a parameter's or gradient's
        storage moves: keep gradients allocated (`zero_grad(set_to_none=False)`) in a steady-state loop."""
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = _lib.load()
        stream = torch.cuda.current_stream().cuda_stream
        pairs = ema.pairs(model) if ema is not None else []
        ema_of = {m.data_ptr(): e for m, e in pairs}
        key = (tuple((p.data_ptr(), p.grad.data_ptr() if p.grad is not None else 0) for g in self.param_groups for p in g["params"]),
               tuple(ema_of))
        if self._plan is None or self._plan[0] != key:
            plans = []
            for gi, group in enumerate(self.param_groups):
                entries = self._entries(group, ema_of)
                if gi == 0 and ema is not None:   # state-dict entries without a gradient (BatchNorm statistics, counters, frozen weights)
                    with_grad = {e[0].data_ptr() for g in self.param_groups for e in self._entries(g, ema_of)}
                    entries += [(m, None, None, e) for m, e in pairs if m.data_ptr() not in with_grad]
                if not entries:
                    plans.append(None)
                    continue
                dev = entries[0][0].device
                table, n = _table(entries, dev)
                plans.append((table, n, torch.zeros(4, dtype=torch.float32, device=dev),
                              {"vals": None, "pinned": torch.zeros(4, dtype=torch.float32).pin_memory()}))
            self._plan = (key, plans)
        self.push_hyper(ema, cur_itrs)
        for group, plan in zip(self.param_groups, self._plan[1]):
            if plan is None:
                continue
            table, n, hyper = plan[:3]
            _lib.check(lib.rtseg_sgd_ema_step(table.data_ptr(), n, hyper.data_ptr(), int(ema.use_ema) if ema is not None else 0,
                                              stream), "rtseg_sgd_ema_step")
        return loss

    def push_hyper(self, ema: Optional[FusedModelEma] = None, cur_itrs: int = 0):
        """Writes (lr, momentum, weight_decay, decay) of every group to its device buffer when they changed.  step()
        calls it; under CUDA-graph capture nothing is written (the captured kernel reads the buffer at replay time), so
        a loop that replays a captured step calls push_hyper() itself after the scheduler moved the learning rate."""
        if self._plan is None or torch.cuda.is_current_stream_capturing():
            return
        for group, plan in zip(self.param_groups, self._plan[1]):
            if plan is None:
                continue
            vals = (float(group["lr"]), float(group["momentum"]), float(group["weight_decay"]),
                    float(ema.decay(cur_itrs)) if ema is not None else 0.0)
            if plan[3].get("vals") != vals:
                plan[3]["vals"] = vals
                plan[3]["pinned"].copy_(torch.tensor(vals, dtype=torch.float32))
                plan[2].copy_(plan[3]["pinned"], non_blocking=True)
