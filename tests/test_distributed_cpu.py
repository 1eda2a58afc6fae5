"""CPU, world_size 2, gloo: the only cross-rank step of the inference path — summing confusion matrices."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import ref_head


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, ret):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from rtseg_b200 import metrics, parallel
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(7)
    pred = rng.integers(0, 19, size=(8, 16, 16))
    target = rng.integers(0, 19, size=(8, 16, 16))
    target[rng.random(target.shape) < 0.1] = 255
    lo, hi = parallel.shard_range(8, rank, world)
    m = metrics.SegMetrics(19, 255)
    # host-side state only: the per-rank matrix comes from the oracle here (no GPU in this test)
    m.confmat = torch.from_numpy(ref_head.confusion_matrix(pred[lo:hi], target[lo:hi], 19, 255))
    iou = m.compute()
    full = ref_head.confusion_matrix(pred, target, 19, 255)
    ok = bool((m.synced_confmat().numpy() == full).all()) and np.array_equal(iou.numpy(), ref_head.iou_from_confmat(full))
    ret[rank] = ok
    dist.destroy_process_group()


def test_confmat_allreduce_world2_gloo():
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, 2, port, ret)) for r in range(2)]
        [p.start() for p in procs]
        [p.join(120) for p in procs]
        assert all(p.exitcode == 0 for p in procs)
        assert ret[0] and ret[1]
