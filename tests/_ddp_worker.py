"""torchrun worker for tests/test_gpu_train.py::test_ddp_syncbn_matches_oracle_on_the_joint_batch (2 ranks, NCCL)."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_nets  # noqa: E402  (weights / inputs only: seeded generators)
from tests.helpers import load_specs  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    import rtseg_b200  # noqa: F401
    from rtseg_b200 import loss as rloss
    from rtseg_b200 import nets
    spec = load_specs()["erfnet"]["spec"]
    sd0 = ref_nets.make_state(spec, 3, True)
    x = ref_nets.make_input(4, 2 * world, 64, 128)
    g = torch.Generator().manual_seed(5)
    labels = torch.randint(0, 19, (2 * world, 64, 128), generator=g)
    labels[torch.rand((2 * world, 64, 128), generator=g) < 0.1] = 255
    m = nets.ERFNet(num_class=19)
    m.load_state_dict(sd0, strict=True)
    m = torch.nn.SyncBatchNorm.convert_sync_batchnorm(m.cuda().train())       # reference utils/parallel.py:29-32
    ddp = torch.nn.parallel.DistributedDataParallel(m, device_ids=[torch.cuda.current_device()])
    lo, hi = 2 * rank, 2 * rank + 2
    loss = rloss.OhemCELoss(0.7, 255)(ddp(x[lo:hi].cuda()), labels[lo:hi].cuda())
    loss.backward()
    losses = [torch.zeros((), device="cuda") for _ in range(world)]
    dist.all_gather(losses, loss.detach())
    if rank == 0:
        out = {"losses": [float(v) for v in losses],
               "grads": {k: p.grad.float().cpu().flatten().tolist() for k, p in m.named_parameters() if p.numel() <= 4096},
               "running": {k: v.float().cpu().flatten().tolist() for k, v in m.state_dict().items() if "running" in k}}
        with open(sys.argv[1], "w") as f:
            json.dump(out, f)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
