"""Multi-rank check of the fused gather's completion protocol (run under torchrun on the GPU box):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/check_gather.py
Each step every rank produces a rank- and step-dependent map; the consumers' gathered tensor must equal the expected
concatenation after EVERY step, with (a) ranks deliberately skewed by device-side sleeps so that arrival really is
late, and (b) producers running ahead of a slow consumer (back-pressure on the double-buffered slots)."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereoformer_b200 import ops                       # noqa: E402
from stereoformer_b200.dist import SignalledGather      # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    B, D, H, W = 2, 24, 17, 120
    for dst in ((0,), tuple(range(world))):
        sg = SignalledGather((B, H, W), dev, dst_ranks=dst)
        out = torch.empty(B, 1, H, W, device=dev)
        gathered = torch.empty(world * B, 1, H, W, device=dev)
        steps = 9
        # every rank can build every rank's input: cv(r, s) from a seeded generator
        def cv_of(r, s):
            g = torch.Generator(device=dev).manual_seed(1000 * r + s)
            return torch.randn(B, D, H, W, device=dev, generator=g)
        for s in range(steps):
            if (s + rank) % 3 == 0:
                torch.cuda._sleep(int(2e7) * (1 + rank))            # skew: this rank's step arrives late
            ops.local_window(cv_of(rank, s), None, mode=ops.SF_WIN_FIXED, radius=2, n_points=20, consider_valid=True,
                             div_mode=ops.SF_DIV_RECIP, out=out, gather=sg)
            if sg.is_consumer:
                if s % 4 == 1:
                    torch.cuda._sleep(int(6e7))                      # slow consumer: producers must not lap it
                sg.wait(gathered)
                want = torch.cat([ops.local_window(cv_of(r, s), None, mode=ops.SF_WIN_FIXED, radius=2, n_points=20,
                                                   consider_valid=True, div_mode=ops.SF_DIV_RECIP)
                                  for r in range(world)], 0)
                assert torch.equal(gathered, want), "rank %d dst %r step %d: gathered maps differ" % (rank, dst, s)
        torch.cuda.synchronize()
        sg.check()
        dist.barrier()
        if rank == 0:
            print("gather to %s: %d steps on %d ranks, every step complete and exact" %
                  ("rank 0" if dst == (0,) else "every rank", steps, world), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
