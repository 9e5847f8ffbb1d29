"""The fused gather's completion protocol on one GPU (world size 1: the rank is its own destination; the kernels and the
signal words are the ones the multi-GPU run uses).  The multi-rank behaviour is exercised by tools/check_gather.py
under torchrun (GPU box) and by bench.py's per-step check during warm-up."""
import os
import tempfile

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pg():
    import torch.distributed as dist
    if dist.is_initialized():
        yield dist.group.WORLD
        return
    torch.cuda.set_device(0)
    f = tempfile.NamedTemporaryFile(delete=False)
    f.close()
    dist.init_process_group("nccl", init_method="file://" + f.name, rank=0, world_size=1,
                            device_id=torch.device("cuda", 0))
    yield dist.group.WORLD
    dist.destroy_process_group()
    os.unlink(f.name)


def test_signalled_gather_single_rank(pg):
    from stereoformer_b200 import ops
    from stereoformer_b200.dist import SignalledGather
    dev = torch.device("cuda", 0)
    B, D, H, W = 2, 24, 12, 120
    try:
        sg = SignalledGather((B, H, W), dev, dst_ranks=(0,))
    except Exception as e:      # noqa: BLE001
        pytest.skip("symmetric memory unavailable: %r" % (e,))
    torch.manual_seed(0)
    out = torch.empty(B, 1, H, W, device=dev)
    gathered = torch.empty(B, 1, H, W, device=dev)
    for step in range(7):                      # both slots, several wrap-arounds
        cv = torch.randn(B, D, H, W, device=dev)
        want = ops.local_window(cv, None, mode=ops.SF_WIN_FIXED, radius=2, n_points=20, consider_valid=True,
                                div_mode=ops.SF_DIV_RECIP)
        ops.local_window(cv, None, mode=ops.SF_WIN_FIXED, radius=2, n_points=20, consider_valid=True,
                         div_mode=ops.SF_DIV_RECIP, out=out, gather=sg)
        sg.wait(gathered)
        torch.cuda.synchronize()
        assert torch.equal(out, want)
        assert torch.equal(gathered, want), "step %d" % step
        sig = sg.signals.cpu()
        assert int(sig[24]) == step + 1                                     # CONSUMED[0]
        assert int(sig[3]) == 0 and int(sig[4]) == 0                        # election counter reset, no error
    sg.check()


def test_signalled_gather_graph_replay(pg):
    """Epochs live in device memory, so one captured step replays correctly any number of times."""
    from stereoformer_b200 import ops
    from stereoformer_b200.dist import SignalledGather
    dev = torch.device("cuda", 0)
    B, D, H, W = 1, 24, 8, 64
    try:
        sg = SignalledGather((B, H, W), dev, dst_ranks=(0,))
    except Exception as e:      # noqa: BLE001
        pytest.skip("symmetric memory unavailable: %r" % (e,))
    cv = torch.randn(B, D, H, W, device=dev)
    out, gathered = torch.empty(B, 1, H, W, device=dev), torch.empty(B, 1, H, W, device=dev)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            ops.local_window(cv, None, mode=ops.SF_WIN_FIXED, radius=2, n_points=20, consider_valid=True,
                             div_mode=ops.SF_DIV_RECIP, out=out, gather=sg)
            sg.wait(gathered)
    for i in range(5):
        cv.normal_()
        gathered.zero_()
        g.replay()
        torch.cuda.synchronize()
        want = ops.local_window(cv, None, mode=ops.SF_WIN_FIXED, radius=2, n_points=20, consider_valid=True,
                                div_mode=ops.SF_DIV_RECIP)
        assert torch.equal(gathered, want), i
    assert int(sg.signals[24].item()) == 5
    sg.check()
