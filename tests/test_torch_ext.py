"""The torch C++-extension front end (stereoformer_b200/csrc/torch_ext.cpp): same C ABI underneath, so it must agree
bit for bit with the ctypes binding; without a GPU: it loads, exports its functions and refuses CPU tensors."""
import pytest
import torch


def test_extension_loads_and_has_no_cpu_path():
    from stereoformer_b200 import ops
    assert ops._ext is not None, "build it with python -m stereoformer_b200.build"
    for name in ("corr_volume", "soft_argmin", "local_window", "upsample_convex"):
        assert hasattr(ops._ext, name)
    with pytest.raises(RuntimeError, match="Not implemented on the CPU"):
        ops._ext.corr_volume(torch.randn(1, 4, 4, 8), torch.randn(1, 4, 4, 8), 4)


@pytest.mark.gpu
def test_extension_matches_ctypes_binding(monkeypatch):
    from stereoformer_b200 import ops
    ext = ops._ext
    assert ext is not None
    torch.manual_seed(3)
    fl, fr = torch.randn(2, 64, 9, 120, device="cuda"), torch.randn(2, 64, 9, 120, device="cuda")
    mask = torch.randn(2, 576, 9, 120, device="cuda")

    def run():
        cv = ops.corr_volume(fl, fr, 24)
        d0 = ops.soft_argmin(cv)
        out = ops.local_window(cv, d0.unsqueeze(1), mode=ops.SF_WIN_FIXED, radius=2, n_points=20, consider_valid=True,
                               div_mode=ops.SF_DIV_RECIP)
        fused = ops.local_window(cv, None, mode=ops.SF_WIN_FIXED, radius=2, n_points=20, consider_valid=True,
                                 div_mode=ops.SF_DIV_RECIP)
        up = ops.upsample_convex(out, mask, 8, mask_scale=0.25)
        return cv, d0, out, fused, up
    a = run()
    monkeypatch.setattr(ops, "_ext", None)          # the ctypes binding
    b = run()
    for x, y in zip(a, b):
        assert torch.equal(x, y)
    monkeypatch.setattr(ops, "_ext", ext)
    with pytest.raises(ValueError):
        ops.local_window(a[0], torch.zeros(3, device="cuda"), mode=ops.SF_WIN_FIXED, radius=2, n_points=20)
    with pytest.raises(ValueError):
        ops.corr_volume(fl, fr[:, :32].contiguous(), 24)
