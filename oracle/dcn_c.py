"""TEST INFRASTRUCTURE: ctypes loader of the plain-C DCN oracle (oracle/dcn_oracle.c)."""
import ctypes
import os
import subprocess

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "_build", "libdcn_oracle.so")


def build():
    src = os.path.join(HERE, "dcn_oracle.c")
    if not os.path.exists(SO) or os.path.getmtime(SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", HERE, "-s", "-B"])
    return SO


_h = None


def dcn_forward(inp, weight, bias, offset, mask, stride, pad, dil, groups, dgroups):
    """CPU float32 tensors in, CPU tensor out (see dcn_oracle.c for the reference lines followed)."""
    global _h
    if _h is None:
        _h = ctypes.CDLL(build())
        _h.dcn_oracle_forward.restype = ctypes.c_int
        _h.dcn_oracle_forward.argtypes = [ctypes.c_void_p] * 6 + [ctypes.c_int] * 15
    ts = [t.detach().cpu().float().contiguous() if t is not None else None for t in (inp, weight, bias, offset, mask)]
    B, Cin, H, W = ts[0].shape
    Cout, _, kh, kw = ts[1].shape
    Ho = (H + 2 * pad[0] - (dil[0] * (kh - 1) + 1)) // stride[0] + 1
    Wo = (W + 2 * pad[1] - (dil[1] * (kw - 1) + 1)) // stride[1] + 1
    out = torch.empty(B, Cout, Ho, Wo)
    rc = _h.dcn_oracle_forward(*[t.data_ptr() if t is not None else None for t in ts], out.data_ptr(), B, Cin, H, W,
                               Cout, kh, kw, stride[0], stride[1], pad[0], pad[1], dil[0], dil[1], groups, dgroups)
    if rc != 0:
        raise RuntimeError("dcn_oracle_forward failed: %d" % rc)
    return out


def dcn_backward(inp, weight, offset, mask, grad_out, stride, pad, dil, groups, dgroups):
    """CPU float32 tensors in -> (grad_input, grad_offset, grad_mask | None, grad_weight, grad_bias), the order
    the reference's DCN.modulated_deform_conv_backward returns (modulated_deform_conv_cuda.cu:277-279)."""
    global _h
    if _h is None:
        dcn_forward(torch.zeros(1, 1, 3, 3), torch.zeros(1, 1, 1, 1), torch.zeros(1), torch.zeros(1, 2, 3, 3), None,
                    (1, 1), (0, 0), (1, 1), 1, 1)
    _h.dcn_oracle_backward.restype = ctypes.c_int
    _h.dcn_oracle_backward.argtypes = [ctypes.c_void_p] * 10 + [ctypes.c_int] * 15
    ts = [t.detach().cpu().float().contiguous() if t is not None else None for t in (inp, weight, offset, mask, grad_out)]
    B, Cin, H, W = ts[0].shape
    Cout, _, kh, kw = ts[1].shape
    gin, gw, gb = torch.empty_like(ts[0]), torch.empty_like(ts[1]), torch.empty(Cout)
    goff = torch.empty_like(ts[2])
    gm = torch.empty_like(ts[3]) if ts[3] is not None else None
    rc = _h.dcn_oracle_backward(*[t.data_ptr() if t is not None else None for t in ts], gin.data_ptr(), gw.data_ptr(),
                                gb.data_ptr(), goff.data_ptr(), gm.data_ptr() if gm is not None else None, B, Cin, H,
                                W, Cout, kh, kw, stride[0], stride[1], pad[0], pad[1], dil[0], dil[1], groups, dgroups)
    if rc != 0:
        raise RuntimeError("dcn_oracle_backward failed: %d" % rc)
    return gin, goff, gm, gw, gb
