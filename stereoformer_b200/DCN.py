"""Drop-in for the reference's pybind extension module ``DCN`` (deform/src/vision.cpp:5-12).

Same six function names and positional signatures as the pybind module; the forward functions run
the sm_100a fused implicit-GEMM kernel through the C ABI (``sf_dcn_forward_f32``), the backward functions
``sf_dcn_backward_f32`` (no column buffer either way).  Register it under
the name the reference imports with ``stereoformer_b200.install.install_dcn()`` (``import DCN`` in
deform/functions/*_func.py:13 then resolves to this module).

Error behaviour mirrors deform/src/modulated_deform_conv.h:28-43 and modulated_deform_conv_cuda.cu:39-73:
CPU tensors -> RuntimeError("Not implemented on the CPU"); non-contiguous input/weight, batch not
divisible by min(batch, im2col_step), channel / kernel-shape mismatches -> RuntimeError.
"""
import torch

from . import ops


def _checks(input, weight, bias, offset, kernel_h, kernel_w, group, im2col_step, mask=None):
    if not input.is_cuda:
        raise RuntimeError("Not implemented on the CPU")
    for t, n in ((weight, "weight"), (bias, "bias"), (offset, "offset")) + (((mask, "mask"),) if mask is not None else ()):
        if not t.is_cuda:
            raise RuntimeError("%s must be a CUDA tensor" % n)
    if not input.is_contiguous():
        raise RuntimeError("input tensor has to be contiguous")
    if not weight.is_contiguous():
        raise RuntimeError("weight tensor has to be contiguous")
    if input.dtype not in (torch.float32, torch.float64):   # AT_DISPATCH_FLOATING_TYPES, modulated_deform_conv_cuda.cu:93
        raise RuntimeError("stereoformer_b200 DCN: float32 and float64 are implemented (got %s)" % input.dtype)
    batch, channels = input.shape[0], input.shape[1]
    step = min(batch, im2col_step)
    if batch % step != 0:  # modulated_deform_conv_cuda.cu:58-60 (im2col_step otherwise never changes results)
        raise RuntimeError("batch(%d) must divide im2col_step(%d)" % (batch, step))
    cout, ck, kh_, kw_ = weight.shape
    if channels % group != 0 or cout % group != 0:
        raise RuntimeError("channels(%d) and channels_out(%d) must divide group(%d)" % (channels, cout, group))
    if kh_ != kernel_h or kw_ != kernel_w:
        raise RuntimeError("Input shape and kernel shape wont match: (%d x %d vs %d x %d)." % (kernel_h, kernel_w, kh_, kw_))
    if channels != ck * group:
        raise RuntimeError("Input shape and kernel channels wont match: (%d vs %d)." % (channels, ck * group))


def modulated_deform_conv_forward(input, weight, bias, offset, mask, kernel_h, kernel_w, stride_h, stride_w, pad_h,
                                  pad_w, dilation_h, dilation_w, group, deformable_group, im2col_step):
    """deform/src/modulated_deform_conv.h:10-44 (note the C++ argument order: weight, bias before offset, mask)."""
    _checks(input, weight, bias, offset, kernel_h, kernel_w, group, im2col_step, mask)
    return ops.dcn_forward(input, weight, bias, offset, mask, (stride_h, stride_w), (pad_h, pad_w),
                           (dilation_h, dilation_w), group, deformable_group)


def deform_conv_forward(input, weight, bias, offset, kernel_h, kernel_w, stride_h, stride_w, pad_h, pad_w,
                        dilation_h, dilation_w, group, deformable_group, im2col_step):
    """deform/src/deform_conv.h:10-43 (DCNv1: no modulation mask)."""
    _checks(input, weight, bias, offset, kernel_h, kernel_w, group, im2col_step)
    return ops.dcn_forward(input, weight, bias, offset, None, (stride_h, stride_w), (pad_h, pad_w),
                           (dilation_h, dilation_w), group, deformable_group)


def _not_built(name, why):
    def f(*a, **k):
        raise NotImplementedError("DCN.%s: %s" % (name, why))
    f.__name__ = name
    return f


def _check_grad(input, weight, grad_output, stride, pad, dil):
    """modulated_deform_conv_cuda.cu:163-192: grad_output must match the forward's output shape."""
    if not grad_output.is_cuda:
        raise RuntimeError("grad_output must be a CUDA tensor")
    kh, kw = weight.shape[2:]
    ho = (input.shape[2] + 2 * pad[0] - (dil[0] * (kh - 1) + 1)) // stride[0] + 1
    wo = (input.shape[3] + 2 * pad[1] - (dil[1] * (kw - 1) + 1)) // stride[1] + 1
    if grad_output.shape[0] != input.shape[0]:
        raise RuntimeError("Input shape and grad_out batch wont match: (%d vs %d)." % (input.shape[0], grad_output.shape[0]))
    if grad_output.shape[1] != weight.shape[0]:
        raise RuntimeError("Input shape and grad_out channels_out wont match: (%d vs %d)." % (weight.shape[0], grad_output.shape[1]))
    if tuple(grad_output.shape[2:]) != (ho, wo):
        raise RuntimeError("Input shape and grad_out shape wont match: (%d x %d vs %d x %d)." %
                           (ho, wo, grad_output.shape[2], grad_output.shape[3]))


def modulated_deform_conv_backward(input, weight, bias, offset, mask, grad_output, kernel_h, kernel_w, stride_h,
                                   stride_w, pad_h, pad_w, dilation_h, dilation_w, group, deformable_group, im2col_step):
    """deform/src/modulated_deform_conv.h:46-92 -> [grad_input, grad_offset, grad_mask, grad_weight, grad_bias]
    (modulated_deform_conv_cuda.cu:277-279)."""
    _checks(input, weight, bias, offset, kernel_h, kernel_w, group, im2col_step, mask)
    _check_grad(input, weight, grad_output, (stride_h, stride_w), (pad_h, pad_w), (dilation_h, dilation_w))
    return list(ops.dcn_backward(input, weight, offset, mask, grad_output.contiguous(), (stride_h, stride_w),
                                 (pad_h, pad_w), (dilation_h, dilation_w), group, deformable_group))


def deform_conv_backward(input, weight, bias, offset, grad_output, kernel_h, kernel_w, stride_h, stride_w, pad_h, pad_w,
                         dilation_h, dilation_w, group, deformable_group, im2col_step):
    """deform/src/deform_conv.h:45-88 -> [grad_input, grad_offset, grad_weight, grad_bias] (DCNv1)."""
    _checks(input, weight, bias, offset, kernel_h, kernel_w, group, im2col_step)
    _check_grad(input, weight, grad_output, (stride_h, stride_w), (pad_h, pad_w), (dilation_h, dilation_w))
    g_in, g_off, _, g_w, g_b = ops.dcn_backward(input, weight, offset, None, grad_output.contiguous(),
                                                (stride_h, stride_w), (pad_h, pad_w), (dilation_h, dilation_w), group,
                                                deformable_group)
    return [g_in, g_off, g_w, g_b]


deform_psroi_pooling_forward = _not_built("deform_psroi_pooling_forward",
                                          "deformable PS-RoI pooling is out of scope (detection op, no stereo caller)")
deform_psroi_pooling_backward = _not_built("deform_psroi_pooling_backward",
                                           "deformable PS-RoI pooling is out of scope (detection op, no stereo caller)")
