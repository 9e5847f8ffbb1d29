"""nn.Module mirror of deform/modules/*.py and autograd mirror of deform/functions/*_func.py.

Same constructor / forward signatures and parameter names (``weight``, ``bias``, ``conv_offset_mask.*``,
``conv_offset.*``) so reference checkpoints load unchanged.  The forward runs the sm_100a kernel through
``stereoformer_b200.DCN`` (the pybind drop-in), the backward ``sf_dcn_backward_f32``.
"""
import math

import torch
from torch import nn
from torch.autograd import Function
from torch.autograd.function import once_differentiable
from torch.nn import init
from torch.nn.modules.utils import _pair

from . import DCN


class ModulatedDeformConvFunction(Function):
    """deform/functions/modulated_deform_conv_func.py:15-56 (argument order: input, offset, mask, weight, bias)."""

    @staticmethod
    def forward(ctx, input, offset, mask, weight, bias, stride, padding, dilation, groups, deformable_groups,
                im2col_step):
        ctx.stride, ctx.padding, ctx.dilation = _pair(stride), _pair(padding), _pair(dilation)
        ctx.kernel_size = _pair(weight.shape[2:4])
        ctx.groups, ctx.deformable_groups, ctx.im2col_step = groups, deformable_groups, im2col_step
        s, p, d, k = ctx.stride, ctx.padding, ctx.dilation, ctx.kernel_size
        output = DCN.modulated_deform_conv_forward(input, weight, bias, offset, mask, k[0], k[1], s[0], s[1], p[0],
                                                   p[1], d[0], d[1], groups, deformable_groups, im2col_step)
        ctx.save_for_backward(input, offset, mask, weight, bias)
        return output

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_output):
        input, offset, mask, weight, bias = ctx.saved_tensors
        s, p, d, k = ctx.stride, ctx.padding, ctx.dilation, ctx.kernel_size
        grad_input, grad_offset, grad_mask, grad_weight, grad_bias = DCN.modulated_deform_conv_backward(
            input, weight, bias, offset, mask, grad_output, k[0], k[1], s[0], s[1], p[0], p[1], d[0], d[1], ctx.groups,
            ctx.deformable_groups, ctx.im2col_step)
        return grad_input, grad_offset, grad_mask, grad_weight, grad_bias, None, None, None, None, None, None


class DeformConvFunction(Function):
    """deform/functions/deform_conv_func.py:15-75."""

    @staticmethod
    def forward(ctx, input, offset, weight, bias, stride, padding, dilation, group, deformable_groups, im2col_step):
        ctx.stride, ctx.padding, ctx.dilation = _pair(stride), _pair(padding), _pair(dilation)
        ctx.kernel_size = _pair(weight.shape[2:4])
        ctx.group, ctx.deformable_groups, ctx.im2col_step = group, deformable_groups, im2col_step
        s, p, d, k = ctx.stride, ctx.padding, ctx.dilation, ctx.kernel_size
        output = DCN.deform_conv_forward(input, weight, bias, offset, k[0], k[1], s[0], s[1], p[0], p[1], d[0], d[1],
                                         group, deformable_groups, im2col_step)
        ctx.save_for_backward(input, offset, weight, bias)
        return output

    @staticmethod
    @once_differentiable
    def backward(ctx, grad_output):
        input, offset, weight, bias = ctx.saved_tensors
        s, p, d, k = ctx.stride, ctx.padding, ctx.dilation, ctx.kernel_size
        grad_input, grad_offset, grad_weight, grad_bias = DCN.deform_conv_backward(
            input, weight, bias, offset, grad_output, k[0], k[1], s[0], s[1], p[0], p[1], d[0], d[1], ctx.group,
            ctx.deformable_groups, ctx.im2col_step)
        return grad_input, grad_offset, grad_weight, grad_bias, None, None, None, None, None, None


class _DeformBase(nn.Module):
    def __init__(self, in_channels, out_channels, kernel_size, stride, padding, dilation=1, groups=1,
                 deformable_groups=1, im2col_step=64, bias=True):
        super().__init__()
        if in_channels % groups != 0:
            raise ValueError('in_channels {} must be divisible by groups {}'.format(in_channels, groups))
        if out_channels % groups != 0:
            raise ValueError('out_channels {} must be divisible by groups {}'.format(out_channels, groups))
        self.in_channels = in_channels
        self.out_channels = out_channels
        self.kernel_size = _pair(kernel_size)
        self.stride = _pair(stride)
        self.padding = _pair(padding)
        self.dilation = _pair(dilation)
        self.groups = groups
        self.deformable_groups = deformable_groups
        self.im2col_step = im2col_step
        self.use_bias = bias
        self.weight = nn.Parameter(torch.Tensor(out_channels, in_channels // groups, *self.kernel_size))
        self.bias = nn.Parameter(torch.Tensor(out_channels))
        self.reset_parameters()
        if not self.use_bias:
            self.bias.requires_grad = False

    def reset_parameters(self):
        init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        fan_in, _ = init._calculate_fan_in_and_fan_out(self.weight)
        bound = 1 / math.sqrt(fan_in)
        init.uniform_(self.bias, -bound, bound)

    def _taps(self):
        return self.deformable_groups * self.kernel_size[0] * self.kernel_size[1]


class ModulatedDeformConv(_DeformBase):
    """deform/modules/modulated_deform_conv.py:14-64."""

    def forward(self, input, offset, mask):
        assert 2 * self._taps() == offset.shape[1]
        assert self._taps() == mask.shape[1]
        return ModulatedDeformConvFunction.apply(input, offset, mask, self.weight, self.bias, self.stride, self.padding,
                                                 self.dilation, self.groups, self.deformable_groups, self.im2col_step)


_ModulatedDeformConv = ModulatedDeformConvFunction.apply


class ModulatedDeformConvPack(ModulatedDeformConv):
    """deform/modules/modulated_deform_conv.py:68-103: offsets and mask from a zero-initialised conv."""

    def __init__(self, in_channels, out_channels, kernel_size, stride, padding, dilation=1, groups=1,
                 deformable_groups=1, im2col_step=64, bias=True, lr_mult=0.1, offset_channels=None):
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, deformable_groups,
                         im2col_step, bias)
        self.conv_offset_mask = nn.Conv2d(self.in_channels if offset_channels is None else offset_channels,
                                          3 * self._taps(), kernel_size=self.kernel_size, stride=self.stride,
                                          padding=self.padding, bias=True)
        self.conv_offset_mask.lr_mult = lr_mult
        self.init_offset()

    def init_offset(self):
        self.conv_offset_mask.weight.data.zero_()
        self.conv_offset_mask.bias.data.zero_()

    def _offset_mask(self, feature):
        out = self.conv_offset_mask(feature)
        o1, o2, mask = torch.chunk(out, 3, dim=1)
        return torch.cat((o1, o2), dim=1), torch.sigmoid(mask)

    def forward(self, input):
        offset, mask = self._offset_mask(input)
        return ModulatedDeformConvFunction.apply(input, offset, mask, self.weight, self.bias, self.stride, self.padding,
                                                 self.dilation, self.groups, self.deformable_groups, self.im2col_step)


class ModulatedDeformNormal(ModulatedDeformConvPack):
    """deform/modules/modulated_deform_conv.py:106-151: offsets predicted from a second (normal) feature."""

    def __init__(self, offset_channel, in_channels, out_channels, kernel_size, stride, padding, dilation=1, groups=1,
                 deformable_groups=1, im2col_step=64, bias=True, lr_mult=0.1):
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, deformable_groups,
                         im2col_step, bias, lr_mult, offset_channels=offset_channel)

    def forward(self, input_feature, normal_feature):
        offset, mask = self._offset_mask(normal_feature)
        return ModulatedDeformConvFunction.apply(input_feature, offset, mask, self.weight, self.bias, self.stride,
                                                 self.padding, self.dilation, self.groups, self.deformable_groups,
                                                 self.im2col_step)


class DeformConv(_DeformBase):
    """deform/modules/deform_conv.py:14-60."""

    def forward(self, input, offset):
        assert 2 * self._taps() == offset.shape[1]
        return DeformConvFunction.apply(input, offset, self.weight, self.bias, self.stride, self.padding, self.dilation,
                                        self.groups, self.deformable_groups, self.im2col_step)


_DeformConv = DeformConvFunction.apply


class DeformConvPack(DeformConv):
    """deform/modules/deform_conv.py:65-101."""

    def __init__(self, in_channels, out_channels, kernel_size, stride, padding, dilation=1, groups=1,
                 deformable_groups=1, im2col_step=64, bias=True, lr_mult=0.1):
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, deformable_groups,
                         im2col_step, bias)
        self.conv_offset = nn.Conv2d(self.in_channels, 2 * self._taps(), kernel_size=self.kernel_size,
                                     stride=self.stride, padding=self.padding, bias=True)
        self.conv_offset.lr_mult = lr_mult
        self.init_offset()

    def init_offset(self):
        self.conv_offset.weight.data.zero_()
        self.conv_offset.bias.data.zero_()

    def forward(self, input):
        offset = self.conv_offset(input)
        return DeformConvFunction.apply(input, offset, self.weight, self.bias, self.stride, self.padding, self.dilation,
                                        self.groups, self.deformable_groups, self.im2col_step)
