"""DLC-GRU refinement step and its 12-iteration loop (the caller of the window sampler).

Mirror of models/LocalCostVolume/Attempts/update.py (ConvGRU :16-49, BasicGuidanceNet :52-90,
OffsetPredictionHead :94-105, DisparityUpdateDLC :110-180, DisparityUpdateDLCWithMask :243-317) with the same
class names, constructor arguments, parameter names (state dicts load unchanged) and forward signatures.  The
convolutions stay cuDNN; everything between them is a hand-written kernel:

  * guidance prologue (soft-max -> uncertainty volume, warp, error map)          sf_guidance_f32
  * ConvGRU gates (sigmoid, product, concatenation / sigmoid, tanh, blend)       sf_gru_reset_f32, sf_gru_blend_f32
  * OffsetPredictionHead's last ReLU + chunk + window sampling + expectation     sf_local_window_f32(SF_WIN_HEAD)
  * `.25 * mask` + convex x8 upsampling                                          sf_upsample_convex_scaled_f32

`DlcGruRefiner` runs the loop of baseline_dynamic_gru.py:128-139 with the loop invariants hoisted (image
resizing) and, for inference, the whole loop captured in one CUDA graph (no tracing compiler involved: plain
stream capture of the launches above plus cuDNN's).
"""
import torch
import torch.nn as nn
import torch.nn.functional as F
import torch.nn.init as init

from . import ops
from ._lib import SF_WIN_HEAD
from . import modules as _m


class ConvGRU(nn.Module):
    """Attempts/update.py:16-49.  conv_zz / conv_bb / conv_gg are kept so reference state dicts load strictly."""

    def __init__(self, nb_channel, softsign):
        super().__init__()
        self.conv_z = nn.Conv2d(nb_channel * 2, nb_channel, 3, 1, 1)
        self.conv_b = nn.Conv2d(nb_channel * 2, nb_channel, 3, 1, 1)
        self.conv_g = nn.Conv2d(nb_channel * 2, nb_channel, 3, 1, 1)
        for c in (self.conv_z, self.conv_b, self.conv_g):
            init.orthogonal_(c.weight)
            init.constant_(c.bias, 0.)
        self.conv_zz = nn.Sequential(self.conv_z, nn.Sigmoid())
        self.conv_bb = nn.Sequential(self.conv_b, nn.Sigmoid())
        self.conv_gg = nn.Sequential(self.conv_g, nn.Softsign() if softsign else nn.Tanh())
        self.softsign = bool(softsign)
        self.nb_channel = nb_channel

    def forward(self, input, prev_h):
        C = self.nb_channel
        if prev_h is None:
            x1 = torch.cat((input, input.new_zeros(input.size(0), C, input.size(2), input.size(3))), 1)
        else:
            x1 = torch.cat((input, prev_h), 1)
        # both gates read x1: one convolution with the two filter banks stacked
        zb = F.conv2d(x1, torch.cat((self.conv_z.weight, self.conv_b.weight), 0),
                      torch.cat((self.conv_z.bias, self.conv_b.bias), 0), 1, 1)
        s = _m._run(lambda b, h, x: ops.gru_reset(b, h, x), "ConvGRU", zb[:, C:], prev_h, input,
                    bwd=lambda a, gs, n: ops.gru_reset_backward(a[0], a[1], a[2].shape, gs, n))
        g = self.conv_g(s)
        softsign = getattr(self, "softsign", None)
        if softsign is None:                 # a reference instance carrying this forward (install.py)
            softsign = isinstance(self.conv_gg[1], nn.Softsign)
        return _m._run(lambda z, gg, h: ops.gru_blend(z, gg, h, softsign), "ConvGRU", zb[:, :C], g, prev_h,
                       bwd=lambda a, gh, n: ops.gru_blend_backward(a[0], a[1], a[2], gh, softsign, n))


class BasicGuidanceNet(nn.Module):
    """Attempts/update.py:52-90."""

    def __init__(self, input_channels, hidden_layer=32):
        super().__init__()
        self.input_channels = input_channels
        self.hidden_layers = hidden_layer
        self.uncertain_encoder = nn.Sequential(
            nn.Conv2d(self.input_channels, self.hidden_layers, kernel_size=3, stride=1, padding=1, bias=False),
            nn.BatchNorm2d(self.hidden_layers), nn.ReLU(inplace=True))
        self.disparity_error_encoder = nn.Sequential(
            nn.Conv2d(3, self.hidden_layers, kernel_size=3, stride=1, padding=1, bias=False),
            nn.BatchNorm2d(self.hidden_layers), nn.ReLU(inplace=True))

    @staticmethod
    def resize_images(cur_disp, left_image, right_image):
        """The loop-invariant part of forward (update.py:74-81): images at the disparity resolution."""
        if cur_disp.size(-1) != left_image.size(-1):
            size = [cur_disp.size(-2), cur_disp.size(-1)]
            return (F.interpolate(left_image, size=size, mode='bilinear', align_corners=False),
                    F.interpolate(right_image, size=size, mode='bilinear', align_corners=False))
        return left_image, right_image

    def forward(self, old_cost_volume, cur_disp, left_image, right_image):
        cur_left, cur_right = BasicGuidanceNet.resize_images(cur_disp, left_image, right_image)
        error_map, uncertainty_volume = _m.guidance_inputs(old_cost_volume, cur_disp, cur_left, cur_right)
        error_feature = self.disparity_error_encoder(error_map)
        uncertainty_feature = self.uncertain_encoder(uncertainty_volume)
        return torch.cat([error_feature, uncertainty_feature], dim=1)


class OffsetPredictionHead(nn.Module):
    """Attempts/update.py:94-105."""

    def __init__(self, input_dim, hidden_dim):
        super().__init__()
        self.conv1 = nn.Conv2d(input_dim, hidden_dim, 3, padding=1)
        self.conv2 = nn.Conv2d(hidden_dim, 2, 3, padding=1)
        self.relu1 = nn.ReLU(inplace=True)
        self.relu2 = nn.ReLU(inplace=True)

    def forward_pre(self, x):
        """conv2 output before relu2: the window kernel applies the ReLU while loading the bounds."""
        return self.conv2(self.relu1(self.conv1(x)))

    def forward(self, x):
        return self.relu2(self.forward_pre(x))


def _head_window(old_cost_volume, cur_disp, bounds_pre, sample_points, consider_valid):
    return _m._run(lambda cv, d, bp: ops.local_window(cv, d, lo=bp, mode=SF_WIN_HEAD, n_points=sample_points,
                                                      consider_valid=consider_valid, div_mode=_m.DEFAULT_DIV_MODE),
                   "DisparityUpdateDLC", old_cost_volume, cur_disp, bounds_pre,
                   bwd=lambda a, g, n: ops.local_window_backward(a[0], a[1], a[2], None, g, SF_WIN_HEAD, 0.0,
                                                                 sample_points, consider_valid, _m.DEFAULT_DIV_MODE,
                                                                 (n[0], n[1], n[2], False))[:3])


class DisparityUpdateDLC(nn.Module):
    """Attempts/update.py:110-180 (the class is defined twice there, identically)."""

    def __init__(self, input_channels, hidden_dim, sample_points=10):
        super().__init__()
        self.sample_points = sample_points
        self.hidden_dim = hidden_dim
        self.encoder = BasicGuidanceNet(input_channels=input_channels, hidden_layer=self.hidden_dim)
        self.gru = ConvGRU(nb_channel=self.hidden_dim * 2, softsign=False)
        self.offset = OffsetPredictionHead(input_dim=32, hidden_dim=64)

    def forward(self, old_cost_volume, cur_disp, left_image, right_image, hidden_state=None, consider_valid=True):
        offset_features = self.encoder(old_cost_volume, cur_disp, left_image, right_image)
        hidden_state = self.gru(offset_features, hidden_state)
        disp = _head_window(old_cost_volume, cur_disp, OffsetPredictionHead.forward_pre(self.offset, hidden_state),
                            self.sample_points, consider_valid)
        return disp, hidden_state


class DisparityUpdateDLCWithMask(nn.Module):
    """Attempts/update.py:243-317."""

    MASK_SCALE = .25

    def __init__(self, input_channels, hidden_dim, sample_points=10):
        super().__init__()
        self.sample_points = sample_points
        self.hidden_dim = hidden_dim
        self.encoder = BasicGuidanceNet(input_channels=input_channels, hidden_layer=self.hidden_dim)
        self.gru = ConvGRU(nb_channel=self.hidden_dim * 2, softsign=False)
        self.offset = OffsetPredictionHead(input_dim=self.hidden_dim * 2, hidden_dim=64)
        self.mask = nn.Sequential(nn.Conv2d(self.hidden_dim * 2, 256, 3, padding=1), nn.ReLU(inplace=True),
                                  nn.Conv2d(256, 64 * 9, 1, padding=0))

    def step(self, old_cost_volume, cur_disp, left_image, right_image, hidden_state, consider_valid):
        """One refinement step -> (disp, hidden_state, UNSCALED mask logits)."""
        offset_features = self.encoder(old_cost_volume, cur_disp, left_image, right_image)
        hidden_state = self.gru(offset_features, hidden_state)
        mask = self.mask(hidden_state)
        disp = _head_window(old_cost_volume, cur_disp, OffsetPredictionHead.forward_pre(self.offset, hidden_state),
                            self.sample_points, consider_valid)
        return disp, hidden_state, mask

    def forward(self, old_cost_volume, cur_disp, left_image, right_image, hidden_state=None, consider_valid=True):
        disp, hidden_state, mask = DisparityUpdateDLCWithMask.step(self, old_cost_volume, cur_disp, left_image,
                                                                   right_image, hidden_state, consider_valid)
        return disp, hidden_state, DisparityUpdateDLCWithMask.MASK_SCALE * mask


class DisparityUpdateDLCWithMaskV2(DisparityUpdateDLCWithMask):
    """Attempts/updatev2.py:168-251: the GRU also sees a 256 -> 64 encoding of the left 1/8 feature map."""

    def __init__(self, input_channels, feature_dim, hidden_dim, sample_points=10):
        nn.Module.__init__(self)
        self.sample_points = sample_points
        self.hidden_dim = hidden_dim
        self.feature_dim = feature_dim
        width = self.hidden_dim * 2 + self.feature_dim
        self.encoder = BasicGuidanceNet(input_channels=input_channels, hidden_layer=self.hidden_dim)
        self.feature_encode = nn.Sequential(nn.Conv2d(256, 64, kernel_size=3, stride=1, padding=1, bias=False),
                                            nn.BatchNorm2d(64), nn.ReLU(inplace=True))
        self.gru = ConvGRU(nb_channel=width, softsign=False)
        self.offset = OffsetPredictionHead(input_dim=width, hidden_dim=64)
        self.mask = nn.Sequential(nn.Conv2d(width, 256, 3, padding=1), nn.ReLU(inplace=True),
                                  nn.Conv2d(256, 64 * 9, 1, padding=0))

    def step(self, old_cost_volume, cur_disp, left_image, right_image, hidden_state, left_feature, consider_valid):
        offset_features = self.encoder(old_cost_volume, cur_disp, left_image, right_image)
        input_features = torch.cat((offset_features, self.feature_encode(left_feature)), dim=1)
        hidden_state = self.gru(input_features, hidden_state)
        mask = self.mask(hidden_state)
        disp = _head_window(old_cost_volume, cur_disp, OffsetPredictionHead.forward_pre(self.offset, hidden_state),
                            self.sample_points, consider_valid)
        return disp, hidden_state, mask

    def forward(self, old_cost_volume, cur_disp, left_image, right_image, hidden_state=None, left_feature=None,
                consider_valid=True):
        disp, hidden_state, mask = DisparityUpdateDLCWithMaskV2.step(
            self, old_cost_volume, cur_disp, left_image, right_image, hidden_state, left_feature, consider_valid)
        return disp, hidden_state, DisparityUpdateDLCWithMask.MASK_SCALE * mask


# ----------------------------------------------------------------------------- single-shot DLC (SmallUNet bounds)
class _ResBlock(nn.Module):
    """models/BasicBlocks/resnet.py:20-48 (parameter names kept)."""
    conv2_factory = staticmethod(lambda n: nn.Conv2d(n, n, kernel_size=3, padding=1))

    def __init__(self, n_in, n_out, kernel_size=3, stride=1):
        super().__init__()
        self.conv1 = nn.Conv2d(n_in, n_out, kernel_size=kernel_size, stride=stride, padding=1)
        self.bn1 = nn.BatchNorm2d(n_out)
        self.relu = nn.ReLU(inplace=True)
        self.conv2 = self.conv2_factory(n_out)
        self.bn2 = nn.BatchNorm2d(n_out)
        self.shortcut = None
        if stride != 1 or n_out != n_in:
            self.shortcut = nn.Sequential(nn.Conv2d(n_in, n_out, kernel_size=1, stride=stride), nn.BatchNorm2d(n_out))

    def forward(self, x):
        residual = x if self.shortcut is None else self.shortcut(x)
        out = self.bn2(self.conv2(self.relu(self.bn1(self.conv1(x)))))
        out += residual
        return self.relu(out)


class _DeformBlock(_ResBlock):
    """models/BasicBlocks/resnet.py:51-79: the second convolution is a ModulatedDeformConvPack (a10)."""

    @staticmethod
    def conv2_factory(n):
        from .dcn_modules import ModulatedDeformConvPack
        return ModulatedDeformConvPack(n, n, kernel_size=3, stride=1, padding=1)


class SmallUNet(nn.Module):
    """Attempts/pure_learning_method.py:35-75."""

    def __init__(self, input_channels, hidden_layer=32):
        super().__init__()
        self.input_channels = input_channels
        self.hidden_layers = hidden_layer
        self.uncertain_encoder = nn.Sequential(
            nn.Conv2d(self.input_channels, self.hidden_layers, kernel_size=3, stride=1, padding=1, bias=False),
            nn.BatchNorm2d(self.hidden_layers), nn.ReLU(inplace=True))
        self.disparity_error_encoder = nn.Sequential(
            nn.Conv2d(3, self.hidden_layers, kernel_size=3, stride=1, padding=1, bias=False),
            nn.BatchNorm2d(self.hidden_layers), nn.ReLU(inplace=True))
        self.feature_fusion1 = _ResBlock(self.hidden_layers * 2, self.hidden_layers, 3, 1)
        self.feature_fusion2 = _DeformBlock(self.hidden_layers, self.hidden_layers // 2, 3, 1)
        self.offset_prediction = nn.Sequential(
            nn.Conv2d(self.hidden_layers // 2, 2, kernel_size=3, padding=1, stride=1), nn.ReLU(inplace=True))

    def forward(self, disparity_error, uncertainity_volume):
        info = torch.cat((self.disparity_error_encoder(disparity_error),
                          self.uncertain_encoder(uncertainity_volume)), dim=1)
        offset = self.offset_prediction(self.feature_fusion2(self.feature_fusion1(info)))
        return torch.chunk(offset, 2, dim=1)


class PureLearningLocalCostVolume(nn.Module):
    """Attempts/pure_learning_method.py:78-143: the two SmallUNet outputs ARE the window bounds (:112).
    with_bounds=True is pure_learning_method_w_super.py: bounds = cur_disp -/+ outputs (:111-113), also returned (:145)."""

    with_bounds = False

    def __init__(self, sample_points=10, with_bounds=False):
        super().__init__()
        self.sample_points = sample_points
        self.with_bounds = with_bounds
        self.offset_prediction_network = SmallUNet(input_channels=24, hidden_layer=32)

    def forward(self, old_cost_volume, cur_disp, left_image, right_image, consider_valid=False):
        cur_left, cur_right = BasicGuidanceNet.resize_images(cur_disp, left_image, right_image)
        error_map, uncertainty_volume = _m.guidance_inputs(old_cost_volume, cur_disp, cur_left, cur_right)
        first, second = self.offset_prediction_network(error_map, uncertainty_volume)
        if not getattr(self, "with_bounds", False):     # reference instances carry this forward (install.py)
            return _m.dlc_window(old_cost_volume, cur_disp, first, second, self.sample_points, consider_valid,
                                 absolute=True)
        disp = _m.dlc_window(old_cost_volume, cur_disp, first, second, self.sample_points, consider_valid)
        lower_bound, upper_bound = cur_disp - first, cur_disp + second
        if not consider_valid:                                     # :120-121: the returned bounds are the clamped ones
            lower_bound = torch.clamp(lower_bound, min=0)
            upper_bound = torch.clamp(upper_bound, min=0, max=old_cost_volume.size(1))
        return disp, lower_bound, upper_bound


class PureLearningLocalCostVolumeWithBounds(PureLearningLocalCostVolume):
    """Attempts/pure_learning_method_w_super.py:78-145."""
    with_bounds = True

    def __init__(self, sample_points=10):
        super().__init__(sample_points, with_bounds=True)


class DlcGruRefiner:
    """The refinement loop of LowCNN.forward, baseline_dynamic_gru.py:128-139, around a DisparityUpdateDLCWithMask.

    refine(final_cost, disp0, left, right, iters) -> list of `iters` full-resolution disparities [B,1,8H,8W].
    Loop invariants (image resizing) are computed once; `.25 * mask` is folded into the upsampling kernel.  With
    use_graph=True (inference) the whole loop is stream-captured once per input shape into a CUDA graph and replayed
    on static buffers; the returned tensors are then views of those buffers (valid until the next call).
    """

    def __init__(self, update, use_graph=True):
        if not isinstance(update, DisparityUpdateDLCWithMask):
            raise TypeError("update must be a stereoformer_b200.update.DisparityUpdateDLCWithMask")
        self.update = update
        self.use_graph = use_graph
        self._graphs = {}

    def _loop(self, cost, disp0, left, right, iters, consider_valid=True):
        left_r, right_r = BasicGuidanceNet.resize_images(disp0, left, right)
        disp, hidden, outs = disp0, None, []
        for _ in range(iters):
            disp, hidden, mask = self.update.step(cost, disp, left_r, right_r, hidden, consider_valid)
            outs.append(ops.upsample_convex(disp, mask, 8, mask_scale=self.update.MASK_SCALE))
        return outs

    @torch.no_grad()
    def refine(self, final_cost, disp0, left, right, iters=12):
        if not self.use_graph:
            return self._loop(final_cost, disp0, left, right, iters)
        if self.update.training:
            raise RuntimeError("DlcGruRefiner: graph replay is inference-only; call update.eval() or use_graph=False")
        key = (tuple(final_cost.shape), tuple(left.shape), iters, final_cost.device)
        ent = self._graphs.get(key)
        if ent is None:
            static = [torch.empty_like(t) for t in (final_cost, disp0, left, right)]
            for s, t in zip(static, (final_cost, disp0, left, right)):
                s.copy_(t)
            side = torch.cuda.Stream(device=final_cost.device)
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(2):                       # warm-up: cuDNN algorithm selection must not be captured
                    self._loop(*static, iters)
                side.synchronize()
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, stream=side):
                    outs = self._loop(*static, iters)
            torch.cuda.current_stream().wait_stream(side)
            ent = self._graphs[key] = (graph, static, outs)
        graph, static, outs = ent
        for s, t in zip(static, (final_cost, disp0, left, right)):
            s.copy_(t)
        graph.replay()
        return outs
