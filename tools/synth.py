"""Seeded synthetic stereo inputs: no datasets exist on the GPU box.

left = ImageNet-normalised uniform noise smoothed a little; right[x] = left[x + d(x)] for a smooth random disparity
field d in [0, 0.8 * maxdisp] plus N(0, 0.01) noise, so soft-argmin / the window samplers see realistic, non-degenerate
windows and disp.min() >= 0 holds.
"""
import torch
import torch.nn.functional as F

MEAN = (0.485, 0.456, 0.406)      # train_file/trainer_DLC_GRU.py:24-25
STD = (0.229, 0.224, 0.225)


def stereo_pair(batch, height, width, maxdisp=192, seed=1024, device="cpu"):
    g = torch.Generator(device="cpu").manual_seed(seed)
    low = torch.rand(batch, 3, height // 4, width // 4, generator=g)
    fine = torch.rand(batch, 3, height, width, generator=g)
    left = 0.7 * F.interpolate(low, size=(height, width), mode="bilinear", align_corners=False) + 0.3 * fine
    mean = torch.tensor(MEAN).view(1, 3, 1, 1)
    std = torch.tensor(STD).view(1, 3, 1, 1)
    left = (left - mean) / std
    field = torch.rand(batch, 1, max(2, height // 64), max(2, width // 64), generator=g) * 0.8 * maxdisp
    disp = F.interpolate(field, size=(height, width), mode="bicubic", align_corners=False).clamp_(0, 0.8 * maxdisp)
    xs = torch.arange(width, dtype=torch.float32).view(1, 1, 1, width).expand(batch, 1, height, width)
    ys = torch.arange(height, dtype=torch.float32).view(1, 1, height, 1).expand(batch, 1, height, width)
    # right image content at x is the left content at x + d: sample left at x + d (border-clamped)
    gx = 2.0 * (xs + disp) / (width - 1) - 1.0
    gy = 2.0 * ys / (height - 1) - 1.0
    grid = torch.cat((gx, gy), 1).permute(0, 2, 3, 1)
    right = F.grid_sample(left, grid, mode="bilinear", padding_mode="border", align_corners=True)
    right = right + 0.01 * torch.randn(right.shape, generator=g)
    return left.to(device), right.to(device), disp.to(device)
