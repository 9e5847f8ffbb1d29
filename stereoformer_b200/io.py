"""On-disk formats either side of the hot path: PFM / KITTI-PNG / NPY disparity maps, RGB
images, filename lists and reference checkpoints, so that real SceneFlow / KITTI / Middlebury pairs and the
reference's trained weights can be fed to the drop-ins when a dataset is mounted.  Host-side only (numpy).

Mirrors utils/file_io.py:11-104 (read_img, read_disp, _read_pfm, write_pfm, _read_kitti_disp), utils/utils.py:9-13
(read_text_lines), :73-100 (load_pretrained_net) and the sample-list construction of
dataloader/SceneflowLoader.py:73-102: same function names, arguments, return types and exceptions.
"""
import os
import re
import sys
from collections import OrderedDict

import numpy as np


def read_img(filename):
    """utils/file_io.py:11-14: RGB float32 [H,W,3]."""
    from PIL import Image
    return np.array(Image.open(filename).convert('RGB')).astype(np.float32)


def _read_pfm(file):
    """utils/file_io.py:34-69 -> (data, scale); rows bottom-to-top in the file, returned top-to-bottom."""
    with open(file, 'rb') as f:
        header = f.readline().rstrip().decode("ascii")
        if header == 'PF':
            color = True
        elif header == 'Pf':
            color = False
        else:
            raise Exception('Not a PFM file.')
        dim = re.match(r'^(\d+)\s(\d+)\s$', f.readline().decode("ascii"))
        if not dim:
            raise Exception('Malformed PFM header.')
        width, height = (int(v) for v in dim.groups())
        scale = float(f.readline().decode("ascii").rstrip())
        endian = '<' if scale < 0 else '>'
        data = np.frombuffer(f.read(), dtype=endian + 'f4')
    shape = (height, width, 3) if color else (height, width)
    return np.flipud(np.reshape(data, shape)), abs(scale)


read_pfm = _read_pfm


def write_pfm(file, image, scale=1):
    """utils/file_io.py:72-98 (also utils/common.py:67-96): byte-identical output."""
    if image.dtype.name != 'float32':
        raise Exception('Image dtype must be float32.')
    image = np.flipud(image)
    if len(image.shape) == 3 and image.shape[2] == 3:
        color = True
    elif len(image.shape) == 2 or len(image.shape) == 3 and image.shape[2] == 1:
        color = False
    else:
        raise Exception('Image must have H x W x 3, H x W x 1 or H x W dimensions.')
    endian = image.dtype.byteorder
    if endian == '<' or endian == '=' and sys.byteorder == 'little':
        scale = -scale
    with open(file, 'wb') as f:
        f.write(b'PF\n' if color else b'Pf\n')
        f.write(b'%d %d\n' % (image.shape[1], image.shape[0]))
        f.write(b'%f\n' % scale)
        f.write(np.ascontiguousarray(image).tobytes())


def _read_kitti_disp(filename):
    """utils/file_io.py:101-104: 16-bit PNG / 256."""
    from PIL import Image
    return np.array(Image.open(filename)).astype(np.float32) / 256.


def read_disp(filename, subset=False):
    """utils/file_io.py:17-31 -> [H,W] float32."""
    if filename.endswith('pfm'):
        disp = np.ascontiguousarray(_read_pfm(filename)[0])
        if subset:
            disp = -disp
    elif filename.endswith('png'):
        disp = _read_kitti_disp(filename)
    elif filename.endswith('npy'):
        disp = np.load(filename)
    else:
        raise Exception('Invalid disparity file format!')
    return disp


def read_text_lines(filepath):
    """utils/utils.py:9-13."""
    with open(filepath, 'r') as f:
        return [line.rstrip() for line in f.readlines()]


def load_samples(data_dir, list_file, save_filename=False, load_pseudo_gt=False):
    """The sample list of StereoDataset.__init__, dataloader/SceneflowLoader.py:73-102: one dict per line of
    `left right [disp]` with keys left / right / disp / pseudo_disp (/ left_name)."""
    samples = []
    for line in read_text_lines(list_file):
        splits = line.split()
        left_img, right_img = splits[:2]
        gt_disp = None if len(splits) == 2 else splits[2]
        sample = dict()
        if save_filename:
            sample['left_name'] = left_img.split('/', 1)[1]
        sample['left'] = os.path.join(data_dir, left_img)
        sample['right'] = os.path.join(data_dir, right_img)
        sample['disp'] = os.path.join(data_dir, gt_disp) if gt_disp is not None else None
        if load_pseudo_gt and sample['disp'] is not None:
            if 'disp_occ_0' in sample['disp']:
                sample['pseudo_disp'] = sample['disp'].replace('disp_occ_0', 'disp_occ_0_pseudo_gt')
            elif 'disp_occ' in sample['disp']:
                sample['pseudo_disp'] = sample['disp'].replace('disp_occ', 'disp_occ_pseudo_gt')
            else:
                raise NotImplementedError
        else:
            sample['pseudo_disp'] = None
        samples.append(sample)
    return samples


def load_pair(sample, subset=False):
    """StereoDataset.__getitem__ without the resize/transform, SceneflowLoader.py:104-121."""
    out = {'img_left': read_img(sample['left']), 'img_right': read_img(sample['right'])}
    if sample.get('disp') is not None:
        out['gt_disp'] = read_disp(sample['disp'], subset=subset)
    if sample.get('pseudo_disp') is not None:
        out['pseudo_disp'] = read_disp(sample['pseudo_disp'], subset=subset)
    return out


def strip_module_prefix(weights, resume=False):
    """utils/utils.py:86-88: DataParallel checkpoints carry a `module.` prefix (dropped unless resuming)."""
    new_state_dict = OrderedDict()
    for k, v in weights.items():
        new_state_dict[k[7:] if 'module' in k and not resume else k] = v
    return new_state_dict


def load_pretrained_net(net, pretrained_path, return_epoch_iter=False, resume=False, no_strict=False):
    """utils/utils.py:73-100."""
    import torch
    if pretrained_path is None:
        return None
    state = torch.load(pretrained_path, map_location='cuda' if torch.cuda.is_available() else 'cpu')
    weights = state['state_dict'] if 'state_dict' in state.keys() else state
    net.load_state_dict(strip_module_prefix(weights, resume), strict=not no_strict)
    if return_epoch_iter:
        return tuple(state[k] if k in state.keys() else None for k in ('epoch', 'num_iter', 'best_epe', 'best_epoch'))
    return None
