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


_PFM_KINDS = {b'PF': 3, b'Pf': 1}          # header tag -> channels


def _pfm_header(buf):
    """Parse the three text lines of a PFM header -> (channels, width, height, scale, payload offset)."""
    pos, fields = 0, []
    for _ in range(3):
        end = buf.index(b'\n', pos)
        fields.append(buf[pos:end + 1])
        pos = end + 1
    tag = fields[0].rstrip()
    if tag not in _PFM_KINDS:
        raise Exception('Not a PFM file.')
    dims = re.match(r'^(\d+)\s(\d+)\s$', fields[1].decode("ascii"))
    if dims is None:
        raise Exception('Malformed PFM header.')
    return _PFM_KINDS[tag], int(dims.group(1)), int(dims.group(2)), float(fields[2].decode("ascii").rstrip()), pos


def _read_pfm(file):
    """utils/file_io.py:34-69 -> (data, scale).  The file stores rows bottom-to-top; a negative scale marks
    little-endian floats.  Returned array: top-to-bottom, [H,W] or [H,W,3], in the file's byte order."""
    with open(file, 'rb') as fh:
        buf = fh.read()
    channels, width, height, scale, start = _pfm_header(buf)
    dtype = np.dtype('<f4' if scale < 0 else '>f4')
    shape = (height, width) if channels == 1 else (height, width, channels)
    data = np.frombuffer(buf, dtype=dtype, offset=start).reshape(shape)
    return data[::-1], abs(scale)


read_pfm = _read_pfm


def write_pfm(file, image, scale=1):
    """utils/file_io.py:72-98 (also utils/common.py:67-96): byte-identical files."""
    if image.dtype.name != 'float32':
        raise Exception('Image dtype must be float32.')
    if image.ndim == 3 and image.shape[2] == 3:
        tag = b'PF'
    elif image.ndim == 2 or (image.ndim == 3 and image.shape[2] == 1):
        tag = b'Pf'
    else:
        raise Exception('Image must have H x W x 3, H x W x 1 or H x W dimensions.')
    little = image.dtype.byteorder == '<' or (image.dtype.byteorder == '=' and sys.byteorder == 'little')
    header = b'%s\n%d %d\n%f\n' % (tag, image.shape[1], image.shape[0], -scale if little else scale)
    with open(file, 'wb') as fh:
        fh.write(header)
        fh.write(np.ascontiguousarray(image[::-1]).tobytes())      # bottom row first


def _read_kitti_disp(filename):
    """utils/file_io.py:101-104: KITTI stores disparity * 256 in a 16-bit PNG."""
    from PIL import Image
    return np.asarray(Image.open(filename), dtype=np.float32) / 256.


_DISP_READERS = (('pfm', lambda f, subset: (-1.0 if subset else 1.0) * np.ascontiguousarray(_read_pfm(f)[0])),
                 ('png', lambda f, subset: _read_kitti_disp(f)),
                 ('npy', lambda f, subset: np.load(f)))


def read_disp(filename, subset=False):
    """utils/file_io.py:17-31 -> [H,W] float32.  SceneFlow `subset` ground truth is stored negated."""
    for suffix, reader in _DISP_READERS:
        if filename.endswith(suffix):
            return reader(filename, subset)
    raise Exception('Invalid disparity file format!')


def read_text_lines(filepath):
    """utils/utils.py:9-13."""
    with open(filepath, 'r') as f:
        return [line.rstrip() for line in f.readlines()]


def _pseudo_gt_path(disp_path):
    """dataloader/SceneflowLoader.py:90-99: pseudo ground truth exists for the KITTI layouts only."""
    for key in ('disp_occ_0', 'disp_occ'):                 # KITTI 2015 first ('disp_occ' is a prefix of it)
        if key in disp_path:
            return disp_path.replace(key, key + '_pseudo_gt')
    raise NotImplementedError


def load_samples(data_dir, list_file, save_filename=False, load_pseudo_gt=False):
    """The sample list of StereoDataset.__init__, dataloader/SceneflowLoader.py:73-102: one dict per line of
    `left right [disp]` with keys left / right / disp / pseudo_disp (/ left_name)."""
    samples = []
    for line in read_text_lines(list_file):
        names = line.split()
        disp = os.path.join(data_dir, names[2]) if len(names) > 2 else None
        entry = {'left': os.path.join(data_dir, names[0]), 'right': os.path.join(data_dir, names[1]), 'disp': disp,
                 'pseudo_disp': _pseudo_gt_path(disp) if (load_pseudo_gt and disp is not None) else None}
        if save_filename:
            entry['left_name'] = names[0].split('/', 1)[1]
        samples.append(entry)
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
