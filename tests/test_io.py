"""f4: PFM / KITTI-PNG / list-file / checkpoint formats against bytes written and arrays read by the reference's own
utils/file_io.py (tests/golden/file_io.npz, made by oracle/gen_golden.py).  CPU only."""
import os

import numpy as np
import pytest
import torch

from stereoformer_b200 import io as sio

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "file_io.npz"))


def _write(tmp_path, name, key):
    fn = str(tmp_path / name)
    with open(fn, "wb") as f:
        f.write(G[key].tobytes())
    return fn


@pytest.mark.parametrize("name,src,scale", [("grey", "grey", 1), ("color", "color", 2.5), ("grey1", "grey", 1)])
def test_write_pfm_bytes_match_reference(tmp_path, name, src, scale):
    img = G[src][:, :, None] if name == "grey1" else G[src]
    fn = str(tmp_path / (name + ".pfm"))
    sio.write_pfm(fn, img, scale)
    assert open(fn, "rb").read() == G["pfm_" + name].tobytes()


@pytest.mark.parametrize("name", ["grey", "color", "grey1", "big"])
def test_read_pfm_matches_reference(tmp_path, name):
    data, scale = sio.read_pfm(_write(tmp_path, name + ".pfm", "pfm_" + name))
    assert np.array_equal(np.ascontiguousarray(data).astype(np.float32), G["read_" + name])
    if name != "big":
        assert scale == float(G["scale_" + name])
    if name in ("grey", "big"):
        assert np.array_equal(data, G["grey"])               # round trip


def test_read_disp_formats(tmp_path):
    assert np.array_equal(sio.read_disp(_write(tmp_path, "k.png", "png_kitti")), G["read_kitti"])
    fn = _write(tmp_path, "g.pfm", "pfm_grey")
    assert np.array_equal(sio.read_disp(fn), G["grey"])
    assert np.array_equal(sio.read_disp(fn, subset=True), G["read_subset"])
    np.save(str(tmp_path / "d.npy"), G["grey"])
    assert np.array_equal(sio.read_disp(str(tmp_path / "d.npy")), G["grey"])
    with pytest.raises(Exception, match="Invalid disparity file format"):
        sio.read_disp("x.tiff")


def test_pfm_errors(tmp_path):
    bad = tmp_path / "bad.pfm"
    bad.write_bytes(b"P6\n1 1\n-1\n\0\0\0\0")
    with pytest.raises(Exception, match="Not a PFM file"):
        sio.read_pfm(str(bad))
    bad.write_bytes(b"Pf\n1x1\n-1\n\0\0\0\0")
    with pytest.raises(Exception, match="Malformed PFM header"):
        sio.read_pfm(str(bad))
    with pytest.raises(Exception, match="float32"):
        sio.write_pfm(str(bad), np.zeros((2, 2), np.float64))
    with pytest.raises(Exception, match="dimensions"):
        sio.write_pfm(str(bad), np.zeros((2, 2, 2), np.float32))


def test_sample_list_and_pair(tmp_path):
    lst = tmp_path / "list.txt"
    lst.write_text("a/left/0001.png\ta/right/0001.png\tdisp/0001.pfm\n"
                   "training/image_2/000000_10.png training/image_3/000000_10.png training/disp_occ_0/000000_10.png\n"
                   "testing/l.png testing/r.png\n")
    s = sio.load_samples("/data", str(lst), save_filename=True)
    assert [x["disp"] for x in s] == ["/data/disp/0001.pfm", "/data/training/disp_occ_0/000000_10.png", None]
    assert s[0]["left"] == "/data/a/left/0001.png" and s[0]["left_name"] == "left/0001.png"
    assert all(x["pseudo_disp"] is None for x in s)
    with pytest.raises(NotImplementedError):                        # pseudo GT exists for KITTI paths only (:90-99)
        sio.load_samples("/data", str(lst), load_pseudo_gt=True)
    lst.write_text("\n".join(lst.read_text().split("\n")[1:]))
    p = sio.load_samples("/data", str(lst), load_pseudo_gt=True)
    assert p[0]["pseudo_disp"] == "/data/training/disp_occ_0_pseudo_gt/000000_10.png" and p[1]["pseudo_disp"] is None
    # one real pair from disk
    from PIL import Image
    rgb = (np.arange(3 * 4 * 5) % 255).astype(np.uint8).reshape(4, 5, 3)
    Image.fromarray(rgb).save(str(tmp_path / "l.png"))
    Image.fromarray(rgb[:, ::-1].copy()).save(str(tmp_path / "r.png"))
    sio.write_pfm(str(tmp_path / "d.pfm"), G["grey"])
    pair = sio.load_pair({"left": str(tmp_path / "l.png"), "right": str(tmp_path / "r.png"),
                          "disp": str(tmp_path / "d.pfm"), "pseudo_disp": None})
    assert pair["img_left"].dtype == np.float32 and np.array_equal(pair["img_left"], rgb.astype(np.float32))
    assert np.array_equal(pair["gt_disp"], G["grey"])


def test_load_pretrained_net_module_prefix(tmp_path):
    net = torch.nn.Sequential(torch.nn.Conv2d(3, 4, 3), torch.nn.BatchNorm2d(4))
    ref = {("module." + k): v.clone() + 1 for k, v in net.state_dict().items()}
    ck = str(tmp_path / "ck.pth")
    torch.save({"state_dict": ref, "epoch": 7, "best_epe": 0.5}, ck)
    out = sio.load_pretrained_net(net.cpu(), ck, return_epoch_iter=True)
    assert out == (7, None, 0.5, None)
    assert torch.equal(net.state_dict()["0.weight"].cpu(), ref["module.0.weight"].cpu())
    torch.save(ref, ck)                                              # bare state dict, resume keeps the prefix
    with pytest.raises(RuntimeError):
        sio.load_pretrained_net(net, ck, resume=True)
    sio.load_pretrained_net(net, ck, resume=True, no_strict=True)
