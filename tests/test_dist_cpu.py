"""Host-side sharding logic, exercised with world-size-2 gloo on the CPU."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from stereoformer_b200 import dist as sfd


def test_pair_shard_covers_batch():
    for n in (1, 7, 8, 9, 64):
        for w in (1, 2, 3, 4, 8):
            sl = [sfd.pair_shard(n, r, w) for r in range(w)]
            assert sl[0][0] == 0 and sl[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(sl, sl[1:]))
            assert max(b - a for a, b in sl) - min(b - a for a, b in sl) <= 1


def test_row_bands_middlebury():
    # 1984 rows = 62 blocks of 32
    for w, want in ((2, [31, 31]), (4, [16, 16, 15, 15]), (8, [8, 8, 8, 8, 8, 8, 7, 7])):
        bands = sfd.row_bands(1984, w, align=32, halo=224)
        assert [(y1 - y0) // 32 for y0, y1, _, _ in bands] == want
        assert bands[0][0] == 0 and bands[-1][1] == 1984
        assert all(a[1] == b[0] for a, b in zip(bands, bands[1:]))
        assert bands[0][2] == 0 and bands[-1][3] == 1984
        assert all(b[2] == max(0, b[0] - 224) for b in bands)
    assert sfd.row_bands(68, 3) == [(0, 23, 0, 23), (23, 46, 23, 46), (46, 68, 46, 68)]


def _worker(rank, world, port):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        n = 5
        a, b = sfd.pair_shard(n, rank, world)
        full = torch.arange(n * 6, dtype=torch.float32).view(n, 1, 2, 3)
        got = sfd.gather_disparity(full[a:b], sizes=[sfd.pair_shard(n, r, world)[1] - sfd.pair_shard(n, r, world)[0]
                                                     for r in range(world)])
        assert torch.equal(got, full)
        even = torch.full((2, 1, 2, 3), float(rank))
        g2 = sfd.gather_disparity(even)
        assert g2.shape[0] == 2 * world and torch.equal(g2[2 * rank:2 * rank + 2], even)
        bands = sfd.row_bands(7, world)
        img = torch.arange(2 * 7 * 4, dtype=torch.float32).view(2, 1, 7, 4)
        y0, y1, _, _ = bands[rank]
        assert torch.equal(sfd.gather_row_bands(img[:, :, y0:y1], bands), img)
    finally:
        dist.destroy_process_group()


def test_gather_world2_gloo():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker, args=(2, port), nprocs=2, join=True)


def test_row_band_host_logic():
    from stereoformer_b200 import bands as sfb
    H = 248
    for world in (2, 4, 8):
        bs = [sfb.RowBand.of_rank(H, world, r, halo=3) for r in range(world)]
        assert bs[0].ys0 == 0 and bs[-1].ys1 == H and sum(b.rows for b in bs) == H
        full = torch.arange(H * 5, dtype=torch.float32).view(1, 1, H, 5)
        for b in bs:
            assert b.covers(3) and not (b.covers(4) and 0 < b.y0 and b.y1 < H)
            assert torch.equal(b.crop(b.source(full)), b.interior(full))
            s = b.scaled(8)
            assert (s.height, s.y0, s.y1, s.ys0, s.ys1) == (H * 8, b.y0 * 8, b.y1 * 8, b.ys0 * 8, b.ys1 * 8)
        assert torch.equal(torch.cat([b.interior(full) for b in bs], 2), full)
    import pytest
    with pytest.raises(ValueError):
        sfb.RowBand(10, 4, 4)
    with pytest.raises(ValueError):
        sfb.RowBand(10, 4, 8, 5, 8)
    off = torch.zeros(1, 18, 4, 4)
    off[0, 4, 1, 1] = -2.25          # a dy channel
    off[0, 5, 1, 1] = 9.0            # a dx channel: does not widen the row halo
    assert sfb.dcn_halo_rows(off, 3, 1, 1) == 1 + 3 + 1
    with pytest.raises(ValueError):
        sfb.dcn_halo_rows(off, 3, 0, 1)
