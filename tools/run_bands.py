"""BASELINE config 5 on N GPUs: ONE Middlebury-shaped pair (1984x2880 -> 248x360 grid, C = 256, D = 48), split by
epipolar row bands (SURVEY 8e partition 2).  Launch: python tools/run_bands.py (N = 1) or
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/run_bands.py

Every rank builds the same seeded whole-image tensors, keeps only the rows a band-partitioned trunk would hand it
(its band + halo) and runs, per step: cost volume + fused soft-argmin / local window (no halo), the guidance stage
(warp with a 1-row halo), the shipped 16->16 DCN (data-dependent halo) and convex x8 upsampling (1-row halo); the
band's full-resolution disparity rows reach every rank through peer stores fused into the upsampling kernel
(--nccl: one NCCL all-gather per step instead; --root: peer stores to rank 0 only, the reference's gather-to-device-0).  Rank 0 checks the gathered map against
the whole-image computation on its own GPU and prints one JSON line (strong scaling: the pair is fixed)."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereoformer_b200 import bands as sfb  # noqa: E402
from stereoformer_b200 import dist as sfd   # noqa: E402
from stereoformer_b200 import ops           # noqa: E402

H, W, C, D = 248, 360, 256, 48


def whole_image_inputs(dev):
    g = torch.Generator(device=dev).manual_seed(1024)
    r = lambda *s: torch.randn(*s, device=dev, generator=g)
    t = dict(left_f=r(1, C, H, W), right_f=r(1, C, H, W), left_i=r(1, 3, H, W), right_i=r(1, 3, H, W),
             x=r(1, 16, H, W), wt=r(16, 16, 3, 3) * 0.1, bias=r(16), off=r(1, 18, H, W) * 2,
             msk=torch.sigmoid(r(1, 9, H, W)), up_mask=r(1, 576, H, W))
    return t


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    full = whole_image_inputs(dev)
    halo = sfb.dcn_halo_rows(full["off"], 3, 1, 1)
    band = sfb.RowBand.of_rank(H, world, rank, halo=halo)
    bands = sfd.row_bands(H, world, halo=halo)
    mine = {k: (band.interior(v) if k in ("left_f", "right_f", "left_i") else
                v if k in ("wt", "bias") else band.source(v)) for k, v in full.items()}

    # the upsampling reads the 3x3 neighbourhood of the low-resolution disparity: the (row-local) window is computed on
    # band + 1 row, the other outputs are cropped from it
    b1 = sfb.RowBand(H, band.y0, band.y1, max(band.y0 - 1, 0), min(band.y1 + 1, H))
    lf1, rf1, um1 = b1.source(full["left_f"]), b1.source(full["right_f"]), b1.source(full["up_mask"])
    even = all(b[1] - b[0] == bands[0][1] - bands[0][0] for b in bands)
    up_full = torch.empty((1, 1, H * 8, W * 8), device=dev)
    # gather of the full-resolution rows: fused into the upsampling kernel's store (NVLink peer mappings); NCCL
    # all-gather only when symmetric memory is unavailable or --nccl is given
    pg, gather_mode = None, "none"
    if world > 1:
        gather_mode = "NCCL all_gather per step"
        if "--nccl" not in sys.argv:
            try:
                pg = sfd.PeerGather(-(-(H * 8 * W * 8) // world), dev)
            except Exception as e:          # noqa: BLE001
                pg, gather_mode = None, "NCCL all_gather per step (symmetric memory unavailable: %s)" % type(e).__name__
        flags = torch.tensor([1.0 if pg is not None else 0.0], device=dev)
        dist.all_reduce(flags, op=dist.ReduceOp.MIN)
        if float(flags.item()) < 1.0:
            pg = None
        else:
            gather_mode = "P2P stores fused into the upsampling kernel (NVLink peer mappings)" + \
                (", to rank 0 only" if "--root" in sys.argv else ", to every rank")

    def band_pass():
        cv1, disp1 = sfb.lcv_band(lf1, rf1, D)
        cv, disp = b1.crop(cv1).contiguous(), b1.crop(disp1).contiguous()
        err, unc = sfb.guidance_band(cv, disp, mine["left_i"], mine["right_i"], band)
        y = sfb.dcn_band(mine["x"], mine["wt"], mine["bias"], mine["off"], mine["msk"], band, halo=halo)
        up = sfb.upsample_convex_band(disp1, um1, b1, gather=pg, root=0 if "--root" in sys.argv else None)
        return disp, err, unc, y, up

    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        for _ in range(3):
            out = band_pass()
        side.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            out = band_pass()
    torch.cuda.synchronize()
    disp, err, unc, y, up = out
    up_bands = [(a * 8, b * 8, c * 8, d * 8) for a, b, c, d in bands]

    def gather():
        if world == 1:
            return up
        if pg is not None:
            return pg.buffer[:H * 8 * W * 8].view(1, 1, H * 8, W * 8)
        if even:
            dist.all_gather_into_tensor(up_full, up)          # B = 1: rank-major rows are the whole map
            return up_full
        return sfd.gather_row_bands(up, up_bands)

    for _ in range(3):
        graph.replay()
        got = gather()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    steps = 50
    e0.record()
    for _ in range(steps):
        graph.replay()
    e1.record()
    for _ in range(steps):
        graph.replay()
        got = gather()
    e2.record()
    torch.cuda.synchronize()
    if pg is not None:
        pg.sync()
    up_full = got
    ms_kernels = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
    ms = torch.tensor([e1.elapsed_time(e2) / steps], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(ms_kernels, op=dist.ReduceOp.MAX)
        y_full = sfd.gather_row_bands(y, bands)
        disp_full = sfd.gather_row_bands(disp, bands)
        err_full = sfd.gather_row_bands(err, bands)
    else:
        y_full, disp_full, err_full = y, disp, err
    if rank == 0:
        whole = sfb.RowBand(H, 0, H)
        cv, d_ref = sfb.lcv_band(full["left_f"], full["right_f"], D)
        e_ref, _ = ops.guidance(cv, d_ref, full["left_i"], full["right_i"], div_mode=ops.SF_DIV_TRUE)
        y_ref = sfb.dcn_band(full["x"], full["wt"], full["bias"], full["off"], full["msk"], whole, halo=halo)
        up_ref = ops.upsample_convex(d_ref, full["up_mask"])
        res = dict(config="C5 1984x2880 (248x360 grid, C=256, D=48), one pair, %d row band(s), DCN halo %d rows"
                          % (world, halo),
                   n_gpus=world, ms_per_pair=float(ms.item()), pairs_per_s=1000.0 / float(ms.item()),
                   ms_per_pair_without_gather=float(ms_kernels.item()),
                   launch="one CUDA graph per band pass", gather=gather_mode,
                   scaling="strong",
                   disparity_bit_exact=bool(torch.equal(disp_full, d_ref)),
                   warp_error_bit_exact=bool(torch.equal(err_full, e_ref)),
                   upsampled_bit_exact=bool(torch.equal(up_full, up_ref)),
                   dcn_max_abs_diff=float((y_full - y_ref).abs().max().item()))
        print(json.dumps(res))
        ok = res["disparity_bit_exact"] and res["warp_error_bit_exact"] and res["upsampled_bit_exact"] and \
            res["dcn_max_abs_diff"] < 1e-4
        if not ok:
            sys.exit(1)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
