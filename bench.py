#!/usr/bin/env python
"""bench.py -- stereo pairs/s of the StereoFormer hot path on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (BASELINE.json configs[1], named in config.workload): the LocalCostVolume hot path of LowCNN at
SceneFlow 544x960, maxdisp 192 -> 1/8 grid 68x120, 256 feature channels, D=24, batch 8 pairs per GPU
(weak scaling: every rank processes its own 8 pairs).  One step = correlation cost volume (a1) +
fused soft-argmin / local disparity-window sampling (a3+a4) over the batch; with N>1 the disparity maps are
all-gathered over NCCL on a side stream (north_star: NCCL only gathers disparity maps).

value : pairs/s, inputs resident in HBM, steps replayed from CUDA graphs, CUDA-event timed, max over ranks.
e2e   : pairs/s through the public module API with HOST (pinned) feature buffers: H2D of both feature maps
        and D2H of the refined disparity inside the timed region.
roofline : the dominant kernel (corr_volume_tc_kernel) alone, algorithmic bytes / CUDA-event time vs the
        measured HBM peak (MEASURED_PEAKS.json).
cpu_baseline / --impl reference : the oracle's CPU port of the same chain on the host cores.
"""
import argparse
import json
import os
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

B_PER_GPU, C, H8, W8, D = 8, 256, 68, 120, 24
RADIUS, NPTS = 2, 20     # LowCNN baseline.py:38
WORKLOAD = "LowCNN LocalCostVolume hot path (corr volume + soft-argmin + local window r=2 N=20), 544x960, maxdisp 192 " \
           "(1/8 grid 68x120, C=256, D=24), batch 8 pairs per GPU"
ALG_BYTES_CV = (2 * C + D) * H8 * W8 * 4          # per pair, SURVEY 8d
ALG_BYTES_WIN = (D + 2) * H8 * W8 * 4


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def traffic_from_profile():
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get("corr_volume_bytes_per_launch")
        except Exception:
            pass
    return None


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self._stop_evt = index, [], set(), None, threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {"hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4)}
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.002)

    def finish(self):
        self._stop_evt.set()
        self.join(timeout=1.0)
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def make_features(device, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    fl = torch.randn(B_PER_GPU, C, H8, W8, generator=g)
    fr = torch.randn(B_PER_GPU, C, H8, W8, generator=g)
    return fl, fr


def cpu_chain(fl, fr):
    """The reference's CPU path for this chain, as restated by the oracle (checker, bench baseline only)."""
    from oracle import stereo_oracle as O
    cv = O.corr_volume(fl, fr, D)
    d0 = O.soft_argmin(cv).unsqueeze(1)
    return O.local_cost_volume(cv, d0, RADIUS, NPTS, True)


def time_cpu(reps, fl, fr):
    torch.set_num_threads(os.cpu_count() or 1)
    cpu_chain(fl[:1], fr[:1])
    best = float("inf")
    for _ in range(reps):
        t0 = time.perf_counter()
        cpu_chain(fl, fr)
        best = min(best, time.perf_counter() - t0)
    return best


def run_reference(args, rank):
    if rank != 0:
        return
    fl, fr = make_features("cpu", 1024)
    torch.set_num_threads(os.cpu_count() or 1)
    for _ in range(max(1, min(args.warmup, 2))):
        cpu_chain(fl, fr)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_chain(fl, fr)
    dt = time.perf_counter() - t0
    v = args.steps * B_PER_GPU / dt
    cores = torch.get_num_threads()
    line = {"impl": "reference", "metric": "stereo pairs/s", "value": v, "unit": "pairs/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "arm": "oracle CPU port of the reference's PyTorch path (the reference is "
                       "Python and cannot travel to the GPU box)"},
            "cpu_baseline": {"value": v, "unit": "pairs/s", "cores": cores, "kind": "port",
                             "sample": "%d steps x %d pairs, full chain" % (args.steps, B_PER_GPU)},
            "e2e": {"value": v, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch.distributed as dist
    import stereoformer_b200 as sf
    from stereoformer_b200 import ops

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- inputs: 3 rotating feature sets per rank (each step reads 134 MB > L2; rotation rules out L2 reuse)
    NSETS = 3
    host = [make_features("cpu", 1024 + 97 * rank + i) for i in range(NSETS)]
    sets = [(a.to(dev), b.to(dev)) for a, b in host]
    gather_buf = torch.empty(world * B_PER_GPU, 1, H8, W8, device=dev) if world > 1 else None
    side = torch.cuda.Stream(device=dev)

    dbuf = [torch.empty(B_PER_GPU, 1, H8, W8, device=dev) for _ in range(NSETS)]

    # N > 1: the disparity gather (north_star: the only exchange of the path).  Preferred: the window kernel stores
    # its result straight into every rank's gather buffer through NVLink peer mappings (torch symmetric memory), so
    # the gather is part of the step graph and costs no host call; fallback: an NCCL all-gather after each step.
    pg, gather_mode = None, "none"
    if world > 1:
        try:
            from stereoformer_b200.dist import PeerGather
            pg = PeerGather(B_PER_GPU * H8 * W8, dev)
            gather_mode = "P2P stores fused into the window kernel (NVLink peer mappings)"
        except Exception as e:            # noqa: BLE001 -- any failure of the symmetric-memory plumbing
            pg, gather_mode = None, "NCCL all_gather_into_tensor per step (symmetric memory unavailable: %s)" % type(e).__name__
        flags = torch.tensor([1.0 if pg is not None else 0.0], device=dev)
        dist.all_reduce(flags, op=dist.ReduceOp.MIN)
        if float(flags.item()) < 1.0:
            pg = None
            if gather_mode.startswith("P2P"):
                gather_mode = "NCCL all_gather_into_tensor per step (symmetric memory unavailable on a peer)"

    def step(i, cv_out=None):
        fl, fr = sets[i % NSETS]
        cv = ops.corr_volume(fl, fr, D, out=cv_out)
        return window(i, cv)

    def window(i, cv):
        return ops.local_window(cv, None, mode=ops.SF_WIN_FIXED, radius=RADIUS, n_points=NPTS, consider_valid=True,
                                div_mode=ops.SF_DIV_RECIP, out=dbuf[i % NSETS],
                                peers=pg.peer_ptrs if pg is not None else None,
                                peer_offset=pg.offset() if pg is not None else 0)

    # CUDA graphs keep the host launch path (ctypes + tensor-map encoding, ~25 us per step) off the device timeline.
    # A step = 2 kernel launches of this library.  `single[i]`: one step.  `pipe`: GSTEPS consecutive steps whose
    # window kernel (5 us, plus the NVLink drain of its peer stores at N > 1) runs on a second captured stream
    # underneath the NEXT step's cost-volume kernel -- the software pipelining any multi-batch caller gets from two
    # streams; every step still does all of its work, the cost volumes are double-buffered.
    GSTEPS = 4 * NSETS
    NCV = 3                                                  # cost-volume buffers in flight
    cvbuf = [torch.empty(B_PER_GPU, D, H8, W8, device=dev) for _ in range(NCV)]
    cap = torch.cuda.Stream(device=dev)
    win = torch.cuda.Stream(device=dev)
    single = []

    cvs = [torch.cuda.Stream(device=dev) for _ in range(2)]

    def capture_pipe(nsteps):
        """nsteps consecutive steps.  Cost-volume kernels alternate between two captured streams, so the persistent
        CTAs of step j+1 move onto each SM as soon as step j's CTA there has finished its tiles (no idle tail: 4352
        tiles over 148 SMs is 29.4 per SM); the window kernels run on a third stream.  Only data dependencies are
        kept: window(j) after cv(j), cv(j) after the window kernel that last read its buffer."""
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=cap):
            fork = torch.cuda.Event()
            fork.record(cap)
            for st in cvs + [win]:
                st.wait_event(fork)
            done = [None] * NCV                              # window kernel that last read cvbuf[k]
            for j in range(nsteps):
                k, st = j % NCV, cvs[j % 2]
                with torch.cuda.stream(st):
                    if done[k] is not None:
                        st.wait_event(done[k])               # cvbuf[k] is free again
                    fl, fr = sets[j % NSETS]
                    ops.corr_volume(fl, fr, D, out=cvbuf[k])
                    ready = torch.cuda.Event()
                    ready.record(st)
                win.wait_event(ready)
                with torch.cuda.stream(win):
                    window(j, cvbuf[k])
                    done[k] = torch.cuda.Event()
                    done[k].record(win)
            for st in cvs + [win]:
                cap.wait_stream(st)
        return g

    with torch.cuda.stream(cap):
        for i in range(NSETS):
            step(i)
        torch.cuda.synchronize()
        for i in range(NSETS):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=cap):
                step(i)
            single.append(g)
        pipes = [(GSTEPS, capture_pipe(GSTEPS)), (NSETS, capture_pipe(NSETS))]
    pipe = pipes[0][1]
    torch.cuda.synchronize()

    def run_steps(first, count):
        """Steps first .. first+count-1: pipelined graphs while GSTEPS-aligned blocks remain, single steps otherwise."""
        i, end = first, first + count
        while i < end:
            blk = [(n, g) for n, g in pipes if end - i >= n] if (i % NSETS == 0 and (pg is not None or world == 1)) else []
            if blk:
                blk[0][1].replay()
                i += blk[0][0]
            else:
                single[i % NSETS].replay()
                if world > 1 and pg is None:
                    dist.all_gather_into_tensor(gather_buf, dbuf[i % NSETS])
                i += 1

    run_steps(0, max(args.warmup, GSTEPS))
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run_steps(0, args.steps)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.finish()
    t = torch.tensor([ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    value = world * B_PER_GPU * args.steps / (ms * 1e-3)

    if pg is not None:
        # the fused gather must have delivered every rank's last disparities into this rank's buffer
        pg.sync()
        last = (args.steps - 1) % NSETS
        want = torch.empty(world * B_PER_GPU, 1, H8, W8, device=dev)
        dist.all_gather_into_tensor(want, dbuf[last])
        assert torch.equal(pg.buffer.view_as(want), want), "P2P disparity gather does not match the NCCL all-gather"

    # ---- dominant kernel alone (roofline): graph of NSETS cost-volume launches, one per feature set
    kg = torch.cuda.CUDAGraph()
    with torch.cuda.stream(cap):
        with torch.cuda.graph(kg, stream=cap):
            kouts = [ops.corr_volume(*sets[i], D) for i in range(NSETS)]
    for i in range(3):
        kg.replay()
    torch.cuda.synchronize()
    nk = max(7, min(args.steps, 200) // NSETS)
    e0.record()
    for i in range(nk):
        kg.replay()
    e1.record()
    torch.cuda.synchronize()
    k_ms = e0.elapsed_time(e1) / (nk * NSETS)
    # the same step without any overlap (one serial 2-kernel graph per step): the denominator for the kernel's share of
    # the step, comparable with an ncu launch list (which serialises kernels)
    ns = 3 * NSETS
    e0.record()
    for i in range(ns):
        single[i % NSETS].replay()
    e1.record()
    torch.cuda.synchronize()
    serial_ms = e0.elapsed_time(e1) / ns
    peak, peak_src = peaks()
    achieved = B_PER_GPU * ALG_BYTES_CV / (k_ms * 1e-3) / 1e9
    traffic = traffic_from_profile()

    # ---- end to end through the public module API, host buffers
    pin = [(a.pin_memory(), b.pin_memory()) for a, b in host]
    out_host = torch.empty(B_PER_GPU, 1, H8, W8).pin_memory()
    cvm, lc = sf.CostVolume(D, "correlation"), sf.LocalCostVolume(RADIUS, NPTS)
    est = sf.DisparityEstimation(D)

    def e2e_step(i):
        a, b = pin[i % NSETS]
        fl, fr = a.to(dev, non_blocking=True), b.to(dev, non_blocking=True)
        cv = cvm(fl, fr)
        d0 = est(cv).unsqueeze(1)
        out_host.copy_(lc(cv, d0, True), non_blocking=True)

    ne = max(3, min(args.steps, 30))
    for i in range(3):
        e2e_step(i)
    barrier()
    e0.record()
    for i in range(ne):
        e2e_step(i)
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_v = world * B_PER_GPU * ne / (float(t.item()) * 1e-3)
    h2d = 2 * B_PER_GPU * C * H8 * W8 * 4
    d2h = B_PER_GPU * H8 * W8 * 4

    # ---- parity spot check of what was just timed (oracle as checker only) + CPU baseline on rank 0
    cpu_base = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        fl, fr = host[0]
        pipe.replay()                                   # the pipelined path that was timed; dbuf[0] = feature set 0
        torch.cuda.synchronize()
        got = dbuf[0].cpu()
        want = cpu_chain(fl[:2], fr[:2])
        err = (got[:2] - want).abs()
        assert err.median() < 1e-3, "bench parity check failed: median |disp err| %g" % err.median()
        best = time_cpu(5, fl, fr)
        cpu_base = {"value": B_PER_GPU / best, "unit": "pairs/s", "cores": torch.get_num_threads(), "kind": "port",
                    "sample": "best of 5 passes over one batch of %d pairs (oracle CPU port, full chain)" % B_PER_GPU}

    if rank == 0:
        line = {"metric": "stereo pairs/s", "value": value, "unit": "pairs/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": WORKLOAD, "pairs_per_gpu_per_step": B_PER_GPU, "parallelism": "pairs sharded "
                           "across %d rank(s); disparity gather: %s" % (world, gather_mode),
                           "l2": "each step reads 134 MB of features (> 126 MB L2) from one of 3 rotating sets", "launch": "CUDA graphs of %d steps (2 kernels per step); consecutive steps overlap on captured streams (window kernel of step i and the tail of its cost volume under the cost volume of step i+1)" % GSTEPS},
                "roofline": {"kernel": "corr_volume_tc_kernel (tcgen05, bf16 hi/mid split)", "bound": "hbm", "achieved": achieved, "peak": peak,
                             "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                             "alg_bytes_per_launch": B_PER_GPU * ALG_BYTES_CV, "us_per_launch": k_ms * 1e3,
                             "serial_step_us": serial_ms * 1e3, "share_of_serial_step": k_ms / serial_ms},
                "cpu_baseline": cpu_base,
                "e2e": {"value": e2e_v, "unit": "pairs/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "steps": ne},
                "gpu_launches": 2 * args.steps, "clocks": clocks}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
