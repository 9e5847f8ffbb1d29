"""Sharding of the hot path across the GPUs of one box (one process per GPU, torch.distributed).

Two exchange-free partitions: by stereo pair (batch) and, for a single large pair, by epipolar row
band.  The only collective on the inference path is the gather of the disparity maps (NCCL over NVLink on GPUs,
gloo in the CPU tests).  Replaces the reference's single-process ``nn.DataParallel``
(train_file/trainer_DLC_GRU.py:102-106).
"""
import torch
import torch.distributed as dist


def pair_shard(n_pairs, rank, world):
    """Contiguous [start, stop) slice of the batch for ``rank``; the first ``n_pairs % world`` ranks get one more."""
    base, rem = divmod(n_pairs, world)
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def row_bands(height, world, align=1, halo=0):
    """Row bands of a single pair: list of (y0, y1, y0_halo, y1_halo) per rank.

    Bands are multiples of ``align`` rows (32 for the LowCNN trunk: 1/32 scale stages), the remainder goes to the
    first ranks; ``halo`` rows (clipped to the image) are added on both sides for ops that are not row-local
    (1 row at the 1/8 grid for disp_warp; the conv trunk's receptive field at full resolution)."""
    blocks = -(-height // align)
    out = []
    b0 = 0
    for r in range(world):
        nb = blocks // world + (1 if r < blocks % world else 0)
        y0, y1 = min(b0 * align, height), min((b0 + nb) * align, height)
        out.append((y0, y1, max(0, y0 - halo), min(height, y1 + halo)))
        b0 += nb
    return out


def _ragged_gather(t, sizes, dim, group):
    """all_gather of tensors whose extent along ``dim`` differs per rank: pad to the largest, gather once, trim."""
    world = len(sizes)
    m = max(sizes)
    shape = list(t.shape)
    shape[dim] = m
    pad = torch.zeros(shape, dtype=t.dtype, device=t.device)
    pad.narrow(dim, 0, t.shape[dim]).copy_(t)
    out = torch.empty([world] + shape, dtype=t.dtype, device=t.device)
    dist.all_gather_into_tensor(out, pad.unsqueeze(0), group=group)
    return torch.cat([out[r].narrow(dim, 0, sizes[r]) for r in range(world)], dim)


def gather_disparity(disp_local, sizes=None, group=None):
    """All-gather per-rank disparity maps [b_r,1,H,W] (b_r may differ per rank) -> [sum b_r,1,H,W] on every rank."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return disp_local
    if sizes is None:
        out = torch.empty((world * disp_local.shape[0],) + tuple(disp_local.shape[1:]), dtype=disp_local.dtype,
                          device=disp_local.device)
        dist.all_gather_into_tensor(out, disp_local.contiguous(), group=group)
        return out
    return _ragged_gather(disp_local, list(sizes), 0, group)


def gather_row_bands(band_local, bands, group=None):
    """All-gather row bands [B,1,h_r,W] (un-haloed interiors, in rank order) -> the full [B,1,H,W] map."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return band_local
    return _ragged_gather(band_local, [y1 - y0 for (y0, y1, _, _) in bands], 2, group)


class PeerGather:
    """Gather buffer that every rank can write directly: torch symmetric memory (CUDA peer mappings over
    NVLink / NVSwitch).  ``peer_ptrs`` are this process's device pointers to every rank's buffer (its own
    included); kernels store their slice at ``offset(rank)`` of each of them (ops.local_window(..., peers=...)),
    so the disparity gather costs no extra launch and no host call per step.  ``sync()`` orders the stores for
    consumers on every rank (stream synchronise + barrier)."""

    def __init__(self, numel_per_rank, device, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        self.numel_per_rank = int(numel_per_rank)
        self.numel_total = self.world * self.numel_per_rank
        self.buffer = symm_mem.empty(self.world * self.numel_per_rank, dtype=torch.float32, device=device)
        self.handle = symm_mem.rendezvous(self.buffer, self.group)
        self.peer_ptrs = [int(p) for p in self.handle.buffer_ptrs]
        if len(self.peer_ptrs) != self.world or any(p == 0 for p in self.peer_ptrs):
            raise RuntimeError("symmetric memory rendezvous returned no peer mappings")

    def offset(self, rank=None):
        return (self.rank if rank is None else rank) * self.numel_per_rank

    def sync(self):
        torch.cuda.synchronize(self.buffer.device)
        dist.barrier(group=self.group)


class SignalledGather:
    """Gather of per-rank disparity maps with COMPLETION semantics, fused into the producing kernel.

    Producers (every rank) store {value, step flag} words straight into the destinations' slots over NVLink from inside
    the window kernel (``ops.local_window(..., gather=self)``); no fence, no extra launch, no host call.  Consumers call
    ``wait(out)`` once per step on their stream: a small kernel that polls the slot until every word carries the
    step's flag, unpacks it into ``out`` [world * B, 1, H, W] and releases the slot to the producers (back-pressure:
    four slots: a producer is at most four steps ahead of the slowest consumer).  ``dst_ranks``: (0,) =
    gather to rank 0, what the reference's ``nn.DataParallel`` does (train_file/trainer_DLC_GRU.py:102-106); all ranks
    = all-gather.  Everything is stream-ordered and graph-capturable: step counters live in device memory.
    Protocol: include/stereoformer_b200.h (sf_local_window_gather_signal_f32)."""

    SIG_WORDS, SIG_ERROR, SLOTS = 64, 4, 4          # SF_SIG_WORDS, SF_SIG_ERROR, SF_GATHER_SLOTS

    def __init__(self, shape, device, dst_ranks=(0,), group=None):
        import ctypes
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        if self.world > 8:
            raise ValueError("at most 8 ranks (SF_MAX_PEERS)")
        self.shape = tuple(int(v) for v in shape)            # (B, H, W) of every rank's maps: fixes the kernel grids
        if len(self.shape) != 3:
            raise ValueError("shape = (B, H, W) of one rank's disparity maps")
        self.numel_per_rank = self.shape[0] * self.shape[1] * self.shape[2]
        self.dst_ranks = tuple(sorted(set(int(r) for r in dst_ranks)))
        if not self.dst_ranks or min(self.dst_ranks) < 0 or max(self.dst_ranks) >= self.world:
            raise ValueError("bad destination ranks %r" % (dst_ranks,))
        self.slot_stride = self.world * self.numel_per_rank          # 8-byte words per slot
        self.slots = symm_mem.empty(self.SLOTS * self.slot_stride, dtype=torch.int64, device=device)
        self.signals = symm_mem.empty(self.SIG_WORDS, dtype=torch.int64, device=device)
        self.slots.zero_()
        self.signals.zero_()
        torch.cuda.synchronize(device)
        self._h_slots = symm_mem.rendezvous(self.slots, self.group)
        self._h_sig = symm_mem.rendezvous(self.signals, self.group)
        self.data_ptrs = [int(p) for p in self._h_slots.buffer_ptrs]
        self.sig_ptrs = [int(p) for p in self._h_sig.buffer_ptrs]
        if any(p == 0 for p in self.data_ptrs + self.sig_ptrs):
            raise RuntimeError("symmetric memory rendezvous returned no peer mappings")
        dist.barrier(group=self.group)               # every block is zeroed before anyone stores
        nd = len(self.dst_ranks)
        self.c_dst_data = (ctypes.c_void_p * nd)(*[ctypes.c_void_p(self.data_ptrs[r]) for r in self.dst_ranks])
        self.c_dst_sig = (ctypes.c_void_p * nd)(*[ctypes.c_void_p(self.sig_ptrs[r]) for r in self.dst_ranks])
        self.c_dst_ranks = (ctypes.c_int * nd)(*self.dst_ranks)
        self.c_src_sig = (ctypes.c_void_p * self.world)(*[ctypes.c_void_p(p) for p in self.sig_ptrs])
        self.is_consumer = self.rank in self.dst_ranks

    def offset(self):
        return self.rank * self.numel_per_rank

    def wait(self, out):
        """Consumer side of one step (stream-ordered): ``out`` receives the gathered world * B * H * W floats."""
        import ctypes
        from . import _lib
        if not self.is_consumer:
            raise RuntimeError("rank %d is not a destination of this gather" % self.rank)
        n = self.world * self.numel_per_rank
        if (out.dtype != torch.float32 or out.numel() != n or not out.is_contiguous() or
                out.device != self.slots.device):
            raise ValueError("out must be a contiguous fp32 tensor of %d elements on the gather's device" % n)
        with torch.cuda.device(self.slots.device):
            _lib.call("sf_gather_wait_f32", ctypes.c_void_p(self.sig_ptrs[self.rank]), self.c_src_sig, self.world,
                      self.rank, ctypes.c_void_p(self.slots.data_ptr()), ctypes.c_longlong(self.slot_stride),
                      ctypes.c_void_p(out.data_ptr()), ctypes.c_longlong(n),
                      ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        return out

    def check(self):
        """Raises if a spin timed out on this rank (a peer never arrived); one device -> host read."""
        if int(self.signals[self.SIG_ERROR].item()) != 0:
            raise RuntimeError("SignalledGather: a wait timed out on rank %d" % self.rank)
