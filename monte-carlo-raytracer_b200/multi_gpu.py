"""Multi-GPU plumbing of the path tracer: one process per GPU, scene and BVH replicated,
work split by image tiles (A) or by sample index (B), and ONE collective -- a sum of the accumulation
buffers  sum(w*L) (float4[P]) || sum(w) (float[P])  over the ranks -- followed by the division
ReconstructionPass does (reconstruction.cl:50-58).  The reference has no multi-device path; this is
the whole exchange step.

The product path is behind the C ABI: ptb_merge_accumulation (include/ptb.h) issues one ncclReduce /
ncclAllReduce on the context's stream over the single 20 B/pixel allocation and divides in one kernel;
create_comm() below only carries the NCCL unique id between the ranks over the launcher's process group.
merge_accumulation() / PipelinedMerger are the same arithmetic on torch tensors (CPU tensors over gloo in
the tests; they were bench.py's path in r01).
"""
import ctypes

import numpy as np


def partition_pixels(width, height, part, num_parts):
    """Global pixel indices of partition A (host-side mirror of ptb_set_partition, needs no GPU)."""
    from . import pt
    lib = pt._lib()
    lib.ptb_partition_pixels.argtypes = [ctypes.c_int] * 4 + [ctypes.c_void_p, ctypes.c_int64, ctypes.POINTER(ctypes.c_int64)]
    n = ctypes.c_int64()
    pt._check(lib.ptb_partition_pixels(width, height, part, num_parts, None, 0, ctypes.byref(n)))
    out = np.empty(n.value, np.uint32)
    pt._check(lib.ptb_partition_pixels(width, height, part, num_parts, out.ctypes.data_as(ctypes.c_void_p), out.size, ctypes.byref(n)))
    return out


def frames_of_rank(rank, world, steps, first=0):
    """Partition B: the sample (frame) indices rank `rank` renders in `steps` steps."""
    return [first + s * world + rank for s in range(steps)]


def merge_accumulation(weighted, weights, stage=None, image=None, group=None):
    """All-reduce (sum) the two accumulation buffers and divide.

    weighted: float32 tensor (P, 4) -- this rank's  sum(w * clamp(L));  weights: float32 (P,).
    stage: optional preallocated float32 (5P,) scratch; image: optional (P, 4) output.
    The rank-local buffers are left untouched (they keep accumulating); returns the merged image."""
    import torch
    import torch.distributed as dist
    P = weights.numel()
    if stage is None:
        stage = torch.empty(5 * P, dtype=torch.float32, device=weighted.device)
    stage[:4 * P].copy_(weighted.reshape(-1))
    stage[4 * P:].copy_(weights.reshape(-1))
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(stage, op=dist.ReduceOp.SUM, group=group)
    if image is None:
        image = torch.empty(P, 4, dtype=torch.float32, device=weighted.device)
    torch.div(stage[:4 * P].view(P, 4), stage[4 * P:].view(P, 1), out=image.view(P, 4))
    return image


class PipelinedMerger:
    """merge_accumulation split in two so the all-reduce of step k overlaps the rendering of step k+1:
    begin() snapshots the rank's buffers and starts the collective asynchronously (NCCL runs it on its
    own stream once the snapshot is complete); finish() makes the current stream wait for it and writes
    the merged image.  The displayed image therefore lags the samples by one step; call finish() once
    more after the last step."""

    def __init__(self, num_pixels, device, group=None):
        import torch
        self.P, self.group = num_pixels, group
        self.stage = torch.empty(5 * num_pixels, dtype=torch.float32, device=device)
        self.work = None
        self.pending = False

    def begin(self, weighted, weights):
        import torch.distributed as dist
        if self.pending:
            raise RuntimeError("PipelinedMerger.begin() before finish() of the previous step")
        P = self.P
        self.stage[:4 * P].copy_(weighted.reshape(-1))
        self.stage[4 * P:].copy_(weights.reshape(-1))
        if dist.is_initialized() and dist.get_world_size(self.group) > 1:
            self.work = dist.all_reduce(self.stage, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        self.pending = True

    def finish(self, image):
        import torch
        if not self.pending:
            return image
        if self.work is not None:
            self.work.wait()
            self.work = None
        P = self.P
        torch.div(self.stage[:4 * P].view(P, 4), self.stage[4 * P:].view(P, 1), out=image.view(P, 4))
        self.pending = False
        return image


def create_comm(pipeline, rank, world, group=None):
    """An ncclComm_t for ptb_merge_accumulation: rank 0 makes the NCCL unique id (ptb_nccl_unique_id), the
    launcher's torch.distributed group carries its 128 bytes to the other ranks, every rank joins on the device
    of its context (ptb_nccl_comm_create).  Returns a ctypes.c_void_p; free it with destroy_comm()."""
    import torch.distributed as dist
    from . import pt
    lib = pt._lib()
    buf = (ctypes.c_ubyte * 128)()
    if rank == 0:
        pt._check(lib.ptb_nccl_unique_id(buf))
    box = [bytes(buf)]
    dist.broadcast_object_list(box, src=0, group=group)
    ident = (ctypes.c_ubyte * 128).from_buffer_copy(box[0])
    comm = ctypes.c_void_p()
    pt._check(lib.ptb_nccl_comm_create(pipeline.scene._h, ident, int(rank), int(world), ctypes.byref(comm)))
    return comm


def destroy_comm(comm):
    from . import pt
    if comm:
        pt._check(pt._lib().ptb_nccl_comm_destroy(comm))
