"""Worker of tests/test_round2_gpu.py::test_two_rank_nccl_merge_through_the_c_abi (launched by torchrun, one rank
per GPU): sample-index partition of 8 frames over the ranks, one ptb_merge_accumulation per step."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mcrt_b200  # noqa: E402,F401
from mcrt_b200 import multi_gpu, pt, rr, scene_library  # noqa: E402


def main():
    out = sys.argv[1]
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    api = rr.IntersectionApi.Create(local)
    pl = pt.make_pt_pipeline(api, scene_library.cornell_box(), 128, 96, 4)
    pl.ensure_size()
    comm = multi_gpu.create_comm(pl, rank, world)
    for f in multi_gpu.frames_of_rank(rank, world, 8 // world):
        pl.passes[0].m_frameIndex = f
        pl.update()
        pl.merge_accumulation(comm, 0)
    pl.synchronize()
    if rank == 0:
        np.save(out + "_rank0.npy", pl.read_buffer("NextFrameImageCL"))
    else:
        np.save(out + "_rank%d_acc.npy" % rank, pl.read_buffer("AccumulationBuffer"))
    dist.barrier()
    multi_gpu.destroy_comm(comm)
    pl.scene.destroy()
    rr.IntersectionApi.Delete(api)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
