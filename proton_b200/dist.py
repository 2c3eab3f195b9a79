"""Multi-GPU plumbing: one process per GPU, one strip of cell rows per rank.  torch.distributed is used
only for the rendezvous (broadcast of the NCCL unique id); the halo exchange and the scalar all-reduces of the PCG
run inside libproton_b200.so on the library's own stream."""
from ._binding import Context, dist_unique_id, strip_rows


def make_rank_context(mip, quad_kind, rank, world, device, torch_dist=None):
    """Context of `rank` on `device` owning its strip of `mip`; joins the NCCL communicator when world > 1."""
    ctx = Context(device)
    rb, re = strip_rows(mip.Ny, world, rank)
    ctx.mesh(mip, quad_kind, row_begin=rb, row_end=re)
    ctx.row_begin, ctx.row_end = rb, re
    if world > 1:
        box = [dist_unique_id() if rank == 0 else None]
        torch_dist.broadcast_object_list(box, src=0)
        ctx.dist_init(rank, world, box[0])
    return ctx
