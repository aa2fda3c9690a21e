"""Multi-GPU plumbing for the ORB front end: frames and keyframe-pair jobs are independent, so they
are dealt round-robin to the ranks (frame f -> rank f mod G) with NO collective on the data path; the only exchange
is one final all_gather of fixed-size padded result records over NCCL/NVLink (gloo in the CPU tests).

Record layout per frame (cap = ORBextractor.max_keypoints()): n (int32), cap x 28 B keypoints, cap x 32 B descriptors.
"""
import torch
import torch.distributed as dist


def shard_indices(n_items, rank, world):
    """Global indices owned by `rank`: rank, rank + world, ... (static round-robin)."""
    return list(range(rank, n_items, world))


def local_count(n_items, rank, world):
    return (n_items - rank + world - 1) // world if rank < n_items else 0


def gather_frames(kps, desc, n, world, group=None, out=None):
    """kps (B, cap, 28) u8, desc (B, cap, 32) u8, n (B,) i32 on this rank -> the same three tensors for all ranks,
    rank-major: (world * B, ...).  B must be equal on all ranks (pad the last shard).  `out` = preallocated tensors."""
    if world == 1:
        return kps, desc, n
    if out is None:
        out = [torch.empty((world * kps.shape[0],) + tuple(kps.shape[1:]), dtype=kps.dtype, device=kps.device),
               torch.empty((world * desc.shape[0],) + tuple(desc.shape[1:]), dtype=desc.dtype, device=desc.device),
               torch.empty(world * n.shape[0], dtype=n.dtype, device=n.device)]
    dist.all_gather_into_tensor(out[0], kps.contiguous(), group=group)
    dist.all_gather_into_tensor(out[1], desc.contiguous(), group=group)
    dist.all_gather_into_tensor(out[2], n.contiguous(), group=group)
    return out


def global_order(gathered, n_items, world):
    """Undo the round-robin deal: returns index tensor `src` with result_of_frame[f] = gathered[src[f]]
    (rank-major gathered layout, B = ceil(n_items / world) slots per rank)."""
    B = (n_items + world - 1) // world
    f = torch.arange(n_items)
    return (f % world) * B + (f // world)
