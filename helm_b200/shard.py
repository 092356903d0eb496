"""Level sharding across the GPUs of one node.

The gates of one netlist level are independent, so rank r evaluates the contiguous slice [r*ceil(W/N), ...) of the
level's gate list with replicated keys, and ONE exchange step per level makes every rank's freshly written output
rows visible everywhere: an all-gather in place on the level's block of the wire array (the reference's precedent is
the per-level gather of src/circuit.rs:877-892, there through host memory).  Levels narrower than `min_width` stay on
one rank (extra GPUs are replicas only for such circuits).  The functions take the process group / tensors as
arguments so the same code runs over NCCL on GPUs and over gloo in the CPU tests."""
import torch


def level_slice(width, world, rank, min_width=0):
    """Rows [g0, g0 + cnt) of a level of `width` gates that `rank` evaluates.  Equal slices (the all-gather needs
    them); the last ranks' slices are padded by the caller when width % world != 0."""
    if world <= 1 or width < min_width:
        return (0, width) if rank == 0 else (0, 0)
    per = (width + world - 1) // world
    g0 = min(rank * per, width)
    return g0, max(0, min(per, width - g0))


def padded_width(width, world):
    per = (width + world - 1) // world
    return per * world


def exchange_level(dist, block, world, rank, group=None):
    """All-gather of the level block `block` ([padded_width, stride] rows; rank r has written rows
    [r*per, (r+1)*per)) in place.  Returns the work handle's result (None for world 1)."""
    if world <= 1:
        return None
    per = block.shape[0] // world
    assert per * world == block.shape[0], "level block must be padded to a multiple of the world size"
    mine = block[rank * per:(rank + 1) * per]
    if block.is_cuda:
        return dist.all_gather_into_tensor(block, mine, group=group)
    parts = [torch.empty_like(mine) for _ in range(world)]       # gloo has no all_gather_into_tensor
    dist.all_gather(parts, mine.contiguous(), group=group)
    for r, p in enumerate(parts):
        block[r * per:(r + 1) * per].copy_(p)
    return None
