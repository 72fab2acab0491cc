"""Multi-GPU partitioner (north-star item 5): the scene is replicated on every GPU, samples-per-pixel are
split across ranks, every rank accumulates its samples into an fp32 SUM buffer (RtProcessParams.blend_mode = 1: rgb += sum of
sample radiances, alpha += sample count) and ONE reduce(sum) per frame over NCCL/NVLink brings them to the root, which divides by
the total sample count. The reference has no multi-GPU code; the per-rank RNG streams are cuRAND subsequences
[rank * n_states, (rank + 1) * n_states) of the same seed (disjoint, 2^67 draws apart), so rank 0 of any world size renders
exactly the single-GPU stream.

torch.distributed is plumbing only (process group, the NCCL reduce); nothing here touches pixels on the host."""


def partition_spp(total_spp, world_size):
    """Samples per rank: even split, the remainder goes to the lowest ranks. sum == total_spp."""
    if total_spp < 0 or world_size < 1:
        raise ValueError("partition_spp: bad argument")
    base, extra = divmod(total_spp, world_size)
    return [base + (1 if r < extra else 0) for r in range(world_size)]


def rank_first_subsequence(rank, n_states):
    """First cuRAND subsequence of `rank` (n_states = the padded per-frame state count, RayTracing.cpp:215-218)."""
    return rank * n_states


def launches_for(spp, spp_per_launch):
    """Split a rank's samples into kernel launches of at most spp_per_launch samples."""
    out = []
    while spp > 0:
        n = min(spp, spp_per_launch)
        out.append(n)
        spp -= n
    return out


def reduce_frame(sum_buffer, dist, dst=0):
    """One collective per frame: fp32 sum of the per-rank [H, W, 4] sum buffers onto `dst` (in place)."""
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(sum_buffer, dst=dst, op=dist.ReduceOp.SUM)
    return sum_buffer


def finalize_mean(sum_buffer):
    """Root only: rgb / sample count (the alpha channel carries the count), alpha = 1 - the reference's frame semantics."""
    out = sum_buffer.clone()
    n = out[..., 3:4].clamp_min(1.0)
    out[..., :3] = out[..., :3] / n
    out[..., 3] = 1.0
    return out
