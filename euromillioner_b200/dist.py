"""One-process-per-GPU plumbing: rendezvous through torch.distributed (NCCL on GPUs, gloo in the CPU
tests), engine communicator through the C ABI (emrf_ctx_create_dist).  torch is used for the
rendezvous only; the data path (the per-group all-reduce of split records) is NCCL inside libemrf.so."""
import os


def env_rank():
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


def tree_owner(tree, world):
    """Tree-sharded fit: rank that builds tree `tree`'s entry lists, histograms and partitions
    (engine.cu uses the same rule: tree % world == rank)."""
    return tree % world


def trees_of_rank(num_trees, rank, world):
    return [t for t in range(num_trees) if tree_owner(t, world) == rank]


def broadcast_bytes(dist, payload, src=0):
    """Rank `src` passes `payload` (bytes); every rank returns it (used for the 128-byte ncclUniqueId)."""
    box = [payload if dist.get_rank() == src else None]
    dist.broadcast_object_list(box, src=src)
    return box[0]


def make_context(dist=None):
    """Context for this rank: plain when WORLD_SIZE == 1, else an NCCL communicator over all ranks."""
    from .classifier import Context
    rank, local_rank, world = env_rank()
    if world == 1:
        return Context(local_rank)
    uid = broadcast_bytes(dist, Context.nccl_unique_id() if rank == 0 else None)
    return Context(local_rank, rank, world, uid)
