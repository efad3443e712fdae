"""Host-side sharding of query batches across ranks: queries shard trivially, the map is
replicated. One process per GPU; torch.distributed only carries the plumbing (NCCL on GPUs, gloo in the
CPU tests). There is no collective in the query loop; results stay on the producing rank unless the
caller asks for them with gather_to_root()."""
import numpy as np


def shard_range(n, rank, world):
    """[start, stop) of the contiguous block of n items owned by `rank` (sizes differ by at most 1)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world: {rank}/{world}")
    base, rem = divmod(int(n), world)
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def shard_scans(n_scans, rank, world):
    """Whole scans per rank (BASELINE config 5: many scans queried against a replicated map)."""
    a, b = shard_range(n_scans, rank, world)
    return list(range(a, b))


def shard_points(xs, ys, zs, rank, world):
    a, b = shard_range(len(xs), rank, world)
    return xs[a:b], ys[a:b], zs[a:b]


def gather_to_root(local, n_total, dist=None, root=0):
    """Reassembles per-rank result arrays (each covering shard_range(n_total, rank, world) along axis 0)
    on `root`; other ranks get None. `local` is a tuple of numpy arrays."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return tuple(local)
    import torch
    rank, world = dist.get_rank(), dist.get_world_size()
    out = []
    for a in local:
        a = np.ascontiguousarray(a)
        view = a.view(np.int64) if a.dtype == np.uint64 else a  # torch has no uint64 collectives
        t = torch.from_numpy(view)
        if rank == root:
            parts = []
            for r in range(world):
                s, e = shard_range(n_total, r, world)
                parts.append(torch.empty((e - s,) + tuple(t.shape[1:]), dtype=t.dtype))
            parts[root].copy_(t)
            reqs = [dist.irecv(parts[r], src=r) for r in range(world) if r != root]
            for q in reqs:
                q.wait()
            full = torch.cat(parts, 0).numpy()
            out.append(full.view(np.uint64) if a.dtype == np.uint64 else full)
        else:
            dist.send(t, dst=root)
            out.append(None)
    return tuple(out) if rank == root else None


def replicate_map(gmap, dist, device, root=0):
    """Broadcasts `gmap` (built on `root`) to every rank with NCCL through the C ABI (mmg_bcast)."""
    from .maps import Communicator
    rank, world = dist.get_rank(), dist.get_world_size()
    if world == 1:
        return None
    obj = [Communicator.unique_id() if rank == root else None]
    dist.broadcast_object_list(obj, src=root)
    comm = Communicator(obj[0], rank, world, device)
    gmap.broadcast(comm, root)
    return comm
