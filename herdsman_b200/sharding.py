"""Row sharding and key distribution for one-process-per-GPU runs (torchrun): the host-side logic of
SURVEY.md 8e.  Rows are independent units; the reference already deals them to partitions as
rows/P + (k < rows%P) (src/service/storage_service.cpp:121-129,328-331) and never communicates
between map tasks, so the only collective is the one-time broadcast of the key blob."""
import numpy as np


def row_slice(n_rows, world, rank):
    """[start, stop) of this rank's contiguous rows -- same sizes as herdsman's partitions"""
    base, rem = divmod(n_rows, world)
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def broadcast_blob(blob, dist, device=None, src=0):
    """broadcast a host uint8 blob from rank `src` (NCCL on `device`, gloo on CPU); returns a torch tensor"""
    import torch
    size = torch.tensor([0 if blob is None else int(blob.size)], dtype=torch.int64, device=device)
    dist.broadcast(size, src=src)
    if dist.get_rank() == src:
        t = torch.from_numpy(np.ascontiguousarray(blob, np.uint8))
        t = t.to(device) if device is not None else t.clone()
    else:
        t = torch.empty(int(size.item()), dtype=torch.uint8, device=device)
    dist.broadcast(t, src=src)
    return t


def gather_rows(local_rows, n_rows, dist):
    """all ranks contribute their slice (CPU tensors, gloo); rank order == row order"""
    import torch
    world = dist.get_world_size()
    parts = [None] * world
    dist.all_gather_object(parts, np.ascontiguousarray(local_rows))
    out = np.concatenate(parts, axis=0)
    assert out.shape[0] == n_rows
    return out
