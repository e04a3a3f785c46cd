"""Host-side plumbing for the one-process-per-GPU launch (torchrun): ring neighbours, exchange of
the opaque IPC blobs, and the single end-of-run reduction of per-rank av_vels partials.

Mirrors the reference's rank bookkeeping: rank_up = (rank+1) % size, rank_down = (rank-1+size) % size
(allocate_rows, d2q9-bgk.c:501-502) and the one MPI_Reduce(SUM) + scaling of collate (:445-446).
Works with any torch.distributed backend (nccl on GPUs, gloo in the CPU tests); no data-path
collective is involved - halos move GPU-to-GPU inside the kernel.
"""
import numpy as np


def ring_neighbours(rank, size):
    """(rank_up, rank_down) as the reference defines them (d2q9-bgk.c:501-502)."""
    return (rank + 1) % size, ((rank - 1) % size + size) % size


def exchange_blobs(dist, torch, blob, rank, size, device="cpu"):
    """all_gather of every rank's fixed-size byte blob; returns (blob_up, blob_down, all_blobs)."""
    mine = torch.frombuffer(bytearray(blob), dtype=torch.uint8).to(device)
    gathered = [torch.empty_like(mine) for _ in range(size)]
    dist.all_gather(gathered, mine)
    raw = [bytes(t.cpu().numpy()) for t in gathered]
    up, down = ring_neighbours(rank, size)
    return raw[up], raw[down], raw


def reduce_av_vels(dist, torch, partial, inv_free_cells, device="cpu"):
    """Sum of the per-rank un-normalised per-step |u| sums (the reference's MPI_Reduce, :445), then the
    scaling by 1/free_cells in fp32 (:446).  `partial` is this rank's float array."""
    t = torch.from_numpy(np.asarray(partial, np.float64)).to(device)
    dist.all_reduce(t)
    return (t.cpu().numpy().astype(np.float32) * np.float32(inv_free_cells)).astype(np.float32)


def gather_rows(dist, torch, local, row_counts, rank, device="cpu"):
    """Gathers every rank's (9, rows_r, nx) block to rank 0 as the global (9, ny, nx) lattice (the
    reference's collate gather, :397-443).  Returns the global array on rank 0, None elsewhere."""
    size = len(row_counts)
    nx = local.shape[2]
    mx = max(row_counts)
    pad = np.zeros((9, mx, nx), np.float32)
    pad[:, :local.shape[1]] = local
    mine = torch.from_numpy(pad).to(device)
    gathered = [torch.empty_like(mine) for _ in range(size)]
    dist.all_gather(gathered, mine)
    if rank != 0:
        return None
    return np.concatenate([g.cpu().numpy()[:, :row_counts[r]] for r, g in enumerate(gathered)], axis=1)
