"""Multi-GPU layout of the hot path: independent streams are sharded over ranks, one process per
GPU, with NO collective on the data path.  The only exchanges are the host-side
gather of per-stream result records and an optional box-wide strongest-peak reduction."""
import numpy as np


def shard_streams(total_streams, world_size, rank):
    """Contiguous shard of `total_streams` for `rank`: (first_stream, count); sizes differ by <= 1."""
    if world_size < 1 or not 0 <= rank < world_size:
        raise ValueError("bad rank / world size")
    base, extra = divmod(total_streams, world_size)
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


def gather_records(local_records, dist=None):
    """All ranks' [streams, blocks] record arrays concatenated in stream order on every rank."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return local_records
    parts = [None] * dist.get_world_size()
    dist.all_gather_object(parts, local_records)
    return np.concatenate(parts, axis=0)


def strongest_peak(local_records, first_stream, dist=None):
    """Box-wide strongest focus-band peak: (absPeakDb, global stream, block, peakBin).
    A 16-byte record per rank; max-reduced (all_gather + host max keeps it backend-agnostic)."""
    flat = local_records["absPeakDb"]
    s, b = np.unravel_index(int(np.argmax(flat)), flat.shape) if flat.size else (0, 0)
    mine = (float(flat[s, b]) if flat.size else -np.inf, first_stream + int(s), int(b),
            int(local_records["peakBin"][s, b]) if flat.size else -1)
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return mine
    parts = [None] * dist.get_world_size()
    dist.all_gather_object(parts, mine)
    return max(parts, key=lambda t: (t[0], -t[1]))
