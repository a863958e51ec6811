"""Multi-GPU plumbing for the batch decoder: streams are independent, so rank r simply owns a contiguous block of
stream ids and decodes only its own packets — no collective on the data path. torch.distributed (NCCL on the GPU box,
gloo in the CPU tests) carries only per-stream checksums and timing scalars at the end of a run.

Reference: one mousiki::Decoder per stream, no shared mutable state (src/opus_decoder.rs:200-241; static tables
src/celt/modes.rs:84-105), which is what makes the partition legal."""
import numpy as np


def shard_range(n_streams: int, world: int, rank: int):
    """Contiguous block [lo, hi) of stream ids owned by `rank` (sizes differ by at most one)."""
    base, rem = divmod(int(n_streams), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def owner_of(stream_id, n_streams: int, world: int):
    """Inverse of shard_range for scalars or arrays of stream ids."""
    base, rem = divmod(int(n_streams), int(world))
    sid = np.asarray(stream_id, dtype=np.int64)
    split = rem * (base + 1)
    return np.where(sid < split, sid // max(base + 1, 1), rem + (sid - split) // max(base, 1)).astype(np.int64)


def shard_packets(arena, offsets, stream_ids, n_streams: int, world: int, rank: int):
    """Select this rank's packets (order preserved). Returns (arena, offsets, local_stream_ids, packet_index)."""
    offsets = np.asarray(offsets, dtype=np.uint32)
    stream_ids = np.asarray(stream_ids, dtype=np.uint32)
    lo, hi = shard_range(n_streams, world, rank)
    sel = np.nonzero((stream_ids >= lo) & (stream_ids < hi))[0]
    lens = (offsets[1:].astype(np.int64) - offsets[:-1].astype(np.int64))[sel]
    new_off = np.zeros(len(sel) + 1, dtype=np.uint32)
    new_off[1:] = np.cumsum(lens)
    if len(sel):
        idx = np.concatenate([np.arange(offsets[p], offsets[p + 1], dtype=np.int64) for p in sel])
        new_arena = np.ascontiguousarray(np.asarray(arena, dtype=np.uint8)[idx])
    else:
        new_arena = np.zeros(0, dtype=np.uint8)
    return new_arena, new_off, (stream_ids[sel] - lo).astype(np.uint32), sel


def stream_checksums(local_stream_ids, n_local_streams: int, final_range, pcm_i16=None):
    """Per-stream 64-bit checksum: low word = sum of final_range over the stream's packets (mod 2^32), high word =
    FNV-1a style fold of the i16 PCM (optional). Order of packets inside a stream matters for neither term's validity:
    both are computed per packet and combined in packet order."""
    ids = np.asarray(local_stream_ids, dtype=np.int64)
    lo = np.zeros(n_local_streams, dtype=np.uint64)
    np.add.at(lo, ids, np.asarray(final_range, dtype=np.uint64))
    lo &= np.uint64(0xFFFFFFFF)
    hi = np.zeros(n_local_streams, dtype=np.uint64)
    if pcm_i16 is not None:
        p = np.asarray(pcm_i16).astype(np.int64)
        w = (np.arange(p.shape[1], dtype=np.int64) * 2654435761 + 97) & 0xFFFF
        per_packet = ((p & 0xFFFF) * (w | 1)).sum(axis=1).astype(np.uint64) & np.uint64(0xFFFFFFFF)
        # packet order inside a stream: weight by occurrence number
        order = np.zeros(len(ids), dtype=np.uint64)
        seen = {}
        for k, s in enumerate(ids.tolist()):
            order[k] = seen.get(s, 0)
            seen[s] = order[k] + 1
        np.add.at(hi, ids, (per_packet * (order * np.uint64(2) + np.uint64(1))) & np.uint64(0xFFFFFFFF))
        hi &= np.uint64(0xFFFFFFFF)
    return (hi << np.uint64(32)) | lo


def gather_checksums(local: np.ndarray, n_streams: int, world: int, rank: int, device=None):
    """all_gather of the per-stream checksums (the only collective of a run). Every rank gets the full
    [n_streams] vector in global stream order. Works with any initialised torch.distributed backend."""
    import torch
    import torch.distributed as dist

    if world == 1:
        return np.asarray(local, dtype=np.uint64).copy()
    sizes = [shard_range(n_streams, world, r) for r in range(world)]
    cap = max(hi - lo for lo, hi in sizes)
    buf = torch.zeros(cap, dtype=torch.int64, device=device)
    buf[: len(local)] = torch.from_numpy(np.asarray(local, dtype=np.uint64).view(np.int64)).to(buf.device)
    out = [torch.zeros_like(buf) for _ in range(world)]
    dist.all_gather(out, buf)
    parts = [o.cpu().numpy().view(np.uint64)[: hi - lo] for o, (lo, hi) in zip(out, sizes)]
    return np.concatenate(parts)
