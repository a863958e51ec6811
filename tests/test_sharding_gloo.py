"""N>1 host path on CPU: world_size-2 gloo run of the sharding + checksum gather used by bench.py under torchrun.
The per-rank "decode" here is the host emulation of the kernel bodies (test infrastructure); what is under test is
that sharding by stream id + all_gather of per-stream checksums reproduces the single-rank result exactly."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import torch.multiprocessing as mp

import helpers
from mousiki_b200 import sharding

ROOT = Path(__file__).resolve().parent.parent
LANES, TICKS = 13, 4  # odd on purpose: shards of unequal size


def _corpus():
    arena, offsets = helpers.fuzz_grid(LANES, TICKS, (20, 120), lambda t, j, r: helpers.celt_toc(3 - (j % 3), j % 2 == 0), seed=77)
    ids = np.tile(np.arange(LANES, dtype=np.uint32), TICKS)
    return arena, offsets, ids


def _decode_shard(arena, offsets, local_ids, n_local):
    # packets arrive tick-major; the emulation wants a regular [ticks x lanes] grid, which a contiguous block of lanes keeps
    ticks = len(local_ids) // max(n_local, 1)
    pcm, samples, rng, _, _ = helpers.emul_decode_grid(arena, offsets, n_local, ticks, 2, 960, i16=True)
    return pcm, samples, rng


def _worker(rank, world, port, q):
    sys.path.insert(0, str(ROOT))
    sys.path.insert(0, str(ROOT / "tests"))
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    arena, offsets, ids = _corpus()
    a, o, lid, sel = sharding.shard_packets(arena, offsets, ids, LANES, world, rank)
    lo, hi = sharding.shard_range(LANES, world, rank)
    pcm, samples, rng = _decode_shard(a, o, lid, hi - lo)
    local = sharding.stream_checksums(lid, hi - lo, rng, pcm)
    full = sharding.gather_checksums(local, LANES, world, rank)
    if rank == 0:
        q.put(full)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_checksums_match_single_rank():
    arena, offsets, ids = _corpus()
    pcm, samples, rng = _decode_shard(arena, offsets, ids, LANES)
    want = sharding.stream_checksums(ids, LANES, rng, pcm)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert np.array_equal(got, want)


def test_shard_ranges_partition_all_streams():
    for n, w in [(10, 3), (7, 2), (5, 8), (100000, 8), (1, 1)]:
        cover = []
        for r in range(w):
            lo, hi = sharding.shard_range(n, w, r)
            cover.extend(range(lo, hi)) if n < 1000 else cover.append((lo, hi))
        if n < 1000:
            assert cover == list(range(n))
            assert (sharding.owner_of(np.arange(n), n, w) == np.concatenate(
                [np.full(sharding.shard_range(n, w, r)[1] - sharding.shard_range(n, w, r)[0], r) for r in range(w)])).all()
        else:
            assert cover[0][0] == 0 and cover[-1][1] == n and all(a[1] == b[0] for a, b in zip(cover, cover[1:]))
