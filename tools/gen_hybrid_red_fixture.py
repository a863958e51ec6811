#!/usr/bin/env python3
"""Hybrid packets that carry a redundancy frame, for tests/test_mode_switch.py.

In hybrid mode the redundancy flag is one range-coder bit with probability 2^-12 (src/opus_decoder.rs:752), so the
reference's `TOC || random bytes` fuzz construction almost never produces one. This script draws random packets until
the oracle reports a redundancy frame (about 4000 draws per hit) and stores PER_TOC of them for each of the 8 hybrid TOCs
in tests/golden/hybrid_redundancy_packets.npz (`arena`, `offsets`). Seeded; rerun to regenerate."""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tests"))
import helpers  # noqa: E402

PER_TOC = 24


def main():
    lib = helpers.oracle_lib()
    rng = np.random.default_rng(20261105)
    pcm = np.zeros(2880 * 2, dtype=np.float32)
    pkts = []
    for fb in (0, 1):
        for ms20 in (0, 1):
            for stereo in (0, 1):
                toc = ((12 + 2 * fb + ms20) << 3) | (4 if stereo else 0)
                h = lib.orc_create(48000, 2, helpers.QUIRKS_MOUSIKI)
                got = 0
                while got < PER_TOC:
                    ln = int(rng.integers(60, 260))
                    pkt = rng.integers(0, 256, size=ln + 1, dtype=np.uint8)
                    pkt[0] = toc
                    lib.orc_reset(h)
                    r = lib.orc_decode_float(h, helpers.ptr(pkt), len(pkt), helpers.ptr(pcm), 2880, 0)
                    if r > 0 and lib.orc_last_redundancy(h):
                        pkts.append(pkt)
                        got += 1
                lib.orc_destroy(h)
                print(f"toc {toc:#x}: {got}", flush=True)
    offsets = np.zeros(len(pkts) + 1, dtype=np.uint32)
    offsets[1:] = np.cumsum([len(p) for p in pkts])
    np.savez_compressed(ROOT / "tests" / "golden" / "hybrid_redundancy_packets.npz", arena=np.concatenate(pkts), offsets=offsets)


if __name__ == "__main__":
    main()
