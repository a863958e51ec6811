#!/usr/bin/env python3
"""Generates tests/golden/celt_fuzz_v1.npz: seeded `TOC || random bytes` corpora (the reference's fuzz construction,
tests/test_opus_decode.rs:372-407) decoded by the oracle (oracle/, the C++ restatement of mousiki's float decoder —
mousiki itself cannot be built or imported in this image: no Rust toolchain). The fixture freezes the oracle's answers
so that (a) an accidental change of the oracle is caught on CPU, and (b) the GPU box can check the CUDA path against
committed vectors as well as against the live oracle. Run from the repo root: python tools/gen_golden.py"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tests"))
import helpers  # noqa: E402

# (name, lm, stereo packets, decoder channels, payload bytes or range, bandwidth, lanes, ticks, seed)
CASES = [
    ("fb20_stereo_160", 3, True, 2, 160, 3, 8, 5, 101),
    ("fb10_stereo_80", 2, True, 2, 80, 3, 6, 5, 102),
    ("fb5_mono_40", 1, False, 1, 40, 3, 6, 5, 103),
    ("fb2p5_stereo_20", 0, True, 2, 20, 3, 6, 5, 104),
    ("swb20_mono_in_stereo_dec", 3, False, 2, (8, 200), 2, 6, 5, 105),
    ("wb20_stereo_in_mono_dec", 3, True, 1, (8, 200), 1, 6, 5, 106),
    ("nb10_stereo_ragged", 2, True, 2, (2, 120), 0, 6, 5, 107),
]

F32_CASES = ("fb20_stereo_160", "fb5_mono_40", "nb10_stereo_ragged")  # float PCM kept for these (fixture size)


def main():
    out = {}
    for name, lm, stereo, ch, payload, bw, lanes, ticks, seed in CASES:
        arena, offsets = helpers.fuzz_grid(lanes, ticks, payload, lambda t, j, r: helpers.celt_toc(lm, stereo, bw), seed=seed)
        pcm, samples, rng = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 960)
        pcm16, samples16, rng16 = helpers.oracle_decode_grid(arena, offsets, lanes, ticks, ch, 960, i16=True)
        assert np.array_equal(rng, rng16) and np.array_equal(samples, samples16)
        n = int(samples.max()) * ch
        out[name + "/meta"] = np.array([lm, int(stereo), ch, bw, lanes, ticks], dtype=np.int32)
        out[name + "/arena"] = arena
        out[name + "/offsets"] = offsets
        out[name + "/samples"] = samples
        out[name + "/final_range"] = rng
        if name in F32_CASES:
            out[name + "/pcm_f32"] = pcm[:, :n].copy()
        out[name + "/pcm_i16"] = pcm16[:, :n].copy()
    np.savez_compressed(ROOT / "tests" / "golden" / "celt_fuzz_v1.npz", **out)
    print("wrote", ROOT / "tests" / "golden" / "celt_fuzz_v1.npz")


if __name__ == "__main__":
    main()
