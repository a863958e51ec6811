#!/usr/bin/env python3
"""Transcribes the reference's own golden vectors for the SILK / hybrid rows into tests/golden/ref_hybrid_vectors.npz
(data only; runs in the build container, needs /root/reference):

  tests/fixtures/hybrid_decode_vectors.rs  packets + final_range + 960-sample f32 PCM produced by opus-c
                                           (asserted by tests/hybrid_decode.rs with tolerance 1e-4)
  tests/fixtures/hybrid_encode_vectors.rs  three libopus hybrid packets + their encoder-side final ranges
                                           (tests/hybrid_encode.rs:81,95,128; used here as a probe only)
"""
import re
from pathlib import Path

import numpy as np

REF = Path("/root/reference/tests/fixtures")
ROOT = Path(__file__).resolve().parent.parent


def consts(path):
    txt = (REF / path).read_text()
    out = {}
    for m in re.finditer(r"const\s+(\w+)\s*:\s*\[(u8|f32);\s*(\d+)\]\s*=\s*\[(.*?)\];", txt, flags=re.S):
        name, ty, n, body = m.groups()
        toks = [t.strip() for t in body.replace("\n", " ").split(",") if t.strip()]
        if ty == "u8":
            arr = np.array([int(t, 0) for t in toks], dtype=np.uint8)
        else:
            arr = np.array([float(t.replace("_", "").rstrip("f32")) for t in toks], dtype=np.float32)
        assert len(arr) == int(n), (name, len(arr), n)
        out[name] = arr
    for m in re.finditer(r"const\s+(\w+)\s*:\s*u32\s*=\s*([\d_]+);", txt):
        out[m.group(1)] = np.array([int(m.group(2).replace("_", ""))], dtype=np.uint32)
    return out


def main():
    out = {}
    out.update(consts("hybrid_decode_vectors.rs"))
    out.update(consts("hybrid_encode_vectors.rs"))
    dst = ROOT / "tests" / "golden" / "ref_hybrid_vectors.npz"
    np.savez_compressed(dst, **out)
    print("wrote", dst, sorted((k, v.shape) for k, v in out.items()))


if __name__ == "__main__":
    main()
