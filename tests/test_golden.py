"""Committed golden vectors (tests/golden/celt_fuzz_v1.npz, made by tools/gen_golden.py from the oracle):
CPU: the oracle and the host emulation of the kernel bodies still reproduce them; GPU: the CUDA path does."""
from pathlib import Path

import numpy as np
import pytest

import helpers

GOLD = Path(__file__).resolve().parent / "golden" / "celt_fuzz_v1.npz"


def _cases():
    z = np.load(GOLD)
    names = sorted({k.split("/")[0] for k in z.files})
    return z, names


Z, NAMES = _cases()


def _case(name):
    lm, stereo, ch, bw, lanes, ticks = (int(v) for v in Z[name + "/meta"])
    return dict(ch=ch, lanes=lanes, ticks=ticks, arena=Z[name + "/arena"], offsets=Z[name + "/offsets"], samples=Z[name + "/samples"],
                rng=Z[name + "/final_range"], pcm16=Z[name + "/pcm_i16"], pcm=Z[name + "/pcm_f32"] if name + "/pcm_f32" in Z.files else None)


@pytest.mark.parametrize("name", NAMES)
def test_oracle_reproduces_golden(name):
    c = _case(name)
    pcm, samples, rng = helpers.oracle_decode_grid(c["arena"], c["offsets"], c["lanes"], c["ticks"], c["ch"], 960)
    assert np.array_equal(samples, c["samples"]) and np.array_equal(rng, c["rng"])
    if c["pcm"] is not None:
        assert np.array_equal(pcm[:, : c["pcm"].shape[1]], c["pcm"])
    pcm16, _, _ = helpers.oracle_decode_grid(c["arena"], c["offsets"], c["lanes"], c["ticks"], c["ch"], 960, i16=True)
    assert np.array_equal(pcm16[:, : c["pcm16"].shape[1]], c["pcm16"])


@pytest.mark.parametrize("name", NAMES)
def test_emulated_kernels_reproduce_golden(name):
    c = _case(name)
    pcm, samples, rng, _, _ = helpers.emul_decode_grid(c["arena"], c["offsets"], c["lanes"], c["ticks"], c["ch"], 960)
    assert np.array_equal(samples, c["samples"]) and np.array_equal(rng, c["rng"])
    if c["pcm"] is not None:
        assert np.abs(pcm[:, : c["pcm"].shape[1]] - c["pcm"]).max() <= 1.0 / 32768.0


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_gpu_reproduces_golden(name):
    from mousiki_b200 import BatchDecoder

    c = _case(name)
    ids = np.tile(np.arange(c["lanes"], dtype=np.uint32), c["ticks"])
    with BatchDecoder(c["lanes"], 48000, c["ch"]) as dec:
        pcm, samples, rng = dec.decode_float(c["arena"], c["offsets"], ids, 960)
        dec.reset()
        pcm16, samples16, rng16 = dec.decode(c["arena"], c["offsets"], ids, 960)
    assert np.array_equal(samples, c["samples"]) and np.array_equal(rng, c["rng"]), "sample counts / final_range are bit-exact"
    assert np.array_equal(samples16, c["samples"]) and np.array_equal(rng16, c["rng"])
    if c["pcm"] is not None:
        assert np.abs(pcm[:, : c["pcm"].shape[1]] - c["pcm"]).max() <= 1.0 / 32768.0  # north star: 1 LSB at i16
    d = np.abs(pcm16[:, : c["pcm16"].shape[1]].astype(np.int32) - c["pcm16"].astype(np.int32))
    assert d.max() <= 1
