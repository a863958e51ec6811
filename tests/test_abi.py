"""The C-ABI library loads here (no GPU) and exports every symbol include/mousiki_b200.h declares; with no
device present the product fails loudly instead of falling back to a CPU path."""
import ctypes as C
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


def _declared_symbols():
    text = (ROOT / "include" / "mousiki_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(bd_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported():
    import mousiki_b200

    lib = mousiki_b200.load_library()
    syms = _declared_symbols()
    assert len(syms) >= 12
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/mousiki_b200.h but not exported"
    assert sorted(mousiki_b200.EXPORTS) == syms


def test_no_cpu_fallback_without_device():
    import torch

    import mousiki_b200

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(mousiki_b200.DecoderError):
        mousiki_b200.BatchDecoder(4, 48000, 2)


def test_bad_arguments_rejected_before_touching_a_device():
    import mousiki_b200

    lib = mousiki_b200.load_library()
    h = C.c_void_p()
    cfg = mousiki_b200.batch_decoder.BdConfig(C.sizeof(mousiki_b200.batch_decoder.BdConfig), 4, 44100, 2, 0, -1, 0x7F, 0, 0, 0)
    assert lib.bd_create(C.byref(cfg), C.byref(h)) == -1  # OpusDecoderInitError::BadArgument (src/opus_decoder.rs:1550-1556)
    cfg.sample_rate_hz, cfg.channels = 48000, 3
    assert lib.bd_create(C.byref(cfg), C.byref(h)) == -1
    assert lib.bd_create(None, C.byref(h)) == -1


def test_product_does_not_reference_oracle():
    """The oracle is test infrastructure: nothing under mousiki_b200/ or include/ may include, link or import it."""
    for path in list((ROOT / "mousiki_b200").rglob("*")) + list((ROOT / "include").rglob("*")):
        if path.suffix in {".py", ".cu", ".cuh", ".h", ".hpp", ".cpp"} or path.name == "Makefile":
            text = path.read_text(errors="ignore")
            assert "liboracle" not in text and "libemul" not in text, path
            assert not re.search(r"(#\s*include|import|from|CDLL|dlopen)[^\n]*(oracle|emul\.)", text), path


def test_library_is_not_older_than_its_sources():
    """The in-tree .so travels to the GPU box as built: a kernel source edited after the last `make` would be tested there
    against a stale binary."""
    root = Path(__file__).resolve().parent.parent
    lib = root / "mousiki_b200" / "libmousiki_b200.so"
    srcs = list((root / "mousiki_b200" / "csrc").glob("*.cu*")) + list((root / "mousiki_b200" / "csrc").glob("*.h")) + [root / "include" / "mousiki_b200.h"]
    newest = max(srcs, key=lambda p: p.stat().st_mtime)
    assert lib.stat().st_mtime >= newest.stat().st_mtime, f"{newest.name} is newer than libmousiki_b200.so: run `python __graft_entry__.py build`"
