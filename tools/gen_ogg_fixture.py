#!/usr/bin/env python3
"""Copies the reference's 180-byte Ogg Opus test file into tests/golden/ (data, MIT, (c) 2023 The Pion community —
/root/reference/testdata/tiny.ogg.license). Run in the build container only; the GPU box has no /root/reference."""
import pathlib, shutil
src = pathlib.Path("/root/reference/testdata/tiny.ogg")
dst = pathlib.Path(__file__).resolve().parent.parent / "tests" / "golden" / "tiny.ogg"
shutil.copyfile(src, dst)
(dst.parent / "tiny.ogg.license").write_text("SPDX-FileCopyrightText: 2023 The Pion community <https://pion.ly>\nSPDX-License-Identifier: MIT\n")
print(dst, dst.stat().st_size, "bytes")
