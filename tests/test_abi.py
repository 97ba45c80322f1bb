"""The C-ABI shared library: it loads, exports every symbol include/rt_b200.h declares, fails loudly without a
GPU (no CPU fallback), and the CLI keeps the reference's argv / exit-code contract (main.cpp:42-94)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, scene_path


def header_symbols():
    text = open(os.path.join(ROOT, "include", "rt_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(rt):
    L = rt.lib()
    syms = header_symbols()
    assert len(syms) >= 17
    for s in syms:
        assert hasattr(L, s), f"{s} declared in rt_b200.h but not exported"
    assert sorted(rt.ABI_SYMBOLS) == syms, "raytracing-project_b200.ABI_SYMBOLS out of date"


def test_struct_layouts_match_header(rt):
    assert C.sizeof(rt.RtPrim) == 48 and C.sizeof(rt.RtMaterial) == 48 and C.sizeof(rt.RtLight) == 32
    assert C.sizeof(rt.RtPokeball) == 48 and C.sizeof(rt.RtAffine) == 48


def test_no_gpu_fails_loudly(rt):
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    with pytest.raises(rt.RtError) as e:
        rt.init(0)
    assert e.value.code == -5 and "no CPU fallback" in str(e.value)
    s = rt.load_scene_from_json(scene_path("quirks"))           # compiling is host work and still succeeds
    with pytest.raises(rt.RtError):
        rt.Tracer(s).render()


def test_argument_errors(rt):
    out = C.c_void_p()
    assert rt.lib().rt_compile_json_text(None, C.byref(out)) == -1
    assert rt.lib().rt_scene_upload(None, C.byref(out)) == -1
    assert rt.lib().rt_write_png_rgb8(b"/tmp/x.png", None, 4, 4) == -1
    assert b"null" in rt.lib().rt_last_error() or b"bad" in rt.lib().rt_last_error()


def test_png_writer_roundtrip(rt, tmp_path):
    from PIL import Image
    rng = np.random.default_rng(0)
    img = rng.integers(0, 256, (37, 53, 3), dtype=np.uint8)
    p = str(tmp_path / "o.png")
    rt.write_png_rgb8(p, img)
    back = np.asarray(Image.open(p))
    assert back.shape == img.shape and back.dtype == np.uint8 and np.array_equal(back, img)   # 8-bit RGB, no alpha
    with pytest.raises(rt.RtError) as e:
        rt.write_png_rgb8("/nonexistent_dir/o.png", img)
    assert e.value.code == -2


def test_png_writer_parallel_bands_form_one_valid_zlib_stream(rt, tmp_path):
    """Frames above 64 rows are deflated in bands on several threads; the concatenation must be ONE zlib stream (header,
    raw-deflate bands ending on full flushes, combined Adler-32) inside one IDAT chunk with a correct CRC."""
    import struct, zlib
    rng = np.random.default_rng(1)
    H, W = 1100, 700                                   # > 256 KB per band on every core count: several bands
    img = np.zeros((H, W, 3), np.uint8)
    img[:, :, 0] = (np.arange(W) % 256)[None, :]; img[:, :, 1] = (np.arange(H) % 256)[:, None]
    img[::3, ::5, 2] = rng.integers(0, 256, img[::3, ::5, 2].shape, dtype=np.uint8)
    p = str(tmp_path / "big.png")
    rt.write_png_rgb8(p, img)
    raw = open(p, "rb").read()
    assert raw[:8] == b"\x89PNG\r\n\x1a\n"
    pos, chunks = 8, []
    while pos < len(raw):
        n, typ = struct.unpack(">I4s", raw[pos:pos + 8])
        data = raw[pos + 8:pos + 8 + n]
        (crc,) = struct.unpack(">I", raw[pos + 8 + n:pos + 12 + n])
        assert crc == zlib.crc32(typ + data) & 0xFFFFFFFF
        chunks.append((typ, data)); pos += 12 + n
    assert [c[0] for c in chunks] == [b"IHDR", b"IDAT", b"IEND"]
    assert struct.unpack(">IIBBBBB", chunks[0][1]) == (W, H, 8, 2, 0, 0, 0)           # 8-bit RGB, no alpha, no interlace
    rows = np.frombuffer(zlib.decompress(chunks[1][1]), np.uint8).reshape(H, 1 + 3 * W)   # checks the combined Adler-32 too
    assert not rows[:, 0].any() and np.array_equal(rows[:, 1:].reshape(H, W, 3), img)


def test_cli_usage_and_error_exit_codes(rt, tmp_path):
    cli = rt.RAY_CLI
    assert os.path.exists(cli)
    r = subprocess.run([cli], capture_output=True, text=True)
    assert r.returncode == 1 and "Usage:" in r.stderr and "--paper" in r.stderr            # main.cpp:43-47
    r = subprocess.run([cli, "/nonexistent.json", str(tmp_path / "o.png")], capture_output=True, text=True)
    assert r.returncode == 3 and r.stderr.startswith("[error] ") and "Cannot open JSON file" in r.stderr   # main.cpp:65-68
    bad = tmp_path / "bad.json"; bad.write_text("{ invalid json }")
    r = subprocess.run([cli, str(bad), str(tmp_path / "o.png")], capture_output=True, text=True)
    assert r.returncode == 3 and "JSON parse error" in r.stderr
