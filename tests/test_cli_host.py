"""Host-side pieces of the `photonmap` drop-in that need no GPU: the flag grammar (gi_parse_args and the Python
parse_flags against the reference's own ParseArgs, fixture tests/golden/parse_args.json) and the image writers."""
import ctypes
import json
import os
import struct
import zlib

import numpy as np
import pytest

import gi_b200
from gi_testlib import GOLDEN

CASES = json.load(open(os.path.join(GOLDEN, "parse_args.json")))
FIELDS = [f for f, _ in gi_b200.GiParams._fields_ if f not in ("seed", "rank", "world_size")]


class GiCliOptions(ctypes.Structure):
    _fields_ = [("input_scene_name", ctypes.c_char_p), ("output_image_name", ctypes.c_char_p), ("width", ctypes.c_int32),
                ("height", ctypes.c_int32), ("aa", ctypes.c_int32), ("real_material", ctypes.c_int32), ("device", ctypes.c_int32)]


def c_parse(argv):
    L = gi_b200.lib()
    p = gi_b200.default_params()
    o = GiCliOptions()
    L.gi_default_cli_options(ctypes.byref(o))
    full = [b"photonmap", b"in.scn", b"out.png"] + [a.encode() for a in argv]
    arr = (ctypes.c_char_p * len(full))(*full)
    rc = L.gi_parse_args(len(full), arr, ctypes.byref(p), ctypes.byref(o))
    return rc, p, o


@pytest.mark.parametrize("case", CASES, ids=[" ".join(c["argv"]) or "defaults" for c in CASES])
def test_flag_grammar_matches_reference_parseargs(case):
    rc, p, o = c_parse(case["argv"])
    assert rc == 0
    want = case["globals"]
    for f in FIELDS:
        assert getattr(p, f) == pytest.approx(want[f], rel=0, abs=0), f
    assert (o.width, o.height, o.aa, o.real_material) == (want["width"], want["height"], want["aa"], want["real_material"])
    assert o.input_scene_name == b"in.scn" and o.output_image_name == b"out.png"
    # the Python mirror used by the tests and bench agrees too
    pp, extra = gi_b200.parse_flags(case["argv"])
    for f in FIELDS:
        assert getattr(pp, f) == getattr(p, f), f
    assert (extra["width"], extra["height"], extra["aa"], int(extra["real"])) == (o.width, o.height, o.aa, o.real_material)


def test_flag_errors():
    rc, _, _ = c_parse(["-bogus"])
    assert rc != 0 and b"Invalid program argument: -bogus" in gi_b200.lib().gi_last_error()
    rc, _, _ = c_parse(["-md"])
    assert rc != 0
    rc, _, _ = c_parse(["third_positional"])
    assert rc != 0
    L = gi_b200.lib()
    p = gi_b200.default_params(); o = GiCliOptions(); L.gi_default_cli_options(ctypes.byref(o))
    arr = (ctypes.c_char_p * 2)(b"photonmap", b"only_input.scn")
    assert L.gi_parse_args(2, arr, ctypes.byref(p), ctypes.byref(o)) != 0
    assert b"Usage: photonmap inputscenefile outputimagefile" in L.gi_last_error()
    with pytest.raises(ValueError):
        gi_b200.parse_flags(["-bogus"])


def _frame(w, h):
    rng = np.random.default_rng(3)
    return rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8)      # row 0 = bottom row


def _write(path, img):
    L = gi_b200.lib()
    L.gi_write_image.argtypes = [ctypes.c_char_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
    return L.gi_write_image(str(path).encode(), img.ctypes.data, img.shape[1], img.shape[0])


@pytest.mark.parametrize("w,h", [(7, 5), (64, 48), (300, 257)])
def test_png_writer_roundtrip(tmp_path, w, h):
    img = _frame(w, h)
    assert _write(tmp_path / "a.png", img) == 0
    raw = open(tmp_path / "a.png", "rb").read()
    assert raw[:8] == b"\x89PNG\r\n\x1a\n"
    pos, chunks = 8, {}
    while pos < len(raw):
        n, typ = struct.unpack(">I4s", raw[pos:pos + 8])
        body = raw[pos + 8:pos + 8 + n]
        assert struct.unpack(">I", raw[pos + 8 + n:pos + 12 + n])[0] == zlib.crc32(typ + body)
        chunks.setdefault(typ, b"")
        chunks[typ] += body
        pos += 12 + n
    assert struct.unpack(">IIBBBBB", chunks[b"IHDR"]) == (w, h, 8, 2, 0, 0, 0)
    data = np.frombuffer(zlib.decompress(chunks[b"IDAT"]), dtype=np.uint8).reshape(h, 1 + 3 * w)
    assert np.all(data[:, 0] == 0)
    assert np.array_equal(data[:, 1:].reshape(h, w, 3), img[::-1])          # PNG row 0 = top = last frame row


def test_ppm_and_bmp_writers(tmp_path):
    img = _frame(6, 4)
    assert _write(tmp_path / "a.ppm", img) == 0
    tok = open(tmp_path / "a.ppm").read().split()
    assert tok[:4] == ["P3", "6", "4", "255"]
    assert np.array_equal(np.array(tok[4:], dtype=int).reshape(4, 6, 3), img[::-1])
    assert _write(tmp_path / "a.bmp", img) == 0
    raw = open(tmp_path / "a.bmp", "rb").read()
    assert raw[:2] == b"BM" and struct.unpack("<iiHH", raw[18:30]) == (6, 4, 1, 24)
    rows = np.frombuffer(raw[54:], dtype=np.uint8).reshape(4, 20)[:, :18].reshape(4, 6, 3)
    assert np.array_equal(rows[:, :, ::-1], img)
    assert _write(tmp_path / "a.jpg", img) != 0 and b"Unsupported" in gi_b200.lib().gi_last_error()
    assert _write(tmp_path / "noext", img) != 0
