"""Row f3 (host part): the PNG writer behind saveScreenshot (Application.cpp:661-674) and the accumulation file.
The PNG is decoded here with zlib by the format's own rules; the decoded pixels must equal the input bit for bit."""
import ctypes
import struct
import zlib

import numpy as np
import pytest


def _decode_png(path):
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    at, chunks = 8, []
    while at < len(data):
        n, kind = struct.unpack(">I4s", data[at:at + 8])
        body = data[at + 8:at + 8 + n]
        (crc,) = struct.unpack(">I", data[at + 8 + n:at + 12 + n])
        assert zlib.crc32(kind + body) == crc, kind
        chunks.append((kind, body))
        at += 12 + n
    assert [k for k, _ in chunks][0] == b"IHDR" and chunks[-1][0] == b"IEND"
    w, h, depth, colour, comp, filt, lace = struct.unpack(">IIBBBBB", chunks[0][1])
    assert (depth, colour, comp, filt, lace) == (8, 6, 0, 0, 0)
    raw = zlib.decompress(b"".join(b for k, b in chunks if k == b"IDAT"))       # checks the Adler-32 too
    rows = np.frombuffer(raw, np.uint8).reshape(h, w * 4 + 1)
    assert not rows[:, 0].any()                                                 # filter type None on every row
    return rows[:, 1:].reshape(h, w, 4)


@pytest.mark.parametrize("shape", [(1, 1), (7, 5), (64, 48), (270, 480)])
def test_png_round_trip(okrt, tmp_path, shape):
    h, w = shape
    rng = np.random.default_rng(h * 1000 + w)
    img = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    img[h // 3:, : w // 2] = (12, 200, 255, 255)           # flat area: runs against the previous pixel and the previous row
    img[:, w // 2:] = img[:1, w // 2:]                     # identical rows
    path = tmp_path / "render.png"
    okrt.abi.check(okrt.abi.lib().okrt_write_png(str(path).encode(), okrt.abi.ptr(img), w, h))
    assert np.array_equal(_decode_png(path), img)
    if h * w > 1000:
        assert path.stat().st_size < img.nbytes            # the matcher does compress


def test_png_rejects_bad_arguments(okrt, tmp_path):
    L = okrt.abi.lib()
    img = np.zeros((2, 2, 4), np.uint8)
    assert L.okrt_write_png(None, okrt.abi.ptr(img), 2, 2) == 1
    assert L.okrt_write_png(str(tmp_path / "a.png").encode(), okrt.abi.ptr(img), 0, 2) == 1
    assert L.okrt_write_png(str(tmp_path / "no_such_dir" / "a.png").encode(), okrt.abi.ptr(img), 2, 2) == 1


def test_accumulation_file_round_trip(okrt, tmp_path):
    L = okrt.abi.lib()
    acc = np.random.default_rng(3).random((9, 16, 4), dtype=np.float32) * 37.0
    path = str(tmp_path / "state.okrtacc").encode()
    okrt.abi.check(L.okrt_write_accumulation_file(path, okrt.abi.ptr(acc), 16, 9, 1234))
    w, h, n = ctypes.c_uint32(0), ctypes.c_uint32(0), ctypes.c_uint32(0)
    okrt.abi.check(L.okrt_read_accumulation_file(path, None, ctypes.byref(w), ctypes.byref(h), ctypes.byref(n)))
    assert (w.value, h.value, n.value) == (16, 9, 1234)
    back = np.zeros_like(acc)
    okrt.abi.check(L.okrt_read_accumulation_file(path, okrt.abi.ptr(back), ctypes.byref(w), ctypes.byref(h), ctypes.byref(n)))
    assert np.array_equal(back.view(np.uint32), acc.view(np.uint32))
    wrong = ctypes.c_uint32(8)
    assert L.okrt_read_accumulation_file(path, okrt.abi.ptr(back), ctypes.byref(wrong), ctypes.byref(h), ctypes.byref(n)) == 1
    open(tmp_path / "junk", "wb").write(b"not a state file")
    assert L.okrt_read_accumulation_file(str(tmp_path / "junk").encode(), None, ctypes.byref(w), ctypes.byref(h), ctypes.byref(n)) == 1
