"""Host-side I/O of the headless layer: OBJ loader, PNG / PFM writers.  No GPU needed."""
import os
import struct
import zlib

import numpy as np
import pytest

from support import objio, scenes


def test_load_obj_matches_reference_loader_semantics(gpt, tmp_path):
    """uf::loadIndexedTris keeps every `v` in file order and fan-triangulates polygons (utilfun.cpp:454-524)."""
    p = tmp_path / "m.obj"
    p.write_text("# comment\no a\nv 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nvn 0 0 1\nvt 0 0\n"
                 "f 1/1/1 2/1/1 3/1/1 4/1/1\n"                     # quad -> 2 triangles
                 "g b\nv 0 0 -1\nv 1 0 -1.5e0\nv 1 1 -1\n"
                 "f -3 -2 -1\n"                                    # negative (relative) indices
                 "f 1//1 5//1 6//1 7//1 2//1\n")                   # pentagon -> 3 triangles
    v, f = gpt.load_obj(str(p))
    v2, f2 = objio.load_obj(str(p))
    assert v.shape == (7, 3) and f.shape == (6, 3)
    np.testing.assert_array_equal(v, v2)
    np.testing.assert_array_equal(f, f2)
    assert f.tolist()[:3] == [[0, 1, 2], [0, 2, 3], [4, 5, 6]]


def test_load_obj_roundtrip_of_a_fixture_and_errors(gpt, tmp_path):
    sc = scenes.load_scene("bunny_cornell")
    p = tmp_path / "bunny.obj"
    objio.write_obj(str(p), sc["verts"], sc["faces"])
    v, f = gpt.load_obj(str(p))
    np.testing.assert_array_equal(f, sc["faces"])
    np.testing.assert_allclose(v, sc["verts"], rtol=0, atol=0)       # %.9g round-trips float32
    import pytest
    with pytest.raises(gpt.GptError):
        gpt.load_obj(str(tmp_path / "missing.obj"))
    bad = tmp_path / "bad.obj"
    bad.write_text("v 0 0 0\nf 1 2 3\n")                              # index beyond the vertices read so far
    with pytest.raises(gpt.GptError):
        gpt.load_obj(str(bad))


def _png_pixels(path):
    """independent decoder: chunks, CRCs, zlib stream, filter type 0 rows"""
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w, h = 8, b"", 0, 0
    while pos < len(data):
        n, tag = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        crc, = struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])
        assert zlib.crc32(tag + body) & 0xffffffff == crc
        if tag == b"IHDR":
            w, h, bits, ctype = struct.unpack(">IIBB", body[:10])
            assert bits == 8 and ctype == 2
        elif tag == b"IDAT":
            idat += body
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + 3 * w)
    assert (raw[:, 0] == 0).all()
    return raw[:, 1:].reshape(h, w, 3)


def test_png_and_pfm_writers(gpt, tmp_path):
    rng = np.random.default_rng(0)
    for w, h in [(7, 5), (301, 230)]:                                # second one needs several stored-deflate blocks
        px = rng.integers(0, 1 << 24, size=(h, w), dtype=np.uint32)
        p = str(tmp_path / ("t%d.png" % w))
        gpt.write_png(p, px, w, h)
        want = np.stack([px & 0xff, (px >> 8) & 0xff, (px >> 16) & 0xff], -1).astype(np.uint8)[::-1]    # 0x00BBGGRR, row 0 = bottom
        np.testing.assert_array_equal(_png_pixels(p), want)
        try:
            from PIL import Image
            np.testing.assert_array_equal(np.array(Image.open(p)), want)
        except ImportError:
            pass
    rgb = rng.random((5, 7, 3)).astype(np.float32)
    p = str(tmp_path / "t.pfm")
    gpt.write_pfm(p, rgb, 7, 5)
    blob = open(p, "rb").read()
    assert blob.startswith(b"PF\n7 5\n-1.0\n")
    np.testing.assert_array_equal(np.frombuffer(blob[len(b"PF\n7 5\n-1.0\n"):], np.float32).reshape(5, 7, 3), rgb)


def _exr_pixels(path):
    """Minimal reader of what gpt_write_exr writes (OpenEXR file layout: magic, version, attributes, offset table, scan lines)."""
    import struct
    b = open(path, "rb").read()
    assert struct.unpack_from("<ii", b, 0) == (20000630, 2)
    pos, attrs = 8, {}
    while b[pos] != 0:
        e = b.index(b"\0", pos); name = b[pos:e].decode(); pos = e + 1
        e = b.index(b"\0", pos); typ = b[pos:e].decode(); pos = e + 1
        (size,) = struct.unpack_from("<i", b, pos); pos += 4
        attrs[name] = (typ, b[pos:pos + size]); pos += size
    pos += 1
    assert attrs["compression"] == ("compression", b"\0") and attrs["lineOrder"] == ("lineOrder", b"\0")
    x0, y0, x1, y1 = struct.unpack("<4i", attrs["dataWindow"][1])
    assert attrs["displayWindow"][1] == attrs["dataWindow"][1] and (x0, y0) == (0, 0)
    w, h = x1 + 1, y1 + 1
    ch, names, q = attrs["channels"][1], [], 0
    while ch[q] != 0:
        e = ch.index(b"\0", q); names.append(ch[q:e].decode())
        assert struct.unpack_from("<i4Bii", ch, e + 1) == (2, 0, 0, 0, 0, 1, 1)      # FLOAT, pLinear 0, sampling 1 x 1
        q = e + 1 + 16
    assert names == ["B", "G", "R"] and q == len(ch) - 1
    offs = struct.unpack_from("<%dQ" % h, b, pos)
    img = np.zeros((h, w, 3), np.float32)
    for y in range(h):
        yy, nbytes = struct.unpack_from("<ii", b, offs[y])
        assert yy == y and nbytes == 12 * w
        rows = np.frombuffer(b, np.float32, 3 * w, offs[y] + 8).reshape(3, w)
        img[y, :, 2], img[y, :, 1], img[y, :, 0] = rows[0], rows[1], rows[2]
    assert offs[-1] + 8 + 12 * w == len(b)
    return img


def test_exr_writer(gpt, tmp_path):
    rng = np.random.default_rng(1)
    for w, h in [(7, 5), (64, 33)]:
        rgb = (rng.random((h, w, 3)) * 10.0).astype(np.float32)
        rgb[0, 0] = (np.inf, -0.0, 1e-40)                            # float data is stored untouched
        p = str(tmp_path / ("t%d.exr" % w))
        gpt.write_exr(p, rgb, w, h)
        got = _exr_pixels(p)
        np.testing.assert_array_equal(got.view(np.uint32), rgb[::-1].view(np.uint32))     # row 0 of the buffer is the bottom row
    # an independent reader where the image has one: OpenCV's OpenEXR codec (BGR, top row first)
    os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")
    try:
        import cv2
        img = cv2.imread(p, cv2.IMREAD_UNCHANGED)
    except Exception:
        img = None
    if img is not None:
        finite = rgb[::-1][..., ::-1].copy()
        np.testing.assert_array_equal(img[1:], finite[1:])           # row 0 holds the inf / denormal probes
    with pytest.raises(gpt.GptError):
        gpt.write_exr(str(tmp_path / "no_such_dir" / "x.exr"), rgb, w, h)


def test_load_obj_with_mtl(gpt, tmp_path):
    """N4: usemtl per face + Kd / Ke / illum from the .mtl the reference's loader parses and drops."""
    (tmp_path / "m.mtl").write_text("newmtl red\nKd 0.8 0.1 0.1\nKe 0 0 0\nillum 2\n\nnewmtl lamp\nKd 1 1 1\nKe 4 4 3\nillum 2\n"
                                    "newmtl mirror\nKd 0.9 0.9 0.9\nillum 5\nnewmtl glass\nillum 7\nnewmtl unused\nKd 0 0 0\n")
    (tmp_path / "m.obj").write_text("mtllib m.mtl\nv 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nf 1 2 3\nusemtl red\nf 1 2 3 4\nusemtl lamp\nf 1 3 4\n"
                                    "usemtl mirror\nf 1 2 4\nusemtl glass\nf 2 3 4\nusemtl red\nf 1 2 3\nusemtl notinmtl\nf 1 2 3\n")
    v, f, fm, mats = gpt.load_obj_materials(str(tmp_path / "m.obj"))
    assert f.shape == (8, 3) and fm.tolist() == [-1, 0, 0, 1, 2, 3, 0, 4] and len(mats) == 5
    assert np.allclose(list(mats[0].col), [0.8, 0.1, 0.1]) and mats[0].mat == gpt.DIFF
    assert np.allclose(list(mats[1].emi), [4, 4, 3]) and mats[2].mat == gpt.SPEC and mats[3].mat == gpt.REFR
    assert np.allclose(list(mats[4].col), [246 / 256, 246 / 255, 70 / 255]) and mats[4].mat == gpt.DIFF     # BasicScene.cpp:234-236 default


def test_accum_checkpoint_roundtrip(gpt, tmp_path):
    import pytest
    rng = np.random.default_rng(1)
    acc = rng.random((5, 7, 4)).astype(np.float32)
    p = str(tmp_path / "a.gptacc")
    gpt.save_accum(p, acc, 12345678901, 0xDEADBEEFCAFE)
    back, n, seed = gpt.load_accum(p, 7, 5)
    np.testing.assert_array_equal(back, acc)
    assert n == 12345678901 and seed == 0xDEADBEEFCAFE
    with pytest.raises(gpt.GptError):
        gpt.load_accum(p, 8, 5)                                       # size mismatch is an error, not a reinterpretation
