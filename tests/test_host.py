"""CPU: the C++ host shell — scene loaders (reference grammar, CRLF-safe), flattening, PNG output."""
import os
import struct
import zlib

import numpy as np
import pytest

import montecarlopathtracing_b200 as M
from conftest import GOLDEN, SCENES


def _parse_obj_independently(path):
    """Plain-Python OBJ reader used as the checker for the C++ loader."""
    v, vn, vt, faces, mats, cur = [], [], [], [], [], None
    with open(path, "rb") as f:
        for raw in f:
            t = raw.decode().split()
            if not t:
                continue
            if t[0] == "v":
                v.append([float(x) for x in t[1:4]])
            elif t[0] == "vn":
                vn.append([float(x) for x in t[1:4]])
            elif t[0] == "vt":
                vt.append([float(x) for x in t[1:3]])
            elif t[0] == "usemtl":
                cur = t[1]
            elif t[0] == "f":
                faces.append([[int(i) - 1 for i in c.split("/")] for c in t[1:4]])
                mats.append(cur)
    return np.array(v), np.array(vn), np.array(vt), np.array(faces), mats


@pytest.mark.parametrize("name,n_faces,n_mat,n_light", [("cornell-box", 15056, 7, 1), ("veach-mis", 3812, 10, 5)])
def test_loader_matches_independent_parse(scenes, name, n_faces, n_mat, n_light):
    s = scenes[name]
    assert s.n_faces == n_faces and len(s.materials) == n_mat and len(s.lights) == n_light
    v, vn, vt, faces, mats = _parse_obj_independently(SCENES + name + ".obj")
    assert np.array_equal(s.pos, v[faces[:, :, 0]])
    # the reference reads the 2nd index into vn[] and the 3rd into vt[] (MTPC/sceneManagement.cpp:145-170)
    assert np.array_equal(s.nrm, vn[faces[:, :, 1]])
    assert np.array_equal(s.uv, vt[faces[:, :, 2]])
    # material ids follow newmtl order of the .mtl; CRLF must not leak into names (lights would never match)
    names = [l.split()[1] for l in open(SCENES + name + ".mtl", "rb").read().decode().splitlines() if l.startswith("newmtl")]
    assert [names[i] for i in s.mat_id] == mats


def test_cornell_materials_camera_lights(scenes):
    s = scenes["cornell-box"]
    cam = s.camera
    assert (cam.width, cam.height, cam.fovy) == (1024, 1024, 19.5)
    assert list(cam.eye) == [0, 1, 6.8] and list(cam.lookat) == [0, 1, 5.8] and list(cam.up) == [0, 1, 0]
    m = s.materials
    assert list(m[0].kd) == [0.63, 0.065, 0.05] and m[0].tex_id == -1           # LeftWall
    assert list(m[6].kd) == [1, 1, 1] and m[6].tex_id == 0                       # Table, map_Kd
    assert (s.textures[0].height, s.textures[0].width) == (408, 612)
    assert s.lights[0].mat_id == 5 and list(s.lights[0].radiance) == [50, 50, 50]


def test_veach_lights_in_camera_file_order(scenes):
    s = scenes["veach-mis"]
    rad = [l.radiance[0] for l in s.lights]
    assert rad == [901.803, 100.0, 11.1111, 1.23457, 800.0]                      # Light1..Light5, trailing blank tolerated
    assert [l.mat_id for l in s.lights] == [9, 8, 7, 6, 5]
    planes = [m for m in s.materials if any(k != 0 for k in m.ks)]
    assert len(planes) == 4 and sorted(p.Ns for p in planes) == [300, 1000, 10000, 100000]


def test_loader_errors_are_reported(tmp_path):
    with pytest.raises(M.McptError, match="cannot open"):
        M.HostScene(str(tmp_path) + "/", "nope")
    (tmp_path / "a.mtl").write_text("newmtl A\nKd 1 1 1\n")
    (tmp_path / "a.obj").write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nvt 0 0\nusemtl B\nf 1/1/1 2/1/1 3/1/1\n")
    (tmp_path / "a.camera").write_text("eye 0 0 1\n")
    with pytest.raises(M.McptError, match="unknown material"):
        M.HostScene(str(tmp_path) + "/", "a")
    (tmp_path / "a.obj").write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nvt 0 0\nusemtl A\nf 1/1/1 2/1/1 9/1/1\n")
    with pytest.raises(M.McptError, match="out of range"):
        M.HostScene(str(tmp_path) + "/", "a")
    (tmp_path / "a.mtl").write_text("newmtl A\nmap_Kd missing.jpg\n")
    with pytest.raises(M.McptError, match="sidecar"):
        M.HostScene(str(tmp_path) + "/", "a")


def test_tiny_scene_lf_and_tabs(tmp_path):
    (tmp_path / "t.mtl").write_text("newmtl A\nKd 0.5 0.25 1\nKs 0 0 0\nNs 7\nNi 1.5\nnewmtl L\nKd 0 0 0\n")
    (tmp_path / "t.obj").write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nv 0 0 1\nvn 0 0 1\nvn 1 0 0\nvt 0.5 0.25\nvt 1 1\n"
                                    "usemtl A\nf 1/1/1 2/1/1 3/1/1\nusemtl L\nf 1/2/2\t2/2/2 4/2/2\n")
    (tmp_path / "t.camera").write_text("eye 0 0 5\nlookat 0 0 4\nup 0 1 0\nfovy 30\nwidth 8\nheight 4\nmtlname L 1 2 3\n")
    s = M.HostScene(str(tmp_path) + "/", "t")
    assert s.n_faces == 2 and list(s.mat_id) == [0, 1]
    assert s.materials[0].Ns == 7 and s.materials[0].Ni == 1.5
    assert np.array_equal(s.nrm[1], [[1, 0, 0]] * 3) and np.array_equal(s.uv[1], [[1, 1]] * 3)
    assert s.lights[0].mat_id == 1 and list(s.lights[0].radiance) == [1, 2, 3]
    assert (s.camera.width, s.camera.height) == (8, 4)


def _decode_png(path):
    raw = open(path, "rb").read()
    assert raw[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w, h = 8, b"", 0, 0
    while pos < len(raw):
        n, tag = struct.unpack(">I4s", raw[pos:pos + 8])
        body = raw[pos + 8:pos + 8 + n]
        assert struct.unpack(">I", raw[pos + 8 + n:pos + 12 + n])[0] == zlib.crc32(tag + body)
        if tag == b"IHDR":
            w, h, depth, ctype = struct.unpack(">IIBB", body[:10])
            assert (depth, ctype) == (8, 2)
        elif tag == b"IDAT":
            idat += body
        pos += 12 + n
    rows = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + 3 * w)
    assert np.all(rows[:, 0] == 0)
    return rows[:, 1:].reshape(h, w, 3)


def test_png_quantisation_matches_imshow(tmp_path):
    """(uchar)clamp(v*255, 0, 255): truncation, no gamma (MTPC/MTPC.cpp:26-28); file complete and closed."""
    rng = np.random.default_rng(0)
    img = rng.uniform(-0.2, 1.4, size=(37, 301, 3)).astype(np.float32)   # one stored block per scanline
    img[0, 0] = [0.0, 1.0, 0.5]; img[0, 1] = [50.0, np.nan, 254.999 / 255]
    f = str(tmp_path / "x.png")
    M.write_png(f, img)
    got = _decode_png(f)
    want = np.clip(img.astype(np.float64) * 255, 0, 255)
    want = np.where(np.isnan(want), 0, want).astype(np.uint8)
    assert np.array_equal(got, want)
    from PIL import Image
    assert np.array_equal(np.asarray(Image.open(f)), want)
    assert os.path.getsize(f) == M.png_size(301, 37)


PNG_CASES = ["render", "noise_7x5", "one_1x1", "ramp_256x3", "wide_300x2", "tall_1x70"]


@pytest.mark.parametrize("case", PNG_CASES)
def test_png_bytes_equal_the_reference_file(tmp_path, case):
    """mcpt_write_png against files the UNMODIFIED reference's imshow + svpng wrote (tests/golden/png_golden.npz, made by
    make_golden.py --png from oracle/_ref/mcpt_ref): same quantisation, one stored deflate block per scanline, IDAT length
    2 + h*(5 + 3w+1) + 4, Adler-32 and CRCs (MTPC/MTPC.cpp:22-30, MTPC/svpng.inc:77-108) — every byte."""
    g = np.load(os.path.join(GOLDEN, "png_golden.npz"))
    img, want = g[case + "_img"], g[case + "_png"]
    h, w = img.shape[:2]
    f = str(tmp_path / "x.png")
    M.write_png(f, img)
    got = np.fromfile(f, np.uint8)
    assert len(got) == len(want) == M.png_size(w, h) == 8 + 25 + 12 + 2 + h * (5 + 3 * w + 1) + 4 + 12
    assert np.array_equal(got, want), "first differing byte at %d" % int(np.nonzero(got != want)[0][0])


def test_png_size_of_the_shipped_resolutions():
    """svpng's cornell-box file is 3 151 935 bytes (result/cornell-box-SPP*.png), veach-mis 3 245 463."""
    assert M.png_size(1024, 1024) == 3151935 and M.png_size(1200, 900) == 3245463
    with pytest.raises(M.McptError, match="21844"):
        M.write_png("/tmp/never.png", np.zeros((1, 21845, 3), np.float32))
