"""Scene ingest on the GPU (csrc/obj_ingest.cu) against the host loader (host/scene_loader.cpp, the
reference's read_obj grammar, MTPC/sceneManagement.cpp:76-189): the flattened FP64 arrays must be BIT-identical, the
errors the same, and everything downstream (Morton codes, image) unchanged."""
import os
import random

import numpy as np
import pytest

import montecarlopathtracing_b200 as M
from conftest import SCENES

pytestmark = pytest.mark.gpu

MTL = """newmtl red
Kd 0.7 0.1 0.1
newmtl white
Kd 0.8 0.8 0.8
Ks 0.1 0.1 0.1
Ns 20
newmtl lamp
Kd 0 0 0
"""
CAMERA = """eye 0 1 5
lookat 0 1 4
up 0 1 0
fovy 40
width 64
height 48
mtlname lamp 10 10 10
"""


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint64)


def _write_scene(d, obj_text, mtl=MTL, camera=CAMERA, name="s"):
    for ext, text in ((".obj", obj_text), (".mtl", mtl), (".camera", camera)):
        with open(os.path.join(d, name + ext), "wb") as f:
            f.write(text.encode() if isinstance(text, str) else text)
    return str(d) + "/", name


def _assert_same_as_host(ctx, path, name):
    host = M.HostScene(path, name)
    cam, ms = ctx.load_scene_gpu(path, name)
    pos, nrm, uv, mat = ctx.get_scene_arrays()
    assert pos.shape[0] == host.n_faces
    assert np.array_equal(_bits(pos), _bits(host.pos)), "positions differ"
    assert np.array_equal(_bits(nrm), _bits(host.nrm)), "normals differ"
    assert np.array_equal(_bits(uv), _bits(host.uv)), "texcoords differ"
    assert np.array_equal(mat, host.mat_id)
    for k in ("eye", "lookat", "up"):
        assert list(getattr(cam, k)) == list(getattr(host.camera, k))
    assert (cam.fovy, cam.width, cam.height) == (host.camera.fovy, host.camera.width, host.camera.height)
    return host, cam


@pytest.mark.parametrize("name", ["cornell-box", "veach-mis"])
def test_shipped_scenes_bit_identical_and_same_image(gpu_ctx, name):
    host, cam = _assert_same_as_host(gpu_ctx, SCENES, name)
    gpu_ctx.build_bvh()
    codes_a = gpu_ctx.get_morton()
    gpu_ctx.set_seed(11)
    img_a = gpu_ctx.render(cam, 2, 160, 120)
    gpu_ctx.upload_scene(host)
    gpu_ctx.build_bvh()
    codes_b = gpu_ctx.get_morton()
    img_b = gpu_ctx.render(host.camera, 2, 160, 120)
    for a, b in zip(codes_a, codes_b):
        assert np.array_equal(a, b)
    assert np.array_equal(img_a.view(np.uint32), img_b.view(np.uint32))   # light tables, textures, materials: all the same


def _random_obj(seed, n_v, n_f, eol="\n"):
    rng = random.Random(seed)
    fmt = ["%.6f", "%g", "%.9g", "%.17g", "%.3e", "%r", "%.25f", "%.30e", "%d"]
    lines = ["# generated", "", "mtllib s.mtl", "o thing", "usemtl white"]
    nv = nvn = nvt = 0
    def num():
        x = rng.uniform(-3, 3) * 10 ** rng.choice([0, 0, 0, -3, 2, -12, -310, 300])
        f = rng.choice(fmt)
        if abs(x) > 1e6 and f in ("%d", "%.6f", "%.25f"):      # would print hundreds of digits; > 40 is a reported error
            f = "%.17g"
        return f % (int(x) if f == "%d" else x)
    sep = lambda: rng.choice([" ", " ", " ", "  ", "\t", " \t "])
    faces = 0
    while faces < n_f:
        r = rng.random()
        if nv < 3 or r < 0.25:
            lines.append(rng.choice(["", " "]) + "v" + sep() + sep().join(num() for _ in range(rng.choice([3, 3, 3, 2, 4]))) + rng.choice(["", " ", "\t"]))
            nv += 1
            lines.append("vn" + sep() + sep().join(num() for _ in range(3)))
            nvn += 1
            if rng.random() < 0.8 or nvt == 0:
                lines.append("vt" + sep() + sep().join(num() for _ in range(rng.choice([2, 2, 3, 1]))))
                nvt += 1
            if nv >= n_v and nv >= 3:
                pass
        elif r < 0.3:
            lines.append(rng.choice(["usemtl red", "usemtl white", "usemtl  lamp ", "usemtl", "s off", "g grp", "#f 1 2 3", "vp 1 2 3"]))
        else:
            def corner():
                a, b, c = rng.randint(1, nv), rng.randint(1, nvn), rng.randint(1, nvt)
                k = rng.random()
                if k < 0.6:
                    return "%d/%d/%d" % (a, b, c)
                m = min(nv, nvn, nvt)
                a = rng.randint(1, m)
                return str(a) if k < 0.8 else "%d/%d" % (a, rng.randint(1, nvn))
            lines.append("f" + sep() + sep().join(corner() for _ in range(3)))
            faces += 1
    return eol.join(lines) + rng.choice(["", eol])


@pytest.mark.parametrize("seed,eol,n_f", [(1, "\n", 50), (2, "\r\n", 400), (3, "\n", 20000), (4, "\r\n", 60000)])
def test_random_obj_text_matches_host_loader(gpu_ctx, tmp_path, seed, eol, n_f):
    """Every number format OBJ writers use (and some they do not: 25-30 digit mantissas, 1e-310, 1e300 — the 896-bit
    path), blanks and tabs, CRLF, comments, statements the reader ignores, corners given as i, i/j and i/j/k."""
    path, name = _write_scene(tmp_path, _random_obj(seed, 1000, n_f, eol))
    _assert_same_as_host(gpu_ctx, path, name)


def test_lines_longer_than_a_chunk_and_missing_final_newline(gpu_ctx, tmp_path):
    pad = " " * 700
    obj = ("usemtl red\n" + "v 0 0 0" + pad + "\nv" + pad + "1 0 0\nv 0 1 0\n# " + "x" * 1000 + "\nvn 0 0 1\nvt 0.5 0.5\n" +
           "f 1/1/1" + pad + "2/1/1 3/1/1\nusemtl lamp\nf 3/1/1 2/1/1 1/1/1")
    path, name = _write_scene(tmp_path, obj)
    host, _ = _assert_same_as_host(gpu_ctx, path, name)
    assert host.n_faces == 2


BAD = {
    "quad": ("usemtl red\nv 0 0 0\nv 1 0 0\nv 0 1 0\nv 1 1 0\nvn 0 0 1\nvt 0 0\nf 1/1/1 2/1/1 3/1/1 4/1/1\n", "only triangles"),
    "two corners": ("usemtl red\nv 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nvt 0 0\nf 1/1/1 2/1/1\n", "only triangles"),
    "unknown material": ("usemtl nope\nv 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nvt 0 0\nf 1/1/1 2/1/1 3/1/1\n", "unknown material"),
    "no usemtl": ("v 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nvt 0 0\nf 1/1/1 2/1/1 3/1/1\n", "unknown material"),
    "index too large": ("usemtl red\nv 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nvt 0 0\nf 1/1/1 2/1/1 4/1/1\n", "out of range"),
    "forward reference": ("usemtl red\nv 0 0 0\nv 1 0 0\nvn 0 0 1\nvt 0 0\nf 1/1/1 2/1/1 3/1/1\nv 0 1 0\n", "out of range"),
    "negative index": ("usemtl red\nv 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nvt 0 0\nf 1/1/1 2/1/1 -1/1/1\n", "out of range"),
    "empty vt slot": ("usemtl red\nv 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nvt 0 0\nf 1//1 2//1 3//1\n", "out of range"),
    "first error wins": ("usemtl red\nv 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nvt 0 0\nf 1/1/1 2/1/1 9/1/1\nusemtl nope\nf 1/1/1 2/1/1 3/1/1\n"
                         "f 1/1/1 2/1/1 3/1/1 1/1/1\n", "out of range"),
    "material before index": ("usemtl nope\nv 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nvt 0 0\nf 1/1/1 2/1/1 9/1/1\n", "unknown material"),
    "no faces": ("usemtl red\nv 0 0 0\n", "no faces"),
}


@pytest.mark.parametrize("case", sorted(BAD))
def test_errors_match_the_host_loader(gpu_ctx, tmp_path, case):
    obj, what = BAD[case]
    path, name = _write_scene(tmp_path, obj)
    with pytest.raises(M.McptError) as host_err:
        M.HostScene(path, name)
    assert what in str(host_err.value)
    with pytest.raises(M.McptError) as gpu_err:
        gpu_ctx.load_scene_gpu(path, name)
    assert what in str(gpu_err.value), str(gpu_err.value)
    with pytest.raises(M.McptError):              # and the context holds no half-read scene
        gpu_ctx.build_bvh()


def test_numbers_atof_reads_differently_are_reported(gpu_ctx, tmp_path):
    for tok in ("inf", "nan", "0x1p3", "0." + "1234567890" * 5):
        obj = "usemtl red\nv 0 0 0\nv 1 0 %s\nv 0 1 0\nvn 0 0 1\nvt 0 0\nf 1/1/1 2/1/1 3/1/1\n" % tok
        path, name = _write_scene(tmp_path, obj)
        with pytest.raises(M.McptError) as e:
            gpu_ctx.load_scene_gpu(path, name)
        assert "atof" in str(e.value)


def test_render_scene_entry_point_uses_the_gpu_ingest(gpu_ctx, tmp_path, scenes):
    """mcpt_render_scene = ingest on the GPU + build + render + device PNG: same file as the host-loader path."""
    s = scenes["cornell-box"]
    M.render_scene(SCENES, "cornell-box", 1, out_dir=str(tmp_path) + "/")
    gpu_ctx.upload_scene(s)
    gpu_ctx.build_bvh()
    gpu_ctx.set_seed(0)
    f = str(tmp_path / "host.png")
    M.write_png(f, gpu_ctx.render(s.camera, 1))
    assert open(f, "rb").read() == open(str(tmp_path / "cornell-box-SPP1.png"), "rb").read()


@pytest.mark.parametrize("name", ["cornell-box", "veach-mis", "glass-box"])
def test_gpu_ingest_equals_what_the_reference_parsed(gpu_ctx, name):
    """K16-K18 against the arrays the UNMODIFIED reference's read_scene produced (tests/golden/*_scene.npz), bit for bit."""
    from conftest import GOLDEN, SCENES
    g = np.load(os.path.join(GOLDEN, name + "_scene.npz"))
    gpu_ctx.load_scene_gpu(SCENES, name)
    pos, nrm, uv, mat = gpu_ctx.get_scene_arrays()
    assert np.array_equal(pos, g["pos"]) and np.array_equal(nrm, g["nrm"]) and np.array_equal(uv, g["uv"])
    import montecarlopathtracing_b200 as M
    names = M.HostScene(SCENES, name).material_names
    assert [names[i] for i in mat] == [str(g["mat_names"][i]) for i in g["mat_index"]]
