"""TEST INFRASTRUCTURE: ctypes binding of oracle/_build/liboracle.so (the CPU restatement) and helpers to run the
unmodified reference binary oracle/_ref/mcpt_ref.  Imported only by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs."""
import ctypes as C
import os
import re
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_build", "liboracle.so")
REF_BIN = os.path.join(HERE, "_ref", "mcpt_ref")
REF_SCENE_DIR = os.path.join(HERE, "_ref", "scene") + "/"
REF_CWD = os.path.join(HERE, "_ref", "cwd")

_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", HERE, "_build/liboracle.so"])
    if os.path.isdir(os.environ.get("MCPT_REFERENCE", "/root/reference")):
        subprocess.check_call(["bash", os.path.join(HERE, "build_ref.sh")], stdout=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        _lib = C.CDLL(LIB_PATH)
        _lib.orc_build.restype = C.c_void_p
        _lib.orc_build.argtypes = [C.c_void_p]
        _lib.orc_free.argtypes = [C.c_void_p]
        _lib.orc_morton.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        _lib.orc_sort.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        _lib.orc_get_order.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        _lib.orc_light_area.restype = C.c_double
        _lib.orc_light_area.argtypes = [C.c_void_p, C.c_int]
        _lib.orc_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib.orc_render.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64,
                                    C.c_void_p, C.c_void_p]
        _lib.orc_render_flags.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_int,
                                          C.c_void_p, C.c_void_p]
        _lib.orc_philox.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    return _lib


def morton(pos):
    pos = np.ascontiguousarray(pos, np.float64).reshape(-1, 9)
    codes = np.empty(pos.shape[0], np.uint32)
    lib().orc_morton(pos.ctypes.data, pos.shape[0], codes.ctypes.data)
    return codes


def stable_order(codes):
    codes = np.ascontiguousarray(codes, np.uint32)
    perm = np.empty(codes.shape[0], np.int32)
    lib().orc_sort(codes.ctypes.data, codes.shape[0], perm.ctypes.data)
    return perm


def philox(ctr, key):
    c = np.asarray(ctr, np.uint32); k = np.asarray(key, np.uint32); o = np.empty(4, np.uint32)
    lib().orc_philox(c.ctypes.data, k.ctypes.data, o.ctypes.data)
    return o


class OracleScene:
    """orc_build over a mcpt_scene_desc (ctypes struct from the product's host loader or make_desc)."""

    def __init__(self, desc):
        self._desc = desc
        self.n = desc.n_faces
        self._s = lib().orc_build(C.addressof(desc))

    def close(self):
        if self._s:
            lib().orc_free(self._s)
            self._s = None

    def __del__(self):
        self.close()

    def order(self):
        codes, perm = np.empty(self.n, np.uint32), np.empty(self.n, np.int32)
        lib().orc_get_order(self._s, codes.ctypes.data, perm.ctypes.data)
        return codes, perm

    def light_area(self, l):
        return lib().orc_light_area(self._s, l)

    def trace(self, rays):
        rays = np.ascontiguousarray(rays, np.float64).reshape(-1, 6)
        n = rays.shape[0]
        hit, face, leaf = np.empty(n, np.int32), np.empty(n, np.int32), np.empty(n, np.int32)
        rec, cnt = np.empty((n, 7), np.float64), np.zeros(2, np.float64)
        lib().orc_trace(self._s, rays.ctypes.data, n, hit.ctypes.data, face.ctypes.data, leaf.ctypes.data, rec.ctypes.data,
                        cnt.ctypes.data)
        return {"hit": hit, "face": face, "leaf": leaf, "t": rec[:, 0], "p": rec[:, 1:4], "pn": rec[:, 4:7],
                "box_tests": cnt[0], "tri_tests": cnt[1]}

    def render(self, camera, spp, width=0, height=0, s0=0, s1=None, seed=0, flags=0):
        """flags: MCPT_FLAG_JITTER (8) / MCPT_FLAG_ANYHIT_SHADOW (16), the twins of the CUDA integrator's quality extensions."""
        w = width if width > 0 else camera.width
        h = height if height > 0 else camera.height
        img, cnt = np.zeros((h, w, 3), np.float64), np.zeros(3, np.float64)
        lib().orc_render_flags(self._s, C.addressof(camera), w, h, spp, s0, spp if s1 is None else s1, seed, flags, img.ctypes.data,
                               cnt.ctypes.data)
        return img, {"primary": cnt[0], "shadow": cnt[1], "extend": cnt[2]}


# ---- the unmodified reference binary ---------------------------------------------------------------------
def have_ref():
    return os.path.exists(REF_BIN) and os.path.isdir(REF_SCENE_DIR)


def _run_ref(args, threads=None):
    env = dict(os.environ)
    env["MCPT_TEXDIR"] = REF_CWD
    if threads:
        env["OMP_NUM_THREADS"] = str(threads)
    r = subprocess.run([REF_BIN] + [str(a) for a in args], cwd=REF_CWD, env=env, stdout=subprocess.DEVNULL,
                       stderr=subprocess.PIPE, text=True)
    if r.returncode != 0:
        raise RuntimeError("mcpt_ref failed: " + r.stderr[-2000:])
    return r.stderr


def _ms(stderr, tag):
    return float(re.search(tag + r" ([0-9.]+)", stderr).group(1))


def ref_morton(scene_base):
    with tempfile.NamedTemporaryFile(suffix=".bin") as f:
        _run_ref(["morton", scene_base, f.name])
        raw = np.fromfile(f.name, np.uint8)
    n = int(np.frombuffer(raw[:4], np.int32)[0])
    codes = np.frombuffer(raw[4:4 + 4 * n], np.uint32).copy()
    perm = np.frombuffer(raw[4 + 4 * n:4 + 8 * n], np.int32).copy()
    cent = np.frombuffer(raw[4 + 8 * n:4 + 20 * n], np.float32).reshape(n, 3).copy()
    return codes, perm, cent


def ref_trace(scene_base, rays):
    rays = np.ascontiguousarray(rays, np.float64).reshape(-1, 6)
    n = rays.shape[0]
    with tempfile.TemporaryDirectory() as d:
        fin, fout = os.path.join(d, "in.bin"), os.path.join(d, "out.bin")
        with open(fin, "wb") as f:
            f.write(np.int32(n).tobytes()); f.write(rays.tobytes())
        _run_ref(["trace", scene_base, fin, fout])
        raw = np.fromfile(fout, np.uint8)
    rec = np.frombuffer(raw[4:], np.dtype([("hit", "i4"), ("face", "i4"), ("d", "f8", 7)]))
    return {"hit": rec["hit"].copy(), "face": rec["face"].copy(), "t": rec["d"][:, 0].copy(), "p": rec["d"][:, 1:4].copy(),
            "pn": rec["d"][:, 4:7].copy()}


def ref_image(scene_base, spp, width=0, height=0, with_png=False, threads=None):
    """generateImg of the unmodified reference; returns (image float64 HxWx3, wall ms of generateImg) and, with_png, also the
    bytes of the PNG the reference's own imshow + svpng (MTPC/MTPC.cpp:10-33) wrote for that image."""
    with tempfile.TemporaryDirectory() as d:
        out = os.path.join(d, "img.bin")
        args = ["image", scene_base, spp, width, height, out]
        if with_png:
            args.append(os.path.join(d, "ref"))
        err = _run_ref(args, threads=threads)
        raw = np.fromfile(out, np.uint8)
        png = open(os.path.join(d, "ref-SPP%d.png" % spp), "rb").read() if with_png else None
    w, h, _ = np.frombuffer(raw[:12], np.int32)
    img = np.frombuffer(raw[12:], np.float64).reshape(h, w, 3).copy()
    return (img, _ms(err, "REF_IMAGE_MS"), png) if with_png else (img, _ms(err, "REF_IMAGE_MS"))


def ref_png(img):
    """imshow + svpng of the unmodified reference on a caller-supplied image (HxWx3 doubles) -> the PNG file's bytes."""
    img = np.ascontiguousarray(img, np.float64)
    h, w = img.shape[:2]
    with tempfile.TemporaryDirectory() as d:
        img.tofile(os.path.join(d, "in.bin"))
        _run_ref(["png", w, h, os.path.join(d, "in.bin"), os.path.join(d, "ref"), 1])
        return open(os.path.join(d, "ref-SPP1.png"), "rb").read()


def ref_scene(scene_base):
    """What the reference's read_scene parsed (MTPC/sceneManagement.cpp:76-189), per face in OBJ order:
    pos (n,3,3), nrm (n,3,3), uv (n,3,2) float64 and the material names."""
    with tempfile.NamedTemporaryFile(suffix=".bin") as f:
        _run_ref(["scene", scene_base, f.name])
        raw = open(f.name, "rb").read()
    n = int(np.frombuffer(raw[:4], np.int32)[0])
    geo, names, off = np.empty((n, 24)), [], 4
    for i in range(n):
        geo[i] = np.frombuffer(raw[off:off + 192], np.float64); off += 192
        ln = int(np.frombuffer(raw[off:off + 4], np.int32)[0]); off += 4
        names.append(raw[off:off + ln].decode()); off += ln
    return geo[:, 0:9].reshape(n, 3, 3), geo[:, 9:18].reshape(n, 3, 3), geo[:, 18:24].reshape(n, 3, 2), names


def ref_build_ms(scene_base, reps=5):
    return _ms(_run_ref(["buildtime", scene_base, reps]), "REF_BUILD_MS")
