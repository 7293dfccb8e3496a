"""Python mirror of the reference's driver surface (MTPC/MTPC.cpp:35-68) over the C ABI.

HostScene  = scene_data after read_scene (MTPC/sceneManagement.cpp:264), flattened.
Context    = one GPU: upload -> build_bvh (seam 1) -> trace (seam 2) / render (seam 3).
All compute happens in libmcpt.so on the GPU; numpy is only the host buffer type.
"""
import ctypes as C

import numpy as np

from . import _capi as capi
from ._capi import check, lib


def _np_view(ptr, n, dtype):
    if n == 0:
        return np.zeros(0, dtype=dtype)
    return np.ctypeslib.as_array(ptr, shape=(n,)).view(dtype)


class HostScene:
    """Scene files parsed by the C++ host loader (CRLF-safe port of the reference grammar)."""

    def __init__(self, path, name):
        self._h = C.c_void_p()
        check(lib().mcpt_scene_load(path.encode(), name.encode(), C.byref(self._h)))
        self.desc_ptr = lib().mcpt_scene_get_desc(self._h)
        self.desc = self.desc_ptr.contents
        self.camera = lib().mcpt_scene_get_camera(self._h).contents
        d = self.desc
        n = d.n_faces
        self.n_faces = n
        self.pos = _np_view(d.pos, 9 * n, np.float64).reshape(n, 3, 3)
        self.nrm = _np_view(d.nrm, 9 * n, np.float64).reshape(n, 3, 3)
        self.uv = _np_view(d.uv, 6 * n, np.float64).reshape(n, 3, 2)
        self.mat_id = _np_view(d.mat_id, n, np.int32)
        self.materials = [d.materials[i] for i in range(d.n_materials)]
        self.lights = [d.lights[i] for i in range(d.n_lights)]
        self.textures = [d.textures[i] for i in range(d.n_textures)]
        self.material_names = [lib().mcpt_scene_get_material_name(self._h, i).decode() for i in range(d.n_materials)]

    def close(self):
        if self._h:
            lib().mcpt_scene_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def make_desc(pos, nrm, uv, mat_id, materials, lights, textures=()):
    """Build a SceneDesc from numpy arrays (synthetic scenes). Returns (desc, keepalive)."""
    pos = np.ascontiguousarray(pos, dtype=np.float64).reshape(-1)
    nrm = np.ascontiguousarray(nrm, dtype=np.float64).reshape(-1)
    uv = np.ascontiguousarray(uv, dtype=np.float64).reshape(-1)
    mat_id = np.ascontiguousarray(mat_id, dtype=np.int32)
    n = mat_id.shape[0]
    assert pos.size == 9 * n and nrm.size == 9 * n and uv.size == 6 * n
    mats = (capi.Material * max(len(materials), 1))()
    for i, m in enumerate(materials):
        mats[i].kd[:] = m.get("kd", (0, 0, 0)); mats[i].ks[:] = m.get("ks", (0, 0, 0))
        mats[i].Ns = m.get("Ns", 1.0); mats[i].Ni = m.get("Ni", 1.0); mats[i].tex_id = m.get("tex_id", -1)
    lts = (capi.Light * max(len(lights), 1))()
    for i, l in enumerate(lights):
        lts[i].mat_id = l["mat_id"]; lts[i].radiance[:] = l["radiance"]
    texs = (capi.Texture * max(len(textures), 1))()
    tex_keep = []
    for i, t in enumerate(textures):
        a = np.ascontiguousarray(t, dtype=np.uint8)
        tex_keep.append(a)
        texs[i].height, texs[i].width = a.shape[0], a.shape[1]
        texs[i].bgr = a.ctypes.data_as(C.POINTER(C.c_uint8))
    d = capi.SceneDesc()
    d.n_faces, d.n_materials, d.n_textures, d.n_lights = n, len(materials), len(textures), len(lights)
    d.pos = pos.ctypes.data_as(C.POINTER(C.c_double)); d.nrm = nrm.ctypes.data_as(C.POINTER(C.c_double))
    d.uv = uv.ctypes.data_as(C.POINTER(C.c_double)); d.mat_id = mat_id.ctypes.data_as(C.POINTER(C.c_int32))
    d.materials, d.textures, d.lights = mats, texs, lts
    return d, (pos, nrm, uv, mat_id, mats, lts, texs, tex_keep)


class Context:
    """One GPU. Raises McptError if there is no usable B200 (no CPU fallback exists)."""

    def __init__(self, device=0):
        self._c = C.c_void_p()
        check(lib().mcpt_create(int(device), C.byref(self._c)))
        self.n_faces = 0
        self.last_stats = capi.Stats()

    def close(self):
        if self._c:
            lib().mcpt_destroy(self._c)
            self._c = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, cuda_stream_ptr):
        check(lib().mcpt_set_stream(self._c, C.c_void_p(cuda_stream_ptr)))

    def use_own_stream(self):
        check(lib().mcpt_use_own_stream(self._c))

    def set_seed(self, seed):
        check(lib().mcpt_set_seed(self._c, C.c_uint64(seed)))

    def upload_scene(self, desc):
        """desc: HostScene, SceneDesc or pointer to one."""
        if isinstance(desc, HostScene):
            desc = desc.desc
        check(lib().mcpt_upload_scene(self._c, C.byref(desc)))
        self.n_faces = desc.n_faces

    def ingest_obj(self, obj_text, tables, material_names=None):
        """read_obj on the GPU: obj_text = the .obj file's bytes; tables = HostScene or SceneDesc giving materials, textures
        and lights; material_names resolves usemtl (taken from a HostScene).  Returns the device time in ms."""
        if isinstance(tables, HostScene):
            material_names = tables.material_names if material_names is None else material_names
            tables = tables.desc
        names = (C.c_char_p * max(len(material_names), 1))(*[n.encode() for n in material_names])
        buf = np.frombuffer(obj_text, dtype=np.uint8)
        ms = C.c_float()
        self.n_faces = 0
        check(lib().mcpt_ingest_obj(self._c, C.c_void_p(buf.ctypes.data), buf.size, C.byref(tables), names, C.byref(ms)))
        self.n_faces = int(lib().mcpt_scene_face_count(self._c))
        return ms.value

    def load_scene_gpu(self, path, name):
        """read_scene with the OBJ parsed on the GPU.  Returns (camera, ingest_ms)."""
        cam, ms = capi.Camera(), C.c_float()
        self.n_faces = 0
        check(lib().mcpt_scene_load_gpu(self._c, path.encode(), name.encode(), C.byref(cam), C.byref(ms)))
        self.n_faces = int(lib().mcpt_scene_face_count(self._c))
        return cam, ms.value

    def get_scene_arrays(self):
        """The flattened scene resident on the device: pos (n,3,3), nrm (n,3,3), uv (n,3,2) float64, mat_id (n,) int32."""
        n = int(lib().mcpt_scene_face_count(self._c))
        if n < 0:
            check(n)
        pos, nrm, uv, mat = np.empty((n, 3, 3)), np.empty((n, 3, 3)), np.empty((n, 3, 2)), np.empty(n, np.int32)
        check(lib().mcpt_get_scene_arrays(self._c, pos.ctypes.data, nrm.ctypes.data, uv.ctypes.data, mat.ctypes.data))
        return pos, nrm, uv, mat

    def build_bvh(self):
        ms = C.c_float()
        check(lib().mcpt_build_bvh(self._c, C.byref(ms)))
        return ms.value

    def get_morton(self):
        n = self.n_faces
        codes_obj, codes_sorted = np.empty(n, np.uint32), np.empty(n, np.uint32)
        face_sorted = np.empty(n, np.int32)
        check(lib().mcpt_get_morton(self._c, codes_obj.ctypes.data, codes_sorted.ctypes.data, face_sorted.ctypes.data))
        return codes_obj, codes_sorted, face_sorted

    def get_nodes(self):
        nn = max(self.n_faces - 1, 1)
        nodes = np.empty((nn, 16), np.float32)
        root = C.c_int32()
        check(lib().mcpt_get_nodes(self._c, nodes.ctypes.data, C.byref(root)))
        return nodes, root.value

    def get_traversal_tree(self):
        nn = max(self.n_faces - 1, 1)
        nodes = np.empty((nn, 16), np.float32)
        root = C.c_int32()
        check(lib().mcpt_get_traversal_tree(self._c, nodes.ctypes.data, C.byref(root)))
        return nodes, root.value

    def get_wide_tree(self):
        """(nodes (n_wide, 4, 4) uint32, grid origin (3,), grid cell (3,)) of the 4-wide traversal tree."""
        cap = max(self.n_faces - 1, 1)
        nodes = np.empty((cap, 4, 4), np.uint32)
        n_wide = C.c_int32()
        grid = np.empty(8, np.float32)
        check(lib().mcpt_get_wide_tree(self._c, nodes.ctypes.data, C.byref(n_wide), grid.ctypes.data))
        return nodes[:n_wide.value].copy(), grid[0:3].copy(), grid[4:7].copy()

    def trace(self, org_dir):
        """org_dir: (n,6) float32 host array -> structured array of hits (t,u,v,tri,leaf)."""
        rays = np.ascontiguousarray(org_dir, dtype=np.float32).reshape(-1, 6)
        n = rays.shape[0]
        hits = np.empty(n, dtype=np.dtype([("t", "f4"), ("u", "f4"), ("v", "f4"), ("tri", "i4"), ("leaf", "i4")]))
        st = capi.Stats()
        check(lib().mcpt_trace(self._c, rays.ctypes.data, n, hits.ctypes.data, C.byref(st)))
        self.last_stats = st
        return hits

    def trace_device(self, d_org_dir_ptr, n, d_hits_ptr, flags=0):
        st = capi.Stats()
        check(lib().mcpt_trace_device(self._c, C.c_void_p(d_org_dir_ptr), n, C.c_void_p(d_hits_ptr), flags, C.byref(st)))
        self.last_stats = st
        return st

    def render(self, camera, spp, width=0, height=0, mode=capi.MODE_WAVEFRONT, flags=0, out=None):
        """generateImg: mean radiance image (H,W,3) float32 in a HOST buffer."""
        w = width if width > 0 else camera.width
        h = height if height > 0 else camera.height
        if out is None:
            out = np.empty((h, w, 3), np.float32)
        st = capi.Stats()
        check(lib().mcpt_render(self._c, C.byref(camera), width, height, spp, mode, flags, out.ctypes.data, C.byref(st)))
        self.last_stats = st
        return out

    def encode_png(self, d_rgb_ptr, width, height):
        """Device image (height*width*3 floats at d_rgb_ptr) -> PNG bytes (numpy uint8), encoded on the GPU."""
        n = lib().mcpt_png_size(width, height)
        out = np.empty(n, np.uint8)
        check(lib().mcpt_encode_png(self._c, C.c_void_p(d_rgb_ptr), width, height, out.ctypes.data, n))
        return out

    def render_png(self, camera, spp, file, width=0, height=0, mode=capi.MODE_WAVEFRONT, flags=0):
        st = capi.Stats()
        check(lib().mcpt_render_png(self._c, C.byref(camera), width, height, spp, mode, flags, file.encode(), C.byref(st)))
        self.last_stats = st
        return st

    def render_chunks(self, camera, spp, chunk_begin, chunk_end, d_partial_ptr, width=0, height=0,
                      mode=capi.MODE_WAVEFRONT, flags=0):
        st = capi.Stats()
        check(lib().mcpt_render_chunks(self._c, C.byref(camera), width, height, spp, chunk_begin, chunk_end, mode, flags,
                                       C.c_void_p(d_partial_ptr), C.byref(st)))
        self.last_stats = st
        return st

    def sum_partials(self, d_partials_ptr, n, count, scale, d_out_ptr):
        check(lib().mcpt_sum_partials(self._c, C.c_void_p(d_partials_ptr), n, count, scale, C.c_void_p(d_out_ptr)))

    def measure_read_bandwidth(self, working_set_bytes, repeats=8):
        """GB/s of a 16-B-vector, L1-bypassing read sweep over a working set of that size (L2 below 126 MB, HBM far above)."""
        g = C.c_float()
        check(lib().mcpt_measure_read_bandwidth(self._c, int(working_set_bytes), int(repeats), C.byref(g)))
        return g.value

    # ---- multi-GPU, one process per GPU: the exchange and the fixed-order sum live in libmcpt.so ----
    def comm_init(self, world, rank, comm_id):
        """Join this context to the render's NCCL communicator; comm_id = the 128 bytes rank 0 got from comm_unique_id()."""
        check(lib().mcpt_comm_init(self._c, int(world), int(rank), bytes(comm_id)))
        self._rank = int(rank)

    def comm_destroy(self):
        check(lib().mcpt_comm_destroy(self._c))

    def render_dist(self, camera, spp, width=0, height=0, mode=capi.MODE_WAVEFRONT, flags=0, out=None, to_host=True):
        """Collective render over the communicator's ranks.  Rank 0 returns the mean image (H,W,3) float32 (host buffer `out`, or
        None with to_host=False: the image stays on the device, see device_image_ptr); other ranks return None."""
        w = width if width > 0 else camera.width
        h = height if height > 0 else camera.height
        st = capi.Stats()
        ptr = None
        to_host = to_host and getattr(self, "_rank", 0) == 0
        if to_host:
            if out is None:
                out = np.empty((h, w, 3), np.float32)
            ptr = out.ctypes.data
        check(lib().mcpt_render_dist(self._c, C.byref(camera), width, height, spp, mode, flags, ptr, C.byref(st)))
        self.last_stats = st
        return out if to_host else None

    def device_image_ptr(self):
        return lib().mcpt_device_image(self._c)


def comm_unique_id():
    """128-byte NCCL id made by rank 0; broadcast it to every rank (torch.distributed, MPI, a file) and pass it to comm_init."""
    buf = C.create_string_buffer(capi.COMM_ID_BYTES)
    check(lib().mcpt_comm_unique_id(buf))
    return buf.raw


class MultiContext:
    """One process driving n GPUs (mcpt_multi_*): ncclCommInitAll, SPP split, one exchange, fixed-order sum."""

    def __init__(self, n_gpus, devices=None):
        self._m = C.c_void_p()
        devs = (C.c_int * n_gpus)(*devices) if devices is not None else None
        check(lib().mcpt_multi_create(devs, int(n_gpus), C.byref(self._m)))
        self.n_gpus = n_gpus
        self.last_stats = capi.Stats()

    def close(self):
        if self._m:
            lib().mcpt_multi_destroy(self._m)
            self._m = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_seed(self, seed):
        check(lib().mcpt_multi_set_seed(self._m, C.c_uint64(seed)))

    def upload_scene(self, desc):
        if isinstance(desc, HostScene):
            desc = desc.desc
        check(lib().mcpt_multi_upload_scene(self._m, C.byref(desc)))

    def load_scene_gpu(self, path, name):
        cam = capi.Camera()
        check(lib().mcpt_multi_load_scene_gpu(self._m, path.encode(), name.encode(), C.byref(cam)))
        return cam

    def build_bvh(self):
        ms = C.c_float()
        check(lib().mcpt_multi_build_bvh(self._m, C.byref(ms)))
        return ms.value

    def render(self, camera, spp, width=0, height=0, mode=capi.MODE_WAVEFRONT, flags=0, out=None):
        w = width if width > 0 else camera.width
        h = height if height > 0 else camera.height
        if out is None:
            out = np.empty((h, w, 3), np.float32)
        st = capi.Stats()
        check(lib().mcpt_multi_render(self._m, C.byref(camera), width, height, spp, mode, flags, out.ctypes.data, C.byref(st)))
        self.last_stats = st
        return out

    def render_png(self, camera, spp, file, width=0, height=0, mode=capi.MODE_WAVEFRONT, flags=0):
        st = capi.Stats()
        check(lib().mcpt_multi_render_png(self._m, C.byref(camera), width, height, spp, mode, flags, file.encode(), C.byref(st)))
        self.last_stats = st
        return st


def render_scene(path, filename, spp, out_dir=None, device=0, gpus=1):
    """render_scene(path, filename, spp) of MTPC/MTPC.cpp:35: writes <out_dir><filename>-SPP<spp>.png (gpus > 1: SPP split over
    that many GPUs of this process, gpus = 0: every visible GPU; the same PNG bytes)."""
    od = out_dir.encode() if out_dir else None
    if gpus == 1:
        check(lib().mcpt_render_scene(path.encode(), filename.encode(), int(spp), od, device))
    else:
        check(lib().mcpt_render_scene_multi(path.encode(), filename.encode(), int(spp), od, int(gpus)))
    return True


def write_png(file, rgb):
    a = np.ascontiguousarray(rgb, dtype=np.float32)
    h, w = a.shape[0], a.shape[1]
    check(lib().mcpt_write_png(file.encode(), a.ctypes.data_as(C.POINTER(C.c_float)), w, h))


def encode_png_static(d_rgb_ptr, width, height, device=0, cuda_stream=0):
    """mcpt_encode_png_static: device image -> PNG bytes without any allocation on the library's side (no Context needed); the
    device scratch is a static 8 MiB array, so images beyond about 1920 x 1080 raise McptError."""
    n = max(int(lib().mcpt_png_size(width, height)), 1)
    out = np.empty(n, np.uint8)
    check(lib().mcpt_encode_png_static(int(device), C.c_void_p(d_rgb_ptr), width, height, out.ctypes.data, n, C.c_void_p(cuda_stream)))
    return out


def png_size(width, height):
    """Size in bytes of the uncompressed PNG both the host and the device writer produce."""
    return int(lib().mcpt_png_size(width, height))
