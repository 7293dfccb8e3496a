"""ctypes view of include/mcpt.h (the C ABI).  Loads the in-tree libmcpt.so; never falls back to CPU."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MCPT_LIB") or os.path.join(HERE, "libmcpt.so")   # MCPT_LIB: developer A/B builds (build.py MCPT_BUILD_TAG)

NUM_CHUNKS = 64
MODE_MEGAKERNEL, MODE_WAVEFRONT = 0, 1
FLAG_COUNT, FLAG_NO_PRIMARY_REUSE, FLAG_NO_SHADOW_EARLY_OUT, FLAG_JITTER, FLAG_ANYHIT_SHADOW, FLAG_MIS_BALANCE = 1, 2, 4, 8, 16, 32


class Material(C.Structure):
    _fields_ = [("kd", C.c_double * 3), ("ks", C.c_double * 3), ("Ns", C.c_double), ("Ni", C.c_double),
                ("tex_id", C.c_int32), ("_pad", C.c_int32)]


class Texture(C.Structure):
    _fields_ = [("height", C.c_int32), ("width", C.c_int32), ("bgr", C.POINTER(C.c_uint8))]


class Light(C.Structure):
    _fields_ = [("mat_id", C.c_int32), ("_pad", C.c_int32), ("radiance", C.c_double * 3)]


class Camera(C.Structure):
    _fields_ = [("eye", C.c_double * 3), ("lookat", C.c_double * 3), ("up", C.c_double * 3), ("fovy", C.c_double),
                ("width", C.c_int32), ("height", C.c_int32)]


class SceneDesc(C.Structure):
    _fields_ = [("n_faces", C.c_int32), ("n_materials", C.c_int32), ("n_textures", C.c_int32), ("n_lights", C.c_int32),
                ("pos", C.POINTER(C.c_double)), ("nrm", C.POINTER(C.c_double)), ("uv", C.POINTER(C.c_double)),
                ("mat_id", C.POINTER(C.c_int32)), ("materials", C.POINTER(Material)), ("textures", C.POINTER(Texture)),
                ("lights", C.POINTER(Light))]


class Hit(C.Structure):
    _fields_ = [("t", C.c_float), ("u", C.c_float), ("v", C.c_float), ("tri", C.c_int32), ("leaf", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("samples", C.c_double), ("rays", C.c_double), ("node_visits", C.c_double), ("tri_tests", C.c_double),
                ("render_ms", C.c_float), ("build_ms", C.c_float), ("kernel_launches", C.c_int32), ("integrate_ms", C.c_float),
                ("trace_ms", C.c_float), ("trace_launches", C.c_int32), ("build_launches", C.c_int32), ("exchange_ms", C.c_float)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if not k.startswith("_")}


# name -> (restype, argtypes); every symbol include/mcpt.h declares
SIGNATURES = {
    "mcpt_scene_load": (C.c_int, [C.c_char_p, C.c_char_p, C.POINTER(C.c_void_p)]),
    "mcpt_scene_get_desc": (C.POINTER(SceneDesc), [C.c_void_p]),
    "mcpt_scene_get_camera": (C.POINTER(Camera), [C.c_void_p]),
    "mcpt_scene_free": (None, [C.c_void_p]),
    "mcpt_scene_get_material_name": (C.c_char_p, [C.c_void_p, C.c_int32]),
    "mcpt_ingest_obj": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(SceneDesc), C.POINTER(C.c_char_p), C.POINTER(C.c_float)]),
    "mcpt_ingest_obj_file": (C.c_int, [C.c_void_p, C.c_char_p, C.POINTER(SceneDesc), C.POINTER(C.c_char_p), C.POINTER(C.c_float)]),
    "mcpt_scene_load_gpu": (C.c_int, [C.c_void_p, C.c_char_p, C.c_char_p, C.POINTER(Camera), C.POINTER(C.c_float)]),
    "mcpt_scene_face_count": (C.c_int64, [C.c_void_p]),
    "mcpt_get_scene_arrays": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "mcpt_render_scene": (C.c_int, [C.c_char_p, C.c_char_p, C.c_int, C.c_char_p, C.c_int]),
    "mcpt_write_png": (C.c_int, [C.c_char_p, C.POINTER(C.c_float), C.c_int, C.c_int]),
    "mcpt_last_error": (C.c_char_p, []),
    "mcpt_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "mcpt_destroy": (None, [C.c_void_p]),
    "mcpt_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "mcpt_use_own_stream": (C.c_int, [C.c_void_p]),
    "mcpt_set_seed": (C.c_int, [C.c_void_p, C.c_uint64]),
    "mcpt_upload_scene": (C.c_int, [C.c_void_p, C.POINTER(SceneDesc)]),
    "mcpt_build_bvh": (C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    "mcpt_get_morton": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "mcpt_get_nodes": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_int32)]),
    "mcpt_get_traversal_tree": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_int32)]),
    "mcpt_get_wide_tree": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_int32), C.c_void_p]),
    "mcpt_trace": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.POINTER(Stats)]),
    "mcpt_trace_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.POINTER(Stats)]),
    "mcpt_render": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                              C.POINTER(Stats)]),
    "mcpt_png_size": (C.c_int64, [C.c_int, C.c_int]),
    "mcpt_encode_png": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int64]),
    "mcpt_encode_png_static": (C.c_int, [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int64, C.c_void_p]),
    "mcpt_render_png": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p,
                                  C.POINTER(Stats)]),
    "mcpt_render_chunks": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                     C.c_int, C.c_void_p, C.POINTER(Stats)]),
    "mcpt_sum_partials": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_void_p]),
    "mcpt_measure_read_bandwidth": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, C.POINTER(C.c_float)]),
    "mcpt_comm_unique_id": (C.c_int, [C.c_char_p]),
    "mcpt_comm_init": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_char_p]),
    "mcpt_comm_destroy": (C.c_int, [C.c_void_p]),
    "mcpt_render_dist": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                   C.POINTER(Stats)]),
    "mcpt_device_image": (C.c_void_p, [C.c_void_p]),
    "mcpt_multi_create": (C.c_int, [C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_void_p)]),
    "mcpt_multi_destroy": (None, [C.c_void_p]),
    "mcpt_multi_size": (C.c_int, [C.c_void_p]),
    "mcpt_multi_ctx": (C.c_void_p, [C.c_void_p, C.c_int]),
    "mcpt_multi_set_seed": (C.c_int, [C.c_void_p, C.c_uint64]),
    "mcpt_multi_upload_scene": (C.c_int, [C.c_void_p, C.POINTER(SceneDesc)]),
    "mcpt_multi_load_scene_gpu": (C.c_int, [C.c_void_p, C.c_char_p, C.c_char_p, C.POINTER(Camera)]),
    "mcpt_multi_build_bvh": (C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    "mcpt_multi_render": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                    C.POINTER(Stats)]),
    "mcpt_multi_render_png": (C.c_int, [C.c_void_p, C.POINTER(Camera), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p,
                                        C.POINTER(Stats)]),
    "mcpt_render_scene_multi": (C.c_int, [C.c_char_p, C.c_char_p, C.c_int, C.c_char_p, C.c_int]),
}
COMM_ID_BYTES = 128
PNG_STATIC_BYTES = 8 << 20

_lib = None


def lib():
    """The loaded libmcpt.so.  Raises (loudly) if it has not been built: there is no other code path."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libmcpt.so is not built; run `python -m montecarlopathtracing_b200.build` "
                               "(or __graft_entry__.build()). There is no CPU fallback.")
        _lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(_lib, name)
            fn.restype, fn.argtypes = res, args
    return _lib


class McptError(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        raise McptError("mcpt error %d: %s" % (rc, lib().mcpt_last_error().decode("utf-8", "replace")))
