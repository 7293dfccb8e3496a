"""Generates the golden fixtures under tests/golden/ from the UNMODIFIED reference (oracle/_ref/mcpt_ref, built by
oracle/build_ref.sh from /root/reference).  Run in the build container only (the GPU box has no /root/reference):

    python tests/golden/make_golden.py [--images]

Fixtures (per scene S in cornell-box, veach-mis):
  S_morton.npz   codes (OBJ order, uint32), perm (stable-sorted leaf -> OBJ face), centroids (float32)
  S_rays.npz     rays (n,6 float64: camera, bounce and light-directed rays), reference hit/face/t/p/pn
  S_image.npz    two independent reference renders (time-seeded RNG) at reduced resolution + their wall ms
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle_py as O  # noqa: E402
import montecarlopathtracing_b200 as M  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
IMAGE_CFG = {"cornell-box": (256, 256, 64), "veach-mis": (200, 150, 32)}


def camera_rays(cam, w, h, step):
    """Primary rays of generateImg (MTPC/pathTracing.cpp:276-308) for every step-th pixel, in float64."""
    eye, look, up = np.array(cam.eye[:]), np.array(cam.lookat[:]), np.array(cam.up[:])
    up = up / np.linalg.norm(up)
    d = look - eye
    l = np.linalg.norm(d)
    dy = np.tan(cam.fovy / 2 / 180 * 3.1415926) * l
    dx = dy / h * w
    X = np.cross(d, up); X /= np.linalg.norm(X)
    start = look - X * dx + up * dy
    ii, jj = np.meshgrid(np.arange(0, h, step), np.arange(0, w, step), indexing="ij")
    pos = start[None, :] - up[None, :] * (2 * dy / h) * ii.reshape(-1, 1) + X[None, :] * (2 * dx / w) * jj.reshape(-1, 1)
    dirs = pos - eye[None, :]
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    return np.concatenate([np.broadcast_to(eye, dirs.shape), dirs], axis=1)


def make_rays(scene, name, rng):
    cam = scene.camera
    step = max(1, int(np.sqrt(cam.width * cam.height / 6000)))
    prim = camera_rays(cam, cam.width, cam.height, step)
    first = O.ref_trace(O.REF_SCENE_DIR + name, prim)
    hit = first["hit"] == 1
    p = first["p"][hit]
    # bounce rays: random directions from the hit points, origin offset 0.01*d as nextRay does (:132)
    d = rng.normal(size=p.shape); d /= np.linalg.norm(d, axis=1, keepdims=True)
    bounce = np.concatenate([p + 0.01 * d, d], axis=1)
    # light-directed rays (:202-209): toward random points on the light materials' triangles
    light_mats = [l.mat_id for l in scene.lights]
    lf = np.nonzero(np.isin(scene.mat_id, light_mats))[0]
    pick = rng.choice(lf, size=p.shape[0])
    w = rng.random((p.shape[0], 3)); w /= w.sum(axis=1, keepdims=True)
    xl = (scene.pos[pick] * w[:, :, None]).sum(axis=1)
    d2 = xl - p; d2 /= np.linalg.norm(d2, axis=1, keepdims=True)
    shadow = np.concatenate([p + 0.01 * d2, d2], axis=1)
    # a few non-unit directions and rays that start inside/outside the scene box
    o = rng.uniform(-2, 3, size=(500, 3)); dd = rng.normal(size=(500, 3)) * rng.uniform(0.1, 5, size=(500, 1))
    misc = np.concatenate([o, dd], axis=1)
    # float32-representable: mcpt_trace takes FP32 rays, so the reference must see exactly the same rays
    return np.concatenate([prim, bounce, shadow, misc], axis=0).astype(np.float32).astype(np.float64)


def main():
    if not O.have_ref():
        O.build()
    rng = np.random.default_rng(20261104)
    for name in ("cornell-box", "veach-mis"):
        scene = M.HostScene(os.path.join(ROOT, "scenes") + "/", name)
        base = O.REF_SCENE_DIR + name
        codes, perm, cent = O.ref_morton(base)
        np.savez_compressed(os.path.join(OUT, name + "_morton.npz"), codes=codes, perm=perm, centroids=cent)
        rays = make_rays(scene, name, rng)
        ref = O.ref_trace(base, rays)
        np.savez_compressed(os.path.join(OUT, name + "_rays.npz"), rays=rays, hit=ref["hit"].astype(np.int8), face=ref["face"],
                            t=ref["t"], p=ref["p"], pn=ref["pn"])
        print(name, "rays", rays.shape[0], "hit fraction", ref["hit"].mean())
        if "--images" in sys.argv:
            w, h, spp = IMAGE_CFG[name]
            imgs, mss = [], []
            for _ in range(2):
                img, ms = O.ref_image(base, spp, w, h)
                imgs.append(img.astype(np.float32)); mss.append(ms)
                time.sleep(1.1)  # the reference seeds its RNG with time(NULL)
            np.savez_compressed(os.path.join(OUT, name + "_image.npz"), a=imgs[0], b=imgs[1], spp=spp, ms=np.array(mss))
            print(name, "image", w, h, spp, "ms", mss)


if __name__ == "__main__":
    main()
