"""Generates the golden fixtures under tests/golden/ from the UNMODIFIED reference (oracle/_ref/mcpt_ref, built by
oracle/build_ref.sh from /root/reference).  Run in the build container only (the GPU box has no /root/reference):

    python tests/golden/make_golden.py [--base] [--images] [--configs] [--glass] [--png] [--scene-arrays]

Fixtures (per scene S in cornell-box, veach-mis):
  S_morton.npz   codes (OBJ order, uint32), perm (stable-sorted leaf -> OBJ face), centroids (float32)
  S_rays.npz     rays (n,6 float64: camera, bounce and light-directed rays), reference hit/face/t/p/pn
  S_image.npz    two independent reference renders (time-seeded RNG) at reduced resolution + their wall ms
--configs (tens of minutes on 8 cores): the BASELINE.json configs at their stated spp, as point-subsamples of the full image
(the reference shoots one un-jittered ray through a pixel's corner, quirk Q4, so fewer pixels at the same camera are a
subset of the full-resolution image with the same per-pixel estimator):
  cornell-box_cfg2.npz   256 x 256 x 256 spp   (config 2: cornell-box at SPP=256)
  veach-mis_cfg3.npz     300 x 225 x 100 spp   (config 3: veach-mis at SPP=100)
  cornell-box_cfg5.npz   384 x 216 x 64 spp    (config 5's 16:9 aspect override, dx = dy*W/H, MTPC/pathTracing.cpp:280)
--glass:  glass-box_image.npz  scenes/glass-box (Ni 1.5, tests/golden/make_glass_scene.py) 160 x 160 x 128 spp, two renders
--png:    png_golden.npz       images (float32-representable doubles) + the bytes of the PNG the reference's imshow + svpng wrote
--scene-arrays: S_scene.npz    what the reference's read_scene parsed (per-face v/vn/vt in OBJ order, material names)
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


CONFIG_IMAGES = {"cornell-box_cfg2": ("cornell-box", 256, 256, 256), "veach-mis_cfg3": ("veach-mis", 300, 225, 100),
                 "cornell-box_cfg5": ("cornell-box", 384, 216, 64)}


def two_renders(scene, w, h, spp, out_name):
    imgs, mss = [], []
    for _ in range(2):
        img, ms = O.ref_image(O.REF_SCENE_DIR + scene, spp, w, h)
        assert np.array_equal(img, img.astype(np.float32).astype(np.float64))   # generateImg accumulates in float (:301,:322)
        imgs.append(img.astype(np.float32)); mss.append(ms)
        time.sleep(1.1)  # the reference seeds its RNG with time(NULL)
    np.savez_compressed(os.path.join(OUT, out_name + ".npz"), a=imgs[0], b=imgs[1], spp=spp, ms=np.array(mss))
    print(out_name, w, h, spp, "ms", mss, flush=True)


def make_png_golden():
    """Byte-parity fixtures for the output stage: the reference's own imshow + svpng on (1) a small reference render and
    (2) hand-made images with negatives, values > 1, exact k/255 boundaries, 1 x 1 and a width whose 3w+1 crosses 256."""
    rng = np.random.default_rng(7)
    img, _, png = O.ref_image(O.REF_SCENE_DIR + "cornell-box", 4, 96, 64, with_png=True)
    out = {"render_img": img.astype(np.float32), "render_png": np.frombuffer(png, np.uint8)}
    assert O.ref_png(img) == png
    cases = {"noise_7x5": rng.uniform(-0.3, 1.4, size=(5, 7, 3)), "one_1x1": np.array([[[0.5, 1.0, 0.25]]]),
             "ramp_256x3": (np.arange(256 * 3 * 3).reshape(3, 256, 3) % 256) / 255.0,
             "wide_300x2": rng.uniform(0, 1, size=(2, 300, 3)), "tall_1x70": rng.uniform(0, 2, size=(70, 1, 3))}
    for k, v in cases.items():
        v32 = v.astype(np.float32)
        out[k + "_img"] = v32
        out[k + "_png"] = np.frombuffer(O.ref_png(v32.astype(np.float64)), np.uint8)
    np.savez_compressed(os.path.join(OUT, "png_golden.npz"), **out)
    print("png_golden", {k: v.shape for k, v in out.items()})


def main():
    if not O.have_ref():
        O.build()
    if "--png" in sys.argv:
        make_png_golden()
    if "--scene-arrays" in sys.argv:
        for name in ("cornell-box", "veach-mis", "glass-box"):
            pos, nrm, uv, names = O.ref_scene(O.REF_SCENE_DIR + name)
            uniq = sorted(set(names))
            np.savez_compressed(os.path.join(OUT, name + "_scene.npz"), pos=pos, nrm=nrm, uv=uv, mat_names=np.array(uniq),
                                mat_index=np.array([uniq.index(x) for x in names], np.int32))
            print(name, "scene arrays", pos.shape)
    if "--glass" in sys.argv:
        two_renders("glass-box", 160, 160, 128, "glass-box_image")
    if "--configs" in sys.argv:
        for out_name, (scene, w, h, spp) in CONFIG_IMAGES.items():
            two_renders(scene, w, h, spp, out_name)
    if "--base" not in sys.argv and "--images" not in sys.argv:
        return
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
