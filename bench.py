#!/usr/bin/env python
"""bench.py — the hot path's headline benchmark (BASELINE.json: Msamples/s, plus Mrays/s and LBVH build ms).

A step = one pass of the hot path over one scene: GPU LBVH build (seam 1) + path-traced render (seams 2+3).
N=1 workload = BASELINE configs[1]: cornell-box at its .camera resolution (1024x1024), 256 spp.
N>1 = SPP split: every GPU integrates 256 spp (total spp = 256*N, weak scaling), partial buffers gathered to
rank 0 over NCCL and summed in the fixed pairwise tree.

  python bench.py --gpus N --steps K --warmup W             (our arm; torchrun for N>1)
  python bench.py --impl reference --gpus N --steps K --warmup W   (reference CPU code, oracle/_ref)
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (scene, width, height, spp per GPU)  — width/height 0 = the .camera's own
    "cornell-box-1024x1024-spp256": ("cornell-box", 0, 0, 256),
    "veach-mis-1200x900-spp100": ("veach-mis", 0, 0, 100),
    "cornell-box-3840x2160-spp1024": ("cornell-box", 3840, 2160, 1024),
    "synthetic-5M-1920x1080-spp64": ("synthetic-5M", 1920, 1080, 64),
    "synthetic-1M-1920x1080-spp64": ("synthetic-1M", 1920, 1080, 64),
}
# DRAM bytes of ONE launch of the dominant kernel from the committed `ncu --set full` capture (bytes; per launch like
# roofline.achieved).  ~89 MB (ray queue in, hit records out) against ~1.4 GB of algorithmic node+triangle bytes per launch: the
# scene is served from L1/L2.
NCU_TRAFFIC = {("cornell-box-1024x1024-spp256", "wavefront"): 81.129984e6 + 9.479680e6,
               ("cornell-box-1024x1024-spp256", "megakernel"): 6.414592e6 + 64.855552e6}
SYNTHETIC_TRIS = {"synthetic-5M": 5_000_000, "synthetic-1M": 1_000_000}
DEFAULT_WORKLOAD = "cornell-box-1024x1024-spp256"
# bounded CPU sample of the same workload: same scene and camera, fewer pixels/samples (no pixel jitter in the
# reference, so a lower resolution is a point-subsampling of the same image with the same cost per sample)
CPU_SAMPLE = {"cornell-box": (192, 192, 16), "veach-mis": (160, 120, 8)}


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.p, self.index = [], None, index

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.p = None

    def mark(self):
        """Rows read so far belong to the warm-up: drop them (keep the last one so a very short timed region still has a sample)."""
        self.rows = self.rows[-1:]

    def _read(self):
        for line in self.p.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.p:
            self.p.terminate()
        sm = sorted(float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit())
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            for k, nm in enumerate(names):
                if len(r) > 3 + k and r[3 + k].lower().startswith("active"):
                    reasons.add(nm)
        mx = float(self.rows[0][1]) if self.rows and len(self.rows[0]) > 1 else None
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def cpu_reference_sample(scene_name, steps=1, warmup=0):
    """Times the reference's own generateImg (unmodified sources, oracle/_ref) on a bounded sample; falls back to the
    oracle port when the reference binary did not travel.  Returns the cpu_baseline dict (value in Msamples/s)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py as O
    times = []
    if scene_name.startswith("synthetic"):
        # the reference would need the mesh as a ~500 MB .obj; the CPU restatement (same algorithm, same balanced tree,
        # Philox instead of the racy RNG) takes the in-memory scene instead: kind "port"
        from montecarlopathtracing_b200.synthetic import make_scene
        w, h, spp = 160, 90, 4
        sample = "%s %dx%d spp %d (same camera, point-subsampled pixels)" % (scene_name, w, h, spp)
        desc, keep, cam = make_scene(SYNTHETIC_TRIS[scene_name])
        osc = O.OracleScene(desc)
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            osc.render(cam, spp, w, h)
            if i >= warmup:
                times.append((time.perf_counter() - t0) * 1e3)
        ms = sum(times) / len(times)
        return {"value": w * h * spp / ms * 1e-3, "unit": "Msamples/s", "cores": os.cpu_count(), "kind": "port", "sample": sample,
                "ms_per_step": ms, "host_cpus": os.cpu_count()}
    w, h, spp = CPU_SAMPLE[scene_name]
    sample = "%s %dx%d spp %d (same camera, point-subsampled pixels)" % (scene_name, w, h, spp)
    if O.have_ref():
        for i in range(warmup + steps):
            _, ms = O.ref_image(O.REF_SCENE_DIR + scene_name, spp, w, h)
            if i >= warmup:
                times.append(ms)
        kind, cores = "reference", min(spp, 8)   # omp_set_num_threads(min(spp,8)), MTPC/pathTracing.cpp:303
    else:
        import montecarlopathtracing_b200 as M
        s = M.HostScene(os.path.join(ROOT, "scenes") + "/", scene_name)
        osc = O.OracleScene(s.desc)
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            osc.render(s.camera, spp, w, h)
            if i >= warmup:
                times.append((time.perf_counter() - t0) * 1e3)
        kind, cores = "port", os.cpu_count()
    ms = sum(times) / len(times)
    return {"value": w * h * spp / ms * 1e-3, "unit": "Msamples/s", "cores": cores, "kind": kind, "sample": sample,
            "ms_per_step": ms, "host_cpus": os.cpu_count()}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    scene_name = WORKLOADS[args.workload][0]
    cb = cpu_reference_sample(scene_name, steps=args.steps, warmup=args.warmup)
    line = {"impl": "reference", "metric": "path-tracing throughput", "value": cb["value"], "unit": "Msamples/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": cb["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "shipped scene",
            "config": {"workload": args.workload, "sample": cb["sample"]},
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def neighbour_stages(M, ctx, scene_name, w, h):
    """The steps either side of the path, outside the timed region, wall clock, median of 5:
    scene ingest (GPU: .obj file -> flattened device arrays; host: the sequential loader + H2D upload) and the output stage
    (device: quantise + PNG framing on the GPU, D2H of the finished file; host: D2H of the float image + mcpt_write_png)."""
    import tempfile
    import numpy as np
    import torch
    out = {}
    med = lambda f, n=5: float(np.median([f() for _ in range(n)]))

    def wall(fn):
        def run():
            torch.cuda.synchronize(); t0 = time.perf_counter(); fn(); torch.cuda.synchronize()
            return (time.perf_counter() - t0) * 1e3
        return run
    tmp = tempfile.mkdtemp()
    if not scene_name.startswith("synthetic"):
        scenes = os.path.join(ROOT, "scenes") + "/"
        keep = []
        def host():
            s = M.HostScene(scenes, scene_name); ctx.upload_scene(s); keep.append(s)
        out["ingest"] = {"obj_bytes": os.path.getsize(scenes + scene_name + ".obj"),
                         "gpu_ms": med(wall(lambda: ctx.load_scene_gpu(scenes, scene_name))), "host_loader_plus_upload_ms": med(wall(host), 3)}
    img = torch.rand(h * w * 3, device="cuda", dtype=torch.float32)
    host_img = torch.empty(h * w * 3, dtype=torch.float32).pin_memory()
    def dev_png():
        ctx.encode_png(img.data_ptr(), w, h).tofile(os.path.join(tmp, "dev.png"))
    def host_png():
        host_img.copy_(img); M.write_png(os.path.join(tmp, "host.png"), host_img.numpy().reshape(h, w, 3))
    out["output_stage"] = {"png_bytes": int(M.png_size(w, h)), "device_ms": med(wall(dev_png)), "host_ms": med(wall(host_png), 3)}
    same = open(os.path.join(tmp, "dev.png"), "rb").read() == open(os.path.join(tmp, "host.png"), "rb").read()
    out["output_stage"]["byte_identical"] = bool(same)
    return out


def run_ours(args):
    import ctypes
    import numpy as np
    import torch
    import torch.distributed as dist
    import montecarlopathtracing_b200 as M
    from montecarlopathtracing_b200 import _capi
    from montecarlopathtracing_b200.multigpu import gpu_backend, render_split

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d: launch with torchrun --nproc-per-node %d" % (args.gpus, world, args.gpus))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    scene_name, width, height, spp_per_gpu = WORKLOADS[args.workload]
    spp = spp_per_gpu * world
    if scene_name.startswith("synthetic"):
        from montecarlopathtracing_b200.synthetic import make_scene
        src_desc, keep, camera = make_scene(SYNTHETIC_TRIS[scene_name], width=width, height=height)
        arrays, textures, n_faces, n_lights = keep[:4], [], src_desc.n_faces, src_desc.n_lights
    else:
        s = M.HostScene(os.path.join(ROOT, "scenes") + "/", scene_name)
        src_desc, camera, arrays, textures, n_faces, n_lights = s.desc, s.camera, (s.pos, s.nrm, s.uv, s.mat_id), s.textures, s.n_faces, len(s.lights)
    w = width or camera.width
    h = height or camera.height
    # host inputs in pinned memory (the e2e leg copies them to the device every step)
    pinned = [torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in arrays]
    desc = _capi.SceneDesc()
    ctypes.memmove(ctypes.byref(desc), ctypes.byref(src_desc), ctypes.sizeof(desc))
    desc.pos = ctypes.cast(pinned[0].data_ptr(), ctypes.POINTER(ctypes.c_double))
    desc.nrm = ctypes.cast(pinned[1].data_ptr(), ctypes.POINTER(ctypes.c_double))
    desc.uv = ctypes.cast(pinned[2].data_ptr(), ctypes.POINTER(ctypes.c_double))
    desc.mat_id = ctypes.cast(pinned[3].data_ptr(), ctypes.POINTER(ctypes.c_int32))
    tex_bytes = sum(t.height * t.width * 3 for t in textures)
    h2d = sum(p.numel() * p.element_size() for p in pinned) + tex_bytes
    host_img = torch.empty(h * w * 3, dtype=torch.float32).pin_memory()

    ctx = M.Context(local)                                   # raises without a B200: no CPU fallback
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)                       # our kernels and torch's collectives share one stream
    ctx.set_seed(2026)
    ctx.upload_scene(desc)
    mode = M.MODE_WAVEFRONT if args.mode == "wavefront" else M.MODE_MEGAKERNEL
    render_partial, sum_partials, stats = gpu_backend(ctx, camera, spp, w, h, mode=mode)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        """inputs resident in HBM: LBVH build + SPP-split render (+ gather and fixed-order sum)."""
        ctx.build_bvh()
        return render_split(render_partial, sum_partials, spp, rank, world)

    def step_e2e():
        """through the C ABI with HOST buffers: upload (H2D) + build + render + image back to the host (D2H)."""
        ctx.upload_scene(desc)
        ctx.build_bvh()
        img = render_split(render_partial, sum_partials, spp, rank, world)
        if img is not None:
            host_img.copy_(img, non_blocking=False)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()                                      # nvidia-smi takes ~0.3 s to print its first row: start it early,
    for _ in range(args.warmup):                             # keep only the rows of the timed region (mark() below)
        step()
    barrier()
    sampler.mark()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    integ_ms, render_ms, build_ms, rays, trace_ms, trace_n, launches = [], [], [], [], [], [], 0
    ev0.record(stream)
    for _ in range(args.steps):
        del stats[:]
        step()
        st = stats[-1]
        integ_ms.append(st.integrate_ms); render_ms.append(st.render_ms); build_ms.append(st.build_ms); rays.append(st.rays)
        trace_ms.append(st.trace_ms); trace_n.append(st.trace_launches)
        launches += st.build_launches + st.kernel_launches + 1     # build (ours + CUB's) + render_chunks' kernels + the final fixed-tree sum
    ev1.record(stream)
    barrier()
    ms_total = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    # end-to-end leg
    for _ in range(min(args.warmup, 2)):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    t = torch.tensor([ms_total, e2e_ms, float(sum(rays))], dtype=torch.float64, device="cuda")
    if world > 1:
        mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = t.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        ms_total, e2e_ms, rays_total = float(mx[0]), float(mx[1]), float(sm[2])
    else:
        rays_total = float(t[2])
    if rank == 0:
        pk, pk_src = peaks()
        samples = float(w) * h * spp * args.steps
        ms_step = ms_total / args.steps
        # algorithmic bytes of the dominant kernel: 64 B per 4-wide node fetch + 48 B per triangle test, counted
        # by the COUNT variant on the same workload at reduced spp (per-ray averages are spp-independent)
        cspp = max(8, min(spp_per_gpu, 16))
        ctx.render(camera, cspp, w, h, mode=mode, flags=M.FLAG_COUNT)
        cs = ctx.last_stats
        bytes_per_ray = (64.0 * cs.node_visits + 48.0 * cs.tri_tests) / cs.rays
        # dominant kernel: the megakernel (one launch per step), or wf_trace (one launch per wavefront iteration; its
        # launches of a step are timed by one CUDA-event pair each and summed)
        if args.mode == "wavefront":
            kernel, n_k = "wf_trace", sum(trace_n) / len(trace_n)
            k_ms_step = sum(trace_ms) / len(trace_ms)
        else:
            kernel, n_k = "pt_megakernel", 1
            k_ms_step = sum(integ_ms) / len(integ_ms)
        k_ms = k_ms_step / n_k                       # average launch duration
        k_rays = rays[-1] / n_k                      # rays per launch
        achieved = bytes_per_ray * k_rays / (k_ms * 1e-3) / 1e9
        line = {
            "metric": "path-tracing throughput", "value": samples / ms_total * 1e-3, "unit": "Msamples/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": ("synthetic mesh (montecarlopathtracing_b200/synthetic.py, seed 1234)" if scene_name.startswith("synthetic")
                                                               else "shipped scene (scenes/%s.obj)" % scene_name),
            "config": {"workload": args.workload, "scene": scene_name, "width": w, "height": h, "spp_total": spp,
                       "spp_per_gpu": spp_per_gpu, "mode": args.mode, "partition": "spp-split, 64 chunks, fixed-tree sum",
                       "l2_policy": "no flush: the scene (%.1f MB of nodes+triangles) is L1/L2-resident by design and every "
                                    "launch re-reads it millions of times; the wavefront state and ray queues (~%d MB) "
                                    "exceed L2 at this size" % ((0.45 * 64 + 48 + 64) * n_faces / 1e6, int(w * h * (100 + 16 * n_lights + 32 * (n_lights + 1)) / 1e6))},
            "mrays_per_s": rays_total / ms_total * 1e-3, "rays_per_sample": rays_total / samples,
            "lbvh_build_ms": sum(build_ms) / len(build_ms), "render_ms": sum(render_ms) / len(render_ms),
            # seam 1: 640 algorithmic bytes per triangle over the radix-tree pipeline + the 4-wide collapse against the HBM peak
            "lbvh_build_roofline": {"bound": "hbm", "achieved": 640.0 * n_faces / (sum(build_ms) / len(build_ms) * 1e-3) / 1e9,
                                    "peak": pk["hbm_gbs"], "unit": "GB/s",
                                    "frac": 640.0 * n_faces / (sum(build_ms) / len(build_ms) * 1e-3) / 1e9 / pk["hbm_gbs"]},
            "e2e": {"value": samples / e2e_ms * 1e-3, "unit": "Msamples/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": h * w * 3 * 4, "ms_per_step": e2e_ms / args.steps},
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": achieved / pk["hbm_gbs"],
                         "traffic": NCU_TRAFFIC.get((args.workload, args.mode)), "traffic_source": "profiles/r1_wide_wf_trace_ncu.md "
                         "(dram__bytes_read.sum + dram__bytes_write.sum of one mid-render wf_trace launch; null for workloads not captured)",
                         "peak_source": pk_src, "kernel": kernel, "kernel_ms": k_ms, "launches_per_step": n_k,
                         "kernel_ms_per_step": k_ms_step, "integrator_ms_per_step": sum(integ_ms) / len(integ_ms),
                         "bytes_per_ray": bytes_per_ray, "nodes_per_ray": cs.node_visits / cs.rays,
                         "tris_per_ray": cs.tri_tests / cs.rays,
                         "achieved_over_integrator": bytes_per_ray * rays[-1] / (sum(integ_ms) / len(integ_ms) * 1e-3) / 1e9,
                         "issue_bound": {"source": "profiles/r1_wide_wf_trace_1pipe_ncu.md (one mid-render launch owning the GPU)",
                                         "issue_slots_busy_pct": 69.5, "alu_pipe_pct": 60.3, "fma_pipe_pct": 23.3, "lsu_wavefront_pct": 67.4,
                                         "active_lanes_per_instruction": 17.3, "warp_instructions_per_ray": 100},
                         "note": "algorithmic node+triangle bytes (64 B per 4-wide node visit, 48 B per triangle test). The scene is "
                                 "L1/L2-resident, so this is on-chip traffic expressed against the HBM copy peak; the binding limit "
                                 "is issue slots under divergence (DESIGN.md, profiles/). In wavefront mode 3 pipes overlap their "
                                 "wf_trace launches, so the per-launch average shares the GPU; achieved_over_integrator divides by "
                                 "the wall time of the whole shade+trace loop instead."},
        }
        if world == 1:
            line["stages"] = neighbour_stages(M, ctx, scene_name, w, h)
        if world == 1 and not args.no_cpu_baseline:
            cb = cpu_reference_sample(scene_name)
            line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--mode", default="wavefront", choices=["wavefront", "megakernel"],
                    help="integrator: wavefront (default, the faster one) or megakernel")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
