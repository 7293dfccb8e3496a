#!/usr/bin/env python
"""bench.py — the hot path's headline benchmark (BASELINE.json: Msamples/s, plus Mrays/s and LBVH build ms, at 1/2/4/8 B200).

A step = one pass of the hot path over one scene: GPU LBVH build (seam 1) + path-traced render (seams 2+3) + — on more than
one GPU — the one exchange of the partial sums and their fixed-order sum on rank 0.

Default workload at EVERY N = BASELINE.json configs[4]: cornell-box at 3840 x 2160, 1024 spp IN TOTAL, the samples split over
the N GPUs (strong scaling: the work is fixed, `value` = W*H*1024 / time).  The image is the same at every N, bit for bit; the
line carries its sha256 so the driver can see that.  At N = 1 the same run also measures, outside the headline timed region,
sub-records for BASELINE configs[1..3] (cornell-box 1024^2 x 256 spp, veach-mis 1200x900 x 100 spp, the synthetic 5 M-triangle
scene 1920x1080 x 64 spp), the steps either side of the path, and the reference's CPU path on the box's host cores.

  python bench.py --gpus N --steps K --warmup W                    (our arm; torchrun for N > 1)
  python bench.py --impl reference --gpus N --steps K --warmup W   (the UNMODIFIED reference's CPU code, oracle/_ref)
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (scene, width, height, spp in total)  — width/height 0 = the .camera's own
    "cornell-box-3840x2160-spp1024": ("cornell-box", 3840, 2160, 1024),
    "cornell-box-1024x1024-spp256": ("cornell-box", 0, 0, 256),
    "veach-mis-1200x900-spp100": ("veach-mis", 0, 0, 100),
    "synthetic-5M-1920x1080-spp64": ("synthetic-5M", 1920, 1080, 64),
    "synthetic-1M-1920x1080-spp64": ("synthetic-1M", 1920, 1080, 64),
}
SYNTHETIC_TRIS = {"synthetic-5M": 5_000_000, "synthetic-1M": 1_000_000}
DEFAULT_WORKLOAD = "cornell-box-3840x2160-spp1024"
SUB_RECORDS = ["cornell-box-1024x1024-spp256", "veach-mis-1200x900-spp100", "synthetic-5M-1920x1080-spp64"]
# Bounded CPU samples of a workload: same scene and camera, fewer pixels / samples.  The reference shoots ONE un-jittered ray
# through a pixel's corner (quirk Q4), so fewer pixels at the same camera and aspect are a point-subsample of the same image
# with the same cost per sample.  (width, height, spp): ~10-20 s on 8 threads.
CPU_SAMPLE = {"cornell-box-3840x2160-spp1024": (384, 216, 16), "cornell-box-1024x1024-spp256": (192, 192, 32),
              "veach-mis-1200x900-spp100": (160, 120, 12)}
# sources that define the profiled kernels: profiles/kernel_metrics.json is only believed while their hash matches
KERNEL_SOURCES = ["wavefront.cu", "trace.cuh", "shade.cuh", "exact.cuh", "render.cu", "device_types.cuh", "vecmath.cuh", "philox.cuh", "quant.cuh"]


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback (B200_PROFILING.md)"


def kernel_source_hash():
    h = hashlib.sha256()
    for f in KERNEL_SOURCES:
        h.update(open(os.path.join(ROOT, "montecarlopathtracing_b200", "csrc", f), "rb").read())
    return h.hexdigest()[:16]


def profile_metrics():
    """Per-kernel counters from the committed ncu captures (tools/ncu_to_metrics.py -> profiles/kernel_metrics.json), or None
    when the file is missing or was captured from other kernel sources than the ones built now."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "kernel_metrics.json")))
    except Exception:
        return None, "profiles/kernel_metrics.json missing"
    if d.get("source_hash") != kernel_source_hash():
        return None, "profiles/kernel_metrics.json is stale (kernel sources changed since the capture)"
    return d, d.get("captured", "")


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


# ---- the reference's CPU path ---------------------------------------------------------------------------------------
def cpu_reference_sample(workload, steps=1, warmup=0, full=False):
    """Times the reference's own generateImg (unmodified sources, oracle/_ref) on a bounded sample of `workload`; falls back
    to the oracle port when the reference binary did not travel.  Returns the cpu_baseline dict (value in Msamples/s)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py as O
    scene_name = WORKLOADS[workload][0]
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
    if full:   # BASELINE config 1 as the reference runs it: the .camera's own resolution, 16 spp
        w, h, spp = 0, 0, 16
        sample = "full: %s at its .camera resolution, 16 spp (BASELINE.json configs[0])" % scene_name
    else:
        w, h, spp = CPU_SAMPLE[workload]
        sample = "%s %dx%d spp %d (same camera and aspect, point-subsampled pixels: the reference has no pixel jitter)" % (scene_name, w, h, spp)
    if O.have_ref():
        for i in range(warmup + steps):
            img, ms = O.ref_image(O.REF_SCENE_DIR + scene_name, spp, w, h)
            if i >= warmup:
                times.append(ms)
        w, h = img.shape[1], img.shape[0]
        kind, cores = "reference", min(spp, 8)   # omp_set_num_threads(min(spp,8)), MTPC/pathTracing.cpp:303
    else:
        import montecarlopathtracing_b200 as M
        s = M.HostScene(os.path.join(ROOT, "scenes") + "/", scene_name)
        osc = O.OracleScene(s.desc)
        w, h = w or s.camera.width, h or s.camera.height
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
    scene_name, width, height, spp = WORKLOADS[args.workload]
    cb = cpu_reference_sample(args.workload, steps=args.steps, warmup=args.warmup, full=args.cpu_full)
    line = {"impl": "reference", "metric": "path-tracing throughput", "value": cb["value"], "unit": "Msamples/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": cb["ms_per_step"],
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "shipped scene (scenes/%s.obj)" % scene_name,
            "config": {"workload": args.workload, "scene": scene_name, "width": width, "height": height, "spp_total": spp,
                       "sample": cb["sample"]},
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ---- our arm ---------------------------------------------------------------------------------------------------------
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
    torch.cuda.synchronize()
    def dev_png():
        ctx.encode_png(img.data_ptr(), w, h).tofile(os.path.join(tmp, "dev.png"))
    def host_png():
        host_img.copy_(img); M.write_png(os.path.join(tmp, "host.png"), host_img.numpy().reshape(h, w, 3))
    out["output_stage"] = {"png_bytes": int(M.png_size(w, h)), "device_ms": med(wall(dev_png)), "host_ms": med(wall(host_png), 3)}
    same = open(os.path.join(tmp, "dev.png"), "rb").read() == open(os.path.join(tmp, "host.png"), "rb").read()
    out["output_stage"]["byte_identical"] = bool(same)
    return out


def load_workload(M, workload):
    """-> (scene_name, src_desc, camera, (pos, nrm, uv, mat_id), textures, n_faces, n_lights, w, h, spp, keepalive)."""
    scene_name, width, height, spp = WORKLOADS[workload]
    if scene_name.startswith("synthetic"):
        from montecarlopathtracing_b200.synthetic import make_scene
        src_desc, keep, camera = make_scene(SYNTHETIC_TRIS[scene_name], width=width, height=height)
        arrays, textures, n_faces, n_lights = keep[:4], [], src_desc.n_faces, src_desc.n_lights
    else:
        keep = s = M.HostScene(os.path.join(ROOT, "scenes") + "/", scene_name)
        src_desc, camera, arrays, textures, n_faces, n_lights = s.desc, s.camera, (s.pos, s.nrm, s.uv, s.mat_id), s.textures, s.n_faces, len(s.lights)
    return scene_name, src_desc, camera, arrays, textures, n_faces, n_lights, width or camera.width, height or camera.height, spp, keep


def roofline_record(M, ctx, camera, workload, mode, w, h, spp_local, n_faces, rays_step, samples_step, integ_ms, trace_ms, trace_n,
                    clocks_mhz, bw):
    """The roofline object of one workload.  Measured in this run: rays and samples (device counters), node visits and triangle
    tests per ray (the COUNT variant on the same workload at reduced spp: per-ray averages do not depend on spp), the
    integrator's CUDA-event time and the summed per-launch event times of wf_trace, the SM clock under load, the L2 / HBM read
    bandwidth of mcpt_measure_read_bandwidth.  From the committed ncu captures (profiles/kernel_metrics.json, believed only
    while the kernel sources hash the same): warp and thread instructions per ray / per sample, DRAM bytes per launch."""
    pk, pk_src = peaks()
    cspp = max(8, min(spp_local, 16))
    ctx.render(camera, cspp, w, h, mode=mode, flags=M.FLAG_COUNT)
    cs = ctx.last_stats
    nodes_per_ray, tris_per_ray = cs.node_visits / cs.rays, cs.tri_tests / cs.rays
    bytes_per_ray = 64.0 * nodes_per_ray + 48.0 * tris_per_ray          # 64-B wide node, 48-B triangle
    scene_mb = (0.45 * 64 + 48) * n_faces / 1e6
    fits_l2 = scene_mb < 100.0
    mem_peak, mem_kind = (bw["l2_gbs"], "measured L2 read bandwidth (32 MB working set)") if fits_l2 else \
                         (bw["hbm_read_gbs"], "measured HBM read bandwidth (4 GB working set)")
    kernel = "wf_trace" if mode == M.MODE_WAVEFRONT else "pt_megakernel"
    n_k = trace_n if mode == M.MODE_WAVEFRONT else 1
    k_ms_step = trace_ms if mode == M.MODE_WAVEFRONT else integ_ms
    k_ms = k_ms_step / max(n_k, 1)
    mem_achieved_launch = bytes_per_ray * (rays_step / max(n_k, 1)) / (k_ms * 1e-3) / 1e9
    mem_achieved_integ = bytes_per_ray * rays_step / (integ_ms * 1e-3) / 1e9
    f_clk = (clocks_mhz or pk.get("sm_max_mhz", 1965.0)) * 1e6
    issue_peak = 148 * 4 * f_clk                                          # warp instructions / s: 148 SMs x 4 schedulers
    rec = {"kernel": kernel, "launches_per_step": n_k, "kernel_ms": k_ms, "kernel_ms_per_step": k_ms_step,
           "integrator_ms_per_step": integ_ms, "nodes_per_ray": nodes_per_ray, "tris_per_ray": tris_per_ray, "bytes_per_ray": bytes_per_ray,
           "memory": {"achieved": mem_achieved_integ, "achieved_per_launch": mem_achieved_launch, "peak": mem_peak, "unit": "GB/s",
                      "frac": mem_achieved_integ / mem_peak, "peak_kind": mem_kind, "hbm_copy_peak": pk["hbm_gbs"], "hbm_peak_source": pk_src,
                      "frac_of_hbm_copy_peak": mem_achieved_integ / pk["hbm_gbs"]},
           "traffic": None, "issue": None}
    prof, prof_note = profile_metrics()
    key = workload.rsplit("-spp", 1)[0]
    if prof and key in prof.get("workloads", {}):
        p = prof["workloads"][key]
        tr, sh = p.get(kernel), p.get("wf_shade") if mode == M.MODE_WAVEFRONT else None
        if tr:
            inst_trace = tr["warp_inst_per_ray"] * rays_step
            inst_shade = sh["warp_inst_per_sample"] * samples_step if sh else 0.0
            rec["issue"] = {"achieved": (inst_trace + inst_shade) / (integ_ms * 1e-3) / 1e9, "peak": issue_peak / 1e9, "unit": "Gwarp-inst/s",
                            "frac": (inst_trace + inst_shade) / (integ_ms * 1e-3) / issue_peak,
                            "frac_dominant_kernel": inst_trace / (integ_ms * 1e-3) / issue_peak,
                            "warp_inst_per_ray": tr["warp_inst_per_ray"], "lanes_per_inst": tr["lanes_per_inst"],
                            "lane_frac": tr["lanes_per_inst"] / 32.0,
                            "thread_inst_frac": (inst_trace * tr["lanes_per_inst"] + inst_shade * (sh["lanes_per_inst"] if sh else 0)) /
                                                (integ_ms * 1e-3) / (issue_peak * 32.0),
                            "shade_warp_inst_per_sample": sh["warp_inst_per_sample"] if sh else None, "sm_mhz": f_clk / 1e6,
                            "source": "instruction counts: %s (%s); rays, samples, time and clock: this run" % ("profiles/kernel_metrics.json", prof_note)}
            rec["traffic"] = tr.get("dram_bytes_per_launch")
    else:
        rec["issue_note"] = prof_note if not prof else "no ncu capture for " + key
    mem_f = rec["memory"]["frac"]
    iss_f = rec["issue"]["frac"] if rec["issue"] else None
    if iss_f is not None and iss_f >= mem_f:
        rec.update({"bound": "issue", "achieved": rec["issue"]["achieved"], "peak": rec["issue"]["peak"], "unit": "Gwarp-inst/s", "frac": iss_f})
    else:
        rec.update({"bound": "l2" if fits_l2 else "hbm", "achieved": mem_achieved_integ, "peak": mem_peak, "unit": "GB/s", "frac": mem_f})
    rec["note"] = ("achieved = this kernel's algorithmic work per step over the integrator's CUDA-event time: the three wavefront pipes overlap "
                   "their wf_trace launches, so a launch's own duration is not exclusive GPU time (per-launch figures are given beside it). "
                   "bound = the larger of the issue fraction (warp instructions / s against 148 SMs x 4 schedulers x the SM clock under load) "
                   "and the memory fraction (64 B per wide-node visit + 48 B per triangle test against the measured L2 bandwidth when the "
                   "tree fits L2, the measured HBM read bandwidth otherwise), SURVEY.md 8(d).")
    return rec


def measure_workload(M, ctx, workload, mode, steps, warmup, torch, stream, bw):
    """One GPU, scene resident: a sub-record for BASELINE configs 1-3 (same step definition as the headline)."""
    scene_name, src_desc, camera, arrays, textures, n_faces, n_lights, w, h, spp, keep = load_workload(M, workload)
    ctx.upload_scene(src_desc)
    def step():
        ctx.build_bvh()
        ctx.render_dist(camera, spp, w, h, mode=mode, to_host=False)
        return ctx.last_stats
    for _ in range(warmup):
        step()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler = ClockSampler(torch.cuda.current_device()); sampler.start(); time.sleep(0.35); sampler.mark()
    ev0.record(stream)
    sts = [step().as_dict() for _ in range(steps)]
    ev1.record(stream); ev1.synchronize()
    clocks = sampler.stop()
    ms = ev0.elapsed_time(ev1) / steps
    samples = float(w) * h * spp
    avg = lambda k: sum(s[k] for s in sts) / len(sts)
    rec = {"workload": workload, "value": samples / ms * 1e-3, "unit": "Msamples/s", "ms_per_step": ms, "steps": steps, "warmup": warmup,
           "mrays_per_s": avg("rays") / ms * 1e-3, "rays_per_sample": avg("rays") / samples, "lbvh_build_ms": avg("build_ms"),
           "build_launches": sts[-1]["build_launches"], "triangles": n_faces, "clocks": clocks,
           "lbvh_build_roofline": build_roofline(n_faces, avg("build_ms")),
           "roofline": roofline_record(M, ctx, camera, workload, mode, w, h, spp, n_faces, avg("rays"), samples, avg("integrate_ms"),
                                       avg("trace_ms"), avg("trace_launches"), clocks["sm_mhz"], bw)}
    del keep
    return rec


def build_roofline(n_faces, build_ms):
    """Seam 1: 640 algorithmic bytes per triangle over the radix-tree pipeline + the 4-wide collapse against the HBM peak."""
    pk, _ = peaks()
    a = 640.0 * n_faces / (build_ms * 1e-3) / 1e9
    return {"bound": "hbm", "achieved": a, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": a / pk["hbm_gbs"], "bytes_per_triangle": 640}


def run_ours(args):
    import ctypes
    import numpy as np
    import torch
    import torch.distributed as dist
    import montecarlopathtracing_b200 as M
    from montecarlopathtracing_b200 import _capi
    from montecarlopathtracing_b200.multigpu import init_comm

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d: launch with torchrun --nproc-per-node %d" % (args.gpus, world, args.gpus))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    scene_name, src_desc, camera, arrays, textures, n_faces, n_lights, w, h, spp, keep = load_workload(M, args.workload)
    spp_local = spp // world
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
    host_np = host_img.numpy().reshape(h, w, 3)

    ctx = M.Context(local)                                   # raises without a B200: no CPU fallback
    stream = torch.cuda.Stream()                             # ONE stream for the library's kernels, its NCCL exchange and our timing events
    torch.cuda.set_stream(stream)
    ctx.set_stream(stream.cuda_stream)
    ctx.set_seed(2026)
    ctx.upload_scene(desc)
    init_comm(ctx, rank, world)                              # the render's own communicator inside libmcpt.so
    mode = M.MODE_WAVEFRONT if args.mode == "wavefront" else M.MODE_MEGAKERNEL

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        """inputs resident in HBM: LBVH build + SPP-split render + the exchange and the fixed-order sum (image stays on rank 0's GPU)."""
        ctx.build_bvh()
        ctx.render_dist(camera, spp, w, h, mode=mode, to_host=False)
        return ctx.last_stats

    def step_e2e():
        """through the C ABI with HOST buffers: scene upload (H2D) + build + render + the image back in host memory (D2H, rank 0)."""
        ctx.upload_scene(desc)
        ctx.build_bvh()
        ctx.render_dist(camera, spp, w, h, mode=mode, out=host_np, to_host=True)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()                                      # nvidia-smi takes ~0.3 s to print its first row: start it early,
    for _ in range(args.warmup):                             # keep only the rows of the timed region (mark() below)
        step()
    barrier()
    sampler.mark()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sts, launches = [], 0
    ev0.record(stream)
    for _ in range(args.steps):
        st = step().as_dict()
        sts.append(st)
        launches += st["build_launches"] + st["kernel_launches"]      # build (ours + CUB's) + integrator + sums
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
    rays_local = float(sum(s["rays"] for s in sts))
    t = torch.tensor([ms_total, e2e_ms, rays_local], dtype=torch.float64, device="cuda")
    if world > 1:
        mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = t.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        ms_total, e2e_ms, rays_total = float(mx[0]), float(mx[1]), float(sm[2])
    else:
        rays_total = rays_local
    if rank == 0:
        samples = float(w) * h * spp * args.steps
        ms_step = ms_total / args.steps
        avg = lambda k: sum(s[k] for s in sts) / len(sts)
        bw = {"l2_gbs": ctx.measure_read_bandwidth(32 << 20, 64), "hbm_read_gbs": ctx.measure_read_bandwidth(4 << 30, 2)}
        line = {
            "metric": "path-tracing throughput", "value": samples / ms_total * 1e-3, "unit": "Msamples/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": ("synthetic mesh (montecarlopathtracing_b200/synthetic.py, seed 1234)" if scene_name.startswith("synthetic")
                     else "shipped scene (scenes/%s.obj)" % scene_name),
            "config": {"workload": args.workload, "scene": scene_name, "width": w, "height": h, "spp_total": spp,
                       "spp_per_gpu": spp_local, "mode": args.mode,
                       "partition": "spp-split: 64 sample chunks, rank r of R integrates chunks [64r/R, 64(r+1)/R); one grouped "
                                    "ncclSend/ncclRecv of the R partial sums to rank 0; fixed pairwise-tree sum (image independent of R)",
                       "l2_policy": "no flush: the scene (%.1f MB of wide nodes + triangles) is L1/L2-resident by design and every launch "
                                    "re-reads it millions of times; the wavefront state and ray queues (%d MB) exceed L2" %
                                    ((0.45 * 64 + 48) * n_faces / 1e6, int(min(w * h * 64 // world, 6e6 / (n_lights + 1) * 1.4) * (100 + 48 * n_lights) / 1e6))},
            "mrays_per_s": rays_total / ms_total * 1e-3, "rays_per_sample": rays_total / samples,
            "lbvh_build_ms": avg("build_ms"), "build_launches": sts[-1]["build_launches"], "render_ms": avg("render_ms"),
            "exchange_ms": avg("exchange_ms"), "exchange_bytes": (world - 1) * h * w * 3 * 4,
            "image_sha256": hashlib.sha256(host_np.tobytes()).hexdigest(),
            "lbvh_build_roofline": build_roofline(n_faces, avg("build_ms")),
            "e2e": {"value": samples / e2e_ms * 1e-3, "unit": "Msamples/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": h * w * 3 * 4, "ms_per_step": e2e_ms / args.steps},
            "gpu_launches": launches,
            "clocks": clocks,
            "measured_bandwidth": bw,
            "roofline": roofline_record(M, ctx, camera, args.workload, mode, w, h, spp_local, n_faces, avg("rays"), float(w) * h * spp_local,
                                        avg("integrate_ms"), avg("trace_ms"), avg("trace_launches"), clocks["sm_mhz"] if clocks else None, bw),
        }
        if world == 1 and not args.no_sub_records:
            line["sub_records"] = [measure_workload(M, ctx, wl, mode, 3, 3, torch, stream, bw) for wl in SUB_RECORDS if wl != args.workload]
            ctx.upload_scene(desc); ctx.build_bvh()
        if world == 1:
            line["stages"] = neighbour_stages(M, ctx, scene_name, w, h)
        if world == 1 and not args.no_cpu_baseline:
            cb = cpu_reference_sample(args.workload, full=args.cpu_full)
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
    ap.add_argument("--no-sub-records", action="store_true")
    ap.add_argument("--cpu-full", action="store_true",
                    help="cpu_baseline / --impl reference: BASELINE config 1 in full (the scene's .camera resolution, 16 spp; ~2 min) "
                         "instead of the bounded sample")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
