#!/usr/bin/env python
"""bench.py — BASELINE.json's metric ("Sponza 4K Mrays/s, primary + shadow + reflection") on 1..N B200s.

Step   = one frame of the hot path (vkrt_trace [+ NCCL tile gather]) over the resident scene.
value  = rays traced by all ranks / device time (CUDA events on the library's stream), inputs resident in HBM.
e2e    = same metric through the reference-facing calls (Raytracer.render + present): the 256-byte uniform block goes
         host->device and the B8G8R8A8 frame comes back to page-locked host memory every step, through a ring of 3
         images like the reference's swapchain (the copy of frame k overlaps the traversal of frame k+1; the timed
         region ends when the last copy has landed).
--impl reference times the CPU oracle (oracle/, the restatement of the reference's shaders; the reference's own
Vulkan path and Mesa lavapipe are unavailable on this box) on the host cores, on a bounded pixel subset per step.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (width, height, spp, description)
    "config1": (1280, 720, 1, "Sponza 1280x720 1 spp, default camera"),
    "config2": (1920, 1080, 1, "Sponza 1920x1080 1 spp, camera path (W 150 frames, Z 150 frames), BVH built once"),
    "config3": (3840, 2160, 8, "Sponza 3840x2160 8 spp jittered (Halton 2,3), 64x64 interleaved tiles over ranks"),
    "config4": (3840, 2160, 1, "4x4 instanced Sponza grid (16 instances) 3840x2160 1 spp"),
    "config5": (3840, 2160, 1, "procedural triangle soup, full BVH rebuild every frame, 3840x2160 1 spp"),
    # SURVEY 8 f4 (not one of BASELINE.json's configs): the reference's Rasterizer path, single GPU
    "raster": (3840, 2160, 1, "Sponza 3840x2160 through the rasteriser path (shader.vert/.frag, 8x MSAA, depth test, alpha test)"),
}
RASTER_METRIC = "Sponza 4K rasteriser Mpixel/s (8x MSAA)"


def csrc_sha():
    """Hash of the library sources this run was built from (the same function as tools/ncu_traffic.py)."""
    import hashlib

    d = os.path.join(ROOT, "vkrt_b200", "csrc")
    h = hashlib.sha256()
    for name in sorted(os.listdir(d)):
        if name.endswith((".cu", ".cuh", ".hpp")) or name == "Makefile":
            h.update(name.encode())
            h.update(open(os.path.join(d, name), "rb").read())
    return h.hexdigest()[:16]


def read_traffic(workload):
    """ncu counters per launch of the trace kernels (DRAM bytes, warp and thread instructions) from the capture summarised in
    profiles/r02_traffic.json (tools/ncu_traffic.py) - only if it was taken on exactly the library sources this run uses."""
    p = os.path.join(ROOT, "profiles", "r02_traffic.json")
    try:
        e = json.load(open(p)).get(workload)
    except Exception:
        return None, "no capture (profiles/r02_traffic.json missing)"
    if not e:
        return None, f"no capture of {workload} in profiles/r02_traffic.json"
    if e.get("csrc_sha") != csrc_sha():
        return None, f"capture at commit {e.get('commit')} is of other library sources (csrc hash {e.get('csrc_sha')} != {csrc_sha()}): not used"
    return e, f"profiles/r02_traffic.json, ncu capture at commit {e.get('commit')} of these sources"


def read_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        if os.environ.get("VKRT_BENCH_NO_SMI"):  # diagnostic: does the nvidia-smi poller perturb the timed region?
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200", "-i",
                                          str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(len(r) >= 7 and r[3 + k].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def build_scene(workload, with_images=True):
    """with_images=False: a rank that receives the texels from rank 0 (vkrt_textures_broadcast) skips the image decode."""
    from vkrt_b200 import scenes

    if workload == "config5":  # CPU arm only (the GPU arm generates the soup on the device): bounded triangle count
        model = scenes.procedural_soup(2_000_000)
        model["provenance"] = "procedural soup, 2,000,000 triangles (numpy generator; CPU arm is bounded to this size)"
        return model, scenes.identity_instance(1.0)
    model = scenes.sponza_standin(with_images=with_images)
    instances = scenes.instanced_grid_transforms() if workload == "config4" else None
    return model, instances


def set_pose(rt, workload, step):
    """config2 walks the reference's own input path (src/Raytracer.cpp:288-323, dt = 1/60): W for 150 frames then Z."""
    if workload == "config2":
        dt = np.float32(1.0 / 60.0)
        f = step % 300
        if f == 0:
            rt.camera.setPosition((6.3, 4.5, -0.7))
            rt.camera.setRotation((0.0, 1.57, 0.0))
        elif f < 150:
            rt.camera.translate(rt.camera.getForward() * (np.float32(5.0) * dt))
        else:
            rt.camera.rotate((0, 1, 0), np.float32(1.5) * dt)
    elif workload == "config4" and step == 0:
        rt.camera.setPosition((-25.0, 45.0, -20.0))
        rt.camera.setRotation((-0.55, -2.16, 0.0))


def run_reference(args):
    """CPU arm: the oracle (kind "port": the reference has no CPU implementation of this path) on all host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import helpers as H
    from vkrt_b200 import camera, scenes
    from vkrt_b200.raytracer import flatten_model

    W, Hh, spp, desc = WORKLOADS[args.workload]
    model, instances = build_scene(args.workload)
    flat = flatten_model(model)
    inst = instances if instances is not None else scenes.REFERENCE_INSTANCE
    t0 = time.perf_counter()
    o = H.OracleScene(flat, inst)
    build_s = time.perf_counter() - t0
    cores = os.cpu_count() or 1
    # bounded sample: a regular pixel subset of the same frame sized for ~5-10 s of CPU work per step
    target_rays = 6.0e6 * max(1.0, cores / 8.0)
    full_rays = W * Hh * spp * 4.6
    sub = max(1, int(np.ceil(np.sqrt(full_rays / target_rays))))
    u = camera.make_uniforms(W, Hh) if args.workload != "config4" else camera.make_uniforms(W, Hh, (-25.0, 45.0, -20.0), (-0.55, -2.16, 0.0))
    rays = 0
    for _ in range(args.warmup):
        o.render(u, W, Hh, spp=spp, stride=(sub, sub), threads=cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        r = o.render(u, W, Hh, spp=spp, stride=(sub, sub), threads=cores)
        st = r["stats"]
        rays += st.raysPrimary + st.raysShadow + st.raysReflection
    dt = time.perf_counter() - t0
    value = rays / dt / 1e6
    sample = f"every {sub}th pixel in x and y of the {W}x{Hh} {spp} spp frame per step ({rays // max(args.steps,1)} rays/step)"
    line = {
        "impl": "reference", "metric": "Sponza 4K Mrays/s (prim+shadow+refl)", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / max(args.steps, 1) * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {desc}", "scene": model["provenance"],
                   "note": "reference Vulkan path and Mesa lavapipe unavailable on this box; CPU oracle (restatement of the shaders) timed"},
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": cores, "kind": "port", "sample": sample,
                         "bvh_build_mtri_s": len(flat["indices"]) / 3 / max(o.build_seconds(), 1e-9) / 1e6, "setup_s": build_s},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def raster_cpu_baseline(model, W, Hh, repeats=1):
    """The oracle's z-buffer restatement of the reference's Rasterizer on all host cores; a bounded frame size."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import helpers as H
    from vkrt_b200 import camera, scenes
    from vkrt_b200.raytracer import flatten_model

    o = H.OracleScene(flatten_model(model), scenes.REFERENCE_INSTANCE)
    o.generate_mips()
    cam = camera.Camera(np.float32(W) / np.float32(Hh))
    cam.setPosition((-4.0, 2.0, -0.2))  # Rasterizer::setupCamera, src/Rasterizer.cpp:207-211
    cam.setRotation((0.0, 1.51, 0.0))
    wvp = cam.wvp(0.01)
    cores = os.cpu_count() or 1
    t0 = time.perf_counter()
    for _ in range(repeats):
        o.rasterize(wvp, W, Hh, threads=cores)
    dt = (time.perf_counter() - t0) / repeats
    return {"value": W * Hh / dt / 1e6, "unit": "Mpixel/s", "cores": cores, "kind": "port",
            "sample": f"one {W}x{Hh} frame of the same scene and pose ({dt:.2f} s per frame; every band of rows walks all 262,267 triangles)"}


def run_raster(args):
    """--workload raster: one step = one vkrt_rasterize frame.  Single GPU (the rasteriser does not split frames over ranks)."""
    from vkrt_b200 import scenes

    W, Hh, _, desc = WORKLOADS["raster"]
    model = scenes.sponza_standin()
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:
            return
        cw, ch = 1280, 720
        steps = max(args.steps, 1)
        c = raster_cpu_baseline(model, cw, ch, repeats=steps)
        print(json.dumps({"impl": "reference", "metric": RASTER_METRIC, "value": c["value"], "unit": "Mpixel/s", "n_gpus": args.gpus,
                          "steps": steps, "warmup": 0, "ms_per_step": cw * ch / c["value"] / 1e3, "higher_is_better": True, "scaling": "weak",
                          "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                          "config": {"workload": f"raster: {desc}", "scene": model["provenance"],
                                     "note": "the reference's Vulkan rasteriser is unavailable on this box; CPU oracle (z-buffer restatement) timed"},
                          "cpu_baseline": c, "e2e": {"value": c["value"], "unit": "Mpixel/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}),
              flush=True)
        return
    import torch

    from vkrt_b200 import binding as B
    from vkrt_b200.raytracer import Rasterizer

    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist, unique_id = None, None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group(backend="nccl", device_id=torch.device("cuda", local))
        ids = [B.Context.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        unique_id = ids[0]
        if rank != 0:
            model = scenes.sponza_standin(with_images=False)
    rz = Rasterizer(model, W, Hh, device=local, rank=rank, world_size=world, unique_id=unique_id, textures_from_root=0 if world > 1 else None)
    stream = torch.cuda.ExternalStream(rz.ctx.stream, device=local)
    for _ in range(args.warmup):
        rz.render(B.TRACE_NO_AOV)
    rz.ctx.synchronize()
    launches0 = int(rz.ctx.stats().kernelLaunches)
    if dist is not None:
        dist.barrier()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        rz.render(B.TRACE_NO_AOV)
    rz.frame.gather_wait()
    e1.record(stream)
    rz.ctx.synchronize()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop()
    last = rz.readback(B.AOV_BGRA8)
    frame_fnv = f"{int(rz.ctx.lib.vkrt_hash_fnv1a64(last.ctypes.data, last.nbytes)):016x}"
    if dist is not None:
        t = torch.tensor([ms], dtype=torch.float64, device=f"cuda:{local}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    launches = int(rz.ctx.stats().kernelLaunches) - launches0
    # end to end: the 64-byte matrix goes in with every call, the resolved frame comes back through the presentation ring
    # (on rank 0; the other ranks take part in the gather of every step)
    ring = [rz.ctx.host_array((Hh, W, 4)) for _ in range(3)] if rank == 0 else []
    tickets = [None, None, None]

    def e2e_step(k):
        if rank == 0 and tickets[k % 3] is not None:
            rz.frame.present_wait(tickets[k % 3])
        rz.render(B.TRACE_NO_AOV)
        if rank == 0:
            tickets[k % 3] = rz.frame.present(ring[k % 3])

    def drain():
        for i, t in enumerate(tickets):
            if t is not None:
                rz.frame.present_wait(t)
                tickets[i] = None
        rz.ctx.synchronize()

    for k in range(3):
        e2e_step(k)
    drain()
    t0 = time.perf_counter()
    for k in range(args.steps):
        e2e_step(k)
    drain()
    if dist is not None:
        dist.barrier()
    e2e_s = time.perf_counter() - t0
    if rank != 0:
        rz.close()
        dist.destroy_process_group()
        return
    tris = int(rz.blas.info().triangleCount)
    px = W * Hh
    peak, peak_src = read_peaks()
    algo = px * (64 + 64 + 4) + tris * (76 + 96)  # keys cleared + keys resolved + image; set-up read + write
    line = {
        "metric": RASTER_METRIC, "value": px * args.steps / (ms * 1e-3) / 1e6, "unit": "Mpixel/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms / max(args.steps, 1), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "frame_fnv": frame_fnv,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"raster: {desc}", "scene": model["provenance"], "width": W, "height": Hh, "msaa": 8, "triangles": tris,
                   "l2": "the 531 MB visibility-key buffer of a 4K frame exceeds the 126 MB L2; no explicit flush"},
        "mtri_s": tris * args.steps / (ms * 1e-3) / 1e6,
        "e2e": {"value": px * args.steps / e2e_s / 1e6, "unit": "Mpixel/s", "h2d_bytes_per_step": 64, "d2h_bytes_per_step": px * 4},
        "gpu_launches": launches, "clocks": clocks,
        "roofline": {"bound": "hbm", "achieved": algo * args.steps / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                     "frac": algo * args.steps / (ms * 1e-3) / 1e9 / peak, "traffic": None, "peak_source": peak_src,
                     "note": "whole frame, algorithmic bytes 132 B/pixel + 172 B/triangle; ncu (profiles/r01_ncu_raster_kernels_v1.txt) shows the "
                             "dominant kernel k_raster_large latency-bound at low occupancy, DRAM at 14 % of peak"},
    }
    if not args.no_cpu_baseline and world == 1:
        line["cpu_baseline"] = raster_cpu_baseline(model, 1280, 720)
    print(json.dumps(line), flush=True)
    rz.close()
    if dist is not None:
        dist.destroy_process_group()


def main():
    # stdout carries exactly one JSON line: this image exports NCCL_DEBUG=VERSION, which makes NCCL printf its version banner
    # to stdout; other NCCL_DEBUG levels (a user asking for INFO) are left alone but sent to stderr
    if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
        os.environ["NCCL_DEBUG"] = "WARN"
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="vkrt_b200", choices=["vkrt_b200", "reference"])
    ap.add_argument("--workload", default="config3", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--triangles", type=int, default=50_000_000, help="config5 only: triangles in the procedural soup")
    ap.add_argument("--build-flags", type=int, default=-1, help="VkrtBuildFlags; default 8 (PLOC + reinsertion passes, for hierarchies built once), config5 1 (LBVH: faster rebuild and, on the soup, fewer nodes per ray)")
    ap.add_argument("--texture-lod", default="zero", choices=["zero", "cones"],
                    help="zero: every texture() samples level 0 as the reference's ray-tracing shaders do (default, the parity mode); "
                         "cones: ray-cone level of detail over the generated mip chain (SURVEY 8 f3)")
    ap.add_argument("--lane-refill", type=int, default=-1, help="vkrt_trace_set_lane_refill; default 0, config5 20")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup

    if args.workload == "raster":
        run_raster(args)
        return
    if args.impl == "reference":
        run_reference(args)
        return

    import torch

    from vkrt_b200 import binding as B
    from vkrt_b200.raytracer import Raytracer

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    unique_id = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group(backend="nccl", device_id=torch.device("cuda", local))
        ids = [B.Context.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        unique_id = ids[0]

    W, Hh, spp, desc = WORKLOADS[args.workload]
    if args.build_flags < 0:
        # config 5 rebuilds every frame (LBVH); the others build once and trace many frames, like the reference's constructor:
        # VKRT_BUILD_PREFER_FAST_TRACE (reinsertion passes after PLOC, ~10x the build time, fewer node visits per ray)
        args.build_flags = 1 if args.workload == "config5" else 8
    soup = None
    t_ingest0 = time.perf_counter()
    if args.workload == "config5":
        from vkrt_b200 import scenes
        from vkrt_b200.scenes_device import SoupDevice

        soup = SoupDevice(args.triangles, local)
        model = soup.model_stub()
        instances = scenes.identity_instance(1.0)
        flat = {"vertices": soup.vertices, "indices": soup.indices, "ranges": soup.ranges(), "rows": np.zeros((1, 4), dtype=np.int32),
                "images": model["images"]}
        rt = Raytracer(None, W, Hh, device=local, instances=instances, rank=rank, world_size=world, unique_id=unique_id, flat=flat,
                       build_flags=args.build_flags)
    else:
        # scene ingest over N ranks: rank 0 decodes the 69 image files (thread pool) and uploads them; the other ranks build
        # only the geometry and receive the texel pool over NVLink (vkrt_textures_broadcast)
        model, instances = build_scene(args.workload, with_images=(rank == 0))
        t_ingest1 = time.perf_counter()
        rt = Raytracer(model, W, Hh, device=local, instances=instances, rank=rank, world_size=world, unique_id=unique_id,
                       build_flags=args.build_flags, textures_from_root=0 if world > 1 else None)
    rt.ctx.synchronize()
    t_ingest2 = time.perf_counter()
    if soup is not None:
        t_ingest1 = t_ingest0
    # ingest: host side (glTF / stand-in geometry + image decode), then upload + mip chain + first BLAS/TLAS build (cold)
    ingest = {"host_s": t_ingest1 - t_ingest0, "device_s": t_ingest2 - t_ingest1}
    build_ms = []
    lane_refill = args.lane_refill if args.lane_refill >= 0 else (20 if args.workload == "config5" else 0)
    rt.ctx.set_lane_refill(lane_refill)
    rt.set_texture_lod(B.TEXTURE_LOD_RAY_CONES if args.texture_lod == "cones" else B.TEXTURE_LOD_ZERO)
    info = rt.blas.info()
    cold_build_ms = float(info.buildMilliseconds)
    # warm BLAS build: median of 10 vkrt_blas_rebuild calls per hierarchy flavour (device time of the whole kernel chain,
    # CUDA events inside the library; storage and scratch already allocated) - the number "BVH build Mtri/s" is quoted on
    warm = {}
    verts_in = soup.vertices if soup is not None else rt.flat["vertices"]
    for name, fl in (("ploc_optimal", 0), ("lbvh_greedy", 1), ("ploc_reinsert_optimal", 8)):
        if fl == 8 and not (args.build_flags & 8):
            continue
        if soup is not None and fl != 1 and args.triangles > 20_000_000:
            continue  # config 5 at 50 M rebuilds with LBVH every frame; the PLOC scratch would not leave room for the 4K queues
        times = []
        for _ in range(10):
            rt.blas.rebuild(verts_in, fl)
            times.append(float(rt.blas.info().buildMilliseconds))
        warm[name] = {"ms_median": float(np.median(times)), "ms_min": float(np.min(times)),
                      "mtri_s": info.triangleCount / float(np.median(times)) / 1e3, "nodes": int(rt.blas.info().nodeCount),
                      "sah_cost": float(rt.blas.info().sahCost)}
    flavour = "lbvh_greedy" if args.build_flags & 1 else ("ploc_reinsert_optimal" if args.build_flags & 8 else "ploc_optimal")
    rt.rebuild(verts_in, flags=args.build_flags)  # back to the flavour the frames are traced on (+ a fresh TLAS)
    info = rt.blas.info()
    stream = torch.cuda.ExternalStream(rt.ctx.stream, device=local)
    flags = B.TRACE_NO_AOV

    def barrier():
        rt.ctx.synchronize()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()

    # untimed counting pass: mean nodes / triangles per ray for the algorithmic-bytes model
    set_pose(rt, args.workload, 0)
    rt.ctx.stats_enable(True, False)
    rt.ctx.stats_reset()
    rt.render(spp=1, flags=flags, gather=False)
    st = rt.ctx.stats()
    any_rays, closest_rays = max(st.raysShadow, 1), max(st.raysPrimary + st.raysReflection, 1)
    n_node_any, n_tri_any = st.nodesVisitedAny / any_rays, st.trianglesTestedAny / any_rays
    n_node_cl = (st.nodesVisited - st.nodesVisitedAny) / closest_rays
    n_tri_cl = (st.trianglesTested - st.trianglesTestedAny) / closest_rays
    rt.ctx.stats_enable(False, True)

    step_counter = [0]

    def rebuild_step():
        if soup is not None:  # config 5: every frame moves all triangles and rebuilds the whole BVH
            rt.rebuild(soup.set_frame(step_counter[0]), flags=args.build_flags)
            build_ms.append(rt.blas.info().buildMilliseconds)

    def device_step():
        set_pose(rt, args.workload, step_counter[0])
        rebuild_step()
        step_counter[0] += 1
        rt.render(spp=spp, flags=flags, gather=True)

    for _ in range(args.warmup):
        device_step()
    barrier()
    rt.ctx.stats_reset()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        device_step()
    rt.frame.gather_wait()  # the last frame's tile gather (its own stream) belongs to the timed region
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    st = rt.ctx.stats()
    # checksum of the last frame of the timed region (gathered on rank 0): equal at N = 1, 2, 4, 8 <=> same image
    frame_fnv = None
    if rank == 0:
        last = rt.readback(B.AOV_BGRA8)
        frame_fnv = f"{int(rt.ctx.lib.vkrt_hash_fnv1a64(last.ctypes.data, last.nbytes)):016x}"
    wave_samples = max(1, int(st.samplesPerWave))  # the stage times below cover the first wave = this many samples
    rays_local = np.array([st.raysPrimary, st.raysShadow, st.raysReflection], dtype=np.float64)
    launches = int(st.kernelLaunches)
    stage = {"raygen": st.msRaygen, "trace_closest": st.msTraceClosest, "shade": st.msShade, "trace_any": st.msTraceAny,
             "resolve": st.msResolve, "sample_total": st.msTotal}
    stage_each = [float(v) for v in st.msStage]  # raygen, closest d0, surface, any d0, resolve, closest d1, surface, any d1, resolve, sample resolve
    bounce_frac = st.raysShadowBounce / max(st.raysShadow, 1)

    # end to end through the reference-facing call: render + readback into pinned host memory every step
    ring = [rt.ctx.host_array((Hh, W, 4)) for _ in range(3)] if rank == 0 else []
    tickets = [None, None, None]
    rt.ctx.stats_enable(False, False)

    def e2e_step(k):
        set_pose(rt, args.workload, step_counter[0])
        rebuild_step()
        step_counter[0] += 1
        if rank == 0 and tickets[k % 3] is not None:  # acquire: the ring image's previous copy has landed
            rt.present_wait(tickets[k % 3])
        rt.render(spp=spp, flags=flags, gather=True)
        if rank == 0:
            tickets[k % 3] = rt.present(ring[k % 3])

    def e2e_drain():
        for i, t in enumerate(tickets):
            if t is not None:
                rt.present_wait(t)
                tickets[i] = None
        rt.ctx.synchronize()

    # restart the camera path so that the timed e2e steps walk the same frames as the device-timed steps did
    step_counter[0] = (step_counter[0] + 299) // 300 * 300
    for k in range(max(args.warmup, 2)):
        e2e_step(k)
    e2e_drain()
    barrier()
    rt.ctx.stats_reset()
    t0 = time.perf_counter()
    for k in range(args.steps):
        e2e_step(k)
    e2e_drain()
    barrier()
    e2e_s = time.perf_counter() - t0
    st2 = rt.ctx.stats()
    rays2_local = float(st2.raysPrimary + st2.raysShadow + st2.raysReflection)

    t = torch.tensor([ms, e2e_s * 1e3], dtype=torch.float64, device=f"cuda:{local}")
    r = torch.tensor(list(rays_local) + [rays2_local, float(launches)], dtype=torch.float64, device=f"cuda:{local}")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(r, op=dist.ReduceOp.SUM)
    ms, e2e_ms = t.tolist()
    rp, rs, rr, rays2, launches_all = r.tolist()
    total_rays = rp + rs + rr

    if rank == 0:
        peak, peak_src = read_peaks()
        # roofline of the dominant kernel (k_trace_any: shadow rays are ~3/4 of all rays).  Algorithmic bytes per ray:
        # 32 (ray) + 4 (result) + 80 * nodes visited + 48 * triangles tested.
        b_any = 32 + 4 + 80 * n_node_any + 48 * n_tri_any
        b_cl = 32 + 16 + 80 * n_node_cl + 48 * n_tri_cl
        sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
        issue_peak = 148 * 4 * sm_mhz * 1e6  # warp instructions per second: 148 SMs x 4 schedulers x clock
        per_wave = 1.0 / max(args.steps * spp, 1) / world * wave_samples  # whole-job ray totals -> one rank's first wave
        rays_launch = {"closest_d0": rp * per_wave, "any_d0": rs * (1.0 - bounce_frac) * per_wave,
                       "closest_d1": rr * per_wave, "any_d1": rs * bounce_frac * per_wave}
        ms_launch = {"closest_d0": stage_each[1], "any_d0": stage_each[3], "closest_d1": stage_each[5], "any_d1": stage_each[7]}
        capture, capture_note = read_traffic(args.workload)
        launches_out = {}
        for name in ("closest_d0", "any_d0", "closest_d1", "any_d1"):
            ms_l = max(ms_launch[name], 1e-9)
            algo = rays_launch[name] * (b_any if name.startswith("any") else b_cl)
            e = {"rays": rays_launch[name], "ms": ms_launch[name], "algorithmic_bytes": algo,
                 "frac_algorithmic": algo / (ms_l * 1e-3) / 1e9 / peak, "traffic": None, "dram_frac": None, "issue_frac": None,
                 "thread_inst_per_ray": None}
            c = (capture or {}).get("launches", {}).get(name) if world == 1 else None
            if c:
                e["traffic"] = c.get("dram_read_bytes", 0.0) + c.get("dram_write_bytes", 0.0)
                e["dram_frac"] = e["traffic"] / (ms_l * 1e-3) / 1e9 / peak
                e["issue_frac"] = c["warp_inst"] / (ms_l * 1e-3) / issue_peak
                e["thread_inst_per_ray"] = c["thread_inst"] / max(rays_launch[name], 1.0)
                e["warp_inst"] = c["warp_inst"]
            launches_out[name] = e
        dom = launches_out["any_d0"]
        roofline = {"bound": "hbm", "kernel": "k_trace_any, depth-0 launch (shadow rays of the primary hits; one launch per wave of samples)",
                    "achieved": dom["algorithmic_bytes"] / (max(dom["ms"], 1e-9) * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                    "frac": dom["frac_algorithmic"], "frac_algorithmic": dom["frac_algorithmic"],
                    "frac_of_8tbs_nominal": dom["algorithmic_bytes"] / (max(dom["ms"], 1e-9) * 1e-3) / 1e9 / 8000.0,
                    "peak_source": peak_src, "traffic": dom["traffic"], "traffic_source": capture_note,
                    "dram_frac": dom["dram_frac"], "issue_frac": dom["issue_frac"], "thread_inst_per_shadow_ray": dom["thread_inst_per_ray"],
                    "issue_peak_ginst_s": issue_peak / 1e9, "sm_mhz_used": sm_mhz,
                    "algorithmic_bytes_per_launch": dom["algorithmic_bytes"], "launch_ms": dom["ms"],
                    "bytes_per_ray_any": b_any, "bytes_per_ray_closest": b_cl,
                    "n_node_any": n_node_any, "n_tri_any": n_tri_any, "n_node_closest": n_node_cl, "n_tri_closest": n_tri_cl,
                    "launches": launches_out, "samples_per_wave": wave_samples, "lane_refill": lane_refill,
                    "stage_ms_sample0": stage, "stage_ms_each": stage_each,
                    "note": "frac = frac_algorithmic = algorithmic bytes (32 ray + 4 result + 80/node + 48/triangle, nodes and triangles per ray from "
                            "the counting pass, one mean for both depths) / live launch time / HBM peak; it exceeds 1 when the BVH is served from "
                            "L1/L2.  dram_frac = ncu-measured DRAM bytes of the same launch / live time / peak; issue_frac = ncu warp instructions / "
                            "live time / (148 SMs x 4 x SM clock): the roof that binds these kernels.  Both are null unless "
                            "profiles/r02_traffic.json was captured on exactly these library sources (single GPU only)."}
        line = {
            "metric": "Sponza 4K Mrays/s (prim+shadow+refl)", "value": total_rays / (ms * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / max(args.steps, 1), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {desc}", "scene": model["provenance"], "width": W, "height": Hh, "spp": spp,
                       "texture_lod": args.texture_lod, "triangles": int(info.triangleCount), "bvh8_nodes": int(info.nodeCount), "bvh_build_flags": args.build_flags, "bvh_flavour": flavour, "instances": 1 if instances is None else int(np.asarray(instances).size // 12),
                       "parallelism": f"interleaved 64x64 tiles over {world} rank(s), scene replicated, NCCL send/recv gather to rank 0",
                       "l2": "per-frame ray/hit/shading queues (>2 GB at 4K) exceed the 126 MB L2; no explicit flush"},
            "rays_per_step": {"primary": rp / args.steps, "shadow": rs / args.steps, "reflection": rr / args.steps},
            "mrays_s_by_type": {"primary": rp / (ms * 1e-3) / 1e6, "shadow": rs / (ms * 1e-3) / 1e6, "reflection": rr / (ms * 1e-3) / 1e6},
            "frame_fnv": frame_fnv, "ingest": ingest,
            "bvh_build": {"warm": warm, "warm_ms": (warm.get(flavour) or {}).get("ms_median"),
                          "warm_mtri_s": (warm.get(flavour) or {}).get("mtri_s"), "flavour": flavour,
                          "cold_first_build_ms": cold_build_ms, "tlas": {"instances": int(rt.tlas.info().instanceCount),
                                                                          "nodes": int(rt.tlas.info().nodeCount), "ms": float(rt.tlas.info().buildMilliseconds)},
                          "ms": info.buildMilliseconds, "mtri_s": info.triangleCount / max(info.buildMilliseconds, 1e-9) / 1e3,
                          "sah_cost": info.sahCost, "build_flags": args.build_flags, "rebuilds_in_timed_region": len(build_ms),
                          "rebuild_ms_median": float(np.median(build_ms)) if build_ms else None,
                          "rebuild_mtri_s": info.triangleCount / float(np.median(build_ms)) / 1e3 if build_ms else None},
            "e2e": {"value": rays2 / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": 256 * world,
                    "d2h_bytes_per_step": W * Hh * 4, "ms_per_step": e2e_ms / max(args.steps, 1)},
            "gpu_launches": int(launches_all), "clocks": clocks, "roofline": roofline,
        }
        if not args.no_cpu_baseline and world == 1 and soup is None:
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            import helpers as H
            from vkrt_b200 import scenes

            cores = os.cpu_count() or 1
            o = H.OracleScene(rt.flat, instances if instances is not None else scenes.REFERENCE_INSTANCE)
            sub = max(1, int(np.ceil(np.sqrt(W * Hh * spp * 4.6 / (12.0e6 * max(1.0, cores / 8.0))))))
            res = o.render(rt.camera.uniforms(), W, Hh, spp=spp, stride=(sub, sub), threads=cores)
            s = res["stats"]
            cpu_rays = s.raysPrimary + s.raysShadow + s.raysReflection
            line["cpu_baseline"] = {"value": cpu_rays / s.seconds / 1e6, "unit": "Mrays/s", "cores": cores, "kind": "port",
                                    "sample": f"every {sub}th pixel in x and y of the same frame ({cpu_rays} rays, {s.seconds:.1f} s)",
                                    "bvh_build_mtri_s": info.triangleCount / max(o.build_seconds(), 1e-9) / 1e6}
        print(json.dumps(line), flush=True)
    rt.close()
    del soup
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
