#!/usr/bin/env python
"""bench.py — the hot path's headline metric on N B200s of one node.

  python bench.py --gpus N --steps K --warmup W [--config 1|2|3|3b|4|5a|5b] [--technique PT|BPT|VCM|UPG] [--impl reference]
                  [--scaling weak|strong]
  (N > 1: python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...)

Default workload = BASELINE.json configs[3] ("CrytekSponza BPT cam0 ... 1/2/4/8-GPU sample-split scaling"):
models/CrytekSponza.blend is stripped from the reference mount, so the large-mesh stand-in of SURVEY.md §8(d) is used: the
reference's own models/LivingRoom.blend (567 304 triangles after one area light is added the way loader.cpp:434-456 adds lights;
tests/golden/make_large_scene.py), camera 0, 1024 x 1024 (the thesis resolution, GenerateResults.psm1:22), roulette 0.9,
beta 2 (BPT) / 1 (PT).  --config selects the stand-ins SURVEY.md §8(d) gives for the other BASELINE.json configs (CONFIGS below).

A STEP is `--spp` passes of one sample per pixel over the whole window, i.e. `spp` Technique::render calls in the reference's
terms.  Scaling is weak by default: every rank renders `--spp` passes per step of ITS OWN sample indices (rank g of N takes
samples g, g+N, ...) over a replicated scene; after the K timed steps the float64 accumulators are summed with ONE all-reduce
over NCCL, inside the timed region — the exact analogue of `master merge` (Options.cpp:1340-1409).  --scaling strong keeps the
TOTAL at `--spp` passes per step and gives each rank spp / N of them (a fixed sample budget split over the GPUs).

value     = pixel-samples of all ranks / device time (CUDA events on the launching stream, max over ranks),
            scene + BVH already resident in HBM.
e2e       = the same through the host-buffer entry point b2pt_render_host (what GpuTechnique::render calls), one call per step:
            parameters go host->device, the dvec4 film comes device->host and is added into the caller's host buffer, all
            inside the timed region (wall clock around synchronous calls); at N > 1 the per-rank host views are merged.
e2e_1pass = the same at the REFERENCE'S contract: one pass per render() call (Application.cpp:66), spp x K calls.
"""
import argparse
import ctypes as C
import hashlib
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

LARGE = os.path.join(ROOT, "scenes", "_large", "LivingRoom.b2s.xz")
GOLDEN_SCENES = os.path.join(ROOT, "tests", "golden", "scenes")
HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")              # parity build: RNG shim, Embree capped at AVX
HARNESS_STOCK = os.path.join(ROOT, "oracle", "_ref", "ref_harness_stock")  # as the reference ships: std::mt19937, -O2 AVX2+FMA, Embree AVX2
RAY_BYTES_CLOSEST = 48  # 32 B ray in + 16 B hit out
RAY_BYTES_SHADOW = 36   # 32 B ray in + 4 B visibility out

# BASELINE.json configs and the stand-ins SURVEY.md §8(d) gives for the scenes stripped from the mount.
# oracle_mean: image mean (r, g, b) of the reference's own CPU render of the scene (BASELINE.md §2 / tests/golden/oracle_means.json),
# the bench's correctness guard: a film whose mean is off is not a measurement.
CONFIGS = {
    "1": dict(scene="CornellBoxDiffuse", technique="PT", res=512, spp=64, what="BASELINE config 1: CornellBoxDiffuse PT"),
    "2": dict(scene="CornellBoxDiffuse", technique="BPT", res=512, spp=32, what="BASELINE config 2: CornellBoxDiffuse BPT"),
    "3": dict(scene="MetalRings", technique="BPT", res=512, spp=32,
              what="BASELINE config 3 stand-in (Bearings.blend is stripped): MetalRings BPT, Phong gold + 2 lights"),
    "3b": dict(scene="CornellBoxSpecular", technique="BPT", res=512, spp=32,
               what="BASELINE config 3 stand-in (Bearings.blend is stripped): CornellBoxSpecular BPT, mirror / glass"),
    "4": dict(scene="LivingRoom", technique="BPT", res=1024, spp=16,
              what="BASELINE config 4 stand-in (CrytekSponza.blend is stripped): LivingRoom + one area light, 567304 triangles, cam0"),
    "5a": dict(scene="CornellBoxDiffuse", technique="UPG", res=512, spp=8, radius=0.01,
               what="BASELINE config 5 stand-in (BreakfastRoom1.blend is stripped): CornellBoxDiffuse UPG r = 0.010"),
    "5b": dict(scene="CornellBoxDiffuse", technique="VCM", res=512, spp=8, radius=0.025,
               what="BASELINE config 5 stand-in (BreakfastRoom1.blend is stripped): CornellBoxDiffuse VCM r = 0.025, alpha 0.75"),
}


def scene_file(name):
    if name == "LivingRoom":
        return LARGE
    return os.path.join(GOLDEN_SCENES, name + ".b2s")


def plain_scene_file(path):
    """ref_harness reads plain .b2s; unpack the .xz once into the temp dir."""
    if not path.endswith(".xz"):
        return path
    out = os.path.join(tempfile.gettempdir(), "b2pt_" + os.path.basename(path)[:-3])
    if not os.path.exists(out) or os.path.getmtime(out) < os.path.getmtime(path):
        import lzma
        with lzma.open(path, "rb") as f, open(out + ".tmp", "wb") as g:
            g.write(f.read())
        os.replace(out + ".tmp", out)
    return out


def host_has_avx2():
    try:
        flags = open("/proc/cpuinfo").read()
        return all((" " + f + " ") in flags for f in ("avx2", "fma", "bmi2"))
    except Exception:
        return False


def timing_harness():
    """(binary, description) of the reference build that is timed: the stock-ISA build when this host can run it."""
    if os.path.exists(HARNESS_STOCK) and host_has_avx2():
        return HARNESS_STOCK, "stock build (std::mt19937, g++ -O2 -march=x86-64-v3, Embree 2.12 with EMBREE_MAX_ISA=AVX2 as Makefile:30-37 links it)"
    return HARNESS, "parity build (counter-based RNG shim, g++ -O2 -msse4.2, Embree 2.12 capped at the AVX ISA)"


def run_reference(scene_path, technique, width, height, passes, seconds, threads=0, radius=None):
    """The reference's own Technique::render loop (ref_harness time) on the host cores."""
    binary, build = timing_harness()
    extra = ["radius=%g" % radius] if radius is not None else []
    cmd = [binary, "time", scene_path, "technique=" + technique, "width=%d" % width, "height=%d" % height] + extra + [
        "num_samples=%d" % passes, "seconds=%g" % seconds, "threads=%d" % threads,
        "beta=%d" % (1 if technique == "PT" else 2)]
    out = subprocess.run(cmd, check=True, capture_output=True, text=True).stdout
    r = json.loads(out.strip().splitlines()[-1])
    r["build"] = build
    return r


def kernel_source_hash():
    """Hash of the traversal + builder sources: profiler-only figures (DRAM traffic, warp instructions per ray) come from a committed
    capture and are only quoted while the kernels they were captured from are the ones that run."""
    h = hashlib.sha1()
    for f in ("trace.cuh", "bvh8.cuh", "scene.cu", "scene.cuh"):
        with open(os.path.join(ROOT, "master_b200", "csrc", f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons of one GPU, sampled every 100 ms while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.sm, self.reasons, self.max_sm = index, False, [], set(), None

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_sm = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {
                "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
                "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
            }
            while not self.stop_flag:
                self.sm.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
                time.sleep(0.1)
        except Exception as e:  # pragma: no cover - box without NVML
            self.reasons.add("nvml_unavailable: %s" % e)

    def summary(self):
        sm = sorted(self.sm)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_sm, "reasons": sorted(self.reasons),
                "samples": len(sm)}


def physical_gpu_index(local):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local])
        except Exception:
            return local
    return local


_STDOUT_FD = None


def quiet_stdout():
    """Libraries (NCCL's version banner, for one) write to fd 1; the contract is ONE JSON line on stdout, so everything else
    goes to stderr and emit() writes the line to the real stdout."""
    global _STDOUT_FD
    sys.stdout.flush()
    _STDOUT_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    data = (json.dumps(line) + "\n").encode()
    if _STDOUT_FD is None:
        os.write(1, data)
    else:
        os.write(_STDOUT_FD, data)


def oracle_mean(scene, technique):
    try:
        m = json.load(open(os.path.join(ROOT, "tests", "golden", "oracle_means.json")))
        e = m.get("%s/%s" % (scene, technique)) or m.get(scene)
        return e
    except Exception:
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default=os.environ.get("B2PT_BENCH_CONFIG", "4"), choices=sorted(CONFIGS))
    ap.add_argument("--technique", default=os.environ.get("B2PT_BENCH_TECHNIQUE"), choices=["PT", "BPT", "VCM", "UPG"])
    ap.add_argument("--width", type=int, default=None)
    ap.add_argument("--height", type=int, default=None)
    ap.add_argument("--spp", type=int, default=None, help="passes (samples per pixel) per step")
    ap.add_argument("--radius", type=float, default=None, help="VCM / UPG merge radius (thesis values: UPG 0.01, VCM 0.025)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    a = ap.parse_args()
    quiet_stdout()
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cfg = dict(CONFIGS[a.config])
    technique = a.technique or cfg["technique"]
    W = a.width or cfg["res"]
    H = a.height or cfg["res"]
    spp = a.spp or cfg["spp"]
    if a.technique in ("VCM", "UPG") and "radius" not in cfg:
        cfg["radius"] = 0.01 if a.technique == "UPG" else 0.025
    radius = a.radius if a.radius is not None else cfg.get("radius")
    scene_name = cfg["scene"]
    scene_path = scene_file(scene_name)
    if not os.path.exists(scene_path):
        scene_name, scene_path = "MetalRings", scene_file("MetalRings")
        cfg["what"] += " — FALLBACK to MetalRings (30558 tris): %s absent" % os.path.relpath(scene_file(cfg["scene"]), ROOT)
    N = world
    strong = a.scaling == "strong"
    if strong and spp % N:
        raise SystemExit("bench.py: --scaling strong needs --spp divisible by the number of GPUs")
    rank_spp = spp // N if strong else spp  # passes per step this rank renders
    config = {"workload": "%s; %s %dx%d, %d spp/step%s, roulette 0.9, beta %d" %
                          (cfg["what"], technique, W, H, spp, " in total" if strong else " per GPU", 1 if technique == "PT" else 2),
              "config": a.config, "scene": scene_name, "technique": technique, "resolution": [W, H], "spp_per_step": spp,
              "parallelism": "sample-split x%d over a replicated scene, one all-reduce of the dvec4 film" % a.gpus}
    if radius is not None and technique in ("VCM", "UPG"):
        config["radius"] = radius
    ref_radius = radius if technique in ("VCM", "UPG") else None

    # ------------------------------------------------------------ reference arm: the reference's CPU code
    if a.impl == "reference":
        if rank != 0:
            return 0
        cores = os.cpu_count()
        plain = plain_scene_file(scene_path)
        # a step = ONE pass of the same window (bounded sample of the b200 arm's step of `spp` passes)
        run_reference(plain, technique, W, H, max(a.warmup, 1), 0, cores, radius=ref_radius) if a.warmup else None
        r = run_reference(plain, technique, W, H, a.steps, 0, cores, radius=ref_radius)
        val = r["msamples_per_s"]
        sample = "%d passes of 1 spp at %dx%d, %d host threads, the reference's Technique::render; %s" % (
            r["num_samples"], W, H, r["threads"], r["build"])
        line = {"impl": "reference", "metric": "Msamples/s", "value": val, "unit": "Msamples/s", "n_gpus": a.gpus,
                "steps": a.steps, "warmup": a.warmup, "ms_per_step": r["seconds"] / max(r["num_samples"], 1) * 1e3,
                "higher_is_better": True, "scaling": a.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": config, "mrays_per_s": r["mrays_per_s"],
                "cpu_baseline": {"value": val, "unit": "Msamples/s", "cores": r["threads"], "kind": "reference", "sample": sample},
                "e2e": {"value": val, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        emit(line)
        return 0

    # ------------------------------------------------------------ B200 arm
    import numpy as np
    import torch
    import torch.distributed as dist

    from master_b200 import api
    from master_b200.scene import load_b2s
    from master_b200.dist import reduce_film, sample_split
    from master_b200.technique import GpuTechnique

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the B200 path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    os.environ["B2PT_PROFILE_STAGES"] = "1"

    flat = load_b2s(scene_path)
    scene = api.DeviceSceneHandle(flat)
    info = scene.info()
    tech_options = {}
    if technique in ("VCM", "UPG"):
        tech_options["radius"] = radius
    tech = GpuTechnique(scene, technique, **tech_options)
    accum = torch.zeros(H * W, 4, dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream

    def split(i, passes):
        return sample_split(rank, N, i, passes)  # global sample (i * passes + j) * N + rank for pass j

    def step(i):
        first, stride = split(i, rank_spp)
        return tech.render_device(accum.data_ptr(), W, H, passes=rank_spp, stream=stream, first_sample=first, sample_stride=stride)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(a.warmup):
        step(i)
    accum.zero_()
    barrier()
    sampler = ClockSampler(physical_gpu_index(local))
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    tot = {k: 0 for k in ("num_basic_rays", "num_shadow_rays", "kernel_launches", "trace_launches", "shadow_launches",
                          "numeric_errors", "capacity_events")}
    ms = {"trace_ms": 0.0, "shadow_ms": 0.0, "shade_ms": 0.0}
    for i in range(a.steps):
        s = step(a.warmup + i)
        for k in tot:
            tot[k] += getattr(s, k)
        for k in ms:
            ms[k] += getattr(s, k)
    reduce_film(accum)  # the one collective of the path: sum of the per-rank films (no-op at N = 1)
    e1.record()
    barrier()
    sampler.stop_flag = True
    sampler.join()
    elapsed_ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    counts = torch.tensor([tot["num_basic_rays"], tot["num_shadow_rays"], tot["kernel_launches"]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(elapsed_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(counts)
    t_ms = float(elapsed_ms.item())
    samples = float(W) * H * rank_spp * a.steps * N
    value = samples / (t_ms * 1e-3) * 1e-6
    mrays = float(counts[0] + counts[1]) / (t_ms * 1e-3) * 1e-6
    # ---- correctness guard: every pixel-sample of every rank landed in the film (minus the non-finite ones the reference also drops),
    # and the film's mean is the scene's mean as the REFERENCE renders it (a black or wrongly weighted film is not a measurement)
    dropped = torch.tensor([tot["numeric_errors"]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(dropped)
    counted = float(accum[:, 3].sum().item())
    film_mean = (accum[:, :3].sum(dim=0) / max(counted, 1.0)).tolist()
    want_mean = oracle_mean(scene_name, technique)
    film_check = {"finite": bool(torch.isfinite(accum).all().item()), "samples_committed": counted == samples - float(dropped.item()),
                  "mean": film_mean, "oracle_mean": want_mean["mean"] if want_mean else None,
                  "oracle": want_mean["source"] if want_mean else "no oracle mean committed for this scene"}
    tol = 0.03 if technique in ("VCM",) else 0.02  # VCM is biased by design (0.5 % at r = 0.025 on Cornell)
    film_check["mean_within_tol"] = (bool(np.allclose(film_mean, want_mean["mean"], rtol=tol)) if want_mean else None)
    film_check["rtol"] = tol
    film_ok = film_check["finite"] and film_check["samples_committed"] and film_check["mean_within_tol"] is not False

    # ------------------------------------------------------------ end to end through host buffers
    def host_loop(passes_per_call, calls_per_step):
        """K steps through b2pt_render_host; the per-rank host views are merged at N > 1 (the reduce of the e2e arm)."""
        view = np.zeros((H, W, 4), dtype=np.float64)
        tech.render(view, passes=passes_per_call, first_sample=rank, sample_stride=N)  # warm the host path (buffers, threads)
        tech.render(view, passes=0)  # flush: nothing rendered ahead of the timed region survives into it
        view[:] = 0
        barrier()
        t0 = time.perf_counter()
        for i in range(a.steps):
            for c in range(calls_per_step):
                first, stride = split((a.warmup + i) * calls_per_step + c, passes_per_call)
                tech.render(view, passes=passes_per_call, first_sample=first, sample_stride=stride)
        if world > 1:
            dv = torch.from_numpy(view).cuda()
            reduce_film(dv)
            if rank == 0:
                view[:] = dv.cpu().numpy()
        # every call has returned with its passes added into `view` (the device->host copies are complete inside the calls): the
        # clock stops here.  Passes the library is still rendering AHEAD of the caller are not waited for — they are not part of
        # what was asked — and are flushed after the clock is read.
        wall = time.perf_counter() - t0
        tech.render(view, passes=0)
        torch.cuda.synchronize()
        dt = torch.tensor([wall], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        ok = bool(np.isfinite(view).all()) and (rank != 0 or abs(float(view[:, :, 3].sum()) - samples) <= 64)
        return samples / float(dt.item()) * 1e-6, ok

    e2e = e2e_1pass = None
    if not a.no_e2e:
        v, ok = host_loop(rank_spp, 1)
        e2e = {"value": v, "unit": "Msamples/s", "h2d_bytes_per_step": C.sizeof(api.RenderParams), "d2h_bytes_per_step": W * H * 32, "film_ok": ok,
               "note": "b2pt_render_host, one call of %d passes per step: parameters in, dvec4 film out to host memory and added into "
                       "the caller's view" % rank_spp}
        v1, ok1 = host_loop(1, rank_spp)
        e2e_1pass = {"value": v1, "unit": "Msamples/s", "h2d_bytes_per_step": C.sizeof(api.RenderParams) * rank_spp,
                     "d2h_bytes_per_step": W * H * 32 * rank_spp, "calls_per_step": rank_spp, "film_ok": ok1,
                     "note": "the reference's contract: ONE pass per Technique::render call (Application.cpp:66); every call returns "
                             "with its pass added into the caller's host view"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    # dominant kernel = closest-hit traversal; algorithmic bytes = 48 B per ray (rank 0's launches)
    trace_s = ms["trace_ms"] * 1e-3
    ach = tot["num_basic_rays"] * RAY_BYTES_CLOSEST / trace_s * 1e-9 if trace_s > 0 else None
    roofline = {"bound": "hbm", "kernel": "closest-hit traversal stage", "achieved": ach, "peak": peak, "unit": "GB/s",
                "frac": (ach / peak) if ach else None, "traffic": None,
                "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650 GB/s",
                "algorithmic_bytes_per_ray": RAY_BYTES_CLOSEST,
                "avg_launch_ms": ms["trace_ms"] / max(tot["trace_launches"], 1),
                "rays_per_launch": tot["num_basic_rays"] / max(tot["trace_launches"], 1),
                "grays_per_s_in_kernel": tot["num_basic_rays"] / trace_s * 1e-9 if trace_s > 0 else None,
                "stage_ms_rank0": ms,
                "shadow": {"achieved": tot["num_shadow_rays"] * RAY_BYTES_SHADOW / (ms["shadow_ms"] * 1e-3) * 1e-9 if ms["shadow_ms"] > 0 else None,
                           "grays_per_s_in_kernel": tot["num_shadow_rays"] / (ms["shadow_ms"] * 1e-3) * 1e-9 if ms["shadow_ms"] > 0 else None},
                "note": "traversal is issue bound by design (SURVEY.md §8(d)): the HBM fraction is small because the scene is served from "
                        "L2/L1; the `issue` object is the roof that binds; see profiles/"}
    # figures that only a profiler sees (DRAM traffic, warp instructions per ray) come from a committed ncu capture of THESE kernels:
    # the capture carries the hash of the traversal + builder sources it was taken from and is refused when they have changed since
    try:
        cap = json.load(open(os.path.join(ROOT, "profiles", "roofline_inputs.json")))
        now = kernel_source_hash()
        if a.config != "4" or technique != "BPT":
            roofline["traffic_note"] = "the committed ncu capture is of config 4 / BPT; not quoted for this workload"
        elif cap.get("kernel_source_hash") != now:
            roofline["traffic_note"] = "STALE capture refused: %s was taken from kernel sources %s, the sources are now %s — re-capture " \
                                       "(tools/gpu_session.sh full) and regenerate profiles/roofline_inputs.json" % (
                                           cap.get("capture"), cap.get("kernel_source_hash"), now)
        else:
            per_ray = (cap["dram_bytes_read"] + cap["dram_bytes_written"]) / cap["rays_in_launch"]
            roofline["traffic"] = per_ray * roofline["rays_per_launch"]  # bytes per launch, scaled to this run's mean launch size
            roofline["traffic_source"] = "static capture %s @ kernel sources %s" % (cap["capture"], cap["kernel_source_hash"])
            roofline["traffic_note"] = "%.0f B/ray of DRAM traffic in the capture vs %d B/ray algorithmic; L2 hit rate %.0f %%" % (
                per_ray, RAY_BYTES_CLOSEST, cap.get("l2_hit_pct", float("nan")))
            clk = sampler.summary()["sm_mhz"] or 1965
            wi = cap["warp_instructions"] / cap["rays_in_launch"]
            peak_issue = 148 * 4 * clk * 1e6 * 1e-9  # Gwarp-instr/s: 148 SMs x 4 schedulers x 1 instruction per cycle
            ach_issue = wi * roofline["grays_per_s_in_kernel"] if roofline["grays_per_s_in_kernel"] else None
            roofline["issue"] = {"bound": "issue", "achieved": ach_issue, "peak": peak_issue, "unit": "Gwarp-instr/s",
                                 "frac": ach_issue / peak_issue if ach_issue else None, "warp_instr_per_ray": wi,
                                 "active_lanes_per_instr": cap["active_lanes_per_instruction"],
                                 "source": roofline["traffic_source"],
                                 "note": "closest-hit stage while it shares the GPU with the other batch's kernels; warp instructions per ray "
                                         "from the capture, rays/s from this run's CUDA events"}
    except Exception as e:  # pragma: no cover
        roofline["traffic_note"] = "profiles/roofline_inputs.json unreadable: %s" % e

    cpu = None
    if not a.no_cpu_baseline and os.path.exists(HARNESS):
        try:
            r = run_reference(plain_scene_file(scene_path), technique, W, H, 1000, 15.0, os.cpu_count(), radius=ref_radius)
            cpu = {"value": r["msamples_per_s"], "unit": "Msamples/s", "cores": r["threads"], "kind": "reference",
                   "mrays_per_s": r["mrays_per_s"],
                   "sample": "%d passes of 1 spp at %dx%d in %.1f s: the unmodified reference (Embree 2.12 + Technique::render) "
                             "on %d host threads; %s" % (r["num_samples"], W, H, r["seconds"], r["threads"], r["build"])}
        except Exception as e:  # pragma: no cover
            cpu = {"value": None, "unit": "Msamples/s", "cores": 0, "kind": "reference", "sample": "failed: %s" % e}

    l2w = scene.l2_window()
    line = {"metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": N, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": t_ms / a.steps, "higher_is_better": True, "scaling": a.scaling, "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": config, "mrays_per_s": mrays,
            "rays_per_sample": float(counts[0] + counts[1]) / samples,
            "gpu_launches": int(counts[2].item()), "e2e": e2e, "e2e_1pass": e2e_1pass, "roofline": roofline, "cpu_baseline": cpu,
            "clocks": sampler.summary(), "film_ok": film_ok, "film_check": film_check,
            "l2_policy": "inputs larger than L2: scene %.0f MB (BVH %.0f MB + shading frames) + path pools > 126 MB, no flush needed; %s" % (
                (info["bvh_bytes"] + info["triangles"] * 112) / 1e6, info["bvh_bytes"] / 1e6,
                "the BVH block is pinned in persisting L2 (cudaAccessPolicyWindow, B2PT_L2_PERSIST=1), pools stream through the rest"
                if l2w.get("window_bytes") else "no persisting-L2 window (measured slower, profiles/r02_ab_runs.md): plain LRU"),
            "scene_info": info, "l2_window": l2w, "numeric_errors": tot["numeric_errors"], "capacity_events": tot["capacity_events"]}
    emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
