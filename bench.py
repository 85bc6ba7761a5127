#!/usr/bin/env python
"""bench.py — the hot path's headline metric on N B200s of one node.

  python bench.py --gpus N --steps K --warmup W [--technique PT|BPT] [--impl reference]
  (N > 1: python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...)

Workload (BASELINE.json configs[3], "CrytekSponza BPT cam0 ... 1/2/4/8-GPU sample-split scaling"):
models/CrytekSponza.blend is stripped from the reference mount, so the large-mesh stand-in of SURVEY.md
§8(d) is used: the reference's own models/LivingRoom.blend (567 304 triangles after one area light is added
the way loader.cpp:434-456 adds lights; tests/golden/make_large_scene.py), camera 0, 1024 x 1024 (the thesis
resolution, GenerateResults.psm1:22), roulette 0.9, beta 2 (BPT) / 1 (PT).

A STEP is `--spp` passes (default 16) of one sample per pixel over the whole window, i.e. one
Technique::render call per pass in the reference's terms.  Scaling is weak: every rank renders `--spp`
passes per step of ITS OWN sample indices (rank g of N takes samples g, g+N, ...) over a replicated scene;
after the K timed steps the float64 accumulators are summed with ONE all-reduce over NCCL, inside the timed
region — the exact analogue of `master merge` (Options.cpp:1340-1409).

value  = pixel-samples of all ranks / device time (CUDA events on the launching stream, max over ranks),
         scene + BVH already resident in HBM.
e2e    = the same through the host-buffer entry point b2pt_render_host (what GpuTechnique::render calls):
         per step the parameters go host->device and the dvec4 view comes device->host and is added into
         the caller's host buffer, all inside the timed region (wall clock around synchronous calls).
"""
import argparse
import ctypes as C
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
FALLBACK = os.path.join(ROOT, "tests", "golden", "scenes", "MetalRings.b2s")
HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
RAY_BYTES_CLOSEST = 48  # 32 B ray in + 16 B hit out
RAY_BYTES_SHADOW = 36   # 32 B ray in + 4 B visibility out


def workload():
    if os.path.exists(LARGE):
        return LARGE, "LivingRoom+arealight (567304 tris; stand-in for the stripped CrytekSponza.blend), cam0"
    return FALLBACK, "MetalRings (30558 tris) — FALLBACK: scenes/_large/LivingRoom.b2s.xz absent"


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


def run_reference(scene_file, technique, width, height, passes, seconds, threads=0, radius=None):
    """The reference's own Technique::render loop (oracle/_ref/ref_harness time) on the host cores."""
    extra = ["radius=%g" % radius] if radius is not None else []
    cmd = [HARNESS, "time", scene_file, "technique=" + technique, "width=%d" % width, "height=%d" % height] + extra + [
        "num_samples=%d" % passes, "seconds=%g" % seconds, "threads=%d" % threads,
        "beta=%d" % (1 if technique == "PT" else 2)]
    out = subprocess.run(cmd, check=True, capture_output=True, text=True).stdout
    return json.loads(out.strip().splitlines()[-1])


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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--technique", default=os.environ.get("B2PT_BENCH_TECHNIQUE", "BPT"), choices=["PT", "BPT", "VCM", "UPG"])
    ap.add_argument("--width", type=int, default=1024)
    ap.add_argument("--height", type=int, default=1024)
    ap.add_argument("--spp", type=int, default=16, help="passes (samples per pixel) per step")
    ap.add_argument("--radius", type=float, default=None, help="VCM / UPG merge radius (thesis values: UPG 0.01, VCM 0.025)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    a = ap.parse_args()
    quiet_stdout()
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    scene_path, scene_name = workload()
    W, H = a.width, a.height
    config = {"workload": "%s, %s %dx%d, %d spp/step, roulette 0.9, beta %d" %
                          (scene_name, a.technique, W, H, a.spp, 1 if a.technique == "PT" else 2),
              "technique": a.technique, "resolution": [W, H], "spp_per_step": a.spp,
              "parallelism": "sample-split x%d over a replicated scene, one all-reduce of the dvec4 film" % a.gpus}

    ref_radius = None
    if a.technique in ("VCM", "UPG"):
        ref_radius = a.radius if a.radius is not None else (0.01 if a.technique == "UPG" else 0.025)
    # ------------------------------------------------------------ reference arm: the reference's CPU code
    if a.impl == "reference":
        if rank != 0:
            return 0
        cores = os.cpu_count()
        plain = plain_scene_file(scene_path)
        # a step = ONE pass of the same window (bounded sample of the b200 arm's step of `spp` passes)
        run_reference(plain, a.technique, W, H, max(a.warmup, 1), 0, cores, radius=ref_radius) if a.warmup else None
        r = run_reference(plain, a.technique, W, H, a.steps, 0, cores, radius=ref_radius)
        val = r["msamples_per_s"]
        sample = "%d passes of 1 spp at %dx%d, %d host threads, Embree 2.12 + the reference's Technique::render" % (
            r["num_samples"], W, H, r["threads"])
        line = {"impl": "reference", "metric": "Msamples/s", "value": val, "unit": "Msamples/s", "n_gpus": a.gpus,
                "steps": a.steps, "warmup": a.warmup, "ms_per_step": r["seconds"] / max(r["num_samples"], 1) * 1e3,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
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
    if a.technique in ("VCM", "UPG"):
        tech_options["radius"] = a.radius if a.radius is not None else (0.01 if a.technique == "UPG" else 0.025)
        config["radius"] = tech_options["radius"]
    tech = GpuTechnique(scene, a.technique, **tech_options)
    accum = torch.zeros(H * W, 4, dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    N = world

    def step(i):
        first, stride = sample_split(rank, N, i, a.spp)  # global sample (i * spp + j) * N + rank for pass j
        return tech.render_device(accum.data_ptr(), W, H, passes=a.spp, stream=stream, first_sample=first, sample_stride=stride)

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
    samples = float(W) * H * a.spp * a.steps * N
    value = samples / (t_ms * 1e-3) * 1e-6
    mrays = float(counts[0] + counts[1]) / (t_ms * 1e-3) * 1e-6
    # every pixel-sample of every rank landed in the film (minus the non-finite ones the reference also drops)
    dropped = torch.tensor([tot["numeric_errors"]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(dropped)
    film_ok = bool(torch.isfinite(accum).all().item()) and float(accum[:, 3].sum().item()) == samples - float(dropped.item())

    # ------------------------------------------------------------ end to end through host buffers
    e2e = None
    if not a.no_e2e:
        view = np.zeros((H, W, 4), dtype=np.float64)
        tech.render(view, passes=a.spp, first_sample=rank, sample_stride=N)  # warm the host path
        view[:] = 0
        barrier()
        t0 = time.perf_counter()
        for i in range(a.steps):
            tech.render(view, passes=a.spp, first_sample=(a.warmup + i) * a.spp * N + rank, sample_stride=N)
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": samples / float(dt.item()) * 1e-6, "unit": "Msamples/s",
               "h2d_bytes_per_step": C.sizeof(api.RenderParams), "d2h_bytes_per_step": W * H * 32,
               "note": "b2pt_render_host: parameters in, dvec4 film out to pageable host memory and added into the caller's view"}

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
    # figures that only a profiler sees (DRAM traffic, warp instructions per ray) come from the committed ncu capture
    try:
        cap = json.load(open(os.path.join(ROOT, "profiles", "r01_roofline_inputs.json")))
        per_ray = (cap["dram_bytes_read"] + cap["dram_bytes_written"]) / cap["rays_in_launch"]
        roofline["traffic"] = per_ray * roofline["rays_per_launch"]  # bytes per launch, scaled to this run's mean launch size
        roofline["traffic_note"] = "%.0f B/ray of DRAM traffic in the capture (%s) vs %d B/ray algorithmic: the excess is BVH nodes and " \
                                   "triangles that missed the L2" % (per_ray, cap["capture"], RAY_BYTES_CLOSEST)
        clk = sampler.summary()["sm_mhz"] or 1965
        wi = cap["warp_instructions"] / cap["rays_in_launch"]
        peak_issue = 148 * 4 * clk * 1e6 * 1e-9  # Gwarp-instr/s: 148 SMs x 4 schedulers x 1 instruction per cycle
        ach_issue = wi * roofline["grays_per_s_in_kernel"] if roofline["grays_per_s_in_kernel"] else None
        roofline["issue"] = {"bound": "issue", "achieved": ach_issue, "peak": peak_issue, "unit": "Gwarp-instr/s",
                             "frac": ach_issue / peak_issue if ach_issue else None, "warp_instr_per_ray": wi,
                             "active_lanes_per_instr": cap["active_lanes_per_instruction"],
                             "note": "closest-hit stage while it shares the GPU with the other batch's kernels; warp instructions per ray "
                                     "from the capture, rays/s from this run's CUDA events"}
    except Exception as e:  # pragma: no cover
        roofline["traffic_note"] = "profiles/r01_roofline_inputs.json unreadable: %s" % e

    cpu = None
    if not a.no_cpu_baseline and os.path.exists(HARNESS):
        try:
            r = run_reference(plain_scene_file(scene_path), a.technique, W, H, 1000, 15.0, os.cpu_count(), radius=ref_radius)
            cpu = {"value": r["msamples_per_s"], "unit": "Msamples/s", "cores": r["threads"], "kind": "reference",
                   "mrays_per_s": r["mrays_per_s"],
                   "sample": "%d passes of 1 spp at %dx%d in %.1f s: the unmodified reference (Embree 2.12 + Technique::render) "
                             "on %d host threads" % (r["num_samples"], W, H, r["seconds"], r["threads"])}
        except Exception as e:  # pragma: no cover
            cpu = {"value": None, "unit": "Msamples/s", "cores": 0, "kind": "reference", "sample": "failed: %s" % e}

    line = {"metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": N, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": t_ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": config, "mrays_per_s": mrays,
            "rays_per_sample": float(counts[0] + counts[1]) / samples,
            "gpu_launches": int(counts[2].item()), "e2e": e2e, "roofline": roofline, "cpu_baseline": cpu,
            "clocks": sampler.summary(), "film_ok": film_ok,
            "l2_policy": "inputs larger than L2: scene %.0f MB (BVH %.0f MB + shading frames) + path pool > 126 MB" % (
                (info["bvh_bytes"] + info["triangles"] * 112) / 1e6, info["bvh_bytes"] / 1e6),
            "scene": info, "numeric_errors": tot["numeric_errors"], "capacity_events": tot["capacity_events"]}
    emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
