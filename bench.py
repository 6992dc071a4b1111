#!/usr/bin/env python
"""bench.py — BASELINE.json's metric (Msamples/s = paths/s; Mrays/s beside it) on BASELINE config 2
(249,924-tri procedural mesh, 1920x1080, 64 spp, synthetic HDRI env) through the C ABI.

  python bench.py --gpus N --steps K --warmup W            the CUDA path (one process per GPU under torchrun)
  python bench.py --impl reference --gpus N --steps K ...  the reference's own CPU code (oracle/_ref), rank 0 only

A step = one whole job of the hot path: setupKernel + `spp` samples for every pixel of the frame
(er_render_reset + er_render_enqueue).  With N GPUs the job grows with N so that per-GPU work is fixed (weak
scaling): by default the same view is rendered N times as tall (1920 x 1080*N, 64 spp; `--scale-by spp`
keeps the 1920x1080 frame and renders 64*N spp instead); the frame is sharded by interleaved 8-row blocks,
there is no collective in the render loop, only an NCCL reduce at readout (inside the e2e timing).  `value`
is timed on the device with CUDA events with the scene resident in HBM; `e2e` is the same job through the
C ABI from HOST buffers: scene upload (flatten + H2D), render, beauty pass back to host memory.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SPP = 64
ROWS_PER_BLOCK = 8


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), float(d.get("sm_max_mhz", 1965.0)), "measured"
    return 6650.0, 1965.0, "fallback"


class ClockSampler(threading.Thread):
    """samples SM clock + throttle reasons of one GPU through NVML while the timed region runs"""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.sm, self.reasons, self.max_sm = index, False, [], set(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_sm = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception as e:  # noqa: BLE001
            self.nv = None
            log("clock sampler unavailable:", e)

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksEventReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksEventReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksEventReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksEventReasonSwPowerCap: "sw_power_cap",
                 nv.nvmlClocksEventReasonHwPowerBrakeSlowdown: "hw_power_brake"}
        while not self.stop_flag:
            try:
                self.sm.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for bit, n in names.items():
                    if r & bit:
                        self.reasons.add(n)
            except Exception:  # noqa: BLE001
                pass
            time.sleep(0.05)

    def result(self):
        self.stop_flag = True
        if self.nv is None or not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": self.max_sm, "reasons": ["unavailable"]}
        return {"sm_mhz": float(np.median(self.sm)), "sm_max_mhz": self.max_sm, "reasons": sorted(self.reasons),
                "samples": len(self.sm)}


# ---------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the reference's own renderingKernel, per pixel index, on the host cores
# ---------------------------------------------------------------------------------------------------
def build_ref_scene(spec, nodes=None, tri_indices=None):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import common
    return common.to_ref(spec, nodes, tri_indices)


def time_reference(rs, steps, warmup, budget_s=None):
    """each step = renderingKernel once for every pixel (1 spp over the whole frame), all host threads"""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import refharness as R
    threads = R.hardware_threads()
    for _ in range(warmup):
        rs.render(1, threads)
    times = []
    t_all = time.perf_counter()
    for _ in range(steps):
        t0 = time.perf_counter()
        rs.render(1, threads)
        times.append(time.perf_counter() - t0)
        if budget_s is not None and time.perf_counter() - t_all > budget_s:
            break
    return times, threads


def run_reference(args, rank, world):
    if rank != 0:
        return
    from elevenrender_b200 import workloads
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import refharness as R
    spec = workloads.workload(args.workload)
    if args.scale_by == "rows" and world > 1:  # the job of the CUDA arm at this N (same view, N times as tall)
        spec["y_res"] *= world
        spec["desc"] += f" [weak scaling: frame {spec['x_res']}x{spec['y_res']}, same view]"
    if not R.available():
        emit({"impl": "reference", "unavailable": "oracle/_ref/liber_ref.so not built"})
        return
    t0 = time.perf_counter()
    rs = build_ref_scene(spec)  # the reference's own BVH::build
    log(f"reference scene + BVH built in {time.perf_counter() - t0:.1f} s")
    times, threads = time_reference(rs, args.steps, min(args.warmup, 1))
    P = spec["x_res"] * spec["y_res"]
    ms = 1e3 * float(np.mean(times))
    v = P / (ms * 1e-3) / 1e6
    sample = f"1 spp over the full {spec['x_res']}x{spec['y_res']} frame per step ({P} renderingKernel calls)"
    line = {"impl": "reference", "metric": "Msamples/s", "value": v, "unit": "Msamples/s", "n_gpus": args.gpus,
            "steps": len(times), "warmup": min(args.warmup, 1), "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": spec["desc"], "x_res": spec["x_res"], "y_res": spec["y_res"], "spp_per_step": 1,
                       "threads": threads, "weak_scaling_by": args.scale_by},
            "cpu_baseline": {"value": v, "unit": "Msamples/s", "cores": threads, "kind": "reference", "sample": sample},
            "e2e": {"value": v, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


# ---------------------------------------------------------------------------------------------------
# the CUDA arm
# ---------------------------------------------------------------------------------------------------
class _DevView:
    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": "<f4", "data": (int(ptr), False), "version": 2}


def algorithmic_per_sample(c, spec, spp_per_launch):
    """algorithmic bytes and FP32 thread-ops per sample from the kernel's own counters"""
    s = max(1, c["samples"])
    hd = spec["hdri"]
    cdf_probes = int(np.ceil(np.log2(hd[0] * hd[1]))) if len(hd) == 5 else 0
    texels = 2 * c["shaded_hits"] + c["samples"]  # env texel at NEE + pdf lookup, env texel on a miss (upper bound 1)
    a_bytes = (32 * c["box_tests"] + 48 * c["tri_tests"] + 48 * c["candidates"] + 160 * c["shaded_hits"] +
               4 * cdf_probes * c["nee_samples"] + 12 * texels) / s + 144.0 / spp_per_launch
    a_ops = (24 * c["box_tests"] + 53 * c["tri_tests"] + 200 * c["candidates"] + 450 * c["shaded_hits"]) / s + 60
    return a_bytes, a_ops


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from elevenrender_b200 import capi, host, sharding, workloads

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    spec = workloads.workload(args.workload)
    # weak scaling (per-GPU work fixed as N grows), two ways to grow the job:
    #   rows (default): the same view rendered N times as tall (1920 x 1080*N pixels, args.spp samples): every
    #                   GPU gets the N=1 number of pixels, interleaved 8-row blocks
    #   spp:            the 1920x1080 frame at args.spp*N samples: every GPU gets 1/N of the pixels. A pixel's
    #                   samples are sequential, so a GPU then holds most of its shard in flight at once and its
    #                   rays span the whole scene (L1 hit rate 51 % -> 29 %, profiles/README.md)
    if args.scale_by == "rows":
        spec["y_res"] *= world
        spp = args.spp
        if world > 1:
            spec["desc"] += f" [weak scaling: frame {spec['x_res']}x{spec['y_res']}, same view]"
    else:
        spp = args.spp * world
    P = spec["x_res"] * spec["y_res"]
    t0 = time.perf_counter()
    scene = workloads.scene_from_spec(spec)
    scene.buildBVH()
    t_bvh = time.perf_counter() - t0
    pars = host.RenderParameters(spec["x_res"], spec["y_res"], spp, device=local_rank, rng_order=host.ER_RNG_LTR,
                                 spp_per_launch=args.spp_per_launch, kernel=args.kernel)
    # a real (non-default) torch stream: the library launches on it and torch.cuda.Event times it.
    # (handle 0 would mean "use the context's own stream", invisible to events on torch's stream)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    rm = host.RenderingManager(pars, shard=(rank, world, ROWS_PER_BLOCK), stream=stream.cuda_stream)
    rm.upload(scene)
    own_pixels = sharding.owned_pixel_count(rank, world, spec["x_res"], spec["y_res"], ROWS_PER_BLOCK)

    # -- counters (untimed, count_stats build of the same kernel): work per sample -------------------
    rm.pars.count_stats = True
    rm.reset_counters()
    rm.render_samples(2)
    counters = rm.get_counters()
    rm.pars.count_stats = False
    rm.reset()
    rm.reset_counters()
    a_bytes, a_ops = algorithmic_per_sample(counters, spec, min(args.spp_per_launch, spp))
    rays_per_sample = counters["rays"] / max(1, counters["samples"])

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    def step():
        rm.reset()               # setupKernel
        rm.render_enqueue(spp)   # the render launches, no host sync

    for _ in range(max(args.warmup, 3) if args.warmup else 0):
        step()
    torch.cuda.synchronize()
    rm.reset_counters()  # also zeroes the library's own kernel-time accumulation

    sampler = ClockSampler(local_rank)
    sampler.start()
    evs = []
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t_wall0 = time.perf_counter()
    for _ in range(args.steps):
        flush.zero_()  # L2 flush between timed iterations (not timed)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        step()
        e1.record(stream)
        evs.append((e0, e1))
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.result()
    ms_total = sum(a.elapsed_time(b) for a, b in evs)
    kern_ms, kern_launches = rm.kernel_time()  # events around the render kernels only
    t = torch.tensor([ms_total, kern_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, kern_ms_max = float(t[0]), float(t[1])
    ms_per_step = ms_total / args.steps
    value = P * spp / (ms_per_step * 1e-3) / 1e6
    launches_per_step = kern_launches // args.steps
    gpu_launches = args.steps * (1 + launches_per_step)  # k_setup + render kernels

    # -- roofline of the dominant kernel (k_render_wavefront), this rank's launches ---------------------
    hbm_peak, sm_max_mhz, peak_src = measured_peaks()
    samples_per_launch = own_pixels * spp / max(1, launches_per_step)
    launch_s = kern_ms / max(1, kern_launches) * 1e-3
    ach_gbs = a_bytes * samples_per_launch / launch_s / 1e9
    sm_mhz = clocks["sm_mhz"] or sm_max_mhz
    info = capi.device_info(local_rank)
    lanes = info["max_compute_units"] * 4 * 32
    ach_ops = a_ops * samples_per_launch / launch_s
    # DRAM traffic of the kernel: per-sample figure of one `ncu --set full` capture (profiles/ncu_traffic.json),
    # scaled to this launch's samples
    traffic, ncu = None, {}
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(tpath):
        ncu = json.load(open(tpath)).get(args.workload, {})
        if "dram_bytes_per_sample" in ncu:
            traffic = ncu["dram_bytes_per_sample"] * samples_per_launch
    roofline = {"bound": "hbm", "achieved": ach_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": ach_gbs / hbm_peak,
                "traffic": traffic, "traffic_unit": "bytes per launch (ncu dram read+write, scaled by samples)",
                "traffic_source": ncu.get("source"), "peak_source": peak_src, "kernel": "k_render_wavefront",
                "algorithmic_bytes_per_sample": a_bytes, "avg_launch_ms": launch_s * 1e3,
                "note": "algorithmic bytes are served mostly by L1/L2 (traffic << achieved*time): the limiting roofline is SM issue rate, see roofline_issue"}
    roofline_issue = {"bound": "sm_issue", "achieved": ach_ops / 1e12, "unit": "T thread-op/s (algorithmic FP32, no FMA)",
                      "peak_at_max_clock": lanes * sm_max_mhz * 1e6 / 1e12,
                      "peak_at_measured_clock": lanes * sm_mhz * 1e6 / 1e12,
                      "frac": ach_ops / (lanes * sm_mhz * 1e6), "algorithmic_ops_per_sample": a_ops,
                      "issue_slots_busy_ncu": ncu.get("issue_active"), "l1_hit_ncu": ncu.get("l1_hit"),
                      "l2_hit_ncu": ncu.get("l2_hit")}

    # -- e2e: the same job from HOST buffers through the C ABI ------------------------------------
    out_host = torch.empty((spec["y_res"], spec["x_res"], 4), dtype=torch.float32).pin_memory()
    out_np = out_host.numpy()
    d2h = out_np.nbytes
    h2d = (scene.tris.shape[0] * (48 + 48 + 80) + 262143 * 64 + scene.hdri.texture.data.nbytes + scene.hdri.cdf.nbytes +
           sum(t.data.nbytes for t in scene.textures) + len(scene.materials) * 128 + 64)

    def e2e_step():
        rm.upload(scene)  # flatten + H2D + setupKernel
        rm.render_enqueue(spp)
        if world > 1:
            tview = torch.as_tensor(_DevView(rm.get_pass_device("beauty"), (spec["y_res"], spec["x_res"], 4)), device="cuda")
            buf = tview.clone()
            sharding.reduce_pass(buf, dst=0)  # disjoint row blocks, zeros elsewhere: exact
            if rank == 0:
                out_host.copy_(buf, non_blocking=True)
            torch.cuda.synchronize()
        else:
            rm.get_pass("beauty", out_np)

    e2e_step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t[0])
    e2e_value = P * spp * args.steps / e2e_s / 1e6
    finite = bool(np.isfinite(out_np).all()) if rank == 0 else True
    mean_beauty = float(out_np[..., :3].mean()) if rank == 0 else 0.0
    rm.close()

    # -- CPU baseline (rank 0, N=1 only): the reference's own code on the host cores --------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import refharness as R
        if R.available():
            rs = build_ref_scene(spec, scene.nodes, scene.triIndices)  # tree identical to BVH::build (tests)
            times, threads = time_reference(rs, 8, 1, budget_s=12.0)
            v = P / float(np.mean(times)) / 1e6
            cpu = {"value": v, "unit": "Msamples/s", "cores": threads, "kind": "reference",
                   "sample": f"{len(times)} x 1 spp over the full {spec['x_res']}x{spec['y_res']} frame "
                             f"(reference renderingKernel per pixel index, g++ -O2, {threads} threads)"}
            rs.close()
        else:
            cpu = {"value": None, "unit": "Msamples/s", "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}

    if rank == 0:
        line = {"metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3) if args.warmup else 0, "ms_per_step": ms_per_step,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": spec["desc"], "x_res": spec["x_res"], "y_res": spec["y_res"], "spp": spp,
                           "weak_scaling_by": args.scale_by,
                           "parallelism": f"{world} x pixel row-block shards ({ROWS_PER_BLOCK} rows), scene replicated",
                           "spp_per_launch": min(args.spp_per_launch, spp), "max_depth": 5, "rng_order": "ltr",
                           "l2": "256 MiB flush between timed steps; working set (passes+scene) > L2"},
                "mrays_per_s": value * rays_per_sample, "rays_per_sample": rays_per_sample,
                "roofline": roofline, "roofline_issue": roofline_issue, "roofline_limiting": "sm_issue",
                "cpu_baseline": cpu,
                "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": int(h2d),
                        "d2h_bytes_per_step": int(d2h), "ms_per_step": e2e_s / args.steps * 1e3,
                        "includes": "er_upload_scene (flatten+H2D+setup) + render + beauty pass D2H"
                                    + (" + NCCL reduce to rank 0" if world > 1 else "")},
                "gpu_launches": int(gpu_launches), "clocks": clocks,
                "kernel_ms_per_step": kern_ms_max / args.steps, "wall_ms_per_step": t_wall / args.steps * 1e3,
                "counters_per_sample": {k: v / max(1, counters["samples"]) for k, v in counters.items()},
                "bvh_build_s": t_bvh, "output_finite": finite, "beauty_mean": mean_beauty}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


_JSON_FD = None


def emit(line):
    """the ONE JSON line, on the process's original stdout"""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def own_stdout():
    """libraries print to stdout too (NCCL's version banner at N>1): send everything but the JSON line to
    stderr, so stdout carries exactly one line"""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def main():
    own_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--spp", type=int, default=SPP)
    ap.add_argument("--spp-per-launch", type=int, default=0, help="0 = the whole job in one launch")
    ap.add_argument("--kernel", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--scale-by", default="rows", choices=["rows", "spp"], help="how the job grows with --gpus (weak scaling)")
    args = ap.parse_args()
    if args.spp_per_launch <= 0:
        w = max(1, int(os.environ.get("WORLD_SIZE", "1"))) if args.scale_by == "spp" else 1
        args.spp_per_launch = min(65535, args.spp * w)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        if world != args.gpus:
            log(f"note: --gpus {args.gpus} but WORLD_SIZE={world}; using WORLD_SIZE")
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
