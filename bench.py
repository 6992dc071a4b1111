#!/usr/bin/env python
"""bench.py — BASELINE.json's metric (Msamples/s = paths/s; Mrays/s beside it) through the C ABI.

  python bench.py --gpus N --steps K --warmup W            the CUDA path (one process per GPU under torchrun)
  python bench.py --impl reference --gpus N --steps K ...  the reference's own CPU code (oracle/_ref), rank 0 only

Workload: BASELINE config 3, the one the north-star's target is stated on — 10,000,000-triangle displaced-sphere
field, 1920x1080, 256 spp (`--workload c3`, the default; config 2 = 249,924 triangles, 64 spp is measured as well
and reported in the same line under "secondary").

A step = one whole job of the hot path: setupKernel + `spp` samples for every pixel of the frame
(er_render_reset + er_render_enqueue).  With N GPUs the job is THE SAME frame ("scaling": "strong"): the pixels
are sharded over the ranks (cost-balanced contiguous stripes by default, `--shard rows` = interleaved 8-row
blocks), the scene is replicated, there is no collective in the render loop; at readout every rank copies the
rows it owns over its own PCIe link into one page-locked frame in POSIX shared memory (er_get_pass_owned).
`--scale-by rows|spp` grow the job with N instead (weak scaling, kept for comparison with round 1).
`value` is timed on the device with CUDA events with the scene resident in HBM; `e2e` is the same job through
the C ABI from HOST buffers: scene upload (flatten + H2D), render, beauty pass back to host memory.
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


def job_spec(args, world):
    """the job at this N: strong scaling keeps the workload's own frame and spp"""
    from elevenrender_b200 import workloads
    spec = workloads.workload(args.workload)
    spp = args.spp if args.spp > 0 else spec["spp"]
    if args.scale_by == "rows" and world > 1:  # weak: the same view, N times as tall
        spec["y_res"] *= world
        spec["desc"] += f" [weak scaling: frame {spec['x_res']}x{spec['y_res']}, same view]"
    elif args.scale_by == "spp":  # weak: N times the samples
        spp *= world
    return spec, spp


def job_config(args, spec, spp, world):
    return {"workload": spec["desc"], "x_res": spec["x_res"], "y_res": spec["y_res"], "spp": spp,
            "triangles": int(spec["tris"].shape[0]), "max_depth": 5, "rng_order": "ltr",
            "scaling_mode": "strong (same frame at every N)" if args.scale_by == "none" else f"weak by {args.scale_by}"}


# ---------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the reference's own renderingKernel, per pixel index, on the host cores
# ---------------------------------------------------------------------------------------------------
def build_ref_scene(spec, nodes=None, tri_indices=None):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import common
    return common.to_ref(spec, nodes, tri_indices)


def ref_tree(spec):
    """The reference's BVH::build is single-threaded and takes minutes (and several GB) at 10 M triangles, so
    above 1 M triangles the reference renders over the tree of er_bvh_build, whose nodes[] / triIndices[] are
    checked to be bit-identical to BVH::build's own (tests/test_oracle_cpu.py::test_bvh_builder_identical_c2/_c3,
    tests/golden/bvh_big.npz).  Below that the reference builds its own tree."""
    if spec["tris"].shape[0] <= 1_000_000:
        return None, None, "the reference's own BVH::build"
    from elevenrender_b200 import capi
    nodes, idx = capi.bvh_build(spec["tris"], 0)
    return nodes, idx, "tree from er_bvh_build (bit-identical to BVH::build: tests/golden/bvh_big.npz)"


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
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import refharness as R
    spec, spp = job_spec(args, world)
    if not R.available():
        emit({"impl": "reference", "unavailable": "oracle/_ref/liber_ref.so not built"})
        return
    t0 = time.perf_counter()
    nodes, idx, tree_note = ref_tree(spec)
    rs = build_ref_scene(spec, nodes, idx)
    log(f"reference scene + BVH ready in {time.perf_counter() - t0:.1f} s ({tree_note})")
    times, threads = time_reference(rs, args.steps, min(args.warmup, 1), budget_s=150.0)
    P = spec["x_res"] * spec["y_res"]
    ms = 1e3 * float(np.mean(times))
    v = P / (ms * 1e-3) / 1e6
    sample = (f"{len(times)} steps of 1 spp over the full {spec['x_res']}x{spec['y_res']} frame ({P} renderingKernel calls per "
              f"step; throughput metric, the reference has no per-launch amortisation, src/kernel.cpp:689-700); {tree_note}")
    cfg = job_config(args, spec, spp, world)
    cfg.update({"spp_per_step": 1, "threads": threads})
    line = {"impl": "reference", "metric": "Msamples/s", "value": v, "unit": "Msamples/s", "n_gpus": args.gpus,
            "steps": len(times), "warmup": min(args.warmup, 1), "ms_per_step": ms, "higher_is_better": True,
            "scaling": "strong" if args.scale_by == "none" else "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": cfg,
            "cpu_baseline": {"value": v, "unit": "Msamples/s", "cores": threads, "kind": "reference", "sample": sample},
            "e2e": {"value": v, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


# ---------------------------------------------------------------------------------------------------
# the CUDA arm
# ---------------------------------------------------------------------------------------------------
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


def measure(args, spec, spp, rank, world, local_rank, steps, warmup, tag, with_cpu, with_t6):
    """value / e2e / roofline / counters of one workload at this N. Returns a dict (rank 0) or None."""
    import torch
    import torch.distributed as dist
    from elevenrender_b200 import capi, host, sharding

    def barrier():
        if world > 1:
            dist.barrier()

    P = spec["x_res"] * spec["y_res"]
    spp_per_launch = min(65535, args.spp_per_launch if args.spp_per_launch > 0 else spp)
    from elevenrender_b200 import workloads
    t0 = time.perf_counter()
    scene = workloads.scene_from_spec(spec)
    scene.buildBVH()
    t_bvh = time.perf_counter() - t0
    pars = host.RenderParameters(spec["x_res"], spec["y_res"], spp, device=local_rank, rng_order=host.ER_RNG_LTR,
                                 spp_per_launch=spp_per_launch, kernel=args.kernel)
    # a real (non-default) torch stream: the library launches on it and torch.cuda.Event times it.
    # (handle 0 would mean "use the context's own stream", invisible to events on torch's stream)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    rm = host.RenderingManager(pars, shard=(rank, world, ROWS_PER_BLOCK), stream=stream.cuda_stream)
    rm.upload(scene)

    # -- sharding --------------------------------------------------------------------------------------
    bounds = None

    row_cost_all = [None]  # the cost prepass summed over the ranks (per image row)

    def install_stripe(yb, cost_all):
        """this rank's stripe; with ER_BENCH_BLOCK_ORDER=1 its row blocks are handed out costliest first"""
        rm.set_stripe(yb[rank], yb[rank + 1], ROWS_PER_BLOCK)
        # (measured neutral on the 8-way split of C3, profiles/r04s_shards.txt: slowest stripe 519 -> 480 ms with the
        # plain kernel, sum 3397 -> 3408 ms; no effect on the chunked kernel, whose ring reorders the pixels anyway)
        if cost_all is not None and yb[rank + 1] > yb[rank] and os.environ.get("ER_BENCH_BLOCK_ORDER"):
            rm.set_block_order(capi.order_blocks(cost_all, ROWS_PER_BLOCK, yb[rank], yb[rank + 1]))

    def install_shard(y_res):
        """cost-balanced contiguous stripes: 1-spp counting prepass over every 4th row block (each rank a 1/N
        share), summed over the ranks, split evenly.  Returns the boundaries."""
        if world == 1 or args.shard == "rows":
            return None
        cost = rm.row_cost(block_first=4 * rank, block_stride=4 * world)
        yb, cost_all = sharding.balanced_stripes(cost, world, ROWS_PER_BLOCK, return_cost=True)
        install_stripe(yb, cost_all)
        row_cost_all[0] = cost_all
        return yb

    t0 = time.perf_counter()
    bounds = install_shard(spec["y_res"])
    torch.cuda.synchronize()
    t_balance = time.perf_counter() - t0
    if bounds is None:
        own_pixels = sharding.owned_pixel_count(rank, world, spec["x_res"], spec["y_res"], ROWS_PER_BLOCK)
    else:
        own_pixels = (bounds[rank + 1] - bounds[rank]) * spec["x_res"]

    # -- counters (untimed, count_stats build of the same kernel): work per sample -------------------
    rm.pars.count_stats = True
    rm.reset_counters()
    rm.render_samples(2)
    counters = rm.get_counters()
    rm.pars.count_stats = False
    rm.reset()
    rm.reset_counters()
    if world > 1:  # the job's counters = the sum over the shards
        keys = sorted(counters)
        t = torch.tensor([counters[k] for k in keys], dtype=torch.int64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        counters = {k: int(v) for k, v in zip(keys, t.cpu())}
    a_bytes, a_ops = algorithmic_per_sample(counters, spec, spp_per_launch)
    rays_per_sample = counters["rays"] / max(1, counters["samples"])

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    def step():
        rm.reset()               # setupKernel
        rm.render_enqueue(spp)   # the render launches, no host sync

    # Warm-up. With stripes the warm-up frames also feed the balance back: after each of them the ranks exchange the
    # kernel time their stripe took and move the boundaries (er_rebalance_stripes); the timed steps run on the
    # stripes that result.
    rebalance = None
    for i in range(warmup):
        rm.reset_counters()  # also zeroes the library's own kernel-time accumulation
        step()
        torch.cuda.synchronize()
        if bounds is not None:
            my_ms, _ = rm.kernel_time()
            if rebalance is None:
                rebalance = {"stripes_from_cost_prepass": list(bounds), "steps": []}
            t_all = torch.zeros(world, dtype=torch.float64, device="cuda")
            t_all[rank] = my_ms
            dist.all_reduce(t_all, op=dist.ReduceOp.SUM)
            new_bounds = [int(v) for v in capi.rebalance_stripes(bounds, t_all.cpu().numpy(), ROWS_PER_BLOCK)]
            rebalance["steps"].append({"kernel_ms": [round(float(v), 2) for v in t_all.cpu()], "stripes": new_bounds})
            if new_bounds != list(bounds):
                bounds = new_bounds
                install_stripe(bounds, row_cost_all[0])
                own_pixels = (bounds[rank + 1] - bounds[rank]) * spec["x_res"]
    torch.cuda.synchronize()
    rm.reset_counters()

    sampler = ClockSampler(local_rank)
    sampler.start()
    evs = []
    barrier()
    torch.cuda.synchronize()
    t_wall0 = time.perf_counter()
    for _ in range(steps):
        flush.zero_()  # L2 flush between timed iterations (not timed)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        step()
        e1.record(stream)
        evs.append((e0, e1))
    torch.cuda.synchronize()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.result()
    ms_total = sum(a.elapsed_time(b) for a, b in evs)
    kern_ms, kern_launches = rm.kernel_time()  # events around the render kernels only
    t = torch.tensor([ms_total, kern_ms, -ms_total], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max, kern_ms_max, ms_min = float(t[0]), float(t[1]), -float(t[2])
    ms_per_step = ms_max / steps
    value = P * spp / (ms_per_step * 1e-3) / 1e6
    launches_per_step = kern_launches // steps
    gpu_launches = steps * (1 + launches_per_step)  # k_setup + render kernels

    # -- roofline of the dominant kernel (k_render_wavefront), this rank's launches ---------------------
    hbm_peak, sm_max_mhz, peak_src = measured_peaks()
    samples_per_launch = own_pixels * spp / max(1, launches_per_step)
    launch_s = kern_ms / max(1, kern_launches) * 1e-3
    ach_gbs = a_bytes * samples_per_launch / launch_s / 1e9
    sm_mhz = clocks["sm_mhz"] or sm_max_mhz
    info = capi.device_info(local_rank)
    lanes = info["max_compute_units"] * 4 * 32
    ach_ops = a_ops * samples_per_launch / launch_s
    # DRAM traffic and issue-slot utilisation of the kernel come from ONE separate `ncu --set full` capture
    # (profiles/ncu_traffic.json, per-sample figures); they are NOT measured in this run and are labelled so
    ncu = {}
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(tpath):
        ncu = json.load(open(tpath)).get(args.workload if tag == "primary" else tag, {})
    traffic = ncu["dram_bytes_per_sample"] * samples_per_launch if "dram_bytes_per_sample" in ncu else None
    frac_bytes = ach_gbs / hbm_peak
    frac_ops = ach_ops / (lanes * sm_mhz * 1e6)
    roofline = {"bound": "hbm", "achieved": ach_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": frac_bytes,
                "traffic": traffic, "traffic_unit": "bytes per launch",
                "traffic_source": "from_profile: " + str(ncu.get("source")) if traffic is not None else None,
                "peak_source": peak_src + " (MEASURED_PEAKS.json hbm_gbs)" if peak_src == "measured" else peak_src,
                "kernel": "k_render_wavefront", "algorithmic_bytes_per_sample": a_bytes,
                "samples_per_launch": samples_per_launch, "avg_launch_ms": launch_s * 1e3,
                "note": "algorithmic bytes (SURVEY §8d) of the kernel's ACTUAL visit set; they are served mostly by L1/L2, so "
                        "DRAM traffic is far below achieved x time; see roofline_issue for the instruction-issue view"}
    roofline_issue = {"bound": "sm_issue", "achieved": ach_ops / 1e12, "unit": "T thread-op/s (algorithmic FP32, no FMA)",
                      "peak_at_max_clock": lanes * sm_max_mhz * 1e6 / 1e12,
                      "peak_at_measured_clock": lanes * sm_mhz * 1e6 / 1e12,
                      "frac": frac_ops, "algorithmic_ops_per_sample": a_ops,
                      "from_profile": {"issue_slots_busy": ncu.get("issue_active"), "l1_hit": ncu.get("l1_hit"),
                                       "l2_hit": ncu.get("l2_hit"), "source": ncu.get("source")}}
    limiting = "hbm_bytes(algorithmic)" if frac_bytes >= frac_ops else "sm_issue(algorithmic ops)"

    # -- e2e: the same job from HOST buffers through the C ABI ------------------------------------
    frame = sharding.SharedFrame(f"{os.environ.get('MASTER_PORT', '0')}_{tag}", spec["y_res"], spec["x_res"], rank, barrier) \
        if world > 1 else None
    if world > 1:
        out_np = frame.array
    else:
        out_host = torch.empty((spec["y_res"], spec["x_res"], 4), dtype=torch.float32).pin_memory()
        out_np = out_host.numpy()
    d2h = own_pixels * 16
    h2d = (scene.tris.shape[0] * (48 + 48 + 80) + 262143 * 64 + scene.hdri.texture.data.nbytes + scene.hdri.cdf.nbytes +
           sum(t.data.nbytes for t in scene.textures) + len(scene.materials) * 128 + 64)

    def restripe():
        if bounds is not None:
            install_stripe(bounds, row_cost_all[0])

    readout = []

    upload_t = []

    def e2e_step():
        t_u = time.perf_counter()
        rm.upload(scene)  # flatten + H2D + setupKernel (the shard layout stays installed)
        upload_t.append(time.perf_counter() - t_u)  # host side of the upload (the copies may still be in flight)
        rm.render_enqueue(spp)
        rm.sync()
        t_r = time.perf_counter()
        if world > 1:
            rm.get_pass_owned("beauty", out_np)  # this rank's rows, its own PCIe link, straight into the shared frame
        else:
            rm.get_pass("beauty", out_np)
        readout.append(time.perf_counter() - t_r)

    # the caller's triangle array is page-locked for the end-to-end steps (the contract's "inputs in pinned host
    # memory"): er_upload_scene then sends the records as they are and the device builds its layout (k_flatten_tris)
    pinned_inputs = False
    if not os.environ.get("ER_BENCH_PAGEABLE_INPUTS") and scene.tris.flags["C_CONTIGUOUS"] and scene.tris.dtype == np.float32:
        try:
            capi.host_register(scene.tris)
            pinned_inputs = True
        except capi.ErError:
            pinned_inputs = False
    if pinned_inputs and scene.tris.shape[0] >= 65536:  # what crosses PCIe on the device-flatten path
        h2d += scene.tris.shape[0] * (scene.tris.shape[1] * 4 + 4 - (48 + 48 + 80))
        d2h += scene.tris.shape[0] * 4  # the per-slot slack, for the tree the host builds
    e2e_step()
    e2e_step()  # (pageable inputs: the second upload page-locks the library's own staging buffers, er_api.cu stage_flat)
    torch.cuda.synchronize()
    readout.clear()
    upload_t.clear()
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        e2e_step()
    torch.cuda.synchronize()
    barrier()  # the frame is complete when every rank has written its rows
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s, float(np.mean(readout)), float(np.mean(upload_t))], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s, readout_s, upload_s = float(t[0]), float(t[1]), float(t[2])
    e2e_value = P * spp * steps / e2e_s / 1e6
    if pinned_inputs:
        capi.host_unregister(scene.tris)
    finite = bool(np.isfinite(out_np).all()) if rank == 0 else True
    mean_beauty = float(out_np[..., :3].mean()) if rank == 0 else 0.0
    alpha_ok = bool((out_np[..., 3] == 1.0).all()) if rank == 0 else True  # every pixel written by its owner

    # -- T6 on real GPUs: the N-GPU frame is bit-identical to the frame one GPU renders alone -------------
    t6 = None
    if with_t6 and world > 1:
        w6, h6, spp6 = 480, 272, 6
        scene.x_res, scene.y_res = w6, h6
        rm.pars.sampleTarget = spp6
        rm.upload(scene)
        rm.set_shard(rank, world, ROWS_PER_BLOCK)
        b6 = install_shard(h6)
        f6 = sharding.SharedFrame(f"{os.environ.get('MASTER_PORT', '0')}_t6", h6, w6, rank, barrier)
        if rank == 0:
            f6.array[...] = -1.0
        barrier()
        rm.render_samples(spp6)
        rm.get_pass_owned("beauty", f6.array)
        barrier()
        if rank == 0:
            rm.set_shard(0, 1, ROWS_PER_BLOCK)
            rm.render_samples(spp6)
            alone = rm.get_pass("beauty")
            same = bool(np.array_equal(alone.view(np.uint32), f6.array.view(np.uint32)))
            t6 = {"bit_identical_to_1gpu": same, "frame": f"{w6}x{h6}", "spp": spp6, "stripes": b6,
                  "differing_pixels": int((alone.view(np.uint32) != f6.array.view(np.uint32)).any(axis=-1).sum())}
        barrier()
        f6.close()
        scene.x_res, scene.y_res = spec["x_res"], spec["y_res"]
    if frame is not None:
        frame.close()
    rm.close()

    # -- CPU baseline (rank 0, N=1 only): the reference's own code on the host cores --------------------
    cpu = None
    if with_cpu and rank == 0 and world == 1 and not args.no_cpu_baseline:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import refharness as R
        if R.available():
            big = spec["tris"].shape[0] > 1_000_000
            # the tree of er_bvh_build is bit-identical to BVH::build's (tests/golden/bvh_big.npz); building it
            # again with the reference's single-threaded builder would take minutes at 10 M triangles
            rs = build_ref_scene(spec, scene.nodes, scene.triIndices) if big else build_ref_scene(spec)
            times, threads = time_reference(rs, 8, 1, budget_s=15.0)
            v = P / float(np.mean(times)) / 1e6
            cpu = {"value": v, "unit": "Msamples/s", "cores": threads, "kind": "reference",
                   "sample": f"{len(times)} x 1 spp over the full {spec['x_res']}x{spec['y_res']} frame "
                             f"(reference renderingKernel per pixel index, g++ -O2, {threads} threads; "
                             + ("tree from er_bvh_build, bit-identical to BVH::build)" if big else "the reference's own BVH::build)")}
            rs.close()
        else:
            cpu = {"value": None, "unit": "Msamples/s", "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}

    if rank != 0:
        return None
    return {"value": value, "ms_per_step": ms_per_step, "ms_per_step_fastest_rank": ms_min / steps,
            "mrays_per_s": value * rays_per_sample, "rays_per_sample": rays_per_sample,
            "roofline": roofline, "roofline_issue": roofline_issue, "roofline_limiting": limiting,
            "cpu_baseline": cpu,
            "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h), "ms_per_step": e2e_s / steps * 1e3, "readout_ms": readout_s * 1e3,
                    "upload_call_ms": upload_s * 1e3,
                    "inputs": "triangles in page-locked host memory" if pinned_inputs else "pageable host memory",
                    "includes": "er_upload_scene (H2D + device-side layout + setup) + render + beauty pass D2H"
                                + (" (every rank its own rows into one pinned shared-memory frame)" if world > 1 else "")},
            "gpu_launches": int(gpu_launches), "clocks": clocks,
            "kernel_ms_per_step": kern_ms_max / steps, "wall_ms_per_step": t_wall / steps * 1e3,
            "counters_per_sample": {k: v / max(1, counters["samples"]) for k, v in counters.items()},
            "bvh_build_s": t_bvh, "balance_s": t_balance, "stripes": bounds, "rebalance": rebalance, "spp_per_launch": spp_per_launch,
            "output_finite": finite, "output_alpha_all_one": alpha_ok, "beauty_mean": mean_beauty, "t6": t6}


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    spec, spp = job_spec(args, world)
    warmup = max(args.warmup, 3) if args.warmup else 0
    m = measure(args, spec, spp, rank, world, local_rank, args.steps, warmup, "primary", True, True)
    sec = None
    if args.workload == "c3" and not args.no_secondary:
        # BASELINE config 2 (round 1's headline), same protocol, fewer steps: keeps the round-over-round history
        a2 = argparse.Namespace(**vars(args))
        a2.workload, a2.spp, a2.spp_per_launch = "c2", 0, 0
        spec2, spp2 = job_spec(a2, world)
        m2 = measure(a2, spec2, spp2, rank, world, local_rank, 3, 3, "c2", False, False)
        if rank == 0:
            sec = {"c2": {"config": job_config(a2, spec2, spp2, world), "value": m2["value"], "unit": "Msamples/s",
                          "ms_per_step": m2["ms_per_step"], "mrays_per_s": m2["mrays_per_s"], "e2e": m2["e2e"],
                          "roofline": m2["roofline"], "roofline_issue": m2["roofline_issue"],
                          "roofline_limiting": m2["roofline_limiting"], "counters_per_sample": m2["counters_per_sample"],
                          "stripes": m2["stripes"], "steps": 3, "warmup": 3}}
    if rank == 0:
        cfg = job_config(args, spec, spp, world)
        cfg.update({"parallelism": (f"{world} x pixel shards (" + ("cost-balanced contiguous row stripes, boundaries fed back from the kernel times of the warm-up frames" if m["stripes"] else
                                                                  f"interleaved {ROWS_PER_BLOCK}-row blocks") + "), scene replicated")
                    if world > 1 else "1 GPU",
                    "spp_per_launch": m["spp_per_launch"],
                    "l2": "256 MiB flush between timed steps; working set (passes+scene) > L2"})
        line = {"metric": "Msamples/s", "value": m["value"], "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
                "warmup": warmup, "ms_per_step": m["ms_per_step"], "higher_is_better": True,
                "scaling": "strong" if args.scale_by == "none" else "weak", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic", "config": cfg}
        for k in ("mrays_per_s", "rays_per_sample", "roofline", "roofline_issue", "roofline_limiting", "cpu_baseline", "e2e",
                  "gpu_launches", "clocks", "kernel_ms_per_step", "wall_ms_per_step", "ms_per_step_fastest_rank",
                  "counters_per_sample", "bvh_build_s", "balance_s", "stripes", "rebalance", "output_finite", "output_alpha_all_one",
                  "beauty_mean", "t6"):
            line[k] = m[k]
        if sec is not None:
            line["secondary"] = sec
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
    ap.add_argument("--workload", default="c3", help="c3 = BASELINE config 3 (10 M triangles, 1080p, 256 spp), the north-star's")
    ap.add_argument("--spp", type=int, default=0, help="0 = the workload's own spp")
    ap.add_argument("--spp-per-launch", type=int, default=0, help="0 = the whole job in one launch")
    ap.add_argument("--kernel", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the config-2 block")
    ap.add_argument("--scale-by", default="none", choices=["none", "rows", "spp"],
                    help="none = strong scaling (same frame at every N); rows/spp grow the job with N (weak)")
    ap.add_argument("--shard", default="stripes", choices=["stripes", "rows"],
                    help="N>1: cost-balanced contiguous row stripes, or interleaved 8-row blocks")
    args = ap.parse_args()
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
