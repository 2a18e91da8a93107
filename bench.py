#!/usr/bin/env python
"""bench.py -- BASELINE.json's headline metric on synthetic data.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference|reference-gpu]

A step is one fit iteration of BASELINE.json configs[1]: 1920x1080 frame, 50 000 Cholesky
Gaussians, project -> bin/sort -> sum-rasterise -> L2 loss -> backward -> Adan.  One JSON line
on stdout (rank 0).  `value` = iterations/s with everything resident in HBM (L2 flushed between
timed steps, each step timed with CUDA events on the launching stream); `e2e` = the same through
the public API with the target frame in pinned HOST memory, copied every step, and the loss read
back every step.  N > 1 (torchrun): every rank fits its own GOP's frame, no data-path collective,
weak scaling, value = all ranks' iterations / max-over-ranks time.

--impl reference: the CPU restatement of the reference's kernels (oracle/, C + OpenMP, all host
threads) + numpy Adan on the same workload; rank 0 only.
--impl reference-gpu: the reference's own CUDA kernels (oracle/_ref, compiled from /root/reference
for sm_100a) strung together with torch.cumsum/.item()/torch.sort/gather exactly as the reference's
Python does, same loop, same GPU.  --impl b200-dropin: the same loop on this repo's operators.

The default arm's line also carries, measured in the same run but OUTSIDE the timed regions of
`value` / `e2e` (each a short leg of its own):
  dropin / dropin_graph   C2 through the unchanged operator loop (GaussianSplats_Represent.py:191-207)
                          on this repo's autograd operators, eager and as one CUDA-graph replay per
                          iteration;
  reference_gpu           the same loop on the reference's own kernels (kind "oracle/_ref"), run in a
                          CHILD process so that the product's process never maps the checker;
  ratio_vs_reference_gpu  fused / drop-in / drop-in-graph over that;
  c5_single_gpu           BASELINE config 5 (4K, 200k Gaussians) on one GPU, with its own roofline;
  gop                     BASELINE config 4: the 120-frame sequence with its full iteration budgets,
                          key-frame-delimited GOPs over the N ranks (strong scaling);
  spatial (N > 1)         config 5 split by tile rows over the N ranks, with the multi-GPU parity checks.
--no-extra-legs prints the headline line alone.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "fit_iters_per_s_1080p_50k"
UNIT = "iters/s"
H, W, NPTS = 1080, 1920, 50_000
WORKLOAD = "C2: 1920x1080 synthetic frame, 50k Cholesky Gaussians, rasterize_gaussians_sum fwd+bwd fit loop (L2, Adan)"
L2_FLUSH_BYTES = 256 << 20


def _env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


# ------------------------------------------------------------------------------------------
# clocks during the timed region (B200_PROFILING.md recipe)
# ------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i",
                 str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------
# algorithmic bytes per launch of each stage (DESIGN.md "Kernels and rooflines")
# ------------------------------------------------------------------------------------------
def algorithmic_bytes(stage: str, n: int, I: int, hw: int, tiles: int) -> float:
    return {
        "project_fwd": 20 * n + 32 * n,
        "scan_map": 4 * n + 12 * n + 8 * I,
        "sort": 2 * 8 * I,                                   # one digit pass: read + write (key, id)
        "edges_pack": 8 * I + 36 * I + 48 * I + 8 * tiles,   # sorted pairs + gathered params -> packed + bins
        "raster_fwd": 48 * I + 8 * tiles + 12 * hw + 4 * hw,
        "raster_bwd": 48 * I + 16 * hw + 36 * I,
        "project_bwd": 12 * n + 4 * n + 12 * n + 20 * n + 32 * n,
        "fit_tiles": 48 * I + 8 * tiles + 1.5 * hw + 36 * I,   # records + ranges + the I420 frame + one gradient row per pair
        "adan": 9 * n * 4 * 9,
        "memset": 0, "other": 0,
    }[stage]


def ncu_capture(stage: str, key: str = "dram_traffic_bytes"):
    """one number per launch of the stage's kernel from the committed `ncu --set full` capture
    (profiles/r1_fit_kernels_ncu.json; taken with this same bench command), or None.  Default:
    dram__bytes_read.sum + dram__bytes_write.sum"""
    kern = {"fit_tiles": "k_fit_tiles", "adan": "k_fit_update", "edges_pack": "k_fit_scatter"}.get(stage)
    recs = None
    for name in ("r2_fit_kernels_ncu.json", "r1_fit_kernels_ncu.json"):     # the newest committed capture
        try:
            recs = json.load(open(os.path.join(ROOT, "profiles", name)))
            break
        except (OSError, ValueError):
            continue
    if recs is None:
        return None
    for r in recs:
        if kern and kern in r.get("kernel", ""):
            return r.get(key)
    return None


# ------------------------------------------------------------------------------------------
# CPU arm: the oracle port on the host cores
# ------------------------------------------------------------------------------------------
def raw_frame(seed=1):
    """the synthetic frame as the reference's input file would hold it: planar I420, uint8 [H*3/2, W]"""
    import torch
    from gsvc_b200 import synth, yuv
    rgb = (synth.target_frame(H, W, seed=seed)[0] * 255).round().to(torch.uint8).permute(1, 2, 0).contiguous()
    return yuv.rgb_to_i420(rgb).contiguous()


def make_cpu_fit(seed=1):
    import numpy as np
    from gsvc_b200 import synth
    from oracle import yuv as oracle_yuv
    from oracle.fit_cpu import CpuFit
    g = synth.gaussians(NPTS, seed=seed)
    # the same frame every arm fits: the raw I420 frame decoded as the reference's loader does
    gt = oracle_yuv.i420_to_rgb(raw_frame(seed).numpy(), H, W).astype(np.float32) / np.float32(255)
    return CpuFit(H, W, {k: g[k].numpy() for k in ("_xyz", "_cholesky", "_features_dc")}, gt)


def run_cpu(steps: int, warmup: int) -> dict:
    from oracle import oracle as _O
    try:   # every host core this process may run on (torchrun exports OMP_NUM_THREADS=1 to its ranks)
        _O.set_num_threads(len(os.sched_getaffinity(0)))
    except (AttributeError, OSError):
        pass
    fit = make_cpu_fit()
    init_isect = None
    for _ in range(warmup):
        fit.step()
        if init_isect is None:
            init_isect = int(fit.last["num_intersects"])
    t0 = time.perf_counter()
    for _ in range(steps):
        fit.step()
    dt = time.perf_counter() - t0
    return {"value": steps / dt, "unit": UNIT, "cores": fit.cores, "kind": "port", "intersections_at_init": init_isect,
            "sample": f"{steps} full-size fit iterations ({WORKLOAD.split(':')[0]}) after {warmup} warm-up, "
                      f"oracle C kernels with OpenMP on {fit.cores} threads + numpy Adan",
            "seconds": dt}


def main_reference(args, rank: int):
    if rank != 0:
        return
    steps, warmup = min(args.steps, 40), max(args.warmup, 3)      # the warm-up the driver asks for (>= 3, as the GPU arm)
    r = run_cpu(steps, warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": 1e3 / r["value"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": base_config(r["intersections_at_init"], f"host cpu, {r['cores']} threads: oracle C kernels (OpenMP) + numpy Adan", "none (host arm)",
                              "rank 0 only", "not applicable (host arm)"),
        "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------
# GPU arms
# ------------------------------------------------------------------------------------------
def base_config(intersections, path: str, e2e_path: str, parallelism: str, l2: str) -> dict:
    """the `config` object: the same keys on every arm, so that the driver's same_config check compares
    like with like"""
    return {"workload": WORKLOAD, "H": H, "W": W, "gaussians": NPTS, "intersections": intersections, "l2": l2,
            "path": path, "e2e_path": e2e_path, "parallelism": parallelism}


def make_fitter(impl: str, dev, seed: int, trained_like: bool = False, slots: bool = True):
    import torch
    if trained_like:
        from gsvc_b200 import synth
        from gsvc_b200.fit import FusedFrameFit
        return FusedFrameFit(H, W, NPTS, dev, init=synth.gaussians(NPTS, seed=seed, trained_like=True), cap=1 << 22,
                             slots=slots)
    if impl == "reference-gpu":
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import ref_fit
        return ref_fit.RefFrameFit(H, W, NPTS, dev, seed=seed)
    from gsvc_b200.fit import FrameFit, FusedFrameFit
    if impl == "b200-dropin":
        return FrameFit(H, W, NPTS, dev, seed=seed)
    return FusedFrameFit(H, W, NPTS, dev, seed=seed, slots=slots)


# ------------------------------------------------------------------------------------------
# extra legs of the default arm (outside the timed regions of `value` / `e2e`)
# ------------------------------------------------------------------------------------------
def _events(n):
    import torch
    return [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]


def operator_loop_leg(dev, seed: int, K: int, Wm: int, gt, gt_host, flush, graphed: bool) -> dict:
    """C2 through the UNCHANGED operator loop -- GaussianVideo_frame.train_iter of the reference
    (GaussianSplats_Represent.py:191-207: tanh / bound, project_gaussians_2d, rasterize_gaussians_sum,
    clamp, permute, mse_loss, backward, loss.item(), Adan step, zero_grad, StepLR) on this repo's
    autograd operators.  graphed: the same iteration as one CUDA-graph replay (the operators never
    synchronise the host, so torch can capture them; the reference's cannot be captured)."""
    import torch
    from gsvc_b200.fit import FrameFit
    fit = FrameFit(H, W, NPTS, dev, seed=seed)
    step = fit.train_iter_graphed if graphed else fit.train_iter
    for _ in range(Wm + (1 if graphed else 0)):
        step(gt)
    torch.cuda.synchronize(dev)
    ev = _events(K)
    for a, b in ev:
        flush.zero_()
        a.record()
        step(gt)
        b.record()
    torch.cuda.synchronize(dev)
    ms = sum(a.elapsed_time(b) for a, b in ev) / K
    a, b = _events(1)[0]
    a.record()
    for _ in range(K):
        g = gt_host.to(dev, non_blocking=True)
        _loss, psnr = step(g)
    b.record()
    torch.cuda.synchronize(dev)
    e2e_ms = a.elapsed_time(b) / K
    render = fit.forward_graphed if graphed else fit.forward
    with torch.no_grad():
        for _ in range(3):
            render()
        torch.cuda.synchronize(dev)
        a.record()
        for _ in range(K):
            render()
        b.record()
        torch.cuda.synchronize(dev)
    render_ms = a.elapsed_time(b) / K
    return {"value": 1e3 / ms, "unit": UNIT, "ms_per_step": ms,
            "e2e": {"value": 1e3 / e2e_ms, "unit": UNIT, "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": gt_host.numel() * gt_host.element_size(), "d2h_bytes_per_step": 4},
            "render": {"value": 1e3 / render_ms, "unit": "frames/s", "ms_per_frame": render_ms},
            "steps": K, "final_psnr": psnr, "path": fit.path_name + (" -- one CUDA-graph replay per iteration" if graphed else ""),
            "loss_read_back_every_step": True}


def reference_gpu_leg(K: int, Wm: int) -> dict:
    """the reference's own kernels (oracle/_ref) in the same loop, in a CHILD process: the checker is
    never mapped into the process whose numbers are the product's"""
    so = os.path.join(ROOT, "oracle", "_ref", "libgsplat_ref.so")
    if not os.path.exists(so):
        return {"unavailable": "oracle/_ref/libgsplat_ref.so not built (oracle/build_ref.sh needs /root/reference)"}
    cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference-gpu", "--steps", str(K), "--warmup", str(Wm),
           "--no-cpu", "--no-extra-legs"]
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env={**os.environ, "WORLD_SIZE": "1", "RANK": "0",
                                                                                   "LOCAL_RANK": os.environ.get("LOCAL_RANK", "0")})
        d = json.loads(r.stdout.strip().splitlines()[-1])
    except (subprocess.TimeoutExpired, ValueError, IndexError) as e:
        return {"unavailable": f"child run failed: {type(e).__name__}"}
    return {"kind": "oracle/_ref", "value": d["value"], "unit": UNIT, "ms_per_step": d["ms_per_step"],
            "e2e": {k: d["e2e"][k] for k in ("value", "unit", "ms_per_step")}, "render": d["render"], "steps": K,
            "path": d["config"]["path"], "process": "child (python bench.py --impl reference-gpu)"}


def c5_leg(dev, K: int, Wm: int, flush, peak_gbs: float) -> dict:
    """BASELINE config 5 on ONE GPU: 3840x2160, 200k Gaussians, the fused step"""
    import torch
    from gsvc_b200 import _lib, synth
    from gsvc_b200.fit import FusedFrameFit
    h, w, n = 2160, 3840, 200_000
    fit = FusedFrameFit(h, w, n, dev, seed=1)
    gt = (synth.target_frame(h, w)[0] * 255).round().to(torch.uint8).permute(1, 2, 0).contiguous().to(dev)
    fit.train_iters(gt, Wm)
    torch.cuda.synchronize(dev)
    ev = _events(K)
    for a, b in ev:
        flush.zero_()
        a.record()
        fit.train_iters(gt, 1)
        b.record()
    torch.cuda.synchronize(dev)
    ms = sum(a.elapsed_time(b) for a, b in ev) / K
    _lib.prof_enable(True)
    for _ in range(K):
        flush.zero_()
        fit.train_iters(gt, 1)
    torch.cuda.synchronize(dev)
    _lib.prof_enable(False)
    prof = _lib.prof_read()
    stages = {k: {"ms_per_step": v[0] / K, "launches_per_step": v[1] / K} for k, v in prof.items() if v[1]}
    I, _of, mx = fit.status()
    tiles = ((w + 15) // 16) * ((h + 15) // 16)
    out = {"workload": "C5 on one GPU: 3840x2160 synthetic frame, 200k Cholesky Gaussians, fused fit step",
           "ms_per_iter": ms, "iters_per_s": 1e3 / ms, "steps": K, "intersections": I, "max_per_tile": mx, "stages": stages,
           "final_psnr": 10 * math.log10(1.0 / max(fit.loss(), 1e-20))}
    if "fit_tiles" in stages:
        alg = algorithmic_bytes("fit_tiles", n, I, h * w, tiles)
        dur = stages["fit_tiles"]["ms_per_step"] * 1e-3
        out["roofline"] = {"kernel": "fit_tiles", "bound": "hbm", "achieved": alg / dur / 1e9, "peak": peak_gbs, "unit": "GB/s",
                           "frac": alg / dur / 1e9 / peak_gbs, "algorithmic_bytes": alg, "avg_launch_us": dur * 1e6}
    return out


def bind_to_gpu_cores(dev, local_rank: int, world: int) -> dict:
    """Pin this rank to host cores near its GPU BEFORE any pinned memory is allocated (first touch decides the
    NUMA node of the pages the copy engine will read): the GPU's PCI device in sysfs names its NUMA node and
    the node's cpu list; the ranks that share a node split that list evenly.  -> what was done, for the line."""
    import torch
    info = {"numa_node": None, "cpus": None, "bound": False}
    try:
        pr = torch.cuda.get_device_properties(dev)
        bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        base = f"/sys/bus/pci/devices/{bdf}"
        node = int(open(base + "/numa_node").read().strip())
        cpulist = open(base + "/local_cpulist").read().strip()
        cpus = []
        for part in cpulist.split(","):
            a, _, b = part.partition("-")
            cpus.extend(range(int(a), int(b or a) + 1))
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
        info.update(numa_node=node, pci=bdf, node_cpus=len(cpus), allowed_cpus=len(allowed))
        if allowed and world > 1:
            # ranks are spread over the node's cpus by their local index (every rank computes the same split)
            share = max(1, len(allowed) // max(world, 1))
            mine = allowed[(local_rank * share) % len(allowed):][:share] or allowed
            os.sched_setaffinity(0, mine)
            info.update(cpus=f"{mine[0]}-{mine[-1]}", bound=True)
    except (OSError, ValueError, AttributeError) as e:
        info["error"] = f"{type(e).__name__}: {e}"[:120]
    return info


def main_gpu(args, rank: int, world: int, local_rank: int):
    import torch
    import torch.distributed as dist
    from gsvc_b200 import _lib, synth

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    host_binding = bind_to_gpu_cores(dev, local_rank, world)      # before the first pin_memory()
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    mine = args.impl in ("b200", "b200-dropin")
    fused = args.impl == "b200"
    fit = None
    raw = raw_frame(seed=1 + rank)                                       # uint8 [H*3/2, W] planar I420
    if fused:
        # the fused path takes the raw frame of the .yuv file and decodes it inside the tile kernel
        gt_host = raw.pin_memory()
        step_resident = lambda g: fit.train_iters(g, 1)                  # nothing read back
    else:
        # fp32 NCHW in [0,1]: what the reference's loader (cv2 decode + ToTensor) hands its loop
        from gsvc_b200 import yuv as _yuv
        rgb = _yuv.i420_to_rgb(raw.to(dev), H, W)
        gt_host = (rgb.permute(2, 0, 1).float() / 255).unsqueeze(0).contiguous().cpu().pin_memory()
        step_resident = lambda g: fit.train_iter(g)                                   # reads the loss back every step, as the reference
    gt = gt_host.to(dev)
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=dev)
    K, Wm = args.steps, max(args.warmup, 3)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    init_isect = [None]

    def fresh_model(stream_warmup: bool = False):
        """every timed phase starts from the same state: the model as initialised + Wm warm-up
        iterations, so `value`, `l2_warm`, `e2e` and the stage pass time the same K iterations.
        stream_warmup: the warm-up goes through fit_stream (the e2e path: its frame pipe then exists)"""
        nonlocal fit
        fit = make_fitter(args.impl, dev, seed=1 + rank, trained_like=args.trained_like, slots=not args.ranges)
        if stream_warmup:
            for _ in fit.fit_stream(gt_host for _ in range(Wm)):
                pass
        else:
            for w_ in range(Wm):
                step_resident(gt)
                if w_ == 0 and init_isect[0] is None and hasattr(fit, "status"):
                    init_isect[0] = fit.status()[0]      # intersections of the model as initialised (step 1)
        barrier()

    fresh_model()

    # ---- value: HBM-resident inputs, L2 flushed between steps, per-step CUDA events ------------
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    launches0 = _lib.launch_count() if mine else 0
    barrier()
    wall0 = time.perf_counter()
    for a, b in ev:
        flush.zero_()
        a.record()
        step_resident(gt)
        b.record()
    barrier()
    wall = time.perf_counter() - wall0
    launches = (_lib.launch_count() - launches0) if mine else 0
    step_ms = sum(a.elapsed_time(b) for a, b in ev) / K

    # ---- same loop back to back (L2 warm: what a real fit loop sees), one event pair -----------
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    fresh_model()
    a.record()
    for _ in range(K):
        step_resident(gt)
    b.record()
    barrier()
    warm_ms = a.elapsed_time(b) / K

    # ---- e2e: target frame in pinned host memory every step, loss read back every step -------
    fresh_model(stream_warmup=fused)
    a.record()
    psnr = 0.0
    if fused:
        # public streaming API: every step takes its frame from pinned host memory (the copy of
        # frame i+1 overlaps step i on a side stream) and hands its loss back before the next one
        for _loss, psnr in fit.fit_stream(gt_host for _ in range(K)):
            pass
    else:
        for _ in range(K):
            g = gt_host.to(dev, non_blocking=True)
            _loss, psnr = fit.train_iter(g)        # reads the loss back (psnr), like the reference loop
    b.record()
    barrier()
    e2e_ms = a.elapsed_time(b) / K
    clocks = sampler.stop() if rank == 0 else None

    # ---- what the host link gives this frame alone (explains e2e; not part of any metric) ----
    barrier()
    a.record()
    for _ in range(20):
        gt.copy_(gt_host, non_blocking=True)
    b.record()
    barrier()
    h2d_ms = a.elapsed_time(b) / 20

    # ---- render-only fps (config 3), device-resident -----------------------------------------
    with torch.no_grad():
        for _ in range(3):
            fit.forward()
        barrier()
        a.record()
        for _ in range(K):
            fit.forward()
        b.record()
        barrier()
    render_ms = a.elapsed_time(b) / K

    # ---- config 3 as BASELINE.json words it: render from the QUANTISED Gaussians (wire format) ----
    renderq_ms = None
    if fused:
        from gsvc_b200.decode import QuantizedFrame
        books = torch.rand(2, 8, 3, generator=torch.Generator().manual_seed(1)) * torch.tensor([1.0, 0.3]).view(2, 1, 1)
        qf = QuantizedFrame.from_state(fit.frame_state(), books, device=dev)
        with torch.no_grad():
            for _ in range(3):
                qf.render(H, W)
            barrier()
            a.record()
            for _ in range(K):
                qf.render(H, W)
            b.record()
            barrier()
        renderq_ms = a.elapsed_time(b) / K

    # ---- per-stage timing pass (CUDA events inside the library, on the launching stream) ------
    stages = {}
    if mine:
        fresh_model()
        _lib.prof_enable(True)
        for _ in range(K):
            if not args.warm_stages:
                flush.zero_()
            step_resident(gt)
        torch.cuda.synchronize(dev)
        _lib.prof_enable(False)
        prof = _lib.prof_read()
        stages = {k: {"ms_per_step": v[0] / K, "launches_per_step": v[1] / K} for k, v in prof.items() if v[1]}

    # ---- per-rank figures that explain e2e at N > 1: each rank's own copy bandwidth and e2e time ----------
    per_rank = None
    if world > 1:
        mine_t = torch.tensor([h2d_ms, e2e_ms, float(host_binding.get("numa_node") if host_binding.get("numa_node") is not None else -1)],
                              dtype=torch.float64, device=dev)
        allr = [torch.zeros_like(mine_t) for _ in range(world)]
        dist.all_gather(allr, mine_t)
        fb = gt_host.numel() * gt_host.element_size()
        per_rank = {"h2d_copy_alone_gbs": [round(fb / (float(x[0]) * 1e-3) / 1e9, 1) for x in allr],
                    "e2e_ms_per_step": [round(float(x[1]), 4) for x in allr],
                    "numa_node": [int(x[2]) for x in allr]}

    # ---- max over ranks ----------------------------------------------------------------------
    t = torch.tensor([step_ms, warm_ms, e2e_ms, render_ms, renderq_ms or 0.0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    step_ms, warm_ms, e2e_ms, render_ms, renderq_ms = (float(x) for x in t.tolist())

    if rank == 0:
        I = int(getattr(fit, "last_num_intersects", 0)) or None
        line = {
            "metric": METRIC, "value": world * 1e3 / step_ms, "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": Wm, "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": base_config(
                init_isect[0] or I, getattr(fit, "path_name", "autograd drop-in ops + torch glue"),
                "FusedFrameFit.fit_stream: pinned-host raw I420 frame per step (decoded in the tile kernel), copy overlapped with the previous step, loss reported into pinned memory" if fused else "train_iter per step: H2D of the fp32 frame, loss.item()",
                f"{world} independent GOP chains, one per GPU, no collective",
                f"flushed between timed steps ({L2_FLUSH_BYTES >> 20} MiB memset outside the event pair)"),
            "clocks": clocks,
            "e2e": {"value": world * 1e3 / e2e_ms, "unit": UNIT, "h2d_bytes_per_step": gt_host.numel() * gt_host.element_size(),
                    "d2h_bytes_per_step": 24 if fused else 4, "ms_per_step": e2e_ms,
                    "h2d_copy_alone_ms": h2d_ms,
                    "h2d_copy_alone_gbs": gt_host.numel() * gt_host.element_size() / (h2d_ms * 1e-3) / 1e9},
            "host_binding": host_binding, "per_rank": per_rank,
            "gpu_launches": launches,
            "l2_warm": {"value": world * 1e3 / warm_ms, "unit": UNIT, "ms_per_step": warm_ms},
            "render": {"value": world * 1e3 / render_ms, "unit": "frames/s", "ms_per_frame": render_ms},
            "wall_s_timed_region": wall, "final_psnr": psnr,
        }
        if renderq_ms:
            line["render_quantized"] = {"value": world * 1e3 / renderq_ms, "unit": "frames/s", "ms_per_frame": renderq_ms,
                                        "bytes_per_gaussian": 9,
                                        "what": "C3: gs2d_render_quantized from fp16 means + 6-bit Cholesky codes + 2x3-bit VQ colour indices"}
        if args.impl != "b200":
            line["impl"] = args.impl
        if args.trained_like:
            line["config"]["workload"] += " [trained-like init: informational, not the headline config]"
        if hasattr(fit, "status"):
            # (beside `config`, not in it: `config` carries the same keys on every arm, base_config)
            line["workload_detail"] = {"max_per_tile": fit.status()[2], "intersections_last_step": I}
            line["workload_detail"]["target"] = ("resident frame converted once into the tile kernel's order (gs2d_fit_target_prepare), "
                                        "12 B/pixel read per iteration; e2e: raw I420 decoded in the tile kernel"
                                        if getattr(fit, "prepare_targets", False) else "raw frame decoded in the tile kernel")
        if stages:
            line["stages"] = stages
            kern = {k: v for k, v in stages.items() if k != "memset"}
            top = max(kern, key=lambda k: kern[k]["ms_per_step"])
            peaks = {}
            try:
                peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            except (OSError, ValueError):
                pass
            peak = float(peaks.get("hbm_gbs", 6650.0))
            n_l = max(kern[top]["launches_per_step"], 1e-9)
            tiles = ((W + 15) // 16) * ((H + 15) // 16)
            alg = algorithmic_bytes(top, NPTS, I or 124_000, H * W, tiles)
            dur_s = kern[top]["ms_per_step"] / n_l * 1e-3
            line["roofline"] = {"kernel": top, "bound": "hbm", "achieved": alg / dur_s / 1e9, "peak": peak,
                                "unit": "GB/s", "frac": alg / dur_s / 1e9 / peak, "traffic": ncu_capture(top),
                                "algorithmic_bytes": alg, "avg_launch_us": dur_s * 1e6,
                                "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured)" if peaks else "fallback 6650",
                                "traffic_note": "algorithmic bytes count the frame at its information content (I420, 1.5 B/pixel); "
                                                "the resident frame is deliberately kept as prepared fp32 (12 B/pixel = 24.9 MB "
                                                "of the measured traffic) to take its decode out of every iteration: "
                                                "bytes traded for instructions on a kernel at ~10 % of HBM bandwidth"
                                if (top == "fit_tiles" and getattr(fit, "prepare_targets", False)) else None}
            # what actually bounds this kernel: warp instructions issued per second
            # against 4 schedulers x SMs x clock -- instruction count from the committed ncu capture, time live
            inst = ncu_capture(top, "smsp__inst_executed.sum")
            if inst and clocks and clocks.get("sm_mhz"):
                sms = torch.cuda.get_device_properties(dev).multi_processor_count
                issue_peak = 4.0 * sms * clocks["sm_mhz"] * 1e6
                line["roofline"]["issue"] = {"warp_instructions": inst, "achieved_ginst_s": inst / dur_s / 1e9,
                                             "peak_ginst_s": issue_peak / 1e9, "frac": inst / dur_s / issue_peak}
            # shared-memory companion (north_star: "fraction of the HBM and shared-memory rooflines"): LSU
            # wavefronts to shared memory (committed ncu capture) per second against one wavefront per
            # clock per SM
            wf = ncu_capture(top, "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum")
            if wf and clocks and clocks.get("sm_mhz"):
                sms = torch.cuda.get_device_properties(dev).multi_processor_count
                smem_peak = sms * clocks["sm_mhz"] * 1e6
                line["roofline"]["smem"] = {"wavefronts": wf, "achieved_gwavefronts_s": wf / dur_s / 1e9,
                                            "peak_gwavefronts_s": smem_peak / 1e9, "frac": wf / dur_s / smem_peak,
                                            "bytes_at_128_per_wavefront": wf * 128}
        if world == 1 and mine and not args.no_cpu:
            r = run_cpu(args.cpu_steps, 1)
            line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}

    # ---- extra legs: short runs of their own, after and outside every timed region above ----------------
    extra = fused and not args.no_extra_legs and not args.trained_like and not args.ranges
    if extra and rank == 0 and world == 1:
        Kx = max(3, min(K, args.leg_steps))
        del fit
        torch.cuda.empty_cache()
        from gsvc_b200 import yuv as _yuv
        rgb = _yuv.i420_to_rgb(raw.to(dev), H, W)
        gt_f32_host = (rgb.permute(2, 0, 1).float() / 255).unsqueeze(0).contiguous().cpu().pin_memory()
        gt_f32 = gt_f32_host.to(dev)
        line["dropin"] = operator_loop_leg(dev, 1, Kx, Wm, gt_f32, gt_f32_host, flush, graphed=False)
        line["dropin_graph"] = operator_loop_leg(dev, 1, Kx, Wm, gt_f32, gt_f32_host, flush, graphed=True)
        del gt_f32
        torch.cuda.synchronize(dev)
        line["reference_gpu"] = reference_gpu_leg(Kx, Wm)
        ref = line["reference_gpu"]
        if "value" in ref:
            line["ratio_vs_reference_gpu"] = {
                name: {"value": d["value"] / ref["value"], "e2e": d["e2e"]["value"] / ref["e2e"]["value"],
                       "render": d["render"]["value"] / ref["render"]["value"]}
                for name, d in (("fused", line), ("dropin", line["dropin"]), ("dropin_graph", line["dropin_graph"]))}
        peak = line.get("roofline", {}).get("peak", 6650.0)
        line["c5_single_gpu"] = c5_leg(dev, Kx, Wm, flush, peak)
    if extra and world > 1:
        barrier()
        sp = run_spatial(args, rank, world, dev, flush)
        if rank == 0:
            line["spatial"] = sp
    if extra and not args.no_gop_leg:
        barrier()
        gather_group = dist.new_group(backend="gloo") if world > 1 else None
        g = run_gop(args, rank, world, dev, gather_group, frames=args.gop_frames, cut=15, chains=1, detect=False)
        if rank == 0:
            line["gop"] = g
        if world == 1:
            g4 = run_gop(args, rank, world, dev, gather_group, frames=args.gop_frames, cut=15, chains=4, detect=False)
            line["gop_concurrent_chains"] = g4
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_spatial(args, rank: int, world: int, dev, flush):
    """BASELINE config 5 over N ranks: the 4K frame split by tile rows (balanced by intersections per
    row), gradient rows reduced and the update sharded inside the update kernel over NVLink
    (SpatialFrameFit, exchange='sharded').  Rank 0 also fits the frame alone; the multi-GPU parity checks
    of tests/run_spatial_multi_gpu.py run here: parameters bit-identical across ranks, loss equal to the
    single-GPU run.  Timing: CUDA events, max over ranks."""
    import torch
    import torch.distributed as dist
    from gsvc_b200 import synth
    from gsvc_b200.fit import FusedFrameFit, SpatialFrameFit
    h, w, n = 2160, 3840, 200_000
    K, Wm = max(10, min(args.steps, 50)), max(args.warmup, 5)
    init = synth.gaussians(n, seed=1)
    gt = (synth.target_frame(h, w)[0] * 255).round().to(torch.uint8).permute(1, 2, 0).contiguous().to(dev)
    out = {"workload": f"C5: 3840x2160 synthetic frame, 200k Cholesky Gaussians, tile-row split over {world} GPUs",
           "world": world, "steps": K}
    # sharded: reduce-scatter of the gradient rows, Adan on the rank's shard and all-gather of the parameters
    # inside the update kernel over NVLink, every rank scatters its band's records from its local copy;
    # sharded_remote_scatter: the update kernel also writes the next step's records straight into the owning
    # rank's slots (an all-to-all by NVLink atomics + stores: measured slower, kept as an option);
    # nccl: replicated update behind one NCCL all-reduce of the gradient rows
    # (sharded_pull: the update kernel reads every rank's gradient rows instead of the tile kernels pushing them)
    def exchange_bytes(exchange, fit):
        """bytes a rank moves over NVLink per step (algorithmic)"""
        if exchange == "nccl":
            return n * 48 * 2                                          # all-reduce: reduce-scatter + all-gather
        gather = (n // world) * 32 * (world - 1)                      # the shard's new rows stored into every peer
        if getattr(fit, "push", False):                                # pushed: 32 bytes per pair of the band, remote owners
            return int(fit.status()[0] * 32 * (world - 1) / world) + gather
        return (n // world) * 48 * (world - 1) + gather                # pulled: the shard's rows read from every peer

    # "sharded" is the default a user gets (pushed reduction); the pulled one runs beside it
    variants = ("sharded", "sharded_pull", "sharded_remote_scatter", "nccl")
    for exchange in variants:
        fit = SpatialFrameFit(h, w, n, dev, init=init, exchange=exchange.split("_")[0], balance=True,
                              remote_scatter=exchange == "sharded_remote_scatter",
                              push={"sharded_pull": False, "sharded_push": True}.get(exchange))
        fit.train_iters(gt, Wm)
        dist.barrier(); torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fit.train_iters(gt, K)
        e1.record()
        dist.barrier(); torch.cuda.synchronize(dev)
        ms = torch.tensor([e0.elapsed_time(e1) / K], dtype=torch.float64, device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        loss = fit.loss()
        chk = torch.stack([fit._params8.double().sum(), fit._params8.double().abs().sum()])
        lo, hi = chk.clone(), chk.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        out[exchange] = {"ms_per_iter": float(ms), "iters_per_s": 1e3 / float(ms), "loss": loss, "bands": fit.bands,
                         "params_identical_across_ranks": bool(torch.equal(lo, hi)),
                         "reduction": ("push" if fit.push else "pull") if exchange.startswith("sharded") else "all-reduce",
                         "exchange_bytes_per_rank_per_iter": exchange_bytes(exchange, fit)}
        if exchange == "sharded":
            keep = fit
        else:
            del fit
    if rank == 0:
        ref = FusedFrameFit(h, w, n, dev, init=init)
        ref.train_iters(gt, Wm)
        torch.cuda.synchronize(dev)
        e0.record()
        ref.train_iters(gt, K)
        e1.record()
        torch.cuda.synchronize(dev)
        ref_ms = e0.elapsed_time(e1) / K
        ref_loss = ref.loss()
        d = (ref._params8 - keep._params8).abs()
        out["single_gpu_ms_per_iter"] = ref_ms
        out["loss_single_gpu"] = ref_loss
        for exchange in variants:
            out[exchange]["speedup_vs_single_gpu"] = ref_ms / out[exchange]["ms_per_iter"]
            out[exchange]["loss_rel_diff_vs_single_gpu"] = abs(out[exchange]["loss"] - ref_loss) / ref_loss
        out["median_abs_param_diff_vs_single_gpu"] = float(d.median())
        out["parity_ok"] = bool(all(out[x]["params_identical_across_ranks"] and out[x]["loss_rel_diff_vs_single_gpu"] <= 2e-3
                                    for x in variants))
    dist.barrier()
    return out if rank == 0 else None


# ------------------------------------------------------------------------------------------
# --workload gop: BASELINE config 4 (not the headline line; run explicitly)
# ------------------------------------------------------------------------------------------
def run_gop(args, rank: int, world: int, dev, gather_group, frames: int, cut: int, chains: int, detect: bool):
    """BASELINE config 4: a `frames`-frame 1080p sequence with a scene cut every `cut` frames, key frames
    fitted from scratch with the removal control, predicted frames from the previous frame with random
    add + prune (train_video_Represent.py:357-367: a key frame resets the chain); key-frame-delimited
    GOPs sharded over the ranks with no data-path collective (gsvc_b200.gop).  Strong scaling: value =
    frames of the whole sequence / max-over-ranks device time.  Frames are pre-decoded into pinned host
    memory per rank (raw I420); the timed region holds every frame's H2D copy, fit, and the final gather
    of the per-frame Gaussians to rank 0.  -> the leg's dict on rank 0, None elsewhere."""
    import torch
    import torch.distributed as dist
    from gsvc_b200 import _lib, gop

    F = frames
    frame_fn, keyframes = gop.synthetic_sequence(H, W, F, cut_every=cut)
    prepass = None
    if detect:
        # the reference's key-frame detection pre-pass (train_video_Represent.py:312-355): 5000-point
        # probe fits, 500 iterations fresh / 100 from the previous frame, frames sharded with a halo
        from gsvc_b200 import yuv as _y
        from gsvc_b200.fit import FusedFrameFit
        edges = [round(r * F / world) for r in range(world + 1)]
        lo_f, hi_f = max(edges[rank] - 1, 0), edges[rank + 1]
        probe_host = {f: _y.rgb_to_i420((frame_fn(f)[0] * 255).round().to(torch.uint8).permute(1, 2, 0).contiguous())
                      .contiguous().pin_memory() for f in range(lo_f, hi_f)}
        probe = lambda prev: FusedFrameFit(H, W, 5000, dev, init=prev) if prev is not None else FusedFrameFit(H, W, 5000, dev, seed=int(torch.initial_seed() % (1 << 30)))
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        scores = gop.keyframe_scores(lambda f: probe_host[f].to(dev, non_blocking=True), F, probe, 500, 100,
                                     group=gather_group)
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        dt = time.perf_counter() - t0
        found = [sorted(set(gop.keyframes_from_scores(scores)))] if rank == 0 else [None]
        if world > 1:
            dist.broadcast_object_list(found, src=0, group=gather_group)
        prepass = {"seconds": dt, "frames_per_s": F / dt, "probe_points": 5000, "iters": [500, 100],
                   "keyframes_found": found[0], "scene_cuts": keyframes, "matches_scene_cuts": found[0] == keyframes}
        keyframes = found[0]
    gops = gop.gops_from_keyframes(keyframes, F)
    cost = lambda g: args.gop_iters_key + (g[1] - g[0] - 1) * args.gop_iters_pred
    mine = gop.assign_gops(gops, world, cost=cost)[rank]
    from gsvc_b200 import yuv
    # raw I420 frames, as the reference's .yuv input holds them; decoded inside the tile kernel
    host = {f: yuv.rgb_to_i420((frame_fn(f)[0] * 255).round().to(torch.uint8).permute(1, 2, 0).contiguous())
            .contiguous().pin_memory() for g in mine for f in range(*g)}
    get = lambda f: host[f].to(dev, non_blocking=True)
    make = gop.reference_fitters(H, W, NPTS, dev)
    # warm-up: one short key frame + predicted frame chain (allocator, kernels, prune path)
    if mine:
        f0 = mine[0][0]
        gop.fit_gop(lambda f: get(f0), (f0, f0 + 2), make, 120, 120)
    sampler = ClockSampler(dev.index or 0)
    if rank == 0:
        sampler.start()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    launches0 = _lib.launch_count()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    wall0 = time.perf_counter()
    a.record()
    out = gop.fit_sequence(get, F, keyframes, make, args.gop_iters_key, args.gop_iters_pred, group=gather_group,
                           chains=chains)
    b.record()
    torch.cuda.synchronize(dev)
    busy_ms = a.elapsed_time(b)                       # this rank's own device time: imbalance shows as a spread
    if world > 1:
        dist.barrier()
    wall = time.perf_counter() - wall0
    launches = _lib.launch_count() - launches0
    t = torch.tensor([busy_ms], dtype=torch.float64, device=dev)
    per_rank = [torch.zeros_like(t) for _ in range(world)] if world > 1 else [t]
    if world > 1:
        dist.all_gather(per_rank, t)
    ms = max(float(x) for x in per_rank)
    clocks = sampler.stop() if rank == 0 else None
    if rank != 0:
        return None
    iters = sum(r["iters"] for r in out.values())
    psnrs = [out[f]["psnr"] for f in sorted(out)]
    npts = [out[f]["num_points"] for f in sorted(out)]
    return {
        "metric": "gop_frames_per_s_1080p_50k", "value": F / (ms * 1e-3), "unit": "frames/s", "n_gpus": world,
        "steps": F, "warmup": 2, "ms_per_step": ms / F, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"C4: {F}-frame 1080p synthetic sequence, scene cut every {cut} frames, "
                               f"{NPTS} Gaussians, key frames {args.gop_iters_key} iters with removal control, "
                               f"predicted frames {args.gop_iters_pred} iters with random add + prune",
                   "H": H, "W": W, "gaussians": NPTS, "gops": len(gops),
                   "gops_per_rank": [len(x) for x in gop.assign_gops(gops, world, cost=cost)],
                   "parallelism": f"{len(gops)} key-frame-delimited GOPs over {world} GPU(s), no data-path collective; "
                                  f"up to {chains} concurrent chain(s) per GPU"},
        "clocks": clocks, "gpu_launches": launches, "fit_iterations": iters,
        "fit_iters_per_s": iters / (ms * 1e-3), "wall_s_timed_region": wall,
        "rank_busy_s": [round(float(x) * 1e-3, 3) for x in per_rank],
        "psnr_mean": sum(psnrs) / len(psnrs), "psnr_min": min(psnrs),
        "num_points_first_last": [npts[0], npts[-1]], "num_points_max": max(npts),
        "keyframe_prepass": prepass,
        "e2e": {"value": F / wall, "unit": "frames/s", "h2d_bytes_per_step": 3 * H * W // 2,
                "d2h_bytes_per_step": 36 * NPTS},
    }


def main_gop(args, rank: int, world: int, local_rank: int):
    """--workload gop: BASELINE config 4 alone (run_gop)"""
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    gather_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        gather_group = dist.new_group(backend="gloo")       # gather_object of CPU dicts
    line = run_gop(args, rank, world, dev, gather_group, args.gop_frames, args.gop_cut, args.gop_chains, args.gop_detect)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="fit", choices=["fit", "gop"],
                    help="fit: the headline line (C2 + C3 render); gop: BASELINE config 4 (sequence, GOP sharding)")
    ap.add_argument("--gop-frames", type=int, default=120)
    ap.add_argument("--gop-cut", type=int, default=30, help="scene cut (key frame) every this many frames")
    ap.add_argument("--gop-chains", type=int, default=4,
                    help="GOPs a rank fits concurrently (own stream + host thread each); 1 = one after the other")
    ap.add_argument("--gop-detect", action="store_true",
                    help="run the key-frame detection pre-pass first and shard by the key frames it finds")
    ap.add_argument("--gop-iters-key", type=int, default=4100,
                    help="iterations per key frame (the removal control runs to iteration 4000)")
    ap.add_argument("--gop-iters-pred", type=int, default=1000,
                    help="iterations per predicted frame (add at 1, prune 500..900, rebuild at 1000)")
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "b200-dropin", "reference", "reference-gpu"])
    ap.add_argument("--cpu-steps", type=int, default=20, help="fit iterations of the cpu_baseline leg")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extra-legs", action="store_true",
                    help="headline line only: no dropin / reference_gpu / c5 / spatial / gop legs")
    ap.add_argument("--no-gop-leg", action="store_true", help="skip the config-4 sequence leg (the longest one)")
    ap.add_argument("--leg-steps", type=int, default=30, help="steps per extra leg (at most --steps)")
    ap.add_argument("--warm-stages", action="store_true",
                    help="per-stage timing pass without the L2 flush (what the back-to-back loop sees); diagnostics")
    ap.add_argument("--ranges", action="store_true",
                    help="tile ranges (count + scan + scatter, 3 kernels/iter) instead of fixed tile slots (2 kernels/iter)")
    ap.add_argument("--trained-like", action="store_true",
                    help="informational: wide Gaussians (L = 4*rand + bound, ~10x the intersections, tiles beyond 256)")
    args = ap.parse_args()
    rank, world, local_rank = _env_int("RANK", 0), _env_int("WORLD_SIZE", 1), _env_int("LOCAL_RANK", 0)
    if args.workload == "gop":
        main_gop(args, rank, world, local_rank)
    elif args.impl == "reference":
        main_reference(args, rank)
    else:
        main_gpu(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
