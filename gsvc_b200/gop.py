"""GOP-sharded fitting of a frame sequence across GPUs (BASELINE config 4).

In the reference a keyframe resets the chain (`Trained_Model = None`,
train_video_Represent.py:362-363) and every P-frame starts from the previous frame's Gaussians
(:359-367), so keyframe-delimited GOPs are independent sequential chains.  One process per GPU
(torchrun); rank r fits the GOPs `assign_gops` gives it with no data-path collective, and the
per-frame parameter dicts (`_xyz`, `_cholesky`, `_features_dc` -- the reference's
gmodels_state_dict entries, train_video_Represent.py:107-114,379) are gathered to rank 0 once at
the end.  The reference itself is single-GPU (`cuda:0`); this module is new design, not a mirror.

Prune / random-add inside a frame's fit (GaussianSplats_Represent.py:98-172) is the fitter's
business (FusedFrameFit(isremoval=...) for key frames, FusedFrameFit(isdensity=...) for predicted
frames, as train_video_Represent.py:362-366 builds them); the Gaussian count of a frame then
follows from the previous one (`num_gaussian_points`, :365).
"""
from __future__ import annotations

import math
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

Gop = Tuple[int, int]   # [first frame, one past the last frame)


def gops_from_keyframes(keyframes: Sequence[int], num_frames: int) -> List[Gop]:
    """keyframes: sorted 0-based frame indices that start a GOP (frame 0 is always one)"""
    ks = sorted(set(int(k) for k in keyframes if 0 <= int(k) < num_frames) | {0})
    return [(a, b) for a, b in zip(ks, ks[1:] + [num_frames]) if b > a]


def assign_gops(gops: Sequence[Gop], world: int, cost: Optional[Callable[[Gop], float]] = None) -> List[List[Gop]]:
    """Longest-processing-time-first: GOPs sorted by cost (default: frame count, the keyframe
    weighted by `cost`), each to the least-loaded rank; every rank's list is in frame order.
    Deterministic, so every rank computes the same assignment without communication."""
    cost = cost or (lambda g: float(g[1] - g[0]))
    loads = [0.0] * world
    out: List[List[Gop]] = [[] for _ in range(world)]
    for g in sorted(gops, key=lambda g: (-cost(g), g[0])):
        r = min(range(world), key=lambda i: (loads[i], i))
        out[r].append(g)
        loads[r] += cost(g)
    for lst in out:
        lst.sort()
    return out


def _fit_all(fit, gt, iters: int) -> float:
    """`iters` iterations that really happen: a fused fitter whose workspace overflows (a tile beyond
    256 slots, more intersections than its capacity) rewinds and continues by itself
    (FusedFrameFit.train_until); other fitters just iterate"""
    if hasattr(fit, "train_until"):
        return fit.train_until(gt, fit.step_index + iters)
    fit.train_iters(gt, iters)
    return fit.loss()


def fit_gop(get_frame: Callable[[int], torch.Tensor], gop: Gop, make_fitter: Callable[[Optional[dict]], object],
            iters_key: int, iters_pred: int, on_frame: Optional[Callable[[int, object], None]] = None,
            seed: Optional[int] = None, states_on_device: bool = False) -> Dict[int, dict]:
    """Fit one GOP: the keyframe from scratch (`make_fitter(None)`), every following frame from
    the previous frame's parameters (`make_fitter(prev_state)`).  Returns frame -> {"state": the
    three parameter tensors on the CPU (states_on_device: where the fitter left them), "psnr": float,
    "iters": int}."""
    out: Dict[int, dict] = {}
    prev = None
    for f in range(gop[0], gop[1]):
        if seed is not None:
            # every frame draws (fresh Gaussians, the random add) from its own stream, so the result
            # of a frame does not depend on which rank (or which of a rank's concurrent chains) fitted
            # which other frames before it
            torch.manual_seed(int(seed) * 100003 + f)
        fit = make_fitter(prev)
        if seed is not None and hasattr(fit, "rng"):
            # the same stream as a private generator (same engine, same seed => same draws): two chains
            # of one process cannot then interleave their draws from the global generator
            fit.rng = torch.Generator().manual_seed(int(seed) * 100003 + f)
        iters = iters_key if prev is None else iters_pred
        gt = get_frame(f)
        loss = _fit_all(fit, gt, iters)
        # the reference hands on _xyz, _cholesky and the WEIGHTED colours (train_video_Represent.py:107-114)
        state = fit.frame_state() if hasattr(fit, "frame_state") else fit.state_dict()
        prev = {k: v.detach().clone() for k, v in state.items()}
        # (prev is this function's own copy and make_fitter only reads it: it can be handed out as it is)
        out[f] = {"state": prev if states_on_device else {k: v.cpu() for k, v in prev.items()},
                  "psnr": 10 * math.log10(1.0 / max(loss, 1e-20)),
                  "iters": iters, "num_points": int(prev["_xyz"].shape[0])}
        if on_frame is not None:
            on_frame(f, fit)
    return out


def fit_sequence(get_frame: Callable[[int], torch.Tensor], num_frames: int, keyframes: Sequence[int],
                 make_fitter: Callable[[Optional[dict]], object], iters_key: int, iters_pred: int,
                 group=None, seed: Optional[int] = 1, chains: int = 1) -> Optional[Dict[int, dict]]:
    """Every rank fits its share of the GOPs; rank 0 returns {frame: result} for the whole
    sequence (other ranks return None).  Works without torch.distributed (one process).

    chains > 1: a rank with several GOPs fits up to that many of them CONCURRENTLY, each chain on a
    CUDA stream of its own fed by a host thread of its own.  A fit iteration is two dependent kernels
    and the second (the per-Gaussian update) leaves most of the GPU idle; a second independent chain
    fills it.  Results do not depend on `chains` (every frame has its own random stream)."""
    distributed = dist.is_available() and dist.is_initialized()
    world = dist.get_world_size(group) if distributed else 1
    rank = dist.get_rank(group) if distributed else 0
    gops = gops_from_keyframes(keyframes, num_frames)
    # a keyframe costs iters_key, a predicted frame iters_pred
    mine = assign_gops(gops, world, cost=lambda g: iters_key + (g[1] - g[0] - 1) * iters_pred)[rank]
    local: Dict[int, dict] = {}
    cost = lambda g: iters_key + (g[1] - g[0] - 1) * iters_pred
    lanes = [l for l in assign_gops(mine, max(1, min(int(chains), len(mine))), cost=cost) if l] if mine else []
    # The results travel to rank 0 once, at the end.  Pickled through gather_object (the only way over a gloo
    # group) the 1.6 MB of a 1080p frame's parameters cost rank 0 a quarter of a second per 100 frames -- 9 % of
    # an 8-GPU run, the whole gap to linear scaling; when the job's default group is NCCL the parameters stay on
    # the GPU until then and go as ONE packed fp32 [rows, 8] tensor per rank over NVLink (_gather_states).
    by_tensor = (distributed and world > 1 and torch.cuda.is_available() and dist.get_backend() == "nccl"
                 and dist.get_world_size() == world)
    if len(lanes) <= 1 or not torch.cuda.is_available():
        for g in mine:
            local.update(fit_gop(get_frame, g, make_fitter, iters_key, iters_pred, seed=seed, states_on_device=by_tensor))
    else:
        local.update(_fit_lanes(get_frame, lanes, make_fitter, iters_key, iters_pred, seed, states_on_device=by_tensor))
    if not distributed:
        return local
    if by_tensor:
        by_tensor = all(v.is_cuda and v.dtype == torch.float32 for r in local.values() for v in r["state"].values()) \
            and all(tuple(r["state"]) == _STATE_KEYS for r in local.values())
        flags = [None] * world
        dist.all_gather_object(flags, bool(by_tensor), group=group)      # (a stand-in fitter may keep CPU tensors)
        if not all(flags):
            by_tensor = False
            for r in local.values():
                r["state"] = {k: v.cpu() for k, v in r["state"].items()}
    if by_tensor:
        merged = _gather_states(local, rank, world, group)
    else:
        gathered = [None] * world if rank == 0 else None
        dist.gather_object(local, gathered, dst=0, group=group)
        merged = {}
        for part in gathered or []:
            merged.update(part)
    if rank != 0:
        return None
    assert sorted(merged) == list(range(num_frames)), "GOP sharding lost or duplicated frames"
    return merged


_STATE_KEYS = ("_xyz", "_cholesky", "_features_dc")
_STATE_COLS = (2, 3, 3)


def _pack_states(local: Dict[int, dict]):
    """frame -> result  =>  (one flat fp32 tensor: per frame in order xyz [n,2] | cholesky [n,3] | features [n,3];
    {frame: the result's small fields + "rows": n})"""
    frames = sorted(local)
    meta = {f: {k: v for k, v in local[f].items() if k != "state"} for f in frames}
    for f in frames:
        meta[f]["rows"] = int(local[f]["state"]["_xyz"].shape[0])
    if not frames:
        return torch.empty((0,), dtype=torch.float32), meta
    return torch.cat([local[f]["state"][k].reshape(-1) for f in frames for k in _STATE_KEYS]), meta


def _unpack_states(flat: torch.Tensor, meta: Dict[int, dict]) -> Dict[int, dict]:
    """inverse of _pack_states: the states are contiguous views of `flat`"""
    want = 8 * sum(int(m["rows"]) for m in meta.values())
    if flat.numel() != want:
        raise ValueError(f"GOP gather: {flat.numel()} packed values for frames that need {want}")
    out: Dict[int, dict] = {}
    at = 0
    for f in sorted(meta):
        m = dict(meta[f])
        n = m.pop("rows")
        m["state"] = {}
        for k, c in zip(_STATE_KEYS, _STATE_COLS):
            m["state"][k] = flat[at:at + n * c].view(n, c)
            at += n * c
        out[f] = m
    return out


def _bit_digest(t: torch.Tensor) -> int:
    """sum of the tensor's 32-bit words as integers: order-free and exact"""
    return int(t.contiguous().view(torch.int32).to(torch.int64).sum()) if t.numel() else 0


def _gather_states(local: Dict[int, dict], rank: int, world: int, group) -> Optional[Dict[int, dict]]:
    """rank 0 <- every rank's frames: the small per-frame records (psnr, iterations, Gaussian count) as objects
    over `group`, the parameters as one flat fp32 tensor per rank (_pack_states) by one all-gather of the
    default NCCL group; rank 0 hands out CPU views of it."""
    dev = torch.device("cuda", torch.cuda.current_device())
    pack, meta = _pack_states(local)
    pack = pack.to(dev)
    both = [None] * world
    dist.all_gather_object(both, (meta, _bit_digest(pack)), group=group)
    metas, digests = [b[0] for b in both], [b[1] for b in both]
    # ONE all-gather of the job's own NCCL communicator, every rank's share padded to the longest.  Not a gather
    # to rank 0 and not send / recv: NCCL sets up point-to-point connections on their first use, which cost an
    # 8-rank job 1.3 s (measured: more than the pickling this replaces), while an all-gather rides the ring that
    # the job's first barrier already built -- and 25 MB per rank over NVLink is a millisecond either way.
    elems_of = [8 * sum(m["rows"] for m in metas[r].values()) for r in range(world)]
    longest = max(max(elems_of), 1)
    padded = torch.zeros((longest,), dtype=torch.float32, device=dev)
    padded[:pack.numel()] = pack
    everything = torch.empty((world * longest,), dtype=torch.float32, device=dev)
    dist.all_gather_into_tensor(everything, padded)
    if rank != 0:
        return None
    merged: Dict[int, dict] = {}
    for r in range(world):
        buf = everything[r * longest:r * longest + elems_of[r]]
        got = _bit_digest(buf)
        if got != digests[r]:
            raise RuntimeError(f"GOP gather: rank {r}'s parameters arrived damaged (digest {got} != {digests[r]})")
        merged.update(_unpack_states(buf.cpu(), metas[r]))
    return merged


def _fit_lanes(get_frame, lanes, make_fitter, iters_key, iters_pred, seed, states_on_device=False) -> Dict[int, dict]:
    """one host thread + one CUDA stream per lane (a list of GOPs fitted in order); the caller's
    stream waits for all of them, so events recorded around fit_sequence span the whole work"""
    import threading
    dev = torch.cuda.current_device()
    main = torch.cuda.current_stream(dev)
    streams = [torch.cuda.Stream(dev) for _ in lanes]
    results: List[Dict[int, dict]] = [dict() for _ in lanes]
    errors: List[BaseException] = []

    def work(i: int):
        try:
            torch.cuda.set_device(dev)
            streams[i].wait_stream(main)
            with torch.cuda.stream(streams[i]):
                for g in lanes[i]:
                    results[i].update(fit_gop(get_frame, g, make_fitter, iters_key, iters_pred, seed=seed,
                                              states_on_device=states_on_device))
        except BaseException as e:   # noqa: BLE001 -- re-raised on the caller's thread
            errors.append(e)

    threads = [threading.Thread(target=work, args=(i,), name=f"gop-lane-{i}") for i in range(len(lanes))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for s in streams:
        main.wait_stream(s)
    if errors:
        raise errors[0]
    out: Dict[int, dict] = {}
    for r in results:
        out.update(r)
    return out


def synthetic_sequence(H: int, W: int, num_frames: int = 120, cut_every: int = 30, shift_px: float = 2.0,
                       seed: int = 1):
    """BASELINE config 4's sequence: one pattern translated by `shift_px` per frame with
    a hard scene cut (new pattern) every `cut_every` frames.  -> (frame(f) -> [1,3,H,W] fp32,
    key frame indices)."""
    from . import synth

    def frame(f: int) -> torch.Tensor:
        return synth.target_frame(H, W, seed=seed, shift=shift_px * (f % cut_every), pattern=f // cut_every)

    return frame, list(range(0, num_frames, cut_every))


def reference_fitters(H: int, W: int, num_points: int, device, is_rm: bool = True, is_ad: bool = True,
                      removal_rate: float = 0.1, densification_interval: int = 100, lr: float = 1e-3, seed: int = 1,
                      cls=None, **kw):
    """make_fitter for fit_gop / fit_sequence, built the way train_video_Represent.py:358-366 builds
    its trainers: a key frame starts from `num_points` fresh Gaussians with the removal control
    (isremoval = --is_rm); a predicted frame starts from the previous frame's Gaussians (their
    count, their weighted colours) with the adaptive add / prune control (isdensity = --is_ad) and
    max_num_points = num_points."""
    if cls is None:
        from .fit import FusedFrameFit as cls

    def make(prev: Optional[dict]):
        if prev is None:
            return cls(H, W, num_points, device, lr=lr, seed=seed, isremoval=is_rm, removal_rate=removal_rate,
                       max_num_points=num_points, densification_interval=densification_interval, **kw)
        return cls(H, W, int(prev["_xyz"].shape[0]), device, lr=lr, init=prev, isdensity=is_ad,
                   removal_rate=removal_rate, max_num_points=num_points,
                   densification_interval=densification_interval, **kw)

    return make


# ------------------------------------------------------------------------------------------
# Key-frame detection pre-pass (train_video_Represent.py:312-355)
# ------------------------------------------------------------------------------------------
def detect_outliers_mean_diff(values: Sequence[float], window_size: int = 10, threshold: float = 3) -> List[int]:
    """utils.py:214-229 of the reference: index i is an outlier if it exceeds the mean of the window
    [i - window_size, i + window_size) by `threshold` standard deviations, or `threshold` times that mean"""
    import numpy as np
    v = np.asarray(values, dtype=np.float64)
    out = []
    for i in range(len(v)):
        w = v[max(0, i - window_size):min(len(v), i + window_size)]
        mean, std = np.mean(w), np.std(w)
        if (v[i] - mean) > threshold * std or v[i] > mean * threshold:
            out.append(i)
    return out


def keyframes_from_scores(scores: Sequence[float]) -> List[int]:
    """scores[i] = loss_P - loss_K of frame i (scores[0] = 0) -> 0-based key-frame indices, frame 0
    first: min-max normalisation of scores[1:] (:336-341), outlier detection (:347), frame 1
    (here: 0) always a key frame (:349)"""
    import numpy as np
    s = np.asarray(scores, dtype=np.float64)
    rest = s[1:]
    if rest.size == 0:
        return [0]
    lo, hi = float(rest.min()), float(rest.max())
    norm = [float(s[0])] + [float((x - lo) / (hi - lo)) if hi > lo else 0.0 for x in rest]
    ks = [int(i) for i in detect_outliers_mean_diff(norm)]
    return [0] + ks       # np.insert(K_frames, 0, 1) in the reference's 1-based numbering: frame 0 may repeat


def keyframe_scores(get_frame: Callable[[int], torch.Tensor], num_frames: int,
                    make_probe: Callable[[Optional[dict]], object], iters_key: int = 500, iters_pred: int = 100,
                    group=None, seed: Optional[int] = 1) -> Optional[List[float]]:
    """The probe fits of the reference's pre-pass: for every frame a fresh small model fitted for
    `iters_key` iterations (loss_K, :322-323 / :328-329) and a model started from the PREVIOUS
    frame's fresh fit, fitted for `iters_pred` (loss_P, :330-331); score = loss_P - loss_K (:334).
    The fresh fit of frame i depends on frame i alone, so the frames are sharded in contiguous blocks
    over the ranks with a one-frame halo (a rank refits the frame before its block) and no data-path
    collective; rank 0 gets the scores (others None)."""
    distributed = dist.is_available() and dist.is_initialized()
    world = dist.get_world_size(group) if distributed else 1
    rank = dist.get_rank(group) if distributed else 0
    edges = [round(r * num_frames / world) for r in range(world + 1)]
    a, b = edges[rank], edges[rank + 1]
    local: Dict[int, float] = {}

    def fresh(f: int):
        if seed is not None:
            torch.manual_seed(int(seed) * 100003 + f)
        fit = make_probe(None)
        loss = _fit_all(fit, get_frame(f), iters_key)
        state = fit.frame_state() if hasattr(fit, "frame_state") else fit.state_dict()
        return {k: v.detach().clone() for k, v in state.items()}, loss

    prev = fresh(a - 1)[0] if (a > 0 and b > a) else None          # the halo
    for f in range(a, b):
        cur, loss_k = fresh(f)
        if f == 0:
            local[f] = 0.0                                          # :324-326
        else:
            p = make_probe(prev)
            local[f] = _fit_all(p, get_frame(f), iters_pred) - loss_k
        prev = cur
    if not distributed:
        return [local[f] for f in range(num_frames)]
    gathered = [None] * world if rank == 0 else None
    dist.gather_object(local, gathered, dst=0, group=group)
    if rank != 0:
        return None
    merged: Dict[int, float] = {}
    for part in gathered:
        merged.update(part)
    assert sorted(merged) == list(range(num_frames)), "key-frame pre-pass lost or duplicated frames"
    return [merged[f] for f in range(num_frames)]
