"""CPU tests of the multi-GPU host logic (no GPU): GOP partitioning, and the N>1 path of
gsvc_b200.gop.fit_sequence under a real world_size-2 `gloo` process group with a stand-in fitter
(the sharding has no data-path collective: only the final gather is communication)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gsvc_b200 import gop


def test_gops_from_keyframes():
    assert gop.gops_from_keyframes([0, 30, 60, 90], 120) == [(0, 30), (30, 60), (60, 90), (90, 120)]
    assert gop.gops_from_keyframes([30, 60], 100) == [(0, 30), (30, 60), (60, 100)]      # frame 0 always starts one
    assert gop.gops_from_keyframes([0, 5, 5, 200], 10) == [(0, 5), (5, 10)]
    assert gop.gops_from_keyframes([], 1) == [(0, 1)]


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_assignment_is_a_balanced_partition(world):
    gops = gop.gops_from_keyframes([0, 15, 30, 45, 60, 75, 90, 105], 120)
    parts = gop.assign_gops(gops, world)
    flat = sorted(g for p in parts for g in p)
    assert flat == gops                                            # every GOP exactly once
    loads = [sum(b - a for a, b in p) for p in parts]
    assert max(loads) - min(loads) <= 15
    assert gop.assign_gops(gops, world) == parts                   # deterministic: no communication needed
    # uneven GOPs: longest-first keeps the spread below the longest GOP
    uneven = gop.gops_from_keyframes([0, 3, 40, 44, 90], 120)
    loads = [sum(b - a for a, b in p) for p in gop.assign_gops(uneven, min(world, 4))]
    assert max(loads) - min(loads) <= 46


class _FakeFit:
    """stand-in with FusedFrameFit's surface: parameters drift by a fixed amount per iteration, so
    the chain structure (key frame from scratch, predicted frames from the previous one) is visible"""

    def __init__(self, prev):
        base = torch.zeros(4, 2) if prev is None else prev["_xyz"]
        self.p = {"_xyz": base.clone(), "_cholesky": torch.ones(4, 3), "_features_dc": torch.ones(4, 3)}
        self.iters = 0

    def train_iters(self, gt, iters):
        self.p["_xyz"] += float(gt.mean()) * 0 + iters
        self.iters += iters

    def loss(self):
        return 1.0 / (1.0 + self.iters)

    def state_dict(self):
        return self.p


def _expected_xyz(frame, keyframes, ik, ip):
    k = max(x for x in keyframes if x <= frame)
    return ik + (frame - k) * ip


def test_fit_sequence_single_process():
    kf = [0, 4, 7]
    out = gop.fit_sequence(lambda f: torch.full((3, 2, 2), float(f)), 10, kf, _FakeFit, 100, 10)
    assert sorted(out) == list(range(10))
    for f, r in out.items():
        assert float(r["state"]["_xyz"][0, 0]) == _expected_xyz(f, kf, 100, 10)
        assert r["iters"] == (100 if f in kf else 10)


def test_chains_split_is_a_partition_and_cpu_falls_back_to_one_chain():
    """fit_sequence(chains=k): a rank's GOPs are split into k lanes by the same longest-first rule;
    without CUDA the lanes run one after the other -- same result either way"""
    kf = [0, 4, 7, 9]
    a = gop.fit_sequence(lambda f: torch.full((3, 2, 2), float(f)), 12, kf, _FakeFit, 100, 10, chains=1)
    b = gop.fit_sequence(lambda f: torch.full((3, 2, 2), float(f)), 12, kf, _FakeFit, 100, 10, chains=3)
    assert sorted(a) == sorted(b) == list(range(12))
    for f in a:
        assert torch.equal(a[f]["state"]["_xyz"], b[f]["state"]["_xyz"]) and a[f]["iters"] == b[f]["iters"]
    gops = gop.gops_from_keyframes(kf, 12)
    lanes = gop.assign_gops(gops, 3)
    assert sorted(g for lane in lanes for g in lane) == sorted(gops) and all(lane == sorted(lane) for lane in lanes)


def test_options_defaults_are_the_reference_and_override_restores():
    import gsvc_b200 as G
    o = G.options
    assert (o.analytic_projection_grad, o.max_per_tile, o.nd_channels) == (False, 256, False)
    with o.override(max_per_tile=0, nd_channels=True):
        assert o.max_per_tile == 0 and o.nd_channels is True
        with o.override(analytic_projection_grad=True):
            assert o.analytic_projection_grad is True
        assert o.analytic_projection_grad is False
    assert (o.analytic_projection_grad, o.max_per_tile, o.nd_channels) == (False, 256, False)
    try:
        with o.override(no_such_option=1):
            pass
    except AttributeError:
        pass
    else:
        raise AssertionError("unknown option accepted")
    assert not hasattr(o, "no_such_option")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        kf = [0, 5, 9, 12]
        seen = []
        fitter = lambda prev: _FakeFit(prev)
        get = lambda f: (seen.append(f), torch.full((3, 2, 2), float(f)))[1]
        out = gop.fit_sequence(get, 16, kf, fitter, 50, 5)
        q.put((rank, sorted(seen), None if out is None else {f: (float(r["state"]["_xyz"][0, 0]), r["iters"])
                                                              for f, r in out.items()}))
    finally:
        dist.destroy_process_group()


def test_fit_sequence_world_size_2_gloo():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = {}
    for _ in range(world):
        rank, seen, out = q.get(timeout=120)
        res[rank] = (seen, out)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # the two ranks touched disjoint frame sets that cover the sequence
    assert sorted(res[0][0] + res[1][0]) == list(range(16))
    assert not set(res[0][0]) & set(res[1][0])
    assert res[1][1] is None                                   # only rank 0 returns the merged result
    merged = res[0][1]
    kf = [0, 5, 9, 12]
    for f in range(16):
        assert merged[f] == (_expected_xyz(f, kf, 50, 5), 50 if f in kf else 5)


def test_split_tile_rows():
    from gsvc_b200.fit import split_tile_rows
    for rows, world in ((135, 8), (68, 2), (68, 4), (9, 8), (4, 4)):
        bands = split_tile_rows(rows, world)
        assert len(bands) == world and bands[0][0] == 0 and bands[-1][1] == rows
        assert all(a[1] == b[0] for a, b in zip(bands, bands[1:])) and all(b > a for a, b in bands)
        assert max(b - a for a, b in bands) - min(b - a for a, b in bands) <= 1
    # weighted: the heaviest band is as light as a contiguous partition allows
    w = [1, 1, 1, 1, 10, 10, 1, 1, 1, 1]
    bands = split_tile_rows(10, 4, w)
    assert max(sum(w[a:b]) for a, b in bands) == 10
    assert split_tile_rows(6, 2, [0] * 6) == [(0, 3), (3, 6)]
    with pytest.raises(ValueError):
        split_tile_rows(3, 4)


# ---- key-frame detection pre-pass (train_video_Represent.py:312-355) -------------------------
def test_keyframe_decision_matches_the_reference_functions():
    import numpy as np
    G = np.load(os.path.join(os.path.dirname(__file__), "golden", "keyframes.npz"))
    for case in range(6):
        got = gop.keyframes_from_scores(G[f"scores{case}"])
        assert [k + 1 for k in got] == [int(k) for k in G[f"kframes{case}"]], case      # the reference numbers frames from 1


class _ProbeFit:
    """probe stand-in: the loss of a fit depends on the frame, on where it started and on how long
    it ran, so a wrong halo or a wrong pairing of frames shows in the score"""

    def __init__(self, prev):
        self.start = None if prev is None else float(prev["_xyz"][0, 0])
        self.frame_val, self.iters = None, 0

    def train_iters(self, gt, iters):
        self.frame_val = float(gt.flatten()[0])
        self.iters += iters

    def loss(self):
        base = 0.01 * (1 + 0.01 * (self.frame_val % 7))
        if self.start is None:
            return base / self.iters
        return base / self.iters + 0.001 * abs(self.frame_val - self.start)     # far from the previous frame: worse

    def frame_state(self):
        return {"_xyz": torch.full((2, 2), self.frame_val)}


def _frame_values(n):
    vals, v = [], 0.0
    for f in range(n):
        v += 9.0 if f in (7, 19) else 1.0          # two scene cuts
        vals.append(v)
    return vals


def _expected_scores(n, ik, ip):
    vals = _frame_values(n)
    out = [0.0]
    for f in range(1, n):
        base = 0.01 * (1 + 0.01 * (vals[f] % 7))
        out.append(base / ip + 0.001 * abs(vals[f] - vals[f - 1]) - base / ik)
    return out


def test_keyframe_scores_single_process():
    n = 24
    vals = _frame_values(n)
    s = gop.keyframe_scores(lambda f: torch.full((3, 2, 2), vals[f]), n, _ProbeFit, 50, 10)
    assert s == pytest.approx(_expected_scores(n, 50, 10))
    assert gop.keyframes_from_scores(s) == [0, 7, 19]


def _kf_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        n = 24
        vals = _frame_values(n)
        seen = []
        get = lambda f: (seen.append(f), torch.full((3, 2, 2), vals[f]))[1]
        scores = gop.keyframe_scores(get, n, _ProbeFit, 50, 10)
        q.put((rank, sorted(set(seen)), scores))
    finally:
        dist.destroy_process_group()


def test_keyframe_scores_world_size_3_gloo_with_halo():
    world = 3
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_kf_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = {}
    for _ in range(world):
        rank, seen, out = q.get(timeout=120)
        res[rank] = (seen, out)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0][1] == pytest.approx(_expected_scores(24, 50, 10))      # identical to the unsharded pass
    assert res[1][1] is None and res[2][1] is None
    assert res[0][0] == list(range(0, 8))                               # block [0, 8)
    assert res[1][0] == list(range(7, 16))                              # block [8, 16) + the frame before it
    assert res[2][0] == list(range(15, 24))


def test_pack_and_unpack_of_a_rank_s_frames_round_trip():
    """the packing behind the multi-GPU gather (gop._gather_states): frames with different Gaussian counts, out
    of order in the dict, come back bit for bit, as contiguous views of the flat buffer; the digest is order-free"""
    from gsvc_b200 import gop
    g = torch.Generator().manual_seed(3)
    local = {}
    for f, n in ((7, 5), (3, 1), (12, 9)):
        local[f] = {"state": {"_xyz": torch.randn(n, 2, generator=g), "_cholesky": torch.randn(n, 3, generator=g),
                              "_features_dc": torch.randn(n, 3, generator=g)},
                    "psnr": 30.0 + f, "iters": 100 * f, "num_points": n}
    flat, meta = gop._pack_states(local)
    assert flat.dtype == torch.float32 and flat.numel() == 8 * (5 + 1 + 9)
    assert sorted(meta) == [3, 7, 12] and [meta[f]["rows"] for f in sorted(meta)] == [1, 5, 9]
    assert "state" not in meta[7] and meta[7]["psnr"] == 37.0
    back = gop._unpack_states(flat.clone(), meta)
    assert sorted(back) == [3, 7, 12]
    for f in local:
        assert back[f]["psnr"] == local[f]["psnr"] and back[f]["iters"] == local[f]["iters"] and "rows" not in back[f]
        for k in ("_xyz", "_cholesky", "_features_dc"):
            assert back[f]["state"][k].is_contiguous() and torch.equal(back[f]["state"][k], local[f]["state"][k])
    assert gop._bit_digest(flat) == gop._bit_digest(flat.flip(0)) != gop._bit_digest(flat + 1)
    empty, meta0 = gop._pack_states({})
    assert empty.numel() == 0 and meta0 == {} and gop._unpack_states(empty, meta0) == {} and gop._bit_digest(empty) == 0
    with pytest.raises(ValueError):
        gop._unpack_states(flat[:-1], meta)
