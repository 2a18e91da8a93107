"""CPU tests of the prune / add control (BASELINE config 4's "pruning and random add"):
the schedule (gsvc_b200.density) and FrameFit's restatement of the control are replayed against
trajectories recorded from the reference's own GaussianVideo_frame + Adan
(tests/golden/make_golden_control.py -> tests/golden/control.npz)."""
import os

import numpy as np
import pytest
import torch

from tests import surrogate
from gsvc_b200.density import Action, DensityPlan, OptimizerClock
from gsvc_b200.fit import FrameFit

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "control.npz"))
KEYS = ("_xyz", "_cholesky", "_features_dc", "rgb_W")


class _SurrogateFit(FrameFit):
    def _render(self, means2d, L_elements, colors):
        xys, depths, radii, conics, nth = surrogate.project(means2d, L_elements, self.H, self.W, self.tile_bounds)
        return surrogate.rasterize(xys, depths, radii, conics, nth, colors, torch.ones(means2d.shape[0], 1), self.H,
                                   self.W)


def _fit(mode):
    init = {k: torch.from_numpy(G[f"{mode}.init.{k}"]) for k in KEYS}
    n = init["_xyz"].shape[0]
    kw = dict(isremoval=True, densification_interval=500) if mode == "removal" else dict(isdensity=True)
    m = _SurrogateFit(8, 8, n, "cpu", init=init, max_num_points=48, removal_rate=0.25, **kw)
    assert torch.equal(m.rgb_W.detach(), init["rgb_W"])        # 0.01 / ones, as the reference initialises it
    return m


@pytest.mark.parametrize("mode,iters", [("removal", 4100), ("density", 1100)])
def test_framefit_control_reproduces_the_reference_class(mode, iters):
    m = _fit(mode)
    gt = torch.from_numpy(G["gt"])
    torch.set_rng_state(torch.from_numpy(G[f"{mode}.rng_state"]))      # the add draws from the global CPU generator
    ns = G[f"{mode}.n"]
    losses = G[f"{mode}.loss"]
    snaps = sorted(int(k.split(".")[1][2:]) for k in G.files if k.startswith(f"{mode}.it") and k.endswith("._xyz"))
    for it in range(1, iters + 1):
        loss, _ = m.train_iter(gt, it)
        assert m._xyz.shape[0] == ns[it - 1], f"iteration {it}: Gaussian count"
        assert float(loss.detach()) == pytest.approx(losses[it - 1], rel=2e-4, abs=1e-7), f"iteration {it}: loss"
        if it in snaps:
            for k in KEYS:
                want = G[f"{mode}.it{it}.{k}"]
                got = getattr(m, k).detach().numpy()
                assert got.shape == want.shape, (it, k)
                np.testing.assert_allclose(got, want, rtol=2e-3, atol=2e-5, err_msg=f"iteration {it}: {k}")


def test_plan_matches_the_recorded_counts():
    """the schedule alone (no model): Gaussian counts after every iteration"""
    for mode, kw, n0 in (("removal", dict(isremoval=True, densification_interval=500), 48),
                         ("density", dict(isdensity=True, densification_interval=100), 36)):
        plan = DensityPlan(48, 0.25, **kw)
        n = n0
        for it, want in enumerate(G[f"{mode}.n"], start=1):
            a = plan.action(it, n)
            n += a.count if a.kind == "add" else -a.count if a.kind == "prune" else 0
            assert n == want, (mode, it)


def test_plan_actions_at_reference_defaults():
    # train_video_Represent.py defaults: removal_rate 0.1, densification_interval 100
    rm = DensityPlan(50_000, 0.1, 100, isremoval=True)
    assert rm.action(50, 50_000) == Action()
    assert rm.action(100, 50_000) == Action("prune", int(0.1 / 40 * 50_000))        # 125 per control step
    assert rm.action(4000, 50_000 - 39 * 125) == Action("prune", 125, True)           # lands on max*(1-rate)
    assert rm.action(4000, 45_000) == Action("none", 0, True)                         # nothing left to drop: rebuild only
    assert rm.action(4100, 45_000) == Action()
    ad = DensityPlan(50_000, 0.1, 100, isdensity=True)
    assert ad.action(1, 45_000) == Action("add", 5000)
    assert ad.action(100, 50_000) == Action() and ad.action(400, 50_000) == Action()
    assert ad.action(500, 50_000) == Action("prune", 1000)
    assert ad.action(900, 46_000) == Action("prune", 1000)
    assert ad.action(1000, 45_000) == Action("none", 0, True)
    assert ad.action(1100, 45_000) == Action()
    assert DensityPlan(100, 0.1).action(100, 100) == Action()                         # neither mode: rgb_W is a buffer
    assert not DensityPlan(100, 0.1).learns_rgb_w
    with pytest.raises(ZeroDivisionError):                                            # int(500 / 600) == 0, as the reference
        DensityPlan(100, 0.1, 600, isdensity=True).action(600, 100)


def test_optimizer_clock():
    c = OptimizerClock()
    assert c.begin(Action()) == (1, True, False)                   # first step: previous gradient := gradient
    assert c.begin(Action()) == (2, False, False)
    assert c.begin(Action("prune", 3)) == (3, True, True)          # replaced: no update, the step still counts
    assert c.begin(Action()) == (4, True, False)                   # fresh state, but bias corrections continue
    assert c.begin(Action()) == (5, False, False)
    assert c.begin(Action("none", 0, True)) == (1, True, False)    # rebuilt optimizer steps right away
    assert c.lr_frozen
    assert c.begin(Action("prune", 2, True)) == (1, True, True)    # rebuilt and replaced: its first step sees no grads
    assert c.begin(Action()) == (2, True, False)


def test_adan_takes_the_reference_constructor_arguments_and_refuses_what_it_lacks():
    """optimizer.py:65-74 of the reference: Adan(params, lr, betas, eps, weight_decay, max_grad_norm,
    no_prox, foreach, fused)"""
    import pytest
    import torch
    from gsvc_b200.adan import Adan
    p = [torch.nn.Parameter(torch.zeros(4, 2))]
    Adan(p, lr=1e-3, betas=(0.98, 0.92, 0.99), eps=1e-8, weight_decay=0.0, max_grad_norm=0.0, no_prox=False,
         foreach=True, fused=False)
    with pytest.raises(NotImplementedError):
        Adan(p, max_grad_norm=1.0)
    with pytest.raises(NotImplementedError):
        Adan(p, no_prox=True)
    opt = Adan(p, fused=True)
    p[0].grad = torch.ones(4, 2)
    with pytest.raises(RuntimeError):
        opt.step()                      # fused=True insists on CUDA tensors
    opt = Adan(p)                       # default: decides per step -> foreach on the CPU
    opt.step()
    assert torch.allclose(p[0], torch.full((4, 2), -1e-3), atol=1e-6)


def test_step_lr_formula_matches_torch_scheduler_on_cpu():
    """FusedFrameFit.current_lr() against torch.optim.lr_scheduler.StepLR across boundaries, without a GPU
    (the method only reads the step index, the schedule and the optimizer clock: called unbound on a stand-in).
    Iteration `it` runs optimizer.step() BEFORE the it-th scheduler.step() (GaussianSplats_Represent.py:52, 203-206)."""
    from types import SimpleNamespace

    import torch

    from gsvc_b200.fit import FusedFrameFit
    for step_size, gamma in ((5, 0.5), (3, 0.1), (20000, 0.5)):
        stub = SimpleNamespace(lr=1e-2, lr_gamma=gamma, lr_step_size=step_size, step_index=0,
                               clock=SimpleNamespace(lr_frozen=False))
        opt = torch.optim.SGD([torch.nn.Parameter(torch.zeros(1))], lr=1e-2)
        sch = torch.optim.lr_scheduler.StepLR(opt, step_size=step_size, gamma=gamma)
        for it in range(1, 4 * min(step_size, 10) + 2):
            stub.step_index = it                      # as _issue has it while iteration `it` is being issued
            assert FusedFrameFit.current_lr(stub) == pytest.approx(opt.param_groups[0]["lr"], rel=1e-12), (step_size, it)
            opt.step()
            sch.step()
        stub.clock.lr_frozen = True                   # update_optimizer() rebuilt the optimizer without a scheduler (:93-97)
        assert FusedFrameFit.current_lr(stub) == 1e-2
