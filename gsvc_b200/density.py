"""The schedule of the reference's Gaussian prune / add control, as pure host logic.

GaussianVideo_frame.train_iter (GaussianSplats_Represent.py:191-207) calls, between
loss.backward() and optimizer.step():
    adaptive_control(iter)   if (iter == 1 or iter % interval == 0) and isdensity     (:199-200)
    removal_control(iter)    elif iter % interval == 0 and isremoval                  (:201-202)
Both rank the Gaussians by |rgb_W| and drop the lowest ones (:103-114, :146-158); adaptive_control
also appends `int(max_num_points * removal_rate)` random Gaussians at iteration 1 (:134-145).

What that does to the optimizer is implicit in the reference and spelled out here, because the
fused step has to reproduce it:
  * a prune or an add REPLACES the four parameter tensors after backward(): the new tensors carry
    no gradient, so this iteration's optimizer.step() updates nothing (optimizer.py:175-177) but
    still advances group['step'] (:165-168); the new tensors have no optimizer state, so the next
    step starts from zero moments with previous gradient := gradient (:180-185)        -> skip + fresh
  * the last control iteration calls update_optimizer() (:93-97, :127, :171): a NEW optimizer --
    group['step'] restarts at 0, all state is fresh, the learning rate is the constructor's and no
    scheduler is attached to it any more                                               -> rebuild

`DensityPlan.action(iter, n)` returns what happens at iteration `iter` with `n` Gaussians alive.
"""
from __future__ import annotations

from dataclasses import dataclass


@dataclass(frozen=True)
class Action:
    kind: str = "none"      # "none" | "add" | "prune"
    count: int = 0          # Gaussians appended / removed
    rebuild: bool = False   # update_optimizer() runs in this iteration (after the prune, if any)

    @property
    def replaces_parameters(self) -> bool:
        """the parameter tensors are re-created, so this iteration's update is skipped and the
        optimizer state of every parameter starts afresh"""
        return self.kind != "none"


class DensityPlan:
    def __init__(self, max_num_points: int, removal_rate: float = 0.1, densification_interval: int = 100,
                 isdensity: bool = False, isremoval: bool = False, removal_until: int = 4000,
                 density_from: int = 500, density_for: int = 500):
        """removal_until / density_from / density_for are the reference's hard-coded
        iter_threshold_remove = 4000 (:99), iter_threshold_add = 500 and iter_threshold_remove = 500
        (:131-132); they are arguments only so that tests can run short schedules."""
        self.max_num_points = int(max_num_points)
        self.removal_rate = float(removal_rate)
        self.interval = int(densification_interval)
        self.isdensity, self.isremoval = bool(isdensity), bool(isremoval)
        self.removal_until = int(removal_until)
        self.density_from, self.density_for = int(density_from), int(density_for)
        if self.interval < 1:
            raise ValueError("densification_interval must be positive")

    @property
    def learns_rgb_w(self) -> bool:
        """rgb_W is an nn.Parameter only in these modes, else a buffer of ones (:31-36)"""
        return self.isdensity or self.isremoval

    @property
    def rgb_w_init(self) -> float:
        return 0.01 if self.isremoval else 1.0   # :31-34 (isremoval wins when both are set)

    def action(self, it: int, n: int) -> Action:
        on_interval = it % self.interval == 0
        if (it == 1 or on_interval) and self.isdensity:
            return self._adaptive(it, n)
        if on_interval and self.isremoval:
            return self._removal(it, n)
        return Action()

    # GaussianSplats_Represent.py:98-128
    def _removal(self, it: int, n: int) -> Action:
        thr = self.removal_until
        if it > thr:
            return Action()
        per_step = self.removal_rate / int(thr / self.interval)        # ZeroDivisionError as the reference
        if it < thr:
            # the mask indexing runs even when nothing is removed (:106-114): still a replacement
            return Action("prune", min(int(per_step * self.max_num_points), n))
        remove = n - int(self.max_num_points * (1 - self.removal_rate))
        return Action("prune", remove, True) if remove > 0 else Action("none", 0, True)

    # GaussianSplats_Represent.py:130-172
    def _adaptive(self, it: int, n: int) -> Action:
        thr_add, thr_rm = self.density_from, self.density_for
        num = int(self.max_num_points * self.removal_rate)
        if it > thr_add + thr_rm or it < thr_add:
            return Action("add", num) if (it == 1 and num > 0) else Action()
        if it < thr_add + thr_rm:
            return Action("prune", min(int(num / int(thr_rm / self.interval)), n))
        remove = n - int(self.max_num_points * (1 - self.removal_rate))
        return Action("prune", remove, True) if remove > 0 else Action("none", 0, True)


class OptimizerClock:
    """group['step'] / freshness of the reference's Adan under the control above.

        begin(action) -> (adan_step, fresh, skip)   for the step about to be issued
    adan_step: the value the bias corrections use; fresh: previous gradient := gradient;
    skip: parameters are being replaced, the update does not happen."""

    def __init__(self):
        self.step = 0
        self.fresh = True
        self.lr_frozen = False     # after update_optimizer(): new optimizer, no scheduler on it

    def begin(self, action: Action):
        if action.rebuild:
            self.step = 0
            self.fresh = True
            self.lr_frozen = True
        self.step += 1
        if action.replaces_parameters:
            self.fresh = True
            return self.step, True, True
        fresh = self.fresh or self.step == 1
        self.fresh = False
        return self.step, fresh, False
