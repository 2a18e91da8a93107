"""Options beyond the reference.  Every default reproduces the reference bit for
bit / formula for formula; nothing here is read by GSVC's unchanged scripts.

    import gsvc_b200
    gsvc_b200.options.analytic_projection_grad = True   # true gradient of the projection (Q12)
    gsvc_b200.options.max_per_tile = 0                  # render every Gaussian of a tile (Q1)
    gsvc_b200.options.nd_channels = True                # colours with 1..32 channels (Q10)
    with gsvc_b200.options.override(max_per_tile=0): ...

The values are read when an operator's forward runs; the matching backward uses what its forward
saw."""
from __future__ import annotations

from contextlib import contextmanager


class Options:
    #: False: project_gaussians_2d / _scale_rot backward use the reference's formula, which counts
    #: the off-diagonal covariance term twice (backward2d.cu:39-40, 103-105).  True: analytic.
    analytic_projection_grad: bool = False
    #: 256: rasterize_gaussians_sum renders the first 256 Gaussians of a tile like the reference
    #: (forward.cu:566-614).  0: no limit.
    max_per_tile: int = 256
    #: False: colours with channels != 3 raise, as in the reference whose N-channel functions are
    #: not bound (rasterize_sum.py:147-150, ext.cpp:4-24).  True: 1..32 channels are rendered, with
    #: no per-tile limit (the reference's N-channel kernel has none, forward.cu:664).
    nd_channels: bool = False
    #: "ranked": rasterize_gaussians_sum bins by per-tile counts + scatter and orders a tile's batch by
    #: Gaussian id inside the tile kernels (three kernels, nothing the host waits for, CUDA-graph
    #: capturable).  "onesweep": the step-by-step chain scan -> map -> custom onesweep radix sort ->
    #: bin edges, which materialises the sorted key / id arrays like the reference does.  Same results.
    binning: str = "ranked"
    #: True: a forward never waits for the device; its intersection count is read later from pinned
    #: memory (gsvc_b200.rasterize_sum module docstring: capacity, lazy overflow check).  False: every
    #: forward waits for its own count before it returns, like the reference's .item() (utils.py:117).
    sync_free: bool = True

    @contextmanager
    def override(self, **kw):
        old = {k: getattr(self, k) for k in kw}
        for k, v in kw.items():
            if not hasattr(Options, k):
                raise AttributeError(f"unknown option {k}")
            setattr(self, k, v)
        try:
            yield self
        finally:
            for k, v in old.items():
                setattr(self, k, v)


options = Options()
