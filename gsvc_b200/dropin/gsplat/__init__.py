"""Drop-in alias: put `<repo>/gsvc_b200/dropin` on PYTHONPATH and GSVC's
`from gsplat.project_gaussians_2d import project_gaussians_2d` /
`from gsplat.rasterize_sum import rasterize_gaussians_sum` resolve to gsvc_b200."""
import os
import sys

_repo = os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
if _repo not in sys.path:
    sys.path.insert(0, _repo)

import gsvc_b200 as _impl  # noqa: E402

_impl.install_as_gsplat()
