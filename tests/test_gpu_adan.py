"""GPU: the one-kernel Adan step (gs2d_adan_step) against the torch._foreach restatement of the
reference's multi-tensor Adan (optimizer.py:296-362 of the reference) on the same gradients."""
import pytest
import torch

pytestmark = pytest.mark.gpu
cuda = torch.device("cuda:0")


def _params(seed, shapes):
    g = torch.Generator().manual_seed(seed)
    return [torch.nn.Parameter(torch.randn(*s, generator=g).to(cuda)) for s in shapes]


@pytest.mark.parametrize("wd", [0.0, 0.02])
def test_fused_adan_follows_the_foreach_path(wd):
    from gsvc_b200.adan import Adan
    shapes = [(50_001, 2), (50_001, 3), (50_001, 3), (50_001, 1), (7,)]     # ragged tails, a tiny tensor
    a, b = _params(3, shapes), _params(3, shapes)
    oa, ob = Adan(a, lr=1e-3, weight_decay=wd, fused=True), Adan(b, lr=1e-3, weight_decay=wd, fused=False)
    g = torch.Generator().manual_seed(5)
    for it in range(12):
        for pa, pb in zip(a, b):
            gr = torch.randn(pa.shape, generator=g).to(cuda) * (0.1 if it % 3 else 1.0)
            pa.grad, pb.grad = gr.clone(), gr.clone()
        oa.step()
        ob.step()
        for pa, pb in zip(a, b):
            # one step moves a parameter by about lr; the two paths round differently (FMA contraction,
            # reciprocal of the bias correction): a few ulps of the update, far below 1e-6
            assert (pa - pb).abs().max().item() <= 2e-6, it
    for pa, pb in zip(a, b):
        for k in ("m", "n", "d", "prev"):
            sa, sb = oa.state[pa][k], ob.state[pb][k]
            assert (sa - sb).abs().max().item() <= 1e-5 * max(1.0, sb.abs().max().item()), k


def test_fused_adan_after_parameters_were_replaced():
    """the prune / add control swaps the parameter tensors: fresh state, previous gradient := gradient,
    while group['step'] keeps counting (GaussianSplats_Represent.py:115-116)"""
    from gsvc_b200.adan import Adan
    a, b = _params(1, [(1000, 3)]), _params(1, [(1000, 3)])
    oa, ob = Adan(a, fused=True), Adan(b, fused=False)
    g = torch.Generator().manual_seed(2)
    for it in range(6):
        if it == 3:
            for params, opt in ((a, oa), (b, ob)):
                params[0] = torch.nn.Parameter(params[0].detach()[:700].clone())
                opt.param_groups[0]["params"] = [params[0]]
        gr = torch.randn(a[0].shape, generator=g).to(cuda)
        a[0].grad, b[0].grad = gr.clone(), gr.clone()
        oa.step()
        ob.step()
        assert (a[0] - b[0]).abs().max().item() <= 2e-6, it
    assert oa.param_groups[0]["step"] == 6


def test_default_picks_the_kernel_on_cuda_and_counts_one_launch():
    from gsvc_b200 import _lib
    from gsvc_b200.adan import Adan
    a = _params(4, [(5000, 2), (5000, 3), (5000, 3)])
    opt = Adan(a)
    for p in a:
        p.grad = torch.randn_like(p)
    before = _lib.launch_count()
    opt.step()
    assert _lib.launch_count() - before == 1
