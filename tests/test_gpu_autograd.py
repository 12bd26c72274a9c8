"""`-m gpu`: the autograd node around the kernels under call patterns the reference's training code can produce —
chunked calls that accumulate into the same .grad, and interleaved graphs (the gradient buffer is zeroed and
the backward arguments are marshalled during the forward; a later backward must notice when that is stale)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def small():
    import bench
    dev = torch.device("cuda", 0)
    model = bench.build_model(dev, grid=64, mask_grid=64)
    rays, target = bench.synthetic_rays(1024, seed=3)
    return model, rays.to(dev), target.to(dev)


def _grads(model):
    return {k: p.grad.detach().clone() for k, p in model.named_parameters()}


def _zero(model):
    for p in model.parameters():
        p.grad = None


def _loss(model, rays, target, seed):
    torch.manual_seed(seed)
    rgb, depth, nv = model(rays, None, white_bg=True, is_train=True, N_samples=221)
    return ((rgb - target) ** 2).sum()


def _close(a, b, tol=2e-5):
    for k in a:
        d = float((a[k] - b[k]).abs().max() / b[k].abs().max().clamp_min(1e-20))
        assert d < tol, (k, d)


@pytest.mark.parametrize("prepare", [True, False])
def test_two_graphs_accumulate_like_two_single_steps(small, prepare):
    model, rays, target = small
    model._no_prepare = not prepare
    h = rays.shape[0] // 2
    _zero(model); _loss(model, rays[:h], target[:h], 1).backward(); g1 = _grads(model)
    _zero(model); _loss(model, rays[h:], target[h:], 2).backward(); g2 = _grads(model)
    want = {k: g1[k] + g2[k] for k in g1}
    # both forwards first, then one backward through both graphs
    _zero(model)
    la = _loss(model, rays[:h], target[:h], 1)
    lb = _loss(model, rays[h:], target[h:], 2)
    (la + lb).backward()
    _close(_grads(model), want)
    # sequential backward calls accumulating into live .grad
    _zero(model)
    _loss(model, rays[:h], target[:h], 1).backward()
    _loss(model, rays[h:], target[h:], 2).backward()
    _close(_grads(model), want)


def test_stale_prepared_backward_is_detected(small):
    model, rays, target = small
    model._no_prepare = False
    h = rays.shape[0] // 2
    _zero(model); _loss(model, rays[:h], target[:h], 1).backward(); g1 = _grads(model)
    _zero(model)
    la = _loss(model, rays[:h], target[:h], 1)          # prepares (zeroes) the persistent buffer
    lb = _loss(model, rays[h:], target[h:], 2)
    lb.backward()                                       # writes the buffer
    _zero(model)                                        # drops those gradients
    la.backward()                                       # must not add to what lb left behind
    _close(_grads(model), g1)
    # and train_step in between
    _zero(model)
    la = _loss(model, rays[:h], target[:h], 1)
    model.train_step(rays[h:], target[h:], None, white_bg=True, N_samples=221)
    _zero(model)
    la.backward()
    _close(_grads(model), g1)
