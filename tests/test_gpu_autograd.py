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


def test_rng_draws_follow_the_reference_order_and_values():
    """SURVEY 0.8 / tensorBase.py:201, :380 and :183: per-ray jitter first, then the background coin, from the CPU
    generator — a seeded run samples exactly the reference's points (the jitter is drawn straight into the pinned
    staging buffer; that must not change the values)."""
    from tests.util import Golden
    g = Golden("vm_fea_train_mask")
    from tests.gpu_util import model_from_field
    m = model_from_field(g.field(), "cuda")
    for n in (1, 96, 5000):
        torch.manual_seed(123)
        want_j = torch.rand(n, 1).reshape(-1)
        want_coin = bool(torch.rand((1,)) < 0.5)
        after = torch.rand(3)
        torch.manual_seed(123)
        j, add_bg = m._draw_rng(n, 70, True, False, False, torch.device("cuda"))
        assert torch.equal(j.cpu(), want_j) and add_bg == want_coin
        assert torch.equal(torch.rand(3), after)            # the generator is left where the reference leaves it
        torch.manual_seed(123)
        j, add_bg = m._draw_rng(n, 70, True, False, True, torch.device("cuda"))   # white_bg: no coin is drawn
        assert torch.equal(j.cpu(), want_j) and add_bg is True
        nxt = torch.rand(1)
        assert torch.equal(nxt, _nth_rand(123, n))          # exactly n values consumed
    near, far = m.near_far
    torch.manual_seed(7)
    z = torch.linspace(near, far, 70).unsqueeze(0)
    want = (z + torch.rand_like(z) * ((far - near) / 70)).reshape(-1)
    torch.manual_seed(7)
    j, _ = m._draw_rng(5, 70, True, True, True, torch.device("cuda"))
    assert torch.equal(j.cpu(), want)
    j, add_bg = m._draw_rng(5, 70, False, False, False, torch.device("cuda"))
    assert j is None and add_bg is False


def _nth_rand(seed, skip):
    """First value the CPU generator yields after `skip` uniform draws."""
    torch.manual_seed(seed)
    torch.rand(skip)
    return torch.rand(1)
