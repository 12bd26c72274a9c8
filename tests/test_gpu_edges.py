"""`-m gpu` edge cases of the per-ray path, checked against the CPU oracle (tests/util.Golden supplies a pinned
field): empty and ragged batches, single rays, rays that miss the box, S below / above a 32-sample chunk and
above 1024 (several chunks per occupancy flag), list capacities that overflow and grow, chunked versus whole
rendering, host / non-contiguous input."""
import pytest
import torch

from oracle import field_oracle as O
from tests.util import Golden, rel_inf

pytestmark = pytest.mark.gpu
TOL = 1e-4


@pytest.fixture(scope="module")
def pair():
    from tests.gpu_util import model_from_field
    g = Golden("vm_fea_train_mask")
    f = g.field()
    return g, f, model_from_field(f)


def _rays(n, seed, miss=False):
    gen = torch.Generator().manual_seed(seed)
    o = torch.randn(n, 3, generator=gen)
    o = 4.0 * o / o.norm(dim=-1, keepdim=True)
    tgt = (torch.rand(n, 3, generator=gen) - 0.5) * 2.0
    d = tgt - o
    if miss:                                   # start outside the box and walk away from it
        o = 5.0 + torch.rand(n, 3, generator=gen)
        d = 1.0 + torch.rand(n, 3, generator=gen)
    return torch.cat([o, d / d.norm(dim=-1, keepdim=True)], -1)


def _check(f, m, rays, S, is_train=False, seed=0):
    from tests.gpu_util import pin_rng
    jitter = torch.rand(rays.shape[0], 1, generator=torch.Generator().manual_seed(seed)) if is_train else None
    pin_rng(m, jitter, False, True)
    want_rgb, want_depth, want_nv = O.render_chunk(f, rays, None, white_bg=True, is_train=is_train, n_samples=S,
                                                   jitter=jitter, bg_coin=False)[:3]
    with torch.no_grad():
        rgb, depth, nv = m(rays.cuda(), None, white_bg=True, is_train=is_train, N_samples=S)
    assert rgb.shape == (rays.shape[0], 3) and depth.shape == (rays.shape[0],)
    if rays.shape[0]:
        assert rel_inf(rgb.cpu(), want_rgb) < TOL and rel_inf(depth.cpu(), want_depth) < TOL
    assert abs(int(nv) - int(want_nv)) <= max(2, int(want_nv) // 500)
    return rgb, depth, int(nv)


@pytest.mark.parametrize("n", [0, 1, 31, 33, 257])
def test_ragged_batch_sizes(pair, n):
    g, f, m = pair
    _check(f, m, _rays(n, 11 + n), 70)


@pytest.mark.parametrize("S", [1, 7, 32, 33, 1100, 2300])
def test_sample_counts(pair, S):
    g, f, m = pair
    _check(f, m, _rays(48, 5), S, is_train=(S % 2 == 1), seed=S)


def test_rays_missing_the_box(pair):
    g, f, m = pair
    rgb, depth, nv = _check(f, m, _rays(64, 3, miss=True), 70)
    assert nv == 0 and torch.all(rgb == 1.0)            # white background only


def test_overflowing_lists_grow_and_repeat(pair):
    g, f, m = pair
    rays = _rays(300, 21)
    m._cap_hint.clear()
    m._cap_hint[(300, 70, False)] = (8, 4)               # absurdly small guesses: the call must retry, not truncate
    _check(f, m, rays, 70)
    assert m._cap_hint[(300, 70, False)][0] > 8


def test_chunked_render_equals_whole(pair):
    import fewtensorf_b200 as ft
    g, f, m = pair
    from tests.gpu_util import pin_rng
    pin_rng(m, None, False, True)
    rays = _rays(1000, 8)
    whole = ft.OctreeRender_trilinear_fast(rays, m, None, chunk=4096, N_samples=-1, white_bg=True, is_train=False, device="cuda")
    # a host tensor, non-contiguous view, in small chunks
    wide = torch.zeros(1000, 12)
    wide[:, ::2] = rays
    parts = ft.OctreeRender_trilinear_fast(wide[:, ::2], m, None, chunk=96, N_samples=-1, white_bg=True, is_train=False,
                                           device="cuda")
    assert torch.equal(whole[0], parts[0]) and torch.equal(whole[2], parts[2]) and whole[5] == parts[5]


def test_no_alpha_mask(pair):
    from tests.gpu_util import model_from_field
    g, f, m = pair
    f2 = O.Field(f.cfg, {k: v.clone() for k, v in f.params.items()}, None)
    m2 = model_from_field(f2)
    _check(f2, m2, _rays(40, 2), 70, is_train=True, seed=4)
