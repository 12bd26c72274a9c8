"""`-m gpu`: `train_step(graph=True)` — the step's launch sequence captured once in a CUDA graph and replayed — against the
same step enqueued call by call: same loss, sample counts and gradients on the same jitter, step after step (the jitter is
drawn per step on the host and copied inside the graph), for device and host inputs, after the field changed under the
graph (new alphaMask, upsampled factors), and when a replay overflows the list capacities baked into it.  And
`train_step(wait="counters")` — the call returns behind the integration stage — against the step the host waits for."""
import pytest
import torch

from oracle import field_oracle as O
from tests.util import rel_inf

pytestmark = pytest.mark.gpu
S = 160


def _model(model="VM", shading="MLP_Fea", seed=3):
    from tests.gpu_util import model_from_field
    comps = (16, 8, 24) if model == "VM" else (48, 48, 48)
    cfg = O.FieldConfig(model=model, shading=shading, grid=(40, 48, 56), density_n_comp=(16, 8, 24) if model == "VM" else (32, 32, 32),
                        app_n_comp=comps)
    p = O.init_params(cfg, seed=seed)
    O.plant_box(cfg, p)
    m = model_from_field(O.Field(cfg, p), "cuda")
    m.updateAlphaMask((32, 32, 32))
    return m


def _batch(n, seed):
    rays = O.synthetic_rays(n, seed=seed)
    target = torch.rand(n, 3, generator=torch.Generator().manual_seed(seed + 1))
    return rays, target


def _grads(m):
    return {k: p.grad.detach().clone() for k, p in m.named_parameters()}


def _both(m, rays, target, seed, **kw):
    """The step call by call and through the graph on the same host RNG state: (loss, n_valid, gradients) of each."""
    out = []
    for graph in (False, True):
        torch.manual_seed(seed)
        loss, nv = m.train_step(rays, target, None, N_samples=S, graph=graph, **kw)
        torch.cuda.synchronize()
        out.append((float(loss), nv, _grads(m)))
    return out


def _same(a, b, tol=2e-5):
    assert a[1] == b[1] and a[1] > 500
    assert abs(a[0] - b[0]) <= 1e-6 * max(1.0, abs(a[0]))
    for k in a[2]:
        assert rel_inf(b[2][k], a[2][k]) < tol, k


@pytest.mark.parametrize("model,shading", [("VM", "MLP_Fea"), ("VM", "SH"), ("CP", "MLP_Fea")])
def test_replayed_step_equals_the_step_call_by_call(model, shading):
    m = _model(model, shading)
    rays, target = (t.cuda() for t in _batch(1500, 5))
    from fewtensorf_b200._abi import lib
    for i in range(4):                                  # first pass captures, the others replay; new jitter every step
        n0 = lib.ftrf_launch_count()
        eager, graphed = _both(m, rays, target, seed=20 + i)
        n1 = lib.ftrf_launch_count()
        _same(eager, graphed)
        if i:
            assert n1 - n0 == 2 * m.step_graph_info()["launches"]        # a replay counts the kernels it runs
    info = m.step_graph_info()
    assert info["graphs"] == 1 and info["launches"] >= 8 and info["error"] is None
    # white_bg=False: the background coin is a scalar of the call — one graph per outcome, both right
    for i in range(4):
        _same(*_both(m, rays, target, seed=40 + i, white_bg=False))
    assert m.step_graph_info()["graphs"] in (2, 3)


def test_inputs_host_or_device_new_every_step_or_the_resident_buffers():
    m = _model()
    n = 1200
    for i in range(3):                                  # a new device batch every step
        rays, target = (t.cuda() for t in _batch(n, 50 + i))
        _same(*_both(m, rays, target, seed=60 + i))
    rays, target = (t.pin_memory() for t in _batch(n, 70))             # pinned host tensors, copied by the step itself
    _same(*_both(m, rays, target, seed=71))
    rb, tb = m.step_buffers(n)                          # the step's own input buffers, filled by the caller
    r2, t2 = _batch(n, 80)
    rb.copy_(r2)
    tb.copy_(t2)
    torch.manual_seed(81)
    loss, nv = m.train_step(rb, tb, None, N_samples=S, graph=True)
    got = (float(loss), nv, _grads(m))
    torch.manual_seed(81)
    loss, nv = m.train_step(r2.cuda(), t2.cuda(), None, N_samples=S)
    _same((float(loss), nv, _grads(m)), got)
    assert m.step_graph_info()["graphs"] == 1


def test_a_changed_field_captures_anew():
    m = _model()
    rays, target = (t.cuda() for t in _batch(1000, 7))
    _same(*_both(m, rays, target, seed=1))
    _same(*_both(m, rays, target, seed=2))
    m.updateAlphaMask((24, 24, 24))                    # new occupancy bits: another pointer and serial
    _same(*_both(m, rays, target, seed=3))
    m.upsample_volume_grid((56, 64, 72))               # new parameters, new gradient buffer
    _same(*_both(m, rays, target, seed=4))
    _same(*_both(m, rays, target, seed=5))
    with torch.no_grad():                               # in-place parameter updates (an optimizer step) keep the graph
        for p in m.parameters():
            p.mul_(0.97)
    g0 = m.step_graph_info()["graphs"]
    _same(*_both(m, rays, target, seed=6))
    assert m.step_graph_info()["graphs"] == g0


def test_a_replay_that_overflows_its_lists_repeats_call_by_call():
    m = _model()
    n = 600
    miss = O.synthetic_rays(n, seed=9).cuda()
    miss[:, :3] = torch.tensor([40.0, 40.0, 40.0], device="cuda")      # origins far outside, looking away: no samples
    miss[:, 3:6] = torch.nn.functional.normalize(torch.rand(n, 3, device="cuda") + 0.1, dim=-1)
    hit, target = (t.cuda() for t in _batch(n, 10))
    m._cap_hint.clear()
    m._cap_hint[(n, S, True)] = (8, 4)
    torch.manual_seed(1)
    loss, nv = m.train_step(miss, target, None, N_samples=S, graph=True)     # captured with lists of ~1 k entries
    assert nv == 0 and m.step_graph_info()["graphs"] == 1
    caps = m._cap_hint[(n, S, True)]
    assert caps[0] < 4096
    torch.manual_seed(2)
    loss, nv = m.train_step(hit, target, None, N_samples=S, graph=True)      # the replay overflows: repeated, new graph
    got = (float(loss), nv, _grads(m))
    assert m._cap_hint[(n, S, True)][0] > caps[0]
    torch.manual_seed(2)
    loss, nv = m.train_step(hit, target, None, N_samples=S)
    _same((float(loss), nv, _grads(m)), got)
    _same(*_both(m, hit, target, seed=3))


def test_steps_the_host_does_not_wait_for_equal_the_waited_ones():
    m = _model()
    rays, target = (t.cuda() for t in _batch(1500, 5))
    want = []
    for i in range(5):
        torch.manual_seed(90 + i)
        loss, nv = m.train_step(rays, target, None, N_samples=S)
        want.append((float(loss), nv, _grads(m)))
    got = []
    for i in range(5):                                  # enqueued back to back: nothing waits for the end of a step
        torch.manual_seed(90 + i)
        loss, nv = m.train_step(rays, target, None, N_samples=S, wait="counters")
        got.append((loss, nv, {k: p.grad.detach().clone() for k, p in m.named_parameters()}))     # stream-ordered copies
    assert m.check_steps() == 0
    for a, b in zip(want, got):
        _same(a, (float(b[0]), b[1], b[2]))
    # an overflowing step is repeated inside the call, as with wait="step"
    m._cap_hint[(1500, S, True)] = (16, 8)
    torch.manual_seed(94)
    loss, nv = m.train_step(rays, target, None, N_samples=S, wait="counters")
    _same(want[4], (float(loss), nv, _grads(m)))
    assert m._cap_hint[(1500, S, True)][0] > 16
    with pytest.raises(ValueError):
        m.train_step(rays, target, None, N_samples=S, wait="never")
