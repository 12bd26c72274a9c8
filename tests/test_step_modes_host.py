"""CPU: the host-side book-keeping of `train_step(wait="counters")` with a gradient exchange (the exchanged overflow flag
of a step is looked at one call late, from one of two alternating pinned slots) and FusedAdam's guard argument — the
parts that need no GPU.  The kernels and the streams behind them: tests/test_gpu_step_graph.py, test_gpu_multi.py."""
import pytest
import torch

from oracle import field_oracle as O


def _model():
    import fewtensorf_b200 as ft
    from tests.gpu_util import args_from_cfg
    cfg = O.FieldConfig(model="VM", grid=(16, 16, 16), density_n_comp=(4, 4, 4), app_n_comp=(8, 8, 8))
    return ft.TensorVMSplit(args_from_cfg(cfg), torch.tensor(cfg.aabb), list(cfg.grid), list(cfg.near_far), "cpu")


def test_the_previous_steps_flag_is_counted_once():
    m = _model()
    m._tail_pins = torch.tensor([0.0, 3.0])           # slot 1 holds a raised flag (three ranks overflowed)
    m._settle_early()                                  # nothing pending: nothing counted
    assert m.__dict__.get("step_overflows", 0) == 0
    m._early_unchecked = 1
    m._settle_early()
    assert m.step_overflows == 1 and m._early_unchecked is None
    m._settle_early()                                  # looked at once
    assert m.step_overflows == 1
    m._early_unchecked = 0                             # a clean step
    m._settle_early()
    assert m.step_overflows == 1


def test_overflow_flag_needs_a_gradient_buffer_and_points_at_its_tail():
    from fewtensorf_b200.module import GRAD_TAIL
    m = _model()
    assert m.overflow_flag() is None
    buf, views = m.grad_buffer()
    f = m.overflow_flag()
    total = sum(p.numel() for p in m._all_params())
    assert buf.numel() == total + GRAD_TAIL and f.numel() == 1
    assert f.data_ptr() == buf.data_ptr() + total * 4
    f.fill_(1.0)
    assert float(buf[total]) == 1.0 and float(buf[:total].abs().sum()) == 0.0


def test_fused_adam_guard_argument_forms():
    import fewtensorf_b200 as ft
    p = torch.nn.Parameter(torch.zeros(8))
    assert ft.FusedAdam([p])._guard_tensor() is None
    assert ft.FusedAdam([p], guard=lambda: None)._guard_tensor() is None
    assert ft.FusedAdam([p], guard=_model())._guard_tensor() is None             # a model without a gradient buffer yet
    for bad in (torch.zeros(1), torch.zeros(2), lambda: torch.zeros(1, dtype=torch.float64)):     # not CUDA / not one element
        with pytest.raises(RuntimeError):
            ft.FusedAdam([p], guard=bad)._guard_tensor()
