"""`-m gpu`, needs two GPUs (skipped otherwise; run with `gpurun --gpus 2`, log under profiles/):
  (1) SURVEY section 4's ladder: the summed 2-rank gradients (each rank its own rays, pre-scaled by 1/2, exchanged by the
      library's own NVLink kernels over the regions of the buffer that can hold a gradient) equal ONE GPU's gradients of
      the concatenated batch to fp32 summation-order noise (1e-5 of the largest entry per tensor), for the p2p and the
      multimem kernel and for NCCL over the whole buffer;
  (2) the overlapped all-reduce inside train_step gives the same gradients as a plain step + one all-reduce."""
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import torch.distributed as dist
    import bench
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", device_id=dev)
    model = bench.build_model(dev, grid=128, mask_grid=128)
    S, n = 443, 2048
    parts = [bench.synthetic_rays(n, seed=10 + r) for r in range(world)]
    jit = [torch.rand(n, 1, generator=torch.Generator().manual_seed(20 + r)) for r in range(world)]

    def run(rays_h, target_h, jitter, scale, **kw):
        model._draw_rng = lambda n_rays, S_, is_train, ndc, wb, device: (jitter.reshape(-1).to(device), True)
        try:
            model.train_step(rays_h.to(dev), target_h.to(dev), None, white_bg=True, N_samples=S, grad_scale=scale, **kw)
        finally:
            del model._draw_rng

    def grads():
        torch.cuda.synchronize()
        return {k: p.grad.detach().clone() for k, p in model.named_parameters()}

    res = {}
    # plain step + NCCL over the whole buffer, and the overlapped variant inside train_step
    run(parts[rank][0], parts[rank][1], jit[rank], 1.0 / world)
    dist.all_reduce(model._gradbuf)
    res["nccl"] = grads()
    run(parts[rank][0], parts[rank][1], jit[rank], 1.0 / world, allreduce=True)
    res["overlapped"] = grads()
    # the library's own kernels on a symmetric-memory gradient buffer, only the regions that can hold a gradient
    from fewtensorf_b200.dist import SymmetricGradients
    sg = SymmetricGradients.create(model)
    frac = None
    if sg is not None:
        frac = sg.exchanged_floats / sg.storage.numel()
        for kind in ("p2p", "multimem", "p2p_fused", "multimem_fused"):      # *_fused: barriers inside the exchange kernel
            if kind.startswith("multimem") and not int(sg.handle.multicast_ptr):
                continue
            sg.kind = kind
            for _ in range(2):
                run(parts[rank][0], parts[rank][1], jit[rank], 1.0 / world)
                sg.all_reduce()
            assert model._gradbuf.untyped_storage().data_ptr() == sg.storage.untyped_storage().data_ptr()
            res[kind] = grads()
        # the exchange enqueued inside train_step, behind the backward and before the step's host wait
        for _ in range(2):
            run(parts[rank][0], parts[rank][1], jit[rank], 1.0 / world, allreduce=sg)
        res["in_step"] = grads()
        # wait="counters": the call returns behind the integration stage, the steps are enqueued back to back
        for _ in range(3):
            run(parts[rank][0], parts[rank][1], jit[rank], 1.0 / world, allreduce=sg, wait="counters")
        assert model.check_steps() == 0
        res["in_step_early"] = grads()
        # ... and a list overflowing on ONE rank: nobody repeats alone, both ranks count the step, the next one is clean
        if rank == 1:
            model._cap_hint[(n, S, True)] = (64, 1 << 20)
        for _ in range(2):
            run(parts[rank][0], parts[rank][1], jit[rank], 1.0 / world, allreduce=sg, wait="counters")
        assert model.check_steps() == 1, (rank, model.check_steps())
        res["in_step_early_after_overflow"] = grads()
        # ... and with the guarded optimizer behind the steps: the overflowed step is dropped on BOTH ranks, the clean
        # one is applied
        import fewtensorf_b200 as ft
        snap = [p.detach().clone() for p in model.parameters()]
        opt = ft.FusedAdam(list(model.parameters()), lr=1e-3, guard=model)
        if rank == 1:
            model._cap_hint[(n, S, True)] = (64, 1 << 20)
        run(parts[rank][0], parts[rank][1], jit[rank], 1.0 / world, allreduce=sg, wait="counters")
        opt.step()
        torch.cuda.synchronize()
        assert all(torch.equal(a, b) for a, b in zip(snap, model.parameters())), rank
        run(parts[rank][0], parts[rank][1], jit[rank], 1.0 / world, allreduce=sg, wait="counters")
        opt.step()
        torch.cuda.synchronize()
        assert not torch.equal(snap[0], next(model.parameters())), rank
        assert model.check_steps() == 2
        with torch.no_grad():
            for p, q in zip(model.parameters(), snap):
                p.copy_(q)
        # the captured step (one graph launch) with the exchange launched behind it
        jp = jit[rank].reshape(-1).pin_memory()
        model._draw_rng_host = lambda n_rays, S_, is_train, ndc, wb: (jp, True)
        try:
            for _ in range(3):
                model.train_step(parts[rank][0].to(dev), parts[rank][1].to(dev), None, white_bg=True, N_samples=S,
                                 grad_scale=1.0 / world, allreduce=sg, graph=True)
        finally:
            del model._draw_rng_host
        assert model.step_graph_info()["graphs"] == 1 and model.step_graph_info()["error"] is None
        res["in_step_graph"] = grads()
    dist.barrier()
    if rank == 0:
        # one GPU, the concatenated batch (mean over 2n rays = sum of the two pre-scaled rank means)
        run(torch.cat([p[0] for p in parts]), torch.cat([p[1] for p in parts]), torch.cat(jit), 1.0)
        ref = grads()
        errs = {}
        for name, g in res.items():
            errs[name] = max(float((g[k] - ref[k]).abs().max() / ref[k].abs().max().clamp_min(1e-30)) for k in ref)
        out.put((errs, float(sum(v.abs().sum() for v in ref.values())), frac))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_rank_gradients_equal_one_gpu_on_the_concatenated_batch():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, 29533, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    errs, norm, frac = out.get(timeout=10)
    print("gradient error vs one GPU on the concatenated batch:", errs, "| exchanged fraction of the buffer:", frac)
    assert norm > 0
    assert "nccl" in errs and "overlapped" in errs
    assert max(errs.values()) < 1e-5, errs
    if frac is not None:
        assert "p2p" in errs and "p2p_fused" in errs and frac < 0.6
        assert "in_step_early" in errs and "in_step_early_after_overflow" in errs and "in_step_graph" in errs
