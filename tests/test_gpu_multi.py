"""`-m gpu`, needs two GPUs (skipped otherwise): the overlapped gradient all-reduce inside train_step gives the
same gradients as a plain step followed by one all-reduce of the flat buffer."""
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
    rays, target = bench.synthetic_rays(2048, seed=rank)
    rays, target = rays.to(dev), target.to(dev)
    S = 443
    res = []
    for overlapped in (False, True, True):
        torch.manual_seed(5 + rank)
        loss, nv = model.train_step(rays, target, None, white_bg=True, N_samples=S, grad_scale=0.5, allreduce=overlapped)
        if not overlapped:
            dist.all_reduce(model._gradbuf)
        torch.cuda.synchronize()
        res.append(model._gradbuf[:-4].clone())
    # the library's NVLS multicast all-reduce on a symmetric-memory gradient buffer
    from fewtensorf_b200.dist import SymmetricGradients
    sg = SymmetricGradients.create(model)
    if sg is not None:
        for kind in ("p2p", "multimem"):
            if kind == "multimem" and not int(sg.handle.multicast_ptr):
                continue
            sg.kind = kind
            for _ in range(2):
                torch.manual_seed(5 + rank)
                model.train_step(rays, target, None, white_bg=True, N_samples=S, grad_scale=0.5)
                sg.all_reduce()
                torch.cuda.synchronize()
            assert model._gradbuf.untyped_storage().data_ptr() == sg.storage.untyped_storage().data_ptr()
            res.append(model._gradbuf[:-4].clone())
    # atomics make the summation order differ run to run: compare in the L2 norm
    err = [float((res[i] - res[0]).norm() / res[0].norm()) for i in range(1, len(res))]
    if rank == 0:
        out.put((err, float(res[0].norm()), sg is not None))
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_overlapped_allreduce_matches_plain():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, 29533, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    err, norm, nvls = out.get(timeout=10)
    assert norm > 0
    assert max(err) < 1e-5, err
    print("symmetric-memory all-reduce kernels exercised:", nvls)
