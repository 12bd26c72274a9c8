"""World-size-2 gloo test (CPU) of the multi-GPU host logic: ray sharding + gradient scaling + one
all-reduce of the flat gradient buffer must reproduce the single-process gradients of the global
batch; row-tile rendering must reassemble the image.  Gradients come from the CPU oracle here (the CUDA
kernels need a GPU); the same flow runs on NCCL in bench.py --gpus N."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import field_oracle as O


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.set_num_threads(1)
    from fewtensorf_b200 import dist as fd
    r, w = fd.init_from_env("gloo")
    assert (r, w) == (rank, world)
    cfg = O.FieldConfig(grid=(16, 16, 16), density_n_comp=(8, 8, 8), app_n_comp=(12, 12, 12))
    p = O.init_params(cfg, seed=3)
    O.plant_box(cfg, p)
    f = O.Field(cfg, p)
    rays, target = O.synthetic_rays(50, seed=4), torch.rand(50, 3, generator=torch.Generator().manual_seed(5))
    idx = fd.shard_indices(50, rank, world)
    _, _, _, _, grads = O.loss_and_grads(f, rays[idx], target[idx], is_train=False, n_samples=40)
    names = sorted(grads)
    flat = torch.cat([grads[k].reshape(-1) for k in names]) * fd.grad_scale(len(idx), 50)
    fd.all_reduce_gradients(flat)
    # tile-split "render": every rank fills its rows with their row index
    H, W = 37, 5
    rows = fd.row_tiles(H, rank, world, block=4)
    img = fd.gather_rows(rows.float()[:, None, None].expand(-1, W, 3).contiguous(), H, rank, world, block=4)
    if rank == 0:
        torch.save({"flat": flat, "names": names, "img": img}, os.path.join(out_dir, "r0.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gradient_allreduce_and_row_tiles(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    got = torch.load(os.path.join(tmp_path, "r0.pt"))
    cfg = O.FieldConfig(grid=(16, 16, 16), density_n_comp=(8, 8, 8), app_n_comp=(12, 12, 12))
    p = O.init_params(cfg, seed=3)
    O.plant_box(cfg, p)
    f = O.Field(cfg, p)
    rays, target = O.synthetic_rays(50, seed=4), torch.rand(50, 3, generator=torch.Generator().manual_seed(5))
    _, _, _, _, grads = O.loss_and_grads(f, rays, target, is_train=False, n_samples=40)
    ref = torch.cat([grads[k].reshape(-1) for k in got["names"]])
    assert float((got["flat"] - ref).abs().max() / ref.abs().max()) < 1e-5
    assert torch.equal(got["img"][:, 0, 0], torch.arange(37).float())


def test_shard_helpers():
    from fewtensorf_b200 import dist as fd
    parts = [fd.shard_indices(10, r, 4) for r in range(4)]
    assert sorted(torch.cat(parts).tolist()) == list(range(10))
    rows = [fd.row_tiles(20, r, 3, block=2) for r in range(3)]
    assert sorted(torch.cat(rows).tolist()) == list(range(20))
    assert fd.grad_scale(3, 12) == 0.25
