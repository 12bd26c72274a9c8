"""Multi-GPU plumbing for the per-ray path.  One process per GPU; rays shard naturally,
parameters are replicated, and the only collective of a training step is one all-reduce(sum) of the flat
gradient buffer the backward kernels wrote into.  Test-view rendering splits the image by interleaved
rows and needs one all-gather of the finished tiles.  The reference has no multi-GPU code at all.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def init_from_env(backend: str | None = None) -> tuple[int, int]:
    """Join the process group torchrun described; returns (rank, world).  No-op for a single process."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    if world > 1 and not dist.is_initialized():
        backend = backend or ("nccl" if torch.cuda.is_available() else "gloo")
        kw = {}
        if backend == "nccl":
            os.environ.setdefault("NCCL_MIN_CTAS", "64")      # more CTAs for the one large all-reduce of a step
            local = int(os.environ.get("LOCAL_RANK", "0"))
            torch.cuda.set_device(local)
            kw["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend, **kw)
    return rank, world


def shard_indices(n: int, rank: int, world: int) -> torch.Tensor:
    """Rank-strided slice of a global batch of n rays (every rank gets floor/ceil(n/world) rays)."""
    return torch.arange(rank, n, world)


def grad_scale(n_local: int, n_global: int) -> float:
    """The loss is a mean over the GLOBAL batch (train.py:338): a rank that back-propagates the mean over its
    own n_local rays must scale by n_local / n_global before the sum-reduce."""
    return float(n_local) / float(n_global)


def all_reduce_gradients(flat_grad: torch.Tensor, group=None) -> torch.Tensor:
    """One all-reduce(sum) over the flat fp32 gradient buffer (TensorBase.grad_buffer()[0])."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(flat_grad, op=dist.ReduceOp.SUM, group=group)
    return flat_grad


class SymmetricGradients:
    """Flat gradient buffer of `model` in symmetric memory + all-reduce by the library's own kernels over
    NVLink: `ftrf_p2p_allreduce_regions` (rank r sums its share of every GPU's copy with peer loads and stores it
    into every copy) or `ftrf_multimem_allreduce_regions` (the same through the NVSwitch multicast address,
    reduction in the switch), bracketed by two cross-GPU barriers on the signal pads.  Only the regions of the
    buffer that can hold a gradient travel (`TensorBase.grad_regions`: rows / columns of each factor inside the
    alphaMask's texel bounding box; 16 % of the planes in the un-shrunk benchmark field).  The constructor times both and NCCL
    on the real buffer and keeps the fastest (2 x B200, 71 MB: p2p 129 us, NCCL 148-166 us, multimem 220 us).
    `SymmetricGradients.create` returns None where symmetric memory cannot be set up and callers use
    `all_reduce_gradients` (NCCL) instead.

        sg = SymmetricGradients.create(model)          # collective call, once (again after upsample / shrink)
        loss, nv = model.train_step(rays, target, grad_scale=grad_scale(n_local, n_global))
        sg.all_reduce() if sg else all_reduce_gradients(model.grad_buffer()[0])
    """

    def __init__(self, model, group=None):
        from ._abi import lib
        self._lib = lib
        pg = group if group is not None else dist.group.WORLD
        self.rank, self.world = dist.get_rank(pg), dist.get_world_size(pg)
        dev = model.basis_mat.weight.device
        # Phase 1, local only (may fail on one rank alone: torch without symmetric memory, allocation failure).
        # The ranks agree on the outcome BEFORE the first collective of this constructor, so a rank that failed
        # never leaves the others blocked in rendezvous.
        symm_mem, err = None, None
        try:
            import torch.distributed._symmetric_memory as symm_mem
            flat, _ = model.grad_buffer(fresh=True)
            n = (flat.numel() + 127) // 128 * 128
            del flat
            self.storage = symm_mem.empty(n, dtype=torch.float32, device=dev)
            self.storage.zero_()
        except Exception as e:                               # noqa: BLE001
            err = e
        ok = torch.tensor([0 if err is not None else 1], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
        if int(ok) == 0:
            raise RuntimeError(f"symmetric gradient storage unavailable on some rank ({err!r})")
        # Phase 2, collective on every rank
        self.handle = symm_mem.rendezvous(self.storage, pg.group_name)
        import ctypes as C
        self.group = group
        self._peer_ptrs = (C.c_void_p * self.world)(*[int(p) for p in self.handle.buffer_ptrs])
        # flags of the barriers the fused kernels run inside themselves (ftrf_allreduce_regions_fused)
        self._flags = symm_mem.empty(128, dtype=torch.int32, device=dev)
        self._flags.zero_()
        self._flag_handle = symm_mem.rendezvous(self._flags, pg.group_name)
        self._flag_ptrs = (C.c_void_p * self.world)(*[int(p) for p in self._flag_handle.buffer_ptrs])
        self._epoch = 0
        torch.cuda.synchronize()
        dist.barrier(group=group)                        # every rank's flags are zero before anybody signals
        self.model = model
        self._regions_key = None
        self._refresh_regions()
        self.kind = self._autotune()          # "p2p", "multimem" or "nccl", whichever is fastest here
        model.set_grad_storage(self.storage)

    def _refresh_regions(self) -> None:
        """(Re)build the table of buffer regions that can hold a gradient (TensorBase.grad_regions): only those are
        exchanged.  Keyed on the alphaMask / grid serials, so a mask rebuild or an upsample is picked up on the next
        all_reduce; every rank derives the same table from the same (replicated) mask."""
        import ctypes as C
        from ._abi import MAX_REGIONS, FtrfRegion
        m = self.model
        key = (m._mask_serial(), m._host["serial"])
        if key == self._regions_key:
            return
        regs = m.grad_regions(self.storage.numel())
        ok = len(regs) <= MAX_REGIONS and all(r[0] % 4 == 0 and r[1] % 4 == 0 and r[2] % 4 == 0 for r in regs)
        if not ok:                                          # (odd component counts: exchange the whole buffer)
            n = self.storage.numel()
            regs = [(0, n, n, 1)]
        self._regions = (FtrfRegion * len(regs))(*[FtrfRegion(*[int(v) for v in r]) for r in regs])
        self._n_regions = len(regs)
        self.exchanged_floats = sum(int(r[1]) * int(r[3]) for r in regs)
        self._regions_key = key

    def describe(self) -> str:
        return (f"{self.kind} over {self.exchanged_floats * 4 / 1e6:.1f} of {self.storage.numel() * 4 / 1e6:.1f} MB "
                f"(rows/columns inside the alphaMask's texel box + basis/MLP; autotuned, ms: {self.timings_ms})")

    @classmethod
    def create(cls, model, group=None):
        """Collective.  None if any rank cannot set up multicast symmetric memory (all ranks agree)."""
        ok, sg = 1, None
        try:
            sg = cls(model, group)      # raises on every rank together when phase 1 failed anywhere (see __init__)
        except Exception:                                    # noqa: BLE001 — old driver / no NVSwitch / torch without symm_mem
            ok = 0
        # rendezvous / autotune failures (driver without multicast) are symmetric, but agree once more before use
        flag = torch.tensor([ok], device=model.basis_mat.weight.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        if int(flag) == 0:
            if sg is not None:
                model.set_grad_storage(None)
            return None
        return sg

    def _launch(self, kind: str) -> None:
        from ._abi import check, current_stream_handle
        st = current_stream_handle()
        if kind == "nccl":
            dist.all_reduce(self.storage, op=dist.ReduceOp.SUM, group=self.group)
            return
        self._refresh_regions()
        if kind in ("p2p_fused", "multimem_fused"):
            self._epoch += 1
            mc = int(self.handle.multicast_ptr) if kind == "multimem_fused" else None
            check(self._lib.ftrf_allreduce_regions_fused(self._peer_ptrs, mc, self._flag_ptrs, self._regions, self._n_regions,
                                                         self.rank, self.world, self._epoch, st), "allreduce_regions_fused")
            return
        self.handle.barrier(channel=0)                       # every rank's backward has written its gradients
        if kind == "multimem":
            check(self._lib.ftrf_multimem_allreduce_regions(int(self.handle.multicast_ptr), self._regions, self._n_regions,
                                                            self.rank, self.world, st), "multimem_allreduce_regions")
        else:
            check(self._lib.ftrf_p2p_allreduce_regions(self._peer_ptrs, self._regions, self._n_regions, self.rank,
                                                       self.world, st), "p2p_allreduce_regions")
        self.handle.barrier(channel=1)                       # every slice is back in every copy

    def _autotune(self) -> str:
        """Time the candidates on the real buffer (all zeros at this point) and pick the fastest; the choice is
        made from the max over ranks so every rank picks the same one."""
        kinds = ["nccl"]
        if int(self.handle.multicast_ptr):
            kinds.append("multimem")
        if self.world in (2, 4, 8):
            kinds.append("p2p")
            kinds.append("p2p_fused")
            if int(self.handle.multicast_ptr):
                kinds.append("multimem_fused")
        times = []
        for kind in kinds:
            for _ in range(3):
                self._launch(kind)
            torch.cuda.synchronize()
            dist.barrier(group=self.group)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                self._launch(kind)
            e1.record()
            torch.cuda.synchronize()
            times.append(e0.elapsed_time(e1) / 5)
        t = torch.tensor(times, device=self.storage.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        self.timings_ms = dict(zip(kinds, [float(x) for x in t]))
        return kinds[int(torch.argmin(t))]

    def all_reduce(self) -> torch.Tensor:
        """Sum the gradient buffer over the group, in place, on the current stream."""
        self._launch(self.kind)
        return self.storage


def row_tiles(height: int, rank: int, world: int, block: int = 8) -> torch.Tensor:
    """Interleaved row blocks of an image for tile-split rendering: rank r renders blocks r, r+world, ...
    (tiles through the object carry more surviving samples, so contiguous tiles would be imbalanced)."""
    rows = torch.arange(height)
    return rows[(rows // block) % world == rank]


def gather_rows(local: torch.Tensor, height: int, rank: int, world: int, block: int = 8, group=None) -> torch.Tensor:
    """Reassemble [H, W, C] from every rank's `row_tiles` part ([rows_r, W, C])."""
    if world == 1:
        return local
    counts = [int(row_tiles(height, r, world, block).numel()) for r in range(world)]
    pad = max(counts)
    buf = torch.zeros((pad,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    buf[: local.shape[0]] = local
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf, group=group)
    out = torch.empty((height,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    for r in range(world):
        out[row_tiles(height, r, world, block).to(local.device)] = parts[r][: counts[r]]
    return out
