"""Column sharding of a radiation call over the GPUs of one box.

Columns are independent in every stage of the path, so a domain of `ncol` columns is cut into
`world` contiguous shards of the x-fastest flattened domain (reference flattening:
src_test/test_rte_rrtmgp.cpp:217); tables are replicated; there is no collective inside the
computation.  The only exchange is the final gather of the broadband flux slabs to one rank
(NCCL on GPUs; any torch.distributed backend works, the CPU tests use gloo).

A flux field is a column-major (ncol_shard, nlev) array, i.e. a C-contiguous (nlev, ncol_shard)
tensor.  `gather_fields` moves all fields of a rank in ONE collective; `assemble` lays the
shards out as the reference's (ncol, nlev) column-major arrays.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_bounds(ncol: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous shard [c0, c1) of rank `rank`; sizes differ by at most one column."""
    if not (0 <= rank < world) or ncol < 0:
        raise ValueError("shard_bounds: bad rank/world/ncol")
    return rank * ncol // world, (rank + 1) * ncol // world


def pack_fields(fields: Sequence[torch.Tensor]) -> torch.Tensor:
    """Stack equal-sized flat flux fields into one (nfield, nlev*ncol_shard) buffer."""
    return torch.stack([f.reshape(-1) for f in fields])


def gather_fields(packed: torch.Tensor, ncol: int, nlev: int, dst: int = 0,
                  recv: Optional[List[torch.Tensor]] = None) -> Optional[List[torch.Tensor]]:
    """Gather every rank's packed fields on `dst`.  Shards may be ragged (ncol % world != 0):
    buffers are padded to the largest shard.  Returns the per-rank buffers on dst, else None."""
    world, rank = dist.get_world_size(), dist.get_rank()
    nfield = packed.shape[0]
    maxc = max(shard_bounds(ncol, world, r)[1] - shard_bounds(ncol, world, r)[0] for r in range(world))
    c0, c1 = shard_bounds(ncol, world, rank)
    if (c1 - c0) != maxc:
        pad = torch.zeros((nfield, nlev * maxc), dtype=packed.dtype, device=packed.device)
        pad.view(nfield, nlev, maxc)[:, :, :c1 - c0] = packed.view(nfield, nlev, c1 - c0)
        packed = pad
    if rank == dst and recv is None:
        recv = [torch.empty_like(packed) for _ in range(world)]
    dist.gather(packed, recv if rank == dst else None, dst=dst)
    return recv if rank == dst else None


def assemble(recv: Sequence[torch.Tensor], ncol: int, nlev: int) -> torch.Tensor:
    """Per-rank (nfield, nlev*maxc) buffers -> (nfield, nlev, ncol): every field as the
    reference's column-major (ncol, nlev) array."""
    world = len(recv)
    nfield = recv[0].shape[0]
    maxc = recv[0].shape[1] // nlev
    out = torch.empty((nfield, nlev, ncol), dtype=recv[0].dtype, device=recv[0].device)
    for r in range(world):
        c0, c1 = shard_bounds(ncol, world, r)
        out[:, :, c0:c1] = recv[r].view(nfield, nlev, maxc)[:, :, :c1 - c0]
    return out


def domain_mean(profile_sum: torch.Tensor, ncol: int) -> torch.Tensor:
    """All-reduce of per-rank column sums (nlev,) -> domain-mean profile on every rank."""
    dist.all_reduce(profile_sum, op=dist.ReduceOp.SUM)
    return profile_sum / ncol


class ShardPlan:
    """rrtmgp_b200_shard_plan (C ABI): this process's column shard of an `ncol`-column domain plus the NCCL communicator the
    flux gather runs on.  The communicator is created inside the library from a unique id that rank 0 obtains
    (rrtmgp_b200_nccl_unique_id) and torch.distributed only carries to the other ranks; world == 1 needs neither."""

    def __init__(self, ncol: int, world: int = 1, rank: int = 0, device: Optional[torch.device] = None):
        import ctypes as C
        from . import capi
        self._C, self._capi, self.L = C, capi, capi.lib()
        self.ncol, self.world, self.rank = int(ncol), int(world), int(rank)
        idbuf = None
        if world > 1:
            # the library resolves NCCL with dlopen: point it at the copy PyTorch ships unless the caller chose one
            import os
            if "RB200_NCCL_LIB" not in os.environ:
                try:
                    import nvidia.nccl
                    cand = os.path.join(list(nvidia.nccl.__path__)[0], "lib", "libnccl.so.2")
                    if os.path.exists(cand):
                        os.environ["RB200_NCCL_LIB"] = cand
                except Exception:
                    pass
            raw = (C.c_ubyte * 128)()
            if rank == 0:
                capi.check(self.L.rrtmgp_b200_nccl_unique_id(raw))
            t = torch.tensor(list(raw), dtype=torch.uint8, device=device if device is not None else "cuda")
            dist.broadcast(t, src=0)
            idbuf = (C.c_ubyte * 128)(*t.cpu().tolist())
        self.h = C.c_void_p()
        capi.check(self.L.rrtmgp_b200_shard_plan_create(self.ncol, self.world, self.rank, idbuf, None, C.byref(self.h)))

    def bounds(self, rank: int = -1) -> Tuple[int, int]:
        C = self._C
        a, b = C.c_longlong(), C.c_longlong()
        self._capi.check(self.L.rrtmgp_b200_shard_plan_bounds(self.h, int(rank), C.byref(a), C.byref(b)))
        return a.value, b.value

    def register(self, full_fields: Optional[Sequence[torch.Tensor]], nfields: int, root: int = 0) -> int:
        """Collective: peer-memory mode of gather() for these whole-domain arrays of the root (rrtmgp_b200_shard_plan_register).
        Returns the number of registered fields (0: peer mapping unavailable, gather() keeps the NCCL path)."""
        C = self._C
        dst = (C.c_void_p * nfields)(*[f.data_ptr() for f in full_fields]) if full_fields is not None else None
        s = torch.cuda.current_stream()
        self._capi.check(self.L.rrtmgp_b200_shard_plan_register(self.h, int(nfields), dst, int(root), C.c_void_p(s.cuda_stream)))
        return int(self.L.rrtmgp_b200_shard_plan_registered(self.h))

    def gather(self, part_fields: Sequence[torch.Tensor], full_fields: Optional[Sequence[torch.Tensor]], nlev: int,
               part: int = 0, nparts: int = 1, root: int = 0, stream=None) -> None:
        """part_fields[f]: this rank's (count, nlev) column-major slab of field f; full_fields[f] on the root: (ncol, nlev)."""
        C = self._C
        n = len(part_fields)
        dt = part_fields[0].dtype
        src = (C.c_void_p * n)(*[f.data_ptr() for f in part_fields])
        dst = (C.c_void_p * n)(*[f.data_ptr() for f in full_fields]) if full_fields is not None else None
        s = stream if stream is not None else torch.cuda.current_stream()
        self._capi.check(self.L.rrtmgp_b200_gather_fluxes(self.h, self._capi.F32 if dt == torch.float32 else self._capi.F64, int(nlev), n,
                                                          src, dst, int(part), int(nparts), int(root), C.c_void_p(s.cuda_stream)))

    def close(self) -> None:
        if getattr(self, "h", None) is not None and self.h.value:
            self.L.rrtmgp_b200_shard_plan_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
