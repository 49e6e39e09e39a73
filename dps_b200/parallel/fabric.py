"""Symmetric arenas over CUDA IPC (the on-box transport).

One arena per rank holds every peer-visible buffer at the same offsets on every rank, so a
single 64-byte IPC handle per rank is all that has to travel.  The exchange uses
``torch.distributed`` (NCCL/gloo) once, at start-up -- bootstrap only; after that the PS
kernels address peers directly.  Replaces the reference's TCP channel + RegisterWorker
handshake on the box (reference: src/workers/worker.py:199-231, src/parameter_server/server.py:190-211):
rank == worker id == shard id == GPU index, so ids never grow without bound.

``emulate=True`` builds all W "ranks" inside one process on one GPU (no IPC); the kernels then run
with rank_lo=0, rank_hi=W in a single launch.  This is how multi-rank logic is tested on one GPU.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from ..ops import _lib as L

_TYPESTR = {torch.float32: "<f4", torch.float16: "<f2", torch.uint16: "<u2", torch.int16: "<i2",
            torch.uint32: "<u4", torch.int32: "<i4", torch.uint8: "|u1"}


class _RawCuda:
    """Minimal __cuda_array_interface__ carrier for a raw device pointer."""

    def __init__(self, ptr: int, nelem: int, typestr: str):
        self.__cuda_array_interface__ = {"shape": (nelem,), "typestr": typestr, "data": (ptr, False), "version": 2}


def tensor_from_ptr(ptr: int, nelem: int, dtype: torch.dtype, device: torch.device) -> torch.Tensor:
    """Zero-copy torch view of raw device memory (bf16 goes through int16)."""
    if dtype == torch.bfloat16:
        return torch.as_tensor(_RawCuda(ptr, nelem, "<i2"), device=device).view(torch.bfloat16)
    return torch.as_tensor(_RawCuda(ptr, nelem, _TYPESTR[dtype]), device=device)


@dataclass
class Region:
    name: str
    offset: int        # bytes from arena base
    nbytes: int


class ArenaSpec:
    """Byte layout of one rank's arena; identical on every rank."""

    def __init__(self):
        self.regions: Dict[str, Region] = {}
        self.size = 0

    def add(self, name: str, nbytes: int, align: int = 1024) -> Region:
        off = (self.size + align - 1) // align * align
        self.regions[name] = Region(name, off, nbytes)
        self.size = off + nbytes
        return self.regions[name]


class Fabric:
    def __init__(self, rank: int, world: int, device_index: int, spec: ArenaSpec, emulate: bool = False,
                 group=None):
        self.rank, self.world, self.device_index, self.spec = rank, world, device_index, spec
        self.emulate = emulate
        self.device = torch.device("cuda", device_index)
        self.lib = L.lib()
        self.bases: List[int] = [0] * world        # arena base address of every rank, as seen from here
        self._owned: List[int] = []
        self._opened: List[int] = []
        torch.cuda.set_device(self.device)
        size = (spec.size + 4095) // 4096 * 4096
        if emulate:
            for r in range(world):
                self.bases[r] = self._alloc(size)
        else:
            self.bases[rank] = self._alloc(size)
            if world > 1:
                self._exchange(group)

    def _alloc(self, size: int) -> int:
        p = C.c_void_p()
        L.check(self.lib.dps_arena_alloc(self.device_index, size, C.byref(p)), "arena_alloc")
        self._owned.append(p.value)
        return p.value

    def _exchange(self, group) -> None:
        import torch.distributed as dist
        buf = C.create_string_buffer(64)
        L.check(self.lib.dps_ipc_export(C.c_void_p(self.bases[self.rank]), buf), "ipc_export")
        handles: List[Optional[Tuple[int, bytes]]] = [None] * self.world
        dist.all_gather_object(handles, (self.device_index, bytes(buf.raw)), group=group)
        for r, (dev, h) in enumerate(handles):
            if r == self.rank:
                continue
            L.check(self.lib.dps_enable_peer(self.device_index, dev), f"enable_peer {dev}")
            p = C.c_void_p()
            L.check(self.lib.dps_ipc_open(self.device_index, h, C.byref(p)), f"ipc_open rank {r}")
            self.bases[r] = p.value
            self._opened.append(p.value)
        dist.barrier(group=group)

    # ------------------------------------------------------------------
    def ptr(self, rank: int, region: str) -> int:
        return self.bases[rank] + self.spec.regions[region].offset

    def tensor(self, rank: int, region: str, dtype: torch.dtype) -> torch.Tensor:
        reg = self.spec.regions[region]
        itemsize = torch.empty((), dtype=dtype).element_size()
        return tensor_from_ptr(self.ptr(rank, region), reg.nbytes // itemsize, dtype, self.device)

    def local_ranks(self) -> Sequence[int]:
        return range(self.world) if self.emulate else (self.rank,)

    def close(self) -> None:
        torch.cuda.synchronize(self.device)
        for p in self._opened:
            self.lib.dps_ipc_close(C.c_void_p(p))
        for p in self._owned:
            self.lib.dps_arena_free(self.device_index, C.c_void_p(p))
        self._opened, self._owned = [], []


class MulticastArena:
    """The pull destinations (bf16 weight arena + fp32 small prefix) in memory that is ALSO bound to an NVLS
    multicast object, so an owner stores each vector once (``multimem.st``, csrc/ps_kernels.cu: bcast_mc) and the
    NVSwitch replicates it into every worker's replica -- the in-switch form of the reference's "send the whole model
    to every worker" (reference: src/parameter_server/server.py:221-230 -> src/workers/worker.py:233-262).

    Allocation, handle exchange and the multicast binding (cuMemCreate / cuMulticastCreate / cuMulticastBindMem) are
    done by ``torch.distributed._symmetric_memory`` -- bootstrap plumbing, like the IPC handle exchange of ``Fabric``;
    the kernels only see three kinds of address: this rank's replica, the peers' replicas (unicast mappings, still used
    by the per-peer path and the async gather), and the multicast window ``mc_base``.

    ``mc_base == 0`` when the box has no multicast support (no NVSwitch / fabric manager without NVLS): the arena then
    works as plain peer memory and the kernels keep their per-peer stores.
    """

    def __init__(self, rank: int, world: int, device_index: int, spec: ArenaSpec, group=None):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        self.rank, self.world, self.device_index, self.spec = rank, world, device_index, spec
        self.device = torch.device("cuda", device_index)
        torch.cuda.set_device(self.device)
        size = (spec.size + 4095) // 4096 * 4096
        # one whole allocation of our own (not a slice of torch's symmetric-memory pool): base address == tensor address
        from torch._C._distributed_c10d import _SymmetricMemory
        self.buf = _SymmetricMemory.empty_strided_p2p((size,), (1,), torch.uint8, self.device)
        self.buf.zero_()
        torch.cuda.synchronize(self.device)
        self.handle = symm.rendezvous(self.buf, group if group is not None else dist.group.WORLD)
        self.bases: List[int] = [int(p) for p in self.handle.buffer_ptrs]
        if self.bases[rank] != self.buf.data_ptr():
            raise RuntimeError("symmetric memory: the local mapping is not the tensor that was allocated")
        mc = int(self.handle.multicast_ptr or 0)
        # every rank must take the same path: multicast only if all of them got a window
        ok = torch.tensor([1 if mc else 0], dtype=torch.int32, device=self.device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
        self.mc_base = mc if int(ok.item()) else 0

    @staticmethod
    def supported(device_index: int) -> bool:
        try:
            from torch._C._autograd import DeviceType
            from torch._C._distributed_c10d import _SymmetricMemory
            return bool(_SymmetricMemory.has_multicast_support(DeviceType.CUDA, device_index))
        except Exception:
            return False

    def ptr(self, rank: int, region: str) -> int:
        return self.bases[rank] + self.spec.regions[region].offset

    def mc_ptr(self, region: str) -> int:
        return self.mc_base + self.spec.regions[region].offset if self.mc_base else 0

    def tensor(self, rank: int, region: str, dtype: torch.dtype) -> torch.Tensor:
        reg = self.spec.regions[region]
        itemsize = torch.empty((), dtype=dtype).element_size()
        return tensor_from_ptr(self.ptr(rank, region), reg.nbytes // itemsize, dtype, self.device)

    def close(self) -> None:
        torch.cuda.synchronize(self.device)
        self.handle = None
        self.buf = None
