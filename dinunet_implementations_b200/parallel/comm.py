"""Comm substrate: a symmetric device-memory arena per site, mapped into every peer.

``SiteGroup`` is what the fused aggregation kernels run on.  Two back ends:

* **ipc** – one process per GPU (``torchrun``): every rank ``cudaMalloc``s an arena of the same
  layout, exports it with CUDA IPC, exchanges the 64-byte handles once through
  ``torch.distributed`` (bootstrap only — SURVEY §5.8: "the steady-state aggregation paths issue
  no NCCL call") and opens its peers' arenas, so kernels address remote HBM directly over
  NVLink/NVSwitch.
* **symm** – same process layout, but the arena comes from ``torch.distributed._symmetric_memory`` (CUDA VMM
  allocation + handle exchange, bootstrap only): besides the per-peer mappings it yields an **NVLS multicast**
  address of the arena (``mc_base``), through which the aggregation kernels reduce inside the NVSwitch
  (``multimem.ld_reduce``) and publish with one store for all sites (``multimem.st``).  Chosen automatically when the
  box supports it (``DINUNET_ARENA=ipc`` forces the plain IPC arena); the kernels themselves never call NCCL.
* **local** – S simulated sites inside one process on one GPU: S arenas in the same address
  space.  The kernels are identical; only the cross-GPU flag barrier is switched off, because
  kernels that wait for one another must not be launched back-to-back on one device.  This is
  how the multi-site logic is tested on a single B200.

The arena is carved by name (``alloc``) in the same order on every rank, so offsets agree.
"""
from __future__ import annotations

import ctypes
import os

import torch

from ..ops import native


class _CudaView:
    """Minimal ``__cuda_array_interface__`` carrier so torch can alias arena memory."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def _as_tensor(ptr, nbytes, device):
    return torch.as_tensor(_CudaView(ptr, nbytes), device=device)


class SiteGroup:
    def __init__(self, nsites: int, rank: int, arena_bytes: int, device, backend: str):
        self.nsites, self.rank, self.arena_bytes, self.backend = int(nsites), int(rank), int(arena_bytes), backend
        self.device = torch.device(device)
        self.base = [0] * self.nsites          # base pointer of every site's arena in THIS process
        self._owned = []
        self._opened = []
        self._cursor = 0
        self._layout = {}
        self.epoch = 0
        self.mc_base = 0                       # multicast (NVLS) address of the arena, 0 = not available
        self._symm = None                      # (tensor, handle) keeping a symmetric-memory arena alive

    # ------------------------------------------------------------------ construction
    @classmethod
    def local(cls, nsites: int, arena_bytes: int, device="cuda:0"):
        lib = native.require_gpu()
        g = cls(nsites, 0, arena_bytes, device, "local")
        with torch.cuda.device(g.device):
            for s in range(nsites):
                p = ctypes.c_void_p()
                native.check(lib.dnb_arena_alloc(ctypes.c_size_t(arena_bytes), ctypes.byref(p)), "arena_alloc")
                g.base[s] = p.value
                g._owned.append(p.value)
        g._reserve_flags()
        return g

    @classmethod
    def from_torch_distributed(cls, arena_bytes: int, device=None):
        """One arena per rank; handles exchanged once with ``all_gather_object``."""
        import torch.distributed as dist
        lib = native.require_gpu()
        rank, world = dist.get_rank(), dist.get_world_size()
        device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        if world > 1 and os.environ.get("DINUNET_ARENA", "symm") != "ipc":
            g = cls._from_symmetric_memory(world, rank, arena_bytes, device)
            if g is not None:
                return g
        g = cls(world, rank, arena_bytes, device, "ipc" if world > 1 else "local")
        with torch.cuda.device(device):
            p = ctypes.c_void_p()
            native.check(lib.dnb_arena_alloc(ctypes.c_size_t(arena_bytes), ctypes.byref(p)), "arena_alloc")
            g.base[rank] = p.value
            g._owned.append(p.value)
            if world > 1:
                hsz = lib.dnb_ipc_handle_size()
                buf = ctypes.create_string_buffer(hsz)
                native.check(lib.dnb_ipc_get_handle(p, buf), "ipc_get_handle")
                handles = [None] * world
                dist.all_gather_object(handles, (bytes(buf.raw), device.index))
                for r, (h, dev_idx) in enumerate(handles):
                    if r == rank:
                        continue
                    native.check(lib.dnb_enable_peer_access(ctypes.c_int(int(dev_idx))), "enable_peer_access")
                    q = ctypes.c_void_p()
                    native.check(lib.dnb_ipc_open_handle(ctypes.create_string_buffer(h, hsz), ctypes.byref(q)),
                                 "ipc_open_handle")
                    g.base[r] = q.value
                    g._opened.append(q.value)
                dist.barrier()
        g._reserve_flags()
        return g

    @classmethod
    def _from_symmetric_memory(cls, world, rank, arena_bytes, device):
        """VMM-backed arena with peer mappings and (where the fabric has it) an NVLS multicast mapping.  Every rank
        must take the same decision, so success is agreed on with one all-reduce; on any failure the caller falls
        back to the CUDA-IPC arena."""
        import torch.distributed as dist
        ok, t, h = 1, None, None
        try:
            import torch.distributed._symmetric_memory as symm
            nbytes = (int(arena_bytes) + (2 << 20) - 1) // (2 << 20) * (2 << 20)
            with torch.cuda.device(device):
                t = symm.empty(nbytes, dtype=torch.uint8, device=device)
                t.zero_()
                torch.cuda.synchronize(device)
                h = symm.rendezvous(t, dist.group.WORLD.group_name)
            if len(h.buffer_ptrs) != world:
                ok = 0
        except Exception as e:      # noqa: BLE001  (API is private and may be missing / refuse this box)
            from ..utils import logger
            logger.warn(f"symmetric-memory arena unavailable ({type(e).__name__}: {e}); using the CUDA-IPC arena", True)
            ok = 0
        flag = torch.tensor([ok], device=device, dtype=torch.int32)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) != 1:
            return None
        g = cls(world, rank, t.numel(), device, "symm")
        g.base = [int(p) for p in h.buffer_ptrs]
        g.mc_base = int(getattr(h, "multicast_ptr", 0) or 0)
        g._symm = (t, h)
        dist.barrier()
        g._reserve_flags()
        return g

    def _reserve_flags(self):
        self.flags = self.alloc("__flags__", native.SIGNAL_WORDS, torch.int32)

    # ------------------------------------------------------------------ carving
    def alloc(self, name: str, numel: int, dtype: torch.dtype, align: int = 256):
        """Reserve ``numel`` elements at the same offset in every site's arena.

        Returns the list of per-site tensors (index = site); in ipc mode entries for remote
        sites alias peer memory and are meant for kernels, not for host-side torch ops.
        """
        esz = torch.empty(0, dtype=dtype).element_size()
        off = (self._cursor + align - 1) // align * align
        nbytes = int(numel) * esz
        if off + nbytes > self.arena_bytes:
            raise MemoryError(f"arena exhausted: need {off + nbytes} bytes, have {self.arena_bytes}")
        self._cursor = off + nbytes
        self._layout[name] = (off, int(numel), dtype)
        return self.views(name)

    def views(self, name):
        off, numel, dtype = self._layout[name]
        esz = torch.empty(0, dtype=dtype).element_size()
        return [_as_tensor(b + off, numel * esz, self.device).view(dtype) for b in self.base]

    def ptrs(self, name):
        off = self._layout[name][0]
        return [b + off for b in self.base]

    def mc_ptr(self, name):
        """Multicast address of a carved buffer (0 when the arena has no NVLS mapping)."""
        return self.mc_base + self._layout[name][0] if self.mc_base else 0

    def local_view(self, name, site=None):
        return self.views(name)[self.rank if site is None else site]

    def next_epoch(self):
        self.epoch += 1
        return self.epoch

    @property
    def use_flags(self):
        return self.backend in ("ipc", "symm")

    def close(self, barrier=None):
        """Unmap the peers' arenas, then free our own.  In ipc mode pass ``barrier`` (e.g.
        ``torch.distributed.barrier``): an arena must not be freed while a peer still has it mapped
        or a kernel of a peer may still touch it."""
        l = native.lib()
        if self._owned or self._opened:
            torch.cuda.synchronize(self.device)
        if barrier is not None:
            barrier()
        for q in self._opened:
            l.dnb_ipc_close_handle(ctypes.c_void_p(q))
        if barrier is not None:
            barrier()
        for p in self._owned:
            l.dnb_arena_free(ctypes.c_void_p(p))
        self._opened, self._owned = [], []
        if self._symm is not None:
            self._symm = None                  # the allocator unmaps and frees once the tensor and handle are gone
            if barrier is not None:
                barrier()
