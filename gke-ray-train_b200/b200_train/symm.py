"""Symmetric (peer-mapped) device memory and the peer-memory collectives built on it.

Bootstrap only goes through torch.distributed (the process group Ray's TorchTrainer — or our shim — already
initialised, ref: ray-jobs/pytorch_llm_ray.py:362-374 `TorchConfig(backend="nccl")`): the 64-byte CUDA IPC handles
are exchanged with all_gather_object. After that the data path is our own kernels doing loads/stores on the peers'
buffers over NVLink (include/b2train.h, b2_peer_*).
"""
import ctypes
import socket

import torch
import torch.distributed as dist

from . import _lib
from ._lib import B2Error, call, dtype_code, stream_ptr


class _Raw:
    """Exposes a raw device allocation to torch through __cuda_array_interface__ (zero-copy)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def tensor_from_ptr(ptr, nbytes, dtype, device):
    raw = torch.as_tensor(_Raw(ptr, nbytes), device=device)
    return raw.view(dtype)


class SymmetricBuffer:
    """`nbytes` of cudaMalloc'd memory on this rank, mapped into every other rank of `group` (same node)."""

    def __init__(self, nbytes, device, group=None):
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.device = device
        self.nbytes = (nbytes + 255) // 256 * 256
        local = ctypes.c_void_p()
        handle = (ctypes.c_ubyte * 64)()
        with torch.cuda.device(device):
            call("b2_symm_alloc", self.nbytes, ctypes.byref(local), handle)
        self.local_ptr = local.value
        infos = [None] * self.world
        dist.all_gather_object(infos, (socket.gethostname(), bytes(handle), self.nbytes), group=group)
        if len({h for h, _, _ in infos}) != 1:
            raise B2Error("peer-memory collectives need all ranks on one node (NVLink/NVSwitch domain); got hosts "
                          f"{sorted({h for h, _, _ in infos})}")
        if len({n for _, _, n in infos}) != 1:
            raise B2Error("symmetric allocation sizes differ across ranks")
        self.ptrs = []
        for k, (_, hbytes, _) in enumerate(infos):
            if k == self.rank:
                self.ptrs.append(self.local_ptr)
                continue
            p = ctypes.c_void_p()
            hb = (ctypes.c_ubyte * 64).from_buffer_copy(hbytes)
            with torch.cuda.device(device):
                call("b2_symm_open", hb, ctypes.byref(p))
            self.ptrs.append(p.value)
        self._closed = False

    def tensor(self, dtype, numel, byte_offset=0):
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        if byte_offset + nbytes > self.nbytes:
            raise B2Error("symmetric buffer too small for the requested view")
        t = tensor_from_ptr(self.local_ptr + byte_offset, nbytes, dtype, self.device)
        t._b2_symm = (self, byte_offset)          # keeps the allocation alive and findable from the tensor
        return t

    def ptr_table(self, byte_offset=0):
        return (ctypes.c_void_p * self.world)(*[p + byte_offset for p in self.ptrs])

    def close(self):
        if self._closed:
            return
        self._closed = True
        torch.cuda.synchronize(self.device)
        for k, p in enumerate(self.ptrs):
            if k != self.rank:
                _lib.lib.b2_symm_close(p)
        dist.barrier(group=self.group)            # owner frees only after every peer unmapped
        _lib.lib.b2_symm_free(self.local_ptr)


class TorchSymmetricBuffer:
    """Same interface as SymmetricBuffer, but the allocation, the peer mappings and — where the fabric supports it — an
    NVLS multicast address come from torch.distributed._symmetric_memory (VMM allocations instead of cudaMalloc + IPC).
    This is bootstrap plumbing only; the collectives that run on these pointers are still our kernels."""

    def __init__(self, nbytes, device, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.device = device
        self.nbytes = (nbytes + 255) // 256 * 256
        self._t = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=device)
        self._t.zero_()
        self._hdl = symm_mem.rendezvous(self._t, group if group is not None else dist.group.WORLD)
        self.ptrs = [int(p) for p in self._hdl.buffer_ptrs]
        self.local_ptr = self.ptrs[self.rank]
        self.mc_ptr = int(self._hdl.multicast_ptr) or None
        self._closed = False

    def tensor(self, dtype, numel, byte_offset=0):
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        if byte_offset + nbytes > self.nbytes:
            raise B2Error("symmetric buffer too small for the requested view")
        t = self._t[byte_offset:byte_offset + nbytes].view(dtype)
        t._b2_symm = (self, byte_offset)
        return t

    def ptr_table(self, byte_offset=0):
        return (ctypes.c_void_p * self.world)(*[p + byte_offset for p in self.ptrs])

    def close(self):
        if not self._closed:
            self._closed = True
            torch.cuda.synchronize(self.device)
            dist.barrier(group=self.group)
            self._hdl = None
            self._t = None


def _pick_backend():
    """B2_SYMM_BACKEND=ipc | torch. Default: torch's symmetric memory when it reports NVLS multicast support (needed for
    the in-switch reduction), else plain CUDA IPC."""
    import os
    forced = os.environ.get("B2_SYMM_BACKEND")
    if forced in ("ipc", "torch"):
        return forced
    try:
        import torch.distributed._symmetric_memory as symm_mem
        from torch._C._autograd import DeviceType
        if symm_mem._SymmetricMemory.has_multicast_support(DeviceType.CUDA, torch.cuda.current_device()):
            return "torch"
    except Exception:  # noqa: BLE001 - any failure here just means "no multicast allocator"
        pass
    return "ipc"


class PeerComm:
    """Per-process handle for the peer-memory collectives: flag block, epoch counter, slice arithmetic."""

    def __init__(self, device, group=None):
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        if self.world > 8:
            raise B2Error("peer-memory collectives support up to 8 ranks (one NVSwitch domain)")
        self.device = device
        self.backend = _pick_backend()
        # in-switch reduction pays off once there are more than two ranks (B2_PEER_NVLS=0/1 overrides)
        import os
        self.use_nvls = bool(int(os.environ.get("B2_PEER_NVLS", int(self.world > 2))))
        self.flags = SymmetricBuffer(_lib.lib.b2_peer_flags_bytes(), device, group)
        self.epoch = 0
        self._flag_table = self.flags.ptr_table()
        self._parts = self.flags.tensor(torch.float32, 64)[32:32 + self.world]
        self.buffers = []
        dist.barrier(group=group)

    def alloc(self, numel, dtype, device):
        """Allocator hook for flat.flatten_params: tensors that live in symmetric memory."""
        esize = torch.empty((), dtype=dtype).element_size()
        cls = TorchSymmetricBuffer if self.backend == "torch" else SymmetricBuffer
        sb = cls(max(numel * esize, 256), device, self.group)
        self.buffers.append(sb)
        return sb.tensor(dtype, numel)

    @staticmethod
    def _symm_of(t):
        info = getattr(t, "_b2_symm", None)
        if info is None:
            raise B2Error("tensor does not live in symmetric memory (allocate it through PeerComm.alloc)")
        return info

    def slice_bounds(self, n, dtype, rank=None):
        b, e = ctypes.c_size_t(), ctypes.c_size_t()
        call("b2_peer_slice", n, dtype_code(dtype), self.world, self.rank if rank is None else rank,
             ctypes.byref(b), ctypes.byref(e))
        return b.value, e.value

    def allreduce_(self, t, scale=None, want_sumsq=False, scatter_only=False, offset=0, numel=None, max_blocks=0,
                   sumsq_out=None):
        """In-place mean (scale = 1/world by default) over ranks of elements [offset, offset+numel) of the flat
        symmetric tensor `t`. want_sumsq: False / True (restart the sum-of-squares slots) / 2 (keep adding to them).
        sumsq_out: fp32 device tensor of `world` elements that the kernel fills with the completed slots as its last act.
        scatter_only: only this rank's slice of the range (see slice_bounds) is reduced, and it stays local."""
        sb, off = self._symm_of(t)
        self.epoch += 1
        numel = t.numel() - offset if numel is None else numel
        sc = float(1.0 / self.world if scale is None else scale)
        mc = getattr(sb, "mc_ptr", None)
        if sumsq_out is not None and (sumsq_out.dtype != torch.float32 or sumsq_out.numel() < self.world
                                      or not sumsq_out.is_contiguous()):
            raise B2Error("sumsq_out must be a contiguous float32 tensor with one element per rank")
        so = None if sumsq_out is None else sumsq_out.data_ptr()
        with torch.cuda.device(self.device):
            if mc and self.use_nvls:
                call("b2_peer_allreduce_mc", mc + off, sb.ptr_table(off), self._flag_table, self.world, self.rank,
                     int(offset), int(numel), dtype_code(t.dtype), sc, self.epoch, int(want_sumsq),
                     int(bool(scatter_only)), int(max_blocks), so, stream_ptr())
            else:
                call("b2_peer_allreduce", sb.ptr_table(off), self._flag_table, self.world, self.rank, int(offset),
                     int(numel), dtype_code(t.dtype), sc, self.epoch, int(want_sumsq), int(bool(scatter_only)),
                     int(max_blocks), so, stream_ptr())
        return self._parts if want_sumsq else None

    def kernel_name(self, t):
        """Which reduce kernel allreduce_ launches for `t` ("mc" = in-switch NVLS reduction, "ldst" = peer loads/stores)."""
        sb, _ = self._symm_of(t)
        return "mc" if (getattr(sb, "mc_ptr", None) and self.use_nvls) else "ldst"

    def allgather_(self, t, offset=0, numel=None, max_blocks=0):
        """Pushes this rank's slice of elements [offset, offset+numel) of `t` to every peer."""
        sb, off = self._symm_of(t)
        self.epoch += 1
        numel = t.numel() - offset if numel is None else numel
        with torch.cuda.device(self.device):
            call("b2_peer_allgather", sb.ptr_table(off), self._flag_table, self.world, self.rank, int(offset), int(numel),
                 dtype_code(t.dtype), self.epoch, int(max_blocks), stream_ptr())

    def close(self):
        for sb in self.buffers:
            sb.close()
        self.flags.close()
