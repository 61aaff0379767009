"""NVLS symmetric buffers: physical memory per rank (CUDA VMM) that is mapped three ways -

* locally (unicast),
* into every peer (unicast, what the CUDA-IPC path of ``parallel/p2p.py`` gives as well), and
* through ONE multicast object bound to all ranks' copies, so that ``multimem.ld_reduce`` on the multicast address sums the
  bucket inside the NVSwitch and ``multimem.st`` writes all replicas at once (csrc/runtime/p2p.cu, mode 2).

Control plane only (driver API through cuda-python): ``cuMemCreate`` with a POSIX-fd handle type, fds handed to the
peers with ``pidfd_getfd`` (same user, one host), ``cuMulticastCreate`` on rank 0, ``cuMulticastAddDevice`` /
``cuMulticastBindMem`` on every rank.  ``supported()`` says whether the device and driver offer multicast at all; every
failure is reported to the caller, who falls back to the IPC + two-shot path.
"""
from __future__ import annotations

import ctypes
import os
from typing import List

import torch

_SYS_pidfd_open = 434
_SYS_pidfd_getfd = 438


def _drv():
    try:
        from cuda.bindings import driver as drv          # cuda-python >= 12.8
    except Exception:                                     # pragma: no cover - older package layout
        from cuda import cuda as drv
    return drv


def _ck(res, what: str):
    err = res[0]
    if int(err) != 0:
        raise RuntimeError(f"{what} failed: {err!r}")
    return res[1] if len(res) == 2 else res[1:]


def supported(device: int) -> bool:
    try:
        drv = _drv()
        _ck(drv.cuInit(0), "cuInit")
        dev = _ck(drv.cuDeviceGet(device), "cuDeviceGet")
        mc = _ck(drv.cuDeviceGetAttribute(drv.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev), "attr")
        fd = _ck(drv.cuDeviceGetAttribute(
            drv.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev), "attr")
        return bool(mc) and bool(fd)
    except Exception:
        return False


def _dup_fd_from(pid: int, fd: int) -> int:
    """A local duplicate of file descriptor ``fd`` of process ``pid`` (Linux >= 5.6, same user)."""
    if pid == os.getpid():
        return os.dup(fd)
    libc = ctypes.CDLL(None, use_errno=True)
    pidfd = libc.syscall(_SYS_pidfd_open, pid, 0)
    if pidfd < 0:
        raise OSError(ctypes.get_errno(), f"pidfd_open({pid})")
    try:
        got = libc.syscall(_SYS_pidfd_getfd, pidfd, fd, 0)
        if got < 0:
            raise OSError(ctypes.get_errno(), f"pidfd_getfd({pid}, {fd})")
        return got
    finally:
        os.close(pidfd)


class _RawCuda:
    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}


class NvlsBuf:
    """One symmetric buffer: ``ptrs[r]`` = rank r's copy mapped here, ``mc_ptr`` = the multicast mapping, ``tensor`` = a
    uint8 view of the local copy."""

    def __init__(self, local, ptrs, mc_ptr, nbytes, tensor, handles, mc_handle, rank):
        self.local, self.ptrs, self.mc_ptr, self.nbytes, self.tensor = local, ptrs, mc_ptr, nbytes, tensor
        self._handles, self._mc_handle, self._rank = handles, mc_handle, rank


class NvlsAllocator:
    """Collective allocator; ``exchange`` is an object with ``all_gather_object`` / ``broadcast_object`` / ``barrier``
    (the P2P backend's native store)."""

    def __init__(self, rank: int, world: int, device: int, exchange):
        self.rank, self.world, self.device, self.x = rank, world, device, exchange
        self.drv = _drv()
        drv = self.drv
        _ck(drv.cuInit(0), "cuInit")
        self.dev = _ck(drv.cuDeviceGet(device), "cuDeviceGet")
        self.fd_type = drv.CUmemAllocationHandleType.CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR
        prop = drv.CUmemAllocationProp()
        prop.type = drv.CUmemAllocationType.CU_MEM_ALLOCATION_TYPE_PINNED
        prop.location.type = drv.CUmemLocationType.CU_MEM_LOCATION_TYPE_DEVICE
        prop.location.id = device
        prop.requestedHandleTypes = self.fd_type
        self.prop = prop
        mem_gran = _ck(drv.cuMemGetAllocationGranularity(
            prop, drv.CUmemAllocationGranularity_flags.CU_MEM_ALLOC_GRANULARITY_RECOMMENDED), "cuMemGetAllocationGranularity")
        mcp = drv.CUmulticastObjectProp()
        mcp.numDevices = world
        mcp.size = 1 << 21
        mcp.handleTypes = self.fd_type
        mc_gran = _ck(drv.cuMulticastGetGranularity(
            mcp, drv.CUmulticastGranularity_flags.CU_MULTICAST_GRANULARITY_RECOMMENDED), "cuMulticastGetGranularity")
        self.gran = max(int(mem_gran), int(mc_gran))
        self.access = drv.CUmemAccessDesc()
        self.access.location.type = drv.CUmemLocationType.CU_MEM_LOCATION_TYPE_DEVICE
        self.access.location.id = device
        self.access.flags = drv.CUmemAccess_flags.CU_MEM_ACCESS_FLAGS_PROT_READWRITE
        self.bufs: List[NvlsBuf] = []

    def _map(self, handle, size: int) -> int:
        drv = self.drv
        va = _ck(drv.cuMemAddressReserve(size, self.gran, 0, 0), "cuMemAddressReserve")
        _ck(drv.cuMemMap(va, size, 0, handle, 0), "cuMemMap")
        _ck(drv.cuMemSetAccess(va, size, [self.access], 1), "cuMemSetAccess")
        return int(va)

    def alloc(self, nbytes: int) -> NvlsBuf:
        drv = self.drv
        size = (int(nbytes) + self.gran - 1) // self.gran * self.gran
        # 1. my physical memory, exported as a file descriptor
        h_local = _ck(drv.cuMemCreate(size, self.prop, 0), "cuMemCreate")
        fd_local = int(_ck(drv.cuMemExportToShareableHandle(h_local, self.fd_type, 0), "cuMemExportToShareableHandle"))
        # 2. the multicast object: created by rank 0, imported by the others
        mcp = drv.CUmulticastObjectProp()
        mcp.numDevices = self.world
        mcp.size = size
        mcp.handleTypes = self.fd_type
        if self.rank == 0:
            h_mc = _ck(drv.cuMulticastCreate(mcp), "cuMulticastCreate")
            fd_mc = int(_ck(drv.cuMemExportToShareableHandle(h_mc, self.fd_type, 0), "export multicast handle"))
        else:
            h_mc, fd_mc = None, -1
        everyone = self.x.all_gather_object((os.getpid(), fd_local, fd_mc))
        handles = [None] * self.world
        ptrs = [0] * self.world
        opened = []
        for r, (pid, fd, _) in enumerate(everyone):
            if r == self.rank:
                handles[r] = h_local
            else:
                lfd = _dup_fd_from(pid, fd)
                opened.append(lfd)
                handles[r] = _ck(drv.cuMemImportFromShareableHandle(lfd, self.fd_type), f"import rank {r}'s memory")
            ptrs[r] = self._map(handles[r], size)
        if self.rank != 0:
            lfd = _dup_fd_from(everyone[0][0], everyone[0][2])
            opened.append(lfd)
            h_mc = _ck(drv.cuMemImportFromShareableHandle(lfd, self.fd_type), "import multicast handle")
        _ck(drv.cuMulticastAddDevice(h_mc, self.dev), "cuMulticastAddDevice")
        self.x.barrier()                                   # every device added before anybody binds
        _ck(drv.cuMulticastBindMem(h_mc, 0, h_local, 0, size, 0), "cuMulticastBindMem")
        self.x.barrier()                                   # every rank bound before anybody uses the mapping
        mc_ptr = self._map(h_mc, size)
        for fd in opened:
            os.close(fd)
        self.x.barrier()                                   # peers have duplicated my descriptors: they can go
        os.close(fd_local)
        if fd_mc >= 0:
            os.close(fd_mc)
        t = torch.as_tensor(_RawCuda(ptrs[self.rank], size), device=torch.device("cuda", self.device))
        t.zero_()
        torch.cuda.synchronize(self.device)
        buf = NvlsBuf(ptrs[self.rank], ptrs, mc_ptr, size, t, handles, h_mc, self.rank)
        self.bufs.append(buf)
        return buf

    def close(self) -> None:
        """Collective teardown: unmap everything, unbind from the multicast objects, release the physical memory."""
        drv = self.drv
        torch.cuda.synchronize(self.device)
        self.x.barrier()                                   # nobody is still reading a peer's copy
        for b in self.bufs:
            b.tensor = None
            for va in list(b.ptrs) + [b.mc_ptr]:
                if va:
                    drv.cuMemUnmap(va, b.nbytes)
                    drv.cuMemAddressFree(va, b.nbytes)
            drv.cuMulticastUnbind(b._mc_handle, self.dev, 0, b.nbytes)
        self.x.barrier()
        for b in self.bufs:
            for h in b._handles:
                if h is not None:
                    drv.cuMemRelease(h)
            drv.cuMemRelease(b._mc_handle)
        self.bufs = []
