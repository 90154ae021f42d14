"""NVSwitch multicast (NVLS) symmetric memory.

A ``MulticastBuffer`` is a symmetric allocation made with the CUDA virtual-memory API (so that it
can be bound to a multicast object) instead of cudaMalloc + cudaIpc:
  * every rank creates one physical allocation, exports it as a POSIX file descriptor and maps
    every peer's allocation (unicast peer pointers, used by the P2P kernels and the flag pads);
  * rank 0 creates a multicast object for all devices, every rank adds its device and binds its
    allocation, then maps the object: loads from that address with ``multimem.ld_reduce`` return
    the in-switch sum of all ranks' copies, ``multimem.st`` writes all copies at once.
File descriptors travel between the per-GPU processes over Unix-domain sockets (SCM_RIGHTS); the
socket paths are exchanged through the bootstrap process group.

Falls back cleanly: ``multicast_supported()`` is False when the device / driver / container does
not expose multicast, and every caller then stays on the P2P kernels.
"""
from __future__ import annotations

import os
import socket
import struct
import uuid

import torch
import torch.distributed as dist

from .comm import SymmBuffer, _RawCuda

try:  # cuda-python >= 12.8 layout, then the legacy one
    from cuda.bindings import driver as cu
except Exception:  # noqa: BLE001
    try:
        from cuda import cuda as cu
    except Exception:  # noqa: BLE001
        cu = None


def _ck(res):
    err = res[0] if isinstance(res, tuple) else res
    if err != cu.CUresult.CUDA_SUCCESS:
        raise RuntimeError(f"CUDA driver error {err}")
    if isinstance(res, tuple):
        return res[1] if len(res) == 2 else res[1:]
    return None


def multicast_supported(device_index: int | None = None) -> bool:
    if cu is None or not torch.cuda.is_available():
        return False
    if os.environ.get("DPT_DISABLE_NVLS", "0") == "1":
        return False
    try:
        idx = torch.cuda.current_device() if device_index is None else device_index
        dev = _ck(cu.cuDeviceGet(idx))
        a = cu.CUdevice_attribute
        return bool(_ck(cu.cuDeviceGetAttribute(a.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev))) and \
            bool(_ck(cu.cuDeviceGetAttribute(a.CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev)))
    except Exception:  # noqa: BLE001
        return False


def _exchange_fds(rank: int, world: int, group, fd_for_peer) -> dict:
    """All-to-all (or one-to-all) file-descriptor exchange over Unix sockets.
    ``fd_for_peer(p)`` returns the fd this rank sends to peer p (or None).  Returns {src: fd}."""
    path = f"/tmp/dpt_b200_{os.getpid()}_{uuid.uuid4().hex[:8]}.sock"
    srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    srv.bind(path)
    srv.listen(world + 2)
    paths = [None] * world
    dist.all_gather_object(paths, path, group=group)
    expect = [None] * world
    sends = [1 if (p != rank and fd_for_peer(p) is not None) else 0 for p in range(world)]
    allsends = [None] * world
    dist.all_gather_object(allsends, sends, group=group)
    n_expect = sum(allsends[src][rank] for src in range(world))
    for p in range(world):
        if sends[p]:
            c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            c.connect(paths[p])
            socket.send_fds(c, [struct.pack("i", rank)], [fd_for_peer(p)])
            c.close()
    got = {}
    for _ in range(n_expect):
        conn, _addr = srv.accept()
        msg, fds, _flags, _a = socket.recv_fds(conn, 16, 1)
        got[struct.unpack("i", msg[:4])[0]] = fds[0]
        conn.close()
    dist.barrier(group=group)
    srv.close()
    try:
        os.unlink(path)
    except OSError:
        pass
    return got


class MulticastBuffer(SymmBuffer):
    """SymmBuffer (local view + unicast peer pointers) plus ``mc_ptr``, this rank's mapping of
    the multicast object that spans every rank's copy."""

    def __init__(self, comm, nbytes: int) -> None:
        rank, world, group = comm.rank, comm.world, comm.group
        dev_index = comm.device.index
        dev = _ck(cu.cuDeviceGet(dev_index))
        FD = cu.CUmemAllocationHandleType.CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR
        prop = cu.CUmemAllocationProp()
        prop.type = cu.CUmemAllocationType.CU_MEM_ALLOCATION_TYPE_PINNED
        prop.location.type = cu.CUmemLocationType.CU_MEM_LOCATION_TYPE_DEVICE
        prop.location.id = dev_index
        prop.requestedHandleTypes = FD
        mcprop = cu.CUmulticastObjectProp()
        mcprop.numDevices = world
        mcprop.handleTypes = FD
        mcprop.flags = 0
        mcprop.size = max(int(nbytes), 1)
        gran = max(int(_ck(cu.cuMemGetAllocationGranularity(
            prop, cu.CUmemAllocationGranularity_flags.CU_MEM_ALLOC_GRANULARITY_RECOMMENDED))),
            int(_ck(cu.cuMulticastGetGranularity(
                mcprop, cu.CUmulticastGranularity_flags.CU_MULTICAST_GRANULARITY_RECOMMENDED))))
        size = (int(nbytes) + gran - 1) // gran * gran
        mcprop.size = size
        self._size, self._gran = size, gran
        access = cu.CUmemAccessDesc()
        access.location.type = cu.CUmemLocationType.CU_MEM_LOCATION_TYPE_DEVICE
        access.location.id = dev_index
        access.flags = cu.CUmemAccess_flags.CU_MEM_ACCESS_FLAGS_PROT_READWRITE

        def map_handle(h):
            va = _ck(cu.cuMemAddressReserve(size, gran, 0, 0))
            _ck(cu.cuMemMap(va, size, 0, h, 0))
            _ck(cu.cuMemSetAccess(va, size, [access], 1))
            return int(va)

        # ---- local physical allocation + unicast mappings of every peer -------------------
        self._h_local = _ck(cu.cuMemCreate(size, prop, 0))
        local_va = map_handle(self._h_local)
        my_fd = int(_ck(cu.cuMemExportToShareableHandle(self._h_local, FD, 0)))
        peer_fds = _exchange_fds(rank, world, group, lambda p: my_fd) if world > 1 else {}
        peer_ptrs = [0] * world
        peer_ptrs[rank] = local_va
        self._peer_handles = {}
        for src, fd in peer_fds.items():
            h = _ck(cu.cuMemImportFromShareableHandle(fd, FD))
            os.close(fd)
            self._peer_handles[src] = h
            peer_ptrs[src] = map_handle(h)
        os.close(my_fd)
        # ---- multicast object -------------------------------------------------------------
        if rank == 0:
            self._mc = _ck(cu.cuMulticastCreate(mcprop))
            mc_fd = int(_ck(cu.cuMemExportToShareableHandle(self._mc, FD, 0)))
            _exchange_fds(rank, world, group, lambda p: mc_fd)
            os.close(mc_fd)
        else:
            got = _exchange_fds(rank, world, group, lambda p: None)
            self._mc = _ck(cu.cuMemImportFromShareableHandle(got[0], FD))
            os.close(got[0])
        _ck(cu.cuMulticastAddDevice(self._mc, dev))
        dist.barrier(group=group)          # every device must be added before anyone binds
        _ck(cu.cuMulticastBindMem(self._mc, 0, self._h_local, 0, size, 0))
        dist.barrier(group=group)
        self.mc_ptr = map_handle(self._mc)
        local = torch.as_tensor(_RawCuda(local_va, size), device=comm.device)
        local.zero_()
        torch.cuda.synchronize()
        dist.barrier(group=group)
        super().__init__(size, local, peer_ptrs, [])
        self._vas = [p for p in peer_ptrs] + [self.mc_ptr]

    def release(self) -> None:
        for va in getattr(self, "_vas", []):
            try:
                cu.cuMemUnmap(va, self._size)
                cu.cuMemAddressFree(va, self._size)
            except Exception:  # noqa: BLE001
                pass
        self._vas = []
        for h in list(getattr(self, "_peer_handles", {}).values()) + [self._h_local, self._mc]:
            try:
                cu.cuMemRelease(h)
            except Exception:  # noqa: BLE001
                pass
