"""Communicator: symmetric (peer-mapped) device memory + the NVLink peer-memory collectives.

This is the data plane that stands in for ProcessGroupNCCL on the gradient path (reference
train_ddp.py:65 ``init_process_group("nccl")`` -> DDP Reducer -> ncclAllReduce).  The torch
process group is still used, but only to *bootstrap*: CUDA IPC handles of each rank's arenas are
exchanged once through it (``all_gather_object``); after that every collective is one of the
kernels in csrc/comm/p2p_collectives.cu reading / writing peers' memory directly.

Scope: one node, <= 8 GPUs, all peer-accessible (NVSwitch).  Anything else uses the
torch.distributed fallback in ``parallel/ddp.py``.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import torch
import torch.distributed as dist

from ..ops import _lib


class _RawCuda:
    """Expose a raw device allocation through __cuda_array_interface__ (zero-copy torch view)."""

    def __init__(self, ptr: int, nbytes: int) -> None:
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1",
                                         "data": (ptr, False), "version": 2}


@dataclass
class SymmBuffer:
    """One allocation per rank, same size everywhere; ``peer_ptrs[r]`` is this process's mapping
    of rank r's copy (``peer_ptrs[rank]`` is the local allocation)."""
    nbytes: int
    local: torch.Tensor          # uint8 view of the local allocation
    peer_ptrs: list[int]
    _opened: list[int] = field(default_factory=list)

    def ptrs(self, byte_offset: int = 0):
        arr = (C.c_void_p * len(self.peer_ptrs))()
        for i, p in enumerate(self.peer_ptrs):
            arr[i] = p + byte_offset
        return arr

    def view(self, dtype: torch.dtype, byte_offset: int = 0, numel: int | None = None):
        t = self.local[byte_offset:]
        item = torch.empty((), dtype=dtype).element_size()
        n = (t.numel() // item) if numel is None else numel
        return t[: n * item].view(dtype)


class Communicator:
    def __init__(self, group=None, timeout_s: float = 30.0) -> None:
        if not torch.cuda.is_available():
            raise RuntimeError("the native communicator needs CUDA devices")
        self.group = group
        if dist.is_available() and dist.is_initialized():
            self.rank = dist.get_rank(group)
            self.world = dist.get_world_size(group)
        else:
            self.rank, self.world = 0, 1
        if self.world > 8:
            raise RuntimeError("native backend is single-node (<= 8 GPUs); use --backend nccl")
        self.lib = _lib.lib()
        self.device = torch.device("cuda", torch.cuda.current_device())
        self._buffers: list[SymmBuffer] = []
        self._mc_buffers: list = []
        self.flags = self.alloc(int(self.lib.dpt_comm_flag_bytes()))
        handle = C.c_void_p()
        _lib.check(self.lib.dpt_comm_create(self.rank, self.world, self.flags.ptrs(), float(timeout_s),
                                            C.byref(handle)), "dpt_comm_create")
        self.handle = handle
        # tiny symmetric scratch: found_inf word (+ room for packed scalar reductions)
        self.scalars = self.alloc(256)
        self.found_inf = self.scalars.view(torch.float32, 0, 1)[0]
        self.launches = 0
        torch.cuda.synchronize()
        self.host_barrier()

    # ------------------------------------------------------------------ memory
    def alloc(self, nbytes: int) -> SymmBuffer:
        nbytes = (int(nbytes) + 255) // 256 * 256
        ptr = C.c_void_p()
        handle = (C.c_ubyte * 64)()
        _lib.check(self.lib.dpt_symm_alloc(nbytes, C.byref(ptr), handle), "dpt_symm_alloc")
        local = torch.as_tensor(_RawCuda(ptr.value, nbytes), device=self.device)
        peer_ptrs = [0] * self.world
        peer_ptrs[self.rank] = ptr.value
        opened = []
        if self.world > 1:
            mine = (self.device.index, bytes(handle))
            everyone = [None] * self.world
            dist.all_gather_object(everyone, mine, group=self.group)
            for r, (dev, h) in enumerate(everyone):
                if r == self.rank:
                    continue
                p = C.c_void_p()
                buf = (C.c_ubyte * 64).from_buffer_copy(h)
                _lib.check(self.lib.dpt_symm_open(buf, C.byref(p)), f"dpt_symm_open(rank {r})")
                peer_ptrs[r] = p.value
                opened.append(p.value)
        sb = SymmBuffer(nbytes, local, peer_ptrs, opened)
        self._buffers.append(sb)
        return sb

    def alloc_multicast(self, nbytes: int):
        """Symmetric buffer that is additionally bound to an NVSwitch multicast object (NVLS).
        Returns None when multicast is not exposed here (callers fall back to ``alloc``)."""
        from . import nvls
        if self.world < 2 or not nvls.multicast_supported(self.device.index):
            return None
        ok = torch.tensor([1], device=self.device)
        try:
            buf = nvls.MulticastBuffer(self, nbytes)
        except Exception as e:  # noqa: BLE001
            import sys
            sys.stderr.write(f"[dpt_b200] rank {self.rank}: NVLS setup failed ({e}); using P2P kernels\n")
            buf = None
            ok.zero_()
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)
        if int(ok.item()) == 0:
            return None
        self._mc_buffers.append(buf)
        return buf

    def allreduce_nvls_(self, buf, byte_offset: int, numel: int, *, scale: float | None = None,
                        inv_scale: torch.Tensor | None = None, check_finite: bool = False,
                        ctas: int = 0, stream: int | None = None) -> None:
        """In-place fused all-reduce through the multicast mapping of ``buf``."""
        if scale is None:
            scale = 1.0 / self.world
        st = _lib.stream() if stream is None else stream
        fi = self.scalars.ptrs(0) if check_finite else None
        _lib.check(self.lib.dpt_allreduce_nvls(self.handle, buf.mc_ptr + byte_offset, numel, scale,
                                               _lib.ptr(inv_scale), fi, ctas, st),
                   "dpt_allreduce_nvls")
        self.launches += 1
        _lib.count()

    def autotune(self, buf, sizes=(1 << 18, 1 << 20, 1 << 22, 1 << 24), iters: int = 20,
                 ctas: int = 0) -> dict:
        """Init-time micro-benchmark: times the two-shot and, when ``buf`` is a
        multicast buffer, the NVLS all-reduce on ``buf`` for a few message sizes (CUDA events,
        max over ranks) and returns {"table": {bytes: {"twoshot_us", "nvls_us"}},
        "nvls_min_bytes": smallest size from which NVLS is the faster kernel (None if never)}.
        The buffer content is overwritten and zeroed again afterwards."""
        has_nvls = hasattr(buf, "mc_ptr")
        table = {}
        for nbytes in sizes:
            nbytes = min(int(nbytes), buf.nbytes) // 16 * 16
            n = nbytes // 4
            row = {}
            for name in (("twoshot", "nvls") if has_nvls else ("twoshot",)):
                fn = (lambda: self.allreduce_(buf, 0, n, ctas=ctas)) if name == "twoshot" else \
                    (lambda: self.allreduce_nvls_(buf, 0, n, ctas=ctas))
                for _ in range(3):
                    fn()
                torch.cuda.synchronize()
                self.host_barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(iters):
                    fn()
                e1.record()
                torch.cuda.synchronize()
                t = torch.tensor([e0.elapsed_time(e1) / iters * 1e3], device=self.device)
                if self.world > 1:
                    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
                row[name + "_us"] = float(t.item())
            table[nbytes] = row
        nvls_min = None
        if has_nvls:
            for nbytes in sorted(table, reverse=True):
                if table[nbytes]["nvls_us"] < table[nbytes]["twoshot_us"]:
                    nvls_min = nbytes
                else:
                    break
        buf.local.zero_()
        torch.cuda.synchronize()
        self.host_barrier()
        self.check_health()
        return {"table": table, "nvls_min_bytes": nvls_min}

    def host_barrier(self) -> None:
        if self.world > 1:
            dist.barrier(group=self.group)

    # ------------------------------------------------------------------ collectives
    def allreduce_(self, buf: SymmBuffer, byte_offset: int, numel: int, *, scale: float | None = None,
                   inv_scale: torch.Tensor | None = None, check_finite: bool = False,
                   ctas: int = 0, stream: int | None = None, sgd: dict | None = None) -> None:
        """In-place fused all-reduce of ``numel`` fp32 values at ``byte_offset`` of ``buf``."""
        if scale is None:
            scale = 1.0 / self.world
        st = _lib.stream() if stream is None else stream
        fi = self.scalars.ptrs(0) if check_finite else None
        if sgd is not None:
            args = (sgd["param"].ptrs(sgd["byte_offset"]), sgd["shadow"].ptrs(sgd["byte_offset"] // 2),
                    sgd["momentum"], sgd["lr"], sgd["momentum_coef"], sgd["weight_decay"],
                    int(sgd.get("nesterov", False)))
        else:
            args = (None, None, None, 0.0, 0.0, 0.0, 0)
        _lib.check(self.lib.dpt_allreduce_fused(self.handle, buf.ptrs(byte_offset), numel, scale,
                                                _lib.ptr(inv_scale), fi, *args, ctas, st),
                   "dpt_allreduce_fused")
        self.launches += 1
        _lib.count()

    def allreduce_oneshot(self, buf: SymmBuffer, byte_offset: int, out: torch.Tensor, *,
                          scale: float | None = None, ctas: int = 0, stream: int | None = None):
        if scale is None:
            scale = 1.0 / self.world
        st = _lib.stream() if stream is None else stream
        _lib.check(self.lib.dpt_allreduce_oneshot(self.handle, buf.ptrs(byte_offset), out.data_ptr(),
                                                  out.numel(), scale, None, None, ctas, st),
                   "dpt_allreduce_oneshot")
        self.launches += 1
        _lib.count()
        return out

    def allreduce_oneshot_push_(self, grads: torch.Tensor, staging: SymmBuffer, cap: int, *,
                                scale: float | None = None, inv_scale: torch.Tensor | None = None,
                                found_inf: torch.Tensor | None = None, sgd: dict | None = None,
                                ctas: int = 0, stream: int | None = None) -> None:
        """One-shot push all-reduce of the local fp32 tensor ``grads`` (numel % 4 == 0, at most
        ``cap``) through ``staging`` (a SymmBuffer of 2 * world * cap floats).  In place, or with
        ``sgd = {param, shadow, momentum (local tensors), lr, momentum_coef, weight_decay}`` the
        replicated fused update."""
        if scale is None:
            scale = 1.0 / self.world
        st = _lib.stream() if stream is None else stream
        if sgd is not None:
            args = (_lib.ptr(sgd["param"]), _lib.ptr(sgd["shadow"]), _lib.ptr(sgd["momentum"]),
                    sgd["lr"], sgd["momentum_coef"], sgd["weight_decay"],
                    int(sgd.get("nesterov", False)))
        else:
            args = (None, None, None, 0.0, 0.0, 0.0, 0)
        _lib.check(self.lib.dpt_allreduce_oneshot_push(
            self.handle, grads.data_ptr(), staging.ptrs(0), cap, grads.numel(), scale,
            _lib.ptr(inv_scale), _lib.ptr(found_inf), *args, ctas, st), "dpt_allreduce_oneshot_push")
        self.launches += 1
        _lib.count()

    def broadcast_(self, buf: SymmBuffer, byte_offset: int, nbytes: int, src: int = 0,
                   ctas: int = 0, stream: int | None = None) -> None:
        if self.world == 1:
            return
        st = _lib.stream() if stream is None else stream
        _lib.check(self.lib.dpt_broadcast(self.handle, buf.ptrs(byte_offset), nbytes, src, ctas, st),
                   "dpt_broadcast")
        self.launches += 1
        _lib.count()

    def barrier(self, stream: int | None = None) -> None:
        st = _lib.stream() if stream is None else stream
        _lib.check(self.lib.dpt_barrier(self.handle, st), "dpt_barrier")

    # ------------------------------------------------------------------ health
    def check_health(self) -> None:
        """Raises when a device-side flag wait timed out (a peer died or stalled)."""
        code = int(self.lib.dpt_comm_error(self.handle))
        if code:
            peer = (code >> 24) & 0x7F
            raise RuntimeError(f"[rank {self.rank}] peer-memory collective timed out waiting for "
                               f"rank {peer} (epoch {code & 0xFFFFFF}); a rank died or stalled")

    def epochs(self) -> list[int]:
        """Per-CTA barrier epoch counters of this rank (synchronises the device; debug only)."""
        out = (C.c_uint * 64)()
        n = int(self.lib.dpt_comm_epochs(self.handle, out, 64))
        if n < 0:
            raise RuntimeError(f"dpt_comm_epochs failed ({n})")
        return [int(out[i]) for i in range(n)]

    def validate_epochs(self) -> list[int]:
        """Debug check: every rank issues the same collectives with the same CTA
        counts, so at a quiescent point (after a step, all ranks) the epoch counters must agree.
        A mismatch means a rank skipped / repeated a collective — the next barrier would either
        time out or, worse, match a flag from a different collective."""
        mine = self.epochs()
        if self.world == 1:
            return mine
        everyone = [None] * self.world
        dist.all_gather_object(everyone, mine, group=self.group)
        for r, other in enumerate(everyone):
            if other != everyone[0]:
                raise RuntimeError(f"[rank {self.rank}] collective epochs differ between rank 0 and "
                                   f"rank {r}: {everyone[0][:8]}… vs {other[:8]}…")
        return mine

    # ------------------------------------------------------------------ watchdog / heartbeat
    def start_watchdog(self, interval_s: float = 1.0, abort: bool = False, heartbeat: bool = True,
                       stale_s: float = 45.0):
        """Host thread that (a) polls the device-written error word, so a peer that died or
        stalled is reported with (rank, peer, epoch) instead of hanging the job — the stand-in
        for ProcessGroupNCCL's watchdog / heartbeat monitor (ProcessGroupNCCL.hpp:610,690) —
        (b) publishes a per-rank heartbeat key in the torch store and (c) reports a peer whose
        heartbeat is older than ``stale_s`` (its process is gone or wedged), usually before the
        device-side wait times out.  ``abort=True`` ends the process with exit code 70 after the
        report (what TORCH_NCCL_ASYNC_ERROR_HANDLING=1 does under the reference)."""
        import os
        import sys
        import threading
        import time
        if getattr(self, "_wd_thread", None) is not None:
            return
        self._wd_stop = threading.Event()
        self.fault = None
        store = None
        if heartbeat and dist.is_available() and dist.is_initialized():
            try:
                store = dist.distributed_c10d._get_default_store()
            except Exception:  # noqa: BLE001
                store = None
        self._wd_store = store
        self._wd_t0 = time.time()

        def report(msg, fault):
            self.fault = fault
            sys.stderr.write(f"[dpt_b200 watchdog] rank {self.rank}: {msg}\n")
            sys.stderr.flush()
            if abort:
                os._exit(70)

        def run():
            tick = 0
            while not self._wd_stop.wait(interval_s):
                tick += 1
                try:
                    code = int(self.lib.dpt_comm_error(self.handle)) if self.handle else 0
                except Exception:  # noqa: BLE001
                    return
                if code:
                    peer = (code >> 24) & 0x7F
                    report(f"collective timed out waiting for rank {peer} (epoch {code & 0xFFFFFF})",
                           (self.rank, peer, code & 0xFFFFFF))
                    return
                if store is None:
                    continue
                try:
                    store.set(f"dpt_b200/heartbeat/{self.rank}", str(time.time()))
                    if tick % 3 == 0:
                        ages = self.peer_heartbeats()
                        late = [r for r, a in ages.items() if r != self.rank and a > stale_s]
                        if late and not self._wd_stop.is_set():
                            report(f"no heartbeat from rank(s) {late} for more than {stale_s:.0f} s "
                                   "(process gone or wedged)", (self.rank, late[0], -1))
                            return
                except Exception as e:  # noqa: BLE001 - the store lives in rank 0's process
                    if self._wd_stop.is_set():
                        return
                    report(f"rendezvous store unreachable ({type(e).__name__}): rank 0 is gone",
                           (self.rank, 0, -1))
                    return

        self._wd_thread = threading.Thread(target=run, name="dpt-watchdog", daemon=True)
        self._wd_thread.start()

    def peer_heartbeats(self) -> dict:
        """Age in seconds of every rank's last heartbeat (needs start_watchdog on all ranks)."""
        import time
        out = {}
        try:
            store = dist.distributed_c10d._get_default_store()
        except Exception:  # noqa: BLE001
            return out
        now = time.time()
        for r in range(self.world):
            key = f"dpt_b200/heartbeat/{r}"
            if not store.check([key]):      # get() would block on a key that is not there yet
                out[r] = now - getattr(self, "_wd_t0", now)
                continue
            raw = store.get(key).decode()
            out[r] = 0.0 if raw == "done" else now - float(raw)
        return out

    def stop_watchdog(self) -> None:
        if getattr(self, "_wd_thread", None) is not None:
            self._wd_stop.set()
            self._wd_thread.join(timeout=2.0)
            self._wd_thread = None
            st = getattr(self, "_wd_store", None)
            if st is not None:
                try:   # a rank that finished cleanly is not a dead rank
                    st.set(f"dpt_b200/heartbeat/{self.rank}", "done")
                except Exception:  # noqa: BLE001
                    pass

    def close(self) -> None:
        self.stop_watchdog()
        if getattr(self, "handle", None) is None:
            return
        torch.cuda.synchronize()
        self.host_barrier()
        for sb in self._buffers:
            for p in sb._opened:
                self.lib.dpt_symm_close(p)
        self.host_barrier()
        for sb in self._buffers:
            self.lib.dpt_symm_free(sb.peer_ptrs[self.rank])
        self._buffers.clear()
        for mb in self._mc_buffers:
            mb.release()
        self._mc_buffers.clear()
        self.lib.dpt_comm_destroy(self.handle)
        self.handle = None
