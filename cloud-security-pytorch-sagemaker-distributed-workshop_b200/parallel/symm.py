"""Symmetric-memory communicator: bootstrap, arena sub-allocation and collective launch.

This is the runtime under ``backend='b200'``.  Bootstrap (once per process group):

1. ``cuMemCreate`` an arena with a POSIX-fd shareable handle and map it locally (``csrc/symm.cu``);
2. exchange the fds with every peer over ``AF_UNIX`` sockets (``SCM_RIGHTS``); socket paths are
   published through the torch.distributed control group — the only thing torch.distributed does here;
3. import and map every peer's arena -> ``peer[r]`` base pointers (NVLink P2P loads/stores);
4. rank 0 creates an NVSwitch multicast object, everyone adds its device and binds its arena, and the
   multicast VA is mapped -> ``mc`` (``multimem.ld_reduce`` / ``multimem.st``).  Missing multicast
   support only disables the NVLS algorithm.

The first MiB of the arena holds the signal pads; the rest is handed out by a bump allocator
(``alloc_arena``) as torch tensors that alias the mapping, e.g. the flat gradient arena of
``DistributedDataParallel`` and a staging window for ``all_reduce`` of ordinary tensors.
"""
from __future__ import annotations

import ctypes
import os
import socket
import tempfile
import threading
import uuid
from ctypes import byref, c_int, c_size_t, c_uint64, c_void_p
from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

from ..ops import _lib as L
from ..utils.config import CommConfig
from ..utils.logging import get_logger

log = get_logger(__name__)

_ELEM = {torch.float32: 4, torch.bfloat16: 2, torch.uint8: 1, torch.int32: 4, torch.int64: 8, torch.float16: 2}


def _exchange_fds(rank: int, world: int, my_fd: int, tag: str) -> List[int]:
    """All-to-all exchange of one file descriptor per rank (SCM_RIGHTS over unix sockets)."""
    path = os.path.join(tempfile.gettempdir(), "b200ddp_%s_%s_%d.sock" % (tag, uuid.uuid4().hex[:8], rank))
    server = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    server.bind(path)
    server.listen(world)

    def serve():
        for _ in range(world - 1):
            conn, _ = server.accept()
            with conn:
                conn.recv(1)
                socket.send_fds(conn, [b"f"], [my_fd])

    th = threading.Thread(target=serve, daemon=True)
    th.start()
    paths = [None] * world
    dist.all_gather_object(paths, path)
    fds = [-1] * world
    fds[rank] = my_fd
    for r in range(world):
        if r == rank:
            continue
        c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        c.connect(paths[r])
        c.sendall(b"r")
        _, got, _, _ = socket.recv_fds(c, 16, 1)
        c.close()
        fds[r] = got[0]
    th.join()
    server.close()
    dist.barrier()
    try:
        os.unlink(path)
    except OSError:
        pass
    return fds


class SymmComm:
    def __init__(self, rank: int, world_size: int, device: torch.device, arena_bytes: Optional[int] = None):
        if world_size > L.MAX_RANKS:
            raise RuntimeError("backend 'b200' is a single-node NVSwitch backend: world_size <= %d" % L.MAX_RANKS)
        self.rank, self.world, self.device = rank, world_size, device
        self.cfg = CommConfig()
        # algorithm thresholds measured on 2 and 8 B200s (profiles/allreduce_sweep.md): one-shot pulls
        # (N-1) x S per GPU, so its cut-over shrinks with N; with 2 ranks the switch reduction buys nothing
        if "B200DDP_ONESHOT_MAX_BYTES" not in os.environ and world_size > 1:
            self.cfg.oneshot_max_bytes = max(64 * 1024, (2 << 20) // (world_size - 1))
        self.prefer_nvls = world_size > 2 or "B200DDP_NVLS_MIN_BYTES" in os.environ
        self.lib = L.lib()
        dev = device.index if device.index is not None else torch.cuda.current_device()
        self.dev_index = dev
        L.check(self.lib.b200_symm_init(c_int(dev)), "symm_init")

        gran, mcg = c_size_t(0), c_size_t(0)
        L.check(self.lib.b200_symm_granularity(c_int(dev), byref(gran), byref(mcg)), "granularity")
        align = max(gran.value, mcg.value, 2 << 20)
        want = arena_bytes or int(os.environ.get("B200DDP_ARENA_MB", "768")) * (1 << 20)
        self.arena_bytes = (want + align - 1) // align * align
        self.staging_bytes = min(int(os.environ.get("B200DDP_STAGING_MB", "128")) * (1 << 20), self.arena_bytes // 4)

        # ---- 1. local arena
        p, fd, h = c_void_p(0), c_int(-1), c_uint64(0)
        L.check(self.lib.b200_symm_alloc(c_int(dev), c_size_t(self.arena_bytes), byref(p), byref(fd), byref(h)), "symm_alloc")
        self.local_ptr, self.local_fd, self.local_handle = p.value, fd.value, h.value
        self.peer_ptrs = [0] * world_size
        self.peer_handles = [0] * world_size
        self.peer_ptrs[rank] = self.local_ptr
        self.mc_ptr, self.mc_handle = 0, 0

        self.bytes_view = L.tensor_from_ptr(self.local_ptr, self.arena_bytes, device)
        self.bytes_view[: L.SIGNAL_BYTES].zero_()
        torch.cuda.synchronize(device)

        # ---- 2/3. peers
        if world_size > 1:
            fds = _exchange_fds(rank, world_size, self.local_fd, "mem")
            for r in range(world_size):
                if r == rank:
                    continue
                pp, hh = c_void_p(0), c_uint64(0)
                L.check(self.lib.b200_symm_import(c_int(dev), c_int(fds[r]), c_size_t(self.arena_bytes), byref(pp), byref(hh)),
                        "symm_import(peer %d)" % r)
                self.peer_ptrs[r], self.peer_handles[r] = pp.value, hh.value
                self.lib.b200_close_fd(c_int(fds[r]))
            self._setup_multicast(dev)
        self.lib.b200_close_fd(c_int(self.local_fd))

        # ---- error flag + communicator struct
        hp, dp = c_void_p(0), c_void_p(0)
        L.check(self.lib.b200_comm_alloc_errflag(byref(hp), byref(dp)), "errflag")
        self._err_host = ctypes.cast(hp, ctypes.POINTER(c_int))
        cs = L.CommStruct()
        cs.rank, cs.world = rank, world_size
        for r in range(world_size):
            cs.peer[r] = self.peer_ptrs[r]
        cs.mc = self.mc_ptr or None
        cs.err_flag = dp.value
        cs.timeout_ns = int(self.cfg.spin_timeout_ms) * 1_000_000
        self.cs = cs

        # ---- streams / bump allocator
        self.comm_stream = torch.cuda.Stream(device=device, priority=-1)
        self._ready_event = torch.cuda.Event()
        self._done_event = torch.cuda.Event()
        self._overlap_dirty = False
        self._bump = L.SIGNAL_BYTES
        self._staging_off = self.arena_bytes - self.staging_bytes
        self.launch_log: List[Tuple[str, int]] = []
        if world_size > 1:
            dist.barrier()          # every signal pad is zeroed and mapped before the first kernel
        log.debug("b200ddp.SymmComm rank %d/%d: arena %d MiB, NVLS multicast %s", rank, world_size,
                  self.arena_bytes >> 20, "bound" if self.mc_ptr else "unavailable")

    # ------------------------------------------------------------------ multicast (NVLS)
    def _all_ok(self, ok: bool) -> bool:
        """Collective AND over ranks — every bootstrap step is agreed on, so a failure on one rank
        turns NVLS off everywhere instead of leaving the others stuck in a barrier."""
        flags = [None] * self.world
        dist.all_gather_object(flags, bool(ok))
        return all(flags)

    def _setup_multicast(self, dev: int):
        if not self._all_ok(bool(self.cfg.use_nvls) and self.lib.b200_mc_supported(c_int(dev)) == 1):
            return
        mch, fd = c_uint64(0), c_int(-1)
        ok = True
        if self.rank == 0:
            ok = self.lib.b200_mc_create(c_int(self.world), c_size_t(self.arena_bytes), byref(mch), byref(fd)) == 0
            if not ok:
                log.debug("multicast create failed: %s", self.lib.b200_last_error().decode())
        if not self._all_ok(ok):
            return
        my_fd = fd.value if self.rank == 0 else os.open(os.devnull, os.O_RDONLY)
        fds = _exchange_fds(self.rank, self.world, my_fd, "mc")
        if self.rank != 0:
            ok = self.lib.b200_mc_import(c_int(fds[0]), byref(mch)) == 0
        for r in range(self.world):
            self.lib.b200_close_fd(c_int(fds[r]))
        if not self._all_ok(ok):
            log.debug("multicast import failed: %s", self.lib.b200_last_error().decode())
            return
        ok = self.lib.b200_mc_add_device(mch, c_int(dev)) == 0
        if not self._all_ok(ok):            # also the "every device added before anyone binds" barrier
            log.debug("multicast add_device failed: %s", self.lib.b200_last_error().decode())
            return
        mp = c_void_p(0)
        ok = self.lib.b200_mc_bind_map(c_int(dev), mch, c_uint64(self.local_handle), c_size_t(self.arena_bytes), byref(mp)) == 0
        if not ok:
            log.debug("multicast bind/map failed: %s", self.lib.b200_last_error().decode())
        if self._all_ok(ok):
            self.mc_ptr, self.mc_handle = mp.value, mch.value

    @property
    def has_nvls(self) -> bool:
        return bool(self.mc_ptr)

    # ------------------------------------------------------------------ arena
    def alloc_arena(self, numel: int, dtype: torch.dtype) -> Tuple[torch.Tensor, int]:
        """Carve ``numel`` elements out of the symmetric arena.  Every rank must make the same
        sequence of calls (offsets are symmetric).  Returns ``(tensor, byte_offset)``."""
        nbytes = numel * _ELEM[dtype]
        off = (self._bump + 4095) // 4096 * 4096
        end = off + (nbytes + 255) // 256 * 256
        if end > self._staging_off:
            raise RuntimeError("symmetric arena exhausted: need %d bytes, %d free (raise B200DDP_ARENA_MB)"
                               % (nbytes, self._staging_off - off))
        self._bump = end
        t = self.bytes_view[off: off + nbytes].view(dtype)
        t.zero_()
        return t, off

    def offset_of(self, t: torch.Tensor) -> Optional[int]:
        p = t.data_ptr()
        if self.local_ptr <= p < self.local_ptr + self.arena_bytes:
            return p - self.local_ptr
        return None

    # ------------------------------------------------------------------ collectives
    def pick_algo(self, nbytes: int) -> int:
        """The native selector (``b200_pick_algo``): the reducer and ``B200_ALGO_AUTO`` take the same decision."""
        return int(self.lib.b200_pick_algo(c_int(self.world), c_int(int(self.has_nvls and self.prefer_nvls)), c_size_t(nbytes),
                                           c_size_t(self.cfg.oneshot_max_bytes), c_size_t(self.cfg.nvls_min_bytes)))

    def _launch_allreduce(self, offset: int, numel: int, dtype: torch.dtype, scale: float, stream, algo: int = 0,
                          max_ctas: Optional[int] = None):
        nbytes = numel * _ELEM[dtype]
        if algo == L.ALGO_AUTO:
            algo = self.pick_algo(nbytes)
        if self.cfg.debug:
            log.debug("b200ddp allreduce: %d bytes at +%d -> %s", nbytes, offset, L.ALGO_NAMES[algo])
        self.launch_log.append((L.ALGO_NAMES[algo], nbytes))
        L.check(self.lib.b200_allreduce(byref(self.cs), c_size_t(offset), c_size_t(numel), c_int(L.dtype_code(dtype)),
                                        ctypes.c_float(scale), c_int(algo), c_int(max_ctas or self.cfg.max_ctas), L.stream_ptr(stream)),
                "allreduce")

    def all_reduce_arena(self, offset_bytes: int, numel: int, scale: float = 1.0, dtype: torch.dtype = torch.float32,
                         overlap: bool = False, algo: int = 0):
        """All-reduce (sum * scale) a range of the symmetric arena in place.  ``overlap=True`` launches on
        the comm stream after everything queued so far on the current stream; ``join_overlap`` joins."""
        if self.world == 1 and scale == 1.0:
            return
        if overlap:
            cur = torch.cuda.current_stream(self.device)
            self._ready_event.record(cur)
            self.comm_stream.wait_event(self._ready_event)
            from ..ops import streams
            for other in streams.producers(self.device):   # bucket members written on the other stream
                ev = torch.cuda.Event()
                ev.record(other)
                self.comm_stream.wait_event(ev)
            self._launch_allreduce(offset_bytes, numel, dtype, scale, self.comm_stream, algo, self.cfg.overlap_ctas)
            self._overlap_dirty = True
        else:
            self._launch_allreduce(offset_bytes, numel, dtype, scale, torch.cuda.current_stream(self.device), algo)

    def join_overlap(self):
        if self._overlap_dirty:
            self._done_event.record(self.comm_stream)
            torch.cuda.current_stream(self.device).wait_event(self._done_event)
            self._overlap_dirty = False

    def all_reduce_(self, tensor: torch.Tensor, op: str = "sum"):
        """In-place all-reduce of an arbitrary CUDA tensor (``dist.all_reduce`` replacement)."""
        if op not in ("sum", "avg"):
            cpu = tensor.detach().cpu()
            dist.all_reduce(cpu, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.MIN)
            tensor.copy_(cpu)
            return
        scale = 1.0 / self.world if op == "avg" else 1.0
        if self.world == 1:
            return
        dtype = tensor.dtype
        if dtype not in (torch.float32, torch.bfloat16):
            tmp = tensor.float()
            self.all_reduce_(tmp, op)
            tensor.copy_(tmp)
            return
        off = self.offset_of(tensor)
        es = _ELEM[dtype]
        if off is not None and tensor.is_contiguous() and off % 16 == 0 and (tensor.numel() * es) % 16 == 0:
            self._launch_allreduce(off, tensor.numel(), dtype, scale, torch.cuda.current_stream(self.device))
            return
        flat = tensor.contiguous().view(-1)
        per = self.staging_bytes // es // 64 * 64
        stage = self.bytes_view[self._staging_off: self._staging_off + self.staging_bytes].view(dtype)
        for s in range(0, flat.numel(), per):
            n = min(per, flat.numel() - s)
            npad = (n * es + 15) // 16 * 16 // es
            stage[:n].copy_(flat[s: s + n])
            if npad > n:
                stage[n:npad].zero_()
            self._launch_allreduce(self._staging_off, npad, dtype, scale, torch.cuda.current_stream(self.device))
            flat[s: s + n].copy_(stage[:n])
        if flat.data_ptr() != tensor.data_ptr():
            tensor.copy_(flat.view(tensor.shape))

    def broadcast_(self, tensor: torch.Tensor, src: int = 0):
        if self.world == 1:
            return
        stream = torch.cuda.current_stream(self.device)
        off = self.offset_of(tensor)
        nbytes = tensor.numel() * tensor.element_size()
        if off is not None and tensor.is_contiguous() and off % 16 == 0 and nbytes % 16 == 0:
            L.check(self.lib.b200_broadcast(byref(self.cs), c_size_t(off), c_size_t(nbytes), c_int(src),
                                            c_int(self.cfg.max_ctas), L.stream_ptr(stream)), "broadcast")
            return
        flat = tensor.contiguous().view(-1).view(torch.uint8)
        stage = self.bytes_view[self._staging_off: self._staging_off + self.staging_bytes]
        for s in range(0, flat.numel(), self.staging_bytes):
            n = min(self.staging_bytes, flat.numel() - s)
            npad = (n + 15) // 16 * 16
            if self.rank == src:
                stage[:n].copy_(flat[s: s + n])
            L.check(self.lib.b200_broadcast(byref(self.cs), c_size_t(self._staging_off), c_size_t(npad), c_int(src),
                                            c_int(self.cfg.max_ctas), L.stream_ptr(stream)), "broadcast")
            if self.rank != src:
                flat[s: s + n].copy_(stage[:n])
        if flat.data_ptr() != tensor.data_ptr():
            tensor.copy_(flat.view(tensor.dtype).view(tensor.shape))

    def barrier(self):
        L.check(self.lib.b200_barrier(byref(self.cs), L.stream_ptr(torch.cuda.current_stream(self.device))), "barrier")
        torch.cuda.synchronize(self.device)
        self.check_health()

    def check_health(self):
        """Raise if a device-side spin timed out (peer never arrived) — SURVEY §5.3."""
        if self._err_host[0] != 0:
            peer, block = self._err_host[1], self._err_host[2]
            raise RuntimeError("b200ddp: rank %d timed out waiting for rank %d in a collective (CTA %d); "
                               "a peer died or launched a different collective sequence" % (self.rank, peer, block))

    def close(self):
        try:
            torch.cuda.synchronize(self.device)
        except Exception:
            pass
        if self.mc_ptr:
            self.lib.b200_mc_release(c_int(self.dev_index), c_uint64(self.mc_handle), c_void_p(self.mc_ptr), c_size_t(self.arena_bytes))
            self.mc_ptr = 0
        for r in range(self.world):
            if r != self.rank and self.peer_ptrs[r]:
                self.lib.b200_symm_unmap(c_void_p(self.peer_ptrs[r]), c_size_t(self.arena_bytes), c_uint64(self.peer_handles[r]))
                self.peer_ptrs[r] = 0
        # the local arena stays mapped until process exit: torch tensors may still alias it
