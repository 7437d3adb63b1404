"""Communicator over NVLink peer memory.

The reference never calls a collective itself: DDP/NCCL do it implicitly behind
``train.torch.prepare_model`` (reference ``torch/estimator.py:243``).  Here the gradient
all-reduce is a hand-written kernel (``csrc/comm/p2p_allreduce.cu``) over *symmetric memory*:
every rank ``cudaMalloc``s one region, the 64-byte CUDA IPC handles travel through the
``torch.distributed`` store, and each rank maps all peers' regions, after which ranks read
and write each other's HBM directly through NVSwitch.  ``torch.distributed`` (NCCL) stays the
bootstrap and the baseline.

``SymmetricMemory``  -- allocation + handle exchange.
``P2PReducer``       -- ``allreduce_()`` (sum in place) and ``allreduce_step_()`` (sum, scale and
                       Adam/SGD update in the same kernel).
``LocalEmulation``   -- all ranks of a world on ONE GPU in one cooperative launch: how the
                       flag protocol is tested where only a single GPU is available.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence

import torch
import torch.distributed as dist

from raydp_b200 import ops
from raydp_b200.ops._abi import OptArgs

FLAG_BYTES = 256
ALGO = {"auto": 0, "one_shot": 1, "two_shot": 2, "two_shot_fixed": 3, "ll": 4, "bf16": 5}
# "bf16": two-shot with bf16 on the wire and fp32 accumulation (half the NVLink bytes; for link-bound buckets)
LL_MAX_BYTES = 1 << 20          # "auto" uses the one-round LL protocol up to this message size
LL_SOURCE_SLOTS = 8             # receive slots per parity in the LL layout (= kMaxRanks of the kernels)


class CommTimeoutError(RuntimeError):
    """A wait on a peer's flag ran past the deadline: a rank died or never entered the collective."""


_status_host: Optional[torch.Tensor] = None
_status_devices = set()
_sym_counter = 0


def enable_timeouts(seconds: Optional[float] = None) -> None:
    """Arm the bounded peer waits of every communication kernel on the current device (idempotent).
    ``RAYDP_B200_COMM_TIMEOUT_S`` (default 60; 0 = wait forever) sets the deadline."""
    import os
    global _status_host
    dev = torch.cuda.current_device()
    if dev in _status_devices and seconds is None:
        return
    if seconds is None:
        seconds = float(os.environ.get("RAYDP_B200_COMM_TIMEOUT_S", "60"))
    if _status_host is None:
        _status_host = torch.zeros(16, dtype=torch.int32).pin_memory()
    ops._check(ops.lib().rdp_comm_set_timeout(ctypes.c_double(seconds), ctypes.c_void_p(_status_host.data_ptr())))
    _status_devices.add(dev)


def comm_failed() -> int:
    return int(_status_host[0]) if _status_host is not None else 0


def check_status() -> None:
    """Raise if a communication kernel of this process gave up waiting for a peer (call after a stream
    synchronisation; the training engines do so once per epoch and at close)."""
    code = comm_failed()
    if code:
        raise CommTimeoutError(f"a wait on a peer GPU timed out (code 0x{code:x}): a rank of the group died or "
                               "did not reach the collective")


def _host_barrier(tag: str, world: int, timeout_s: float = 15.0) -> bool:
    """Rendezvous through the process group's key-value store, on the host, with a deadline: used in
    teardown, where a collective barrier would hang forever on a peer that is already gone."""
    import time
    try:
        store = dist.distributed_c10d._get_default_store()
        store.add(tag, 1)
        t0 = time.time()
        while int(store.add(tag, 0)) < world:
            if time.time() - t0 > timeout_s:
                return False
            time.sleep(0.005)
        return True
    except Exception:  # noqa: BLE001 - store gone: the group is being torn down
        return False


class _RawCudaBuffer:
    """Exposes a raw device pointer through ``__cuda_array_interface__`` (zero-copy torch view)."""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2, "strides": None}


_TYPESTR = {torch.float32: "<f4", torch.int32: "<i4", torch.uint8: "|u1", torch.int16: "<i2"}


def tensor_from_ptr(ptr: int, shape, dtype: torch.dtype, device: torch.device) -> torch.Tensor:
    """Torch tensor over memory this process did not get from torch (symmetric / IPC regions).
    bf16 has no array-interface typestr: it is viewed through int16."""
    if dtype == torch.bfloat16:
        return torch.as_tensor(_RawCudaBuffer(ptr, shape, "<i2"), device=device).view(torch.bfloat16)
    return torch.as_tensor(_RawCudaBuffer(ptr, shape, _TYPESTR[dtype]), device=device)


def p2p_available() -> bool:
    """True when every GPU of this host's ranks can map every other one."""
    if not torch.cuda.is_available() or not dist.is_initialized():
        return False
    n = torch.cuda.device_count()
    me = torch.cuda.current_device()
    lib = ops.lib()
    ok = all(lib.rdp_comm_can_access_peer(me, d) for d in range(n) if d != me) and n >= dist.get_world_size()
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    return bool(flag.item())


def _exchange_fds(offered: List[int], rank: int, world: int, group, tag: str) -> List[List[int]]:
    """Pass file descriptors between the ranks' processes over unix sockets (``SCM_RIGHTS``).  Every rank
    serves ``offered`` to each peer and collects what the peers offer; the socket paths travel through the
    process group.  Returns ``fds[q]`` = descriptors received from rank q (``offered`` itself for q == rank)."""
    import os
    import socket
    import tempfile
    import threading
    path = os.path.join(tempfile.gettempdir(), f"rdp_fd_{os.getpid()}_{tag}.sock")
    if os.path.exists(path):
        os.unlink(path)
    srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    srv.bind(path)
    srv.listen(world)
    srv.settimeout(60.0)
    errors: List[BaseException] = []

    def serve() -> None:
        try:
            for _ in range(world - 1):
                conn, _ = srv.accept()
                with conn:
                    socket.send_fds(conn, [len(offered).to_bytes(4, "little")], list(offered))
        except BaseException as exc:  # noqa: BLE001
            errors.append(exc)
    thread = threading.Thread(target=serve, name="rdp-fd-serve", daemon=True)
    thread.start()
    paths: List[Optional[str]] = [None] * world
    dist.all_gather_object(paths, path, group=group)
    out: List[List[int]] = []
    try:
        for q in range(world):
            if q == rank:
                out.append(list(offered))
                continue
            with socket.socket(socket.AF_UNIX, socket.SOCK_STREAM) as c:
                c.settimeout(60.0)
                c.connect(paths[q])
                msg, fds, _flags, _addr = socket.recv_fds(c, 4, 8)
                if len(fds) != int.from_bytes(msg, "little"):
                    raise RuntimeError(f"rank {q} announced {int.from_bytes(msg, 'little')} descriptors, got {len(fds)}")
                out.append(list(fds))
        thread.join(90.0)
        if errors:
            raise RuntimeError(f"descriptor exchange failed: {errors[0]}")
    finally:
        srv.close()
        try:
            os.unlink(path)
        except OSError:
            pass
    return out


def symm_support(device: Optional[int] = None) -> int:
    """0: only CUDA IPC; 1: driver-API symmetric memory (cuMemCreate + fd exchange); 2: + NVLS multicast."""
    import os
    if os.environ.get("RAYDP_B200_SYMM", "vmm") == "ipc":
        return 0
    dev = torch.cuda.current_device() if device is None else device
    return int(ops.lib().rdp_symm_supported(dev))


class SymmetricMemory:
    """One region per rank, mapped by all ranks.  ``bases[q]`` is rank q's region in this process.

    Preferred plumbing: every rank ``cuMemCreate``s its region as an allocation
    shareable by POSIX file descriptor, the descriptors cross process boundaries over unix sockets, and each
    rank maps all peers' allocations (``rdp_symm_*`` in ``csrc/comm/symmem.cu``).  With ``multicast=True``
    rank 0 also creates a multicast object over all devices, every rank binds its region to it and maps it:
    ``mc_ptr`` is the address ``multimem.ld_reduce`` / ``multimem.st`` operate on (NVLS).  CUDA IPC
    (``cudaIpcGetMemHandle``) remains as the fallback where the driver lacks the virtual-memory API
    (``RAYDP_B200_SYMM=ipc`` forces it).  Teardown never needs a peer's cooperation."""

    def __init__(self, nbytes: int, group=None, multicast: bool = False):
        global _sym_counter
        self.lib = ops.lib()
        enable_timeouts()
        _sym_counter += 1                          # same creation order on every rank: names keys and sockets
        self._tag = f"rdp_sym_close_{_sym_counter}"
        self._group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.nbytes = int(nbytes)
        self.mc_ptr = 0
        self._handles: List[int] = []
        self._mc_handle = 0
        self.device = torch.cuda.current_device()
        support = symm_support(self.device)
        # all ranks must take the same path
        flag = torch.tensor([support], device="cuda", dtype=torch.int32)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        support = int(flag.item())
        if multicast and support < 2:
            raise RuntimeError("this system exposes no NVLS multicast support")
        self.mode = "vmm" if support >= 1 else "ipc"
        if self.mode == "vmm":
            self._init_vmm(multicast)
        else:
            self._init_ipc()

    # ---- driver virtual-memory API + fd exchange ------------------------------------------------
    def _init_vmm(self, multicast: bool) -> None:
        lib, vp = self.lib, ctypes.c_void_p
        gran = ctypes.c_uint64()
        ops._check(lib.rdp_symm_granularity(self.device, 1 if multicast else 0, self.world, ctypes.byref(gran)))
        size = (self.nbytes + gran.value - 1) // gran.value * gran.value
        self.size = size
        handle, ptr = vp(), vp()
        ops._check(lib.rdp_symm_alloc(size, self.device, ctypes.byref(handle), ctypes.byref(ptr)))
        self.local = ptr.value
        self._own_handle = handle.value
        fd = ctypes.c_int(-1)
        ops._check(lib.rdp_symm_export_fd(vp(self._own_handle), ctypes.byref(fd)))
        offered = [fd.value]
        if multicast and self.rank == 0:
            mc = vp()
            ops._check(lib.rdp_mc_create(size, self.world, ctypes.byref(mc)))
            self._mc_handle = mc.value
            mfd = ctypes.c_int(-1)
            ops._check(lib.rdp_symm_export_fd(vp(self._mc_handle), ctypes.byref(mfd)))
            offered.append(mfd.value)
        fds = _exchange_fds(offered, self.rank, self.world, self._group, f"{_sym_counter}")
        self.bases = []
        for q in range(self.world):
            if q == self.rank:
                self.bases.append(self.local)
                continue
            h, p = vp(), vp()
            ops._check(lib.rdp_symm_import_fd(fds[q][0], size, self.device, ctypes.byref(h), ctypes.byref(p)))
            self._handles.append(h.value)
            self.bases.append(p.value)
        if multicast:
            if self.rank != 0:
                mc = vp()
                ops._check(lib.rdp_mc_import_fd(fds[0][1], size, ctypes.byref(mc)))
                self._mc_handle = mc.value
            ops._check(lib.rdp_mc_add_device(vp(self._mc_handle), self.device))
            if not _host_barrier(self._tag + "_added", self.world, 60.0):
                raise RuntimeError("multicast set-up: not every rank added its device")
            ops._check(lib.rdp_mc_bind(vp(self._mc_handle), vp(self._own_handle)))
            if not _host_barrier(self._tag + "_bound", self.world, 60.0):
                raise RuntimeError("multicast set-up: not every rank bound its memory")
            mptr = vp()
            ops._check(lib.rdp_mc_map(vp(self._mc_handle), self.device, ctypes.byref(mptr)))
            self.mc_ptr = mptr.value
        for per_rank in fds:                       # the mappings keep the memory alive; the fds can go
            for f in per_rank:
                lib.rdp_close_fd(int(f))
        torch.cuda.synchronize()
        if not _host_barrier(self._tag + "_ready", self.world, 60.0):
            raise RuntimeError("symmetric memory set-up: a rank did not finish mapping")

    # ---- CUDA IPC fallback ------------------------------------------------------------------------
    def _init_ipc(self) -> None:
        ptr = ctypes.c_void_p()
        ops._check(self.lib.rdp_comm_alloc(self.nbytes, ctypes.byref(ptr)))
        self.local = ptr.value
        self.size = self.nbytes
        handle = ctypes.create_string_buffer(64)
        ops._check(self.lib.rdp_comm_ipc_handle(ctypes.c_void_p(self.local), handle))
        gathered: List[Optional[bytes]] = [None] * self.world
        dist.all_gather_object(gathered, bytes(handle.raw), group=self._group)
        self.bases: List[int] = []
        self._opened: List[int] = []
        for q, blob in enumerate(gathered):
            if q == self.rank:
                self.bases.append(self.local)
                continue
            out = ctypes.c_void_p()
            buf = ctypes.create_string_buffer(blob, 64)
            ops._check(self.lib.rdp_comm_ipc_open(buf, ctypes.byref(out)))
            self.bases.append(out.value)
            self._opened.append(out.value)
        dist.barrier(self._group)

    def close(self) -> None:
        if self.local is None:
            return
        torch.cuda.synchronize()
        # nobody may still be reading this region when it is freed -- but a peer that died will never answer:
        # rendezvous on the host with a deadline instead of a collective barrier
        if not comm_failed():
            _host_barrier(self._tag, self.world)
        vp = ctypes.c_void_p
        if self.mode == "vmm":
            if self._mc_handle:
                self.lib.rdp_symm_free(vp(self._mc_handle))
            for h in self._handles:
                self.lib.rdp_symm_free(vp(h))
            self.lib.rdp_symm_free(vp(self._own_handle))
        else:
            for ptr in self._opened:
                self.lib.rdp_comm_ipc_close(vp(ptr))
            self.lib.rdp_comm_free(vp(self.local))
        self.local = None


def _opt_args(kind: int = 0, p=None, m=None, v=None, lr=0.0, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0,
              momentum=0.0, dampening=0.0, adamw=False, nesterov=False, lr_dev=None, step_dev=None,
              shadow=None) -> OptArgs:
    """``shadow``: optional bf16 tensor (same numel as ``p``) that receives the updated parameters."""
    ptr = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None  # noqa: E731
    return OptArgs(kind, ptr(p), ptr(m), ptr(v), lr, betas[0], betas[1], eps, weight_decay, momentum,
                   dampening, 1 if adamw else 0, 1 if nesterov else 0, ptr(lr_dev), ptr(step_dev), ptr(shadow))


def _launch(lib, bases: Sequence[int], slot_floats: int, rank: int, world: int, ins, outs, epochs, counters,
            opts: Optional[Sequence[OptArgs]], n: int, scale: float, algo: int) -> None:
    nl = len(ins)
    vp = ctypes.c_void_p
    arr = lambda xs: (vp * len(xs))(*[vp(x) for x in xs])  # noqa: E731
    opt_arr = (OptArgs * nl)(*opts) if opts is not None else None
    if algo == ALGO["bf16"]:
        assert n % 8 == 0, "the bf16-wire all-reduce moves 8 values per 16-byte load: pad the buffer to 8"
        ops._check(lib.rdp_allreduce_bf16(arr(bases), slot_floats, rank, world, nl,
                                          arr([t.data_ptr() for t in ins]),
                                          arr([t.data_ptr() if t is not None else 0 for t in outs]),
                                          arr([t.data_ptr() for t in epochs]), arr([t.data_ptr() for t in counters]),
                                          opt_arr, n, scale, vp(torch.cuda.current_stream().cuda_stream)))
        return
    ops._check(lib.rdp_allreduce(arr(bases), slot_floats, rank, world, nl,
                                 arr([t.data_ptr() for t in ins]),
                                 arr([t.data_ptr() if t is not None else 0 for t in outs]),
                                 arr([t.data_ptr() for t in epochs]), arr([t.data_ptr() for t in counters]),
                                 opt_arr, n, scale, algo,
                                 vp(torch.cuda.current_stream().cuda_stream)))


def _launch_ll(lib, bases: Sequence[int], slot_floats: int, rank: int, world: int, ins, outs, epochs, counters,
               opts: Optional[Sequence[OptArgs]], n: int, scale: float) -> None:
    nl = len(ins)
    vp = ctypes.c_void_p
    arr = lambda xs: (vp * len(xs))(*[vp(x) for x in xs])  # noqa: E731
    opt_arr = (OptArgs * nl)(*opts) if opts is not None else None
    ops._check(lib.rdp_allreduce_ll(arr(bases), slot_floats, rank, world, nl,
                                    arr([t.data_ptr() for t in ins]),
                                    arr([t.data_ptr() if t is not None else 0 for t in outs]),
                                    arr([t.data_ptr() for t in epochs]), arr([t.data_ptr() for t in counters]),
                                    opt_arr, n, scale, vp(torch.cuda.current_stream().cuda_stream)))


class P2PReducer:
    """Gradient all-reduce for one flat fp32 buffer."""

    kind = "p2p"

    def __init__(self, grads: Optional[torch.Tensor], device: torch.device, group=None, algo: str = "auto",
                 n: Optional[int] = None):
        """``grads``: the flat fp32 gradient buffer to reduce, or None with ``n`` given for *registered*
        mode, where ``self.grads`` IS the symmetric send buffer (``algo="two_shot_fixed"``)."""
        self.n = grads.numel() if grads is not None else int(n)
        assert self.n % 4 == 0, "flat buffer must be 16-byte granular"
        self.slot = (self.n + 3) // 4 * 4
        if algo == "auto" and grads is not None and self.n * 4 <= LL_MAX_BYTES:
            # latency-bound: one round, flag carried with the data.  LL moves twice the bytes to every peer, so its
            # range shrinks with the world size (8 GPUs: LL 25 us vs one-shot 30 us at 194 KB, 36 vs 33 us at 512 KB,
            # 60 vs 44 us at 1 MB; 2 GPUs: LL wins up to 1 MB -- profiles/r2/allreduce_{2,8}gpu.json)
            world = dist.get_world_size(group) if dist.is_initialized() else 1
            algo = "ll" if self.n * 4 * world <= 2 * LL_MAX_BYTES else "one_shot"
        self.algo = ALGO[algo]
        self.ll = self.algo == ALGO["ll"]
        if self.ll:
            assert grads is not None, "the LL protocol reduces a caller-owned buffer"
            self.sym = SymmetricMemory(2 * LL_SOURCE_SLOTS * 2 * self.slot * 4, group)
        elif self.algo == ALGO["bf16"]:
            # the four slots keep their offsets but hold bf16: half the slot size is enough
            assert grads is not None and self.n % 8 == 0, "bf16 wire: caller-owned buffer, multiple of 8 values"
            self.slot = (self.n // 2 + 3) // 4 * 4
            self.sym = SymmetricMemory(4 * self.slot * 4 + FLAG_BYTES, group)
        else:
            self.sym = SymmetricMemory(4 * self.slot * 4 + FLAG_BYTES, group)
        if grads is None:
            grads = tensor_from_ptr(self.sym.local, (self.n,), torch.float32, device)
            self.recv0 = tensor_from_ptr(self.sym.local + 2 * self.slot * 4, (self.n,), torch.float32, device)
        assert grads.dtype == torch.float32 and grads.is_contiguous() and grads.data_ptr() % 16 == 0
        self.grads = grads
        self.epoch = torch.zeros(1, dtype=torch.int32, device=device)
        self.counters = torch.zeros(4, dtype=torch.int32, device=device)
        self.world = self.sym.world
        self.rank = self.sym.rank
        self._calls = 0           # host mirror of the device epoch (selects the buffer parity)

    def allreduce_(self, scale: float = 1.0) -> None:
        """In-place sum (times ``scale``) of ``grads`` over all ranks."""
        self._calls += 1
        if self.ll:
            _launch_ll(self.sym.lib, self.sym.bases, self.slot, self.rank, self.world, [self.grads], [self.grads],
                       [self.epoch], [self.counters], None, self.n, scale)
            return
        _launch(self.sym.lib, self.sym.bases, self.slot, self.rank, self.world, [self.grads], [self.grads],
                [self.epoch], [self.counters], None, self.n, scale, self.algo)

    def allreduce_step_(self, opt: OptArgs, scale: float) -> None:
        """All-reduce fused with the optimizer update: parameters are updated in the same kernel
        that pulls the peers' gradients; ``grads`` receives the reduced gradient."""
        self._calls += 1
        if self.ll:
            _launch_ll(self.sym.lib, self.sym.bases, self.slot, self.rank, self.world, [self.grads], [self.grads],
                       [self.epoch], [self.counters], [opt], self.n, scale)
            return
        out = None if self.algo == ALGO["two_shot_fixed"] else self.grads
        _launch(self.sym.lib, self.sym.bases, self.slot, self.rank, self.world, [self.grads], [out],
                [self.epoch], [self.counters], [opt], self.n, scale, self.algo)

    # ---- registered (zero-copy) mode for large buffers, two-shot only -----------------------------
    def next_send_buffer(self) -> torch.Tensor:
        """View of the symmetric send buffer the NEXT ``allreduce_registered_`` call will reduce:
        produce the gradient straight into it and the kernel skips its copy-in pass."""
        b = (self._calls + 1) & 1
        return tensor_from_ptr(self.sym.local + b * self.slot * 4, (self.n,), torch.float32, self.grads.device)

    def allreduce_registered_(self, scale: float = 1.0) -> torch.Tensor:
        """Two-shot all-reduce of ``next_send_buffer()``; returns a view of the receive buffer holding
        the result (valid until the call after next).  No copy-in, no copy-out."""
        src = self.next_send_buffer()
        b = (self._calls + 1) & 1
        dst = tensor_from_ptr(self.sym.local + (2 + b) * self.slot * 4, (self.n,), torch.float32, self.grads.device)
        _launch(self.sym.lib, self.sym.bases, self.slot, self.rank, self.world, [src], [dst],
                [self.epoch], [self.counters], None, self.n, scale, ALGO["two_shot"])
        self._calls += 1
        return dst

    def close(self) -> None:
        self.sym.close()


class LocalEmulation:
    """``world`` ranks on one GPU: separate "symmetric regions" in local memory, one cooperative
    launch whose ``blockIdx.y`` is the rank.  Exercises exactly the device code of the multi-GPU
    path (flags, epochs, buffer parity, fused optimizer) without a second GPU."""

    def __init__(self, world: int, n: int, device: torch.device):
        self.world, self.n = world, n
        self.slot = (n + 3) // 4 * 4
        self.lib = ops.lib()
        words = 4 * self.slot + FLAG_BYTES // 4
        self.regions = [torch.zeros(words, dtype=torch.float32, device=device) for _ in range(world)]
        self.ll_regions = None
        self.epochs = [torch.zeros(1, dtype=torch.int32, device=device) for _ in range(world)]
        self.counters = [torch.zeros(4, dtype=torch.int32, device=device) for _ in range(world)]

    def allreduce(self, ins: Sequence[torch.Tensor], outs: Sequence[torch.Tensor], scale: float = 1.0,
                  algo: str = "auto", opts: Optional[Sequence[OptArgs]] = None) -> None:
        if algo == "ll":
            if self.ll_regions is None:
                words = 2 * LL_SOURCE_SLOTS * 2 * self.slot
                self.ll_regions = [torch.zeros(words, dtype=torch.float32, device=ins[0].device)
                                   for _ in range(self.world)]
            _launch_ll(self.lib, [r.data_ptr() for r in self.ll_regions], self.slot, 0, self.world, list(ins),
                       list(outs), self.epochs, self.counters, opts, self.n, scale)
            return
        _launch(self.lib, [r.data_ptr() for r in self.regions], self.slot, 0, self.world, list(ins), list(outs),
                self.epochs, self.counters, opts, self.n, scale, ALGO[algo])


make_opt_args = _opt_args


class NvlsReducer:
    """K3: all-reduce with the reduction done inside the NVSwitch (``multimem.ld_reduce`` /
    ``multimem.st``).  The buffer is allocated as *symmetric memory with a multicast mapping* through
    ``torch.distributed._symmetric_memory`` (plumbing only: cuMulticast object, handle exchange,
    mappings); the kernel is ``rdp_allreduce_nvls``.  ``buffer`` is the gradient buffer itself, so
    there is no copy-in and the result is in place on every rank."""

    kind = "nvls"

    def __init__(self, n: int, device: torch.device, group=None, wire: str = "fp32"):
        """``wire="bf16"``: the multicast buffer holds bf16 (half the NVLink bytes, fp32 accumulation in the
        switch); ``buffer`` is then a plain fp32 tensor that is rounded into it by the kernel."""
        assert n % 4 == 0 and wire in ("fp32", "bf16")
        self.n = n
        self.wire16 = wire == "bf16"
        assert not self.wire16 or n % 8 == 0, "bf16 wire: pad the buffer to a multiple of 8 values"
        self.lib = ops.lib()
        self.sym = None
        data_bytes = (n * 2 + 15) // 16 * 16 if self.wire16 else n * 4
        nbytes = data_bytes + FLAG_BYTES
        if symm_support(device.index or 0) >= 2:
            # own plumbing: cuMemCreate + fd exchange + cuMulticastCreate/AddDevice/BindMem (csrc/comm/symmem.cu)
            self.sym = SymmetricMemory(nbytes, group, multicast=True)
            self.rank, self.world = self.sym.rank, self.sym.world
            self.mc_ptr = int(self.sym.mc_ptr)
            self.buffer = tensor_from_ptr(self.sym.local, (n,), torch.float32, device) if not self.wire16 else None
            self._wire_ptr = int(self.sym.local)
            self.flag_ptrs = [int(b) + data_bytes for b in self.sym.bases]
            self.plumbing = "own (cuMulticast*)"
        else:
            import torch.distributed._symmetric_memory as symm_mem
            group = group or dist.group.WORLD
            self._t = symm_mem.empty(data_bytes // 4 + FLAG_BYTES // 4, dtype=torch.float32, device=device)
            self._t.zero_()
            self._hdl = symm_mem.rendezvous(self._t, group)
            self.rank, self.world = self._hdl.rank, self._hdl.world_size
            self.mc_ptr = int(self._hdl.multicast_ptr or 0)
            if not self.mc_ptr:
                raise RuntimeError("this system exposes no NVLS multicast mapping")
            self.buffer = self._t[:n] if not self.wire16 else None
            self._wire_ptr = int(self._t.data_ptr())
            self.flag_ptrs = [int(p) + data_bytes for p in self._hdl.buffer_ptrs]
            self.plumbing = "torch symmetric memory"
            torch.cuda.synchronize()
            dist.barrier(group)
        if self.wire16:
            self.buffer = torch.zeros(n, dtype=torch.float32, device=device)
        self.epoch = torch.zeros(1, dtype=torch.int32, device=device)
        self.counters = torch.zeros(4, dtype=torch.int32, device=device)

    def allreduce_(self, scale: float = 1.0, opt: Optional[OptArgs] = None, out: Optional[torch.Tensor] = None,
                   overlap: bool = False) -> None:
        """``overlap=True`` launches the register-capped 256-thread shape that co-resides with GEMM CTAs."""
        vp = ctypes.c_void_p
        flags = (vp * self.world)(*[vp(p) for p in self.flag_ptrs])
        if self.wire16:
            if out is None and opt is None:
                out = self.buffer                     # plain all-reduce: result back in place, as fp32
            ops._check(self.lib.rdp_allreduce_nvls_bf16(vp(self.mc_ptr), vp(self._wire_ptr), flags, self.rank,
                                                        self.world, vp(self.epoch.data_ptr()),
                                                        vp(self.counters.data_ptr()),
                                                        ctypes.byref(opt) if opt is not None else None,
                                                        vp(self.buffer.data_ptr()),
                                                        vp(out.data_ptr()) if out is not None else None, self.n,
                                                        scale, 1 if overlap else 0,
                                                        vp(torch.cuda.current_stream().cuda_stream)))
            return
        ops._check(self.lib.rdp_allreduce_nvls(vp(self.mc_ptr), vp(self.buffer.data_ptr()), flags, self.rank,
                                               self.world, vp(self.epoch.data_ptr()), vp(self.counters.data_ptr()),
                                               ctypes.byref(opt) if opt is not None else None,
                                               vp(out.data_ptr()) if out is not None else None, self.n, scale,
                                               1 if overlap else 0,
                                               vp(torch.cuda.current_stream().cuda_stream)))

    def allreduce_step_(self, opt: OptArgs, scale: float, overlap: bool = False) -> None:
        """Same call shape as :meth:`P2PReducer.allreduce_step_` (all-reduce fused with the optimizer)."""
        self.allreduce_(scale, opt=opt, overlap=overlap)

    @property
    def grads(self) -> torch.Tensor:
        return self.buffer

    def close(self) -> None:
        torch.cuda.synchronize()
        if self.sym is not None:
            self.sym.close()
            self.sym = None


def nvls_available(device: torch.device) -> bool:
    if symm_support(device.index or 0) >= 2:
        return True
    try:
        from torch._C._autograd import DeviceType
        from torch._C._distributed_c10d import _SymmetricMemory
        return bool(_SymmetricMemory.has_multicast_support(DeviceType.CUDA, device.index or 0))
    except Exception:  # noqa: BLE001
        return False


class P2PAllToAll:
    """All-to-all of fp32 chunks by direct peer stores (``rdp_alltoall``).  Every rank owns two receive
    buffers ``[world, floats_per_pair]`` in symmetric memory, used alternately, so a call costs ONE flag round;
    the returned view is valid until the call after next."""

    def __init__(self, floats_per_pair: int, device: torch.device, group=None):
        assert floats_per_pair % 4 == 0
        self.chunk = floats_per_pair
        self.lib = ops.lib()
        world = dist.get_world_size(group)
        self.stride = (world * floats_per_pair + 63) // 64 * 64            # floats between the parity buffers
        self.sym = SymmetricMemory(2 * self.stride * 4 + FLAG_BYTES, group)
        self.rank, self.world = self.sym.rank, self.sym.world
        self._recv = [tensor_from_ptr(self.sym.local + b * self.stride * 4, (world, floats_per_pair), torch.float32,
                                      device) for b in range(2)]
        self._flag_ptrs = [b + 2 * self.stride * 4 for b in self.sym.bases]
        self.epoch = torch.zeros(1, dtype=torch.int32, device=device)
        self.counters = torch.zeros(16, dtype=torch.int32, device=device)
        self._calls = 0                 # host mirror of the device epoch

    @property
    def recv(self) -> torch.Tensor:
        """Receive buffer of the most recent call."""
        return self._recv[self._calls & 1]

    def __call__(self, send: torch.Tensor) -> torch.Tensor:
        assert send.dtype == torch.float32 and send.numel() == self.world * self.chunk and send.is_contiguous()
        vp = ctypes.c_void_p
        arr = lambda xs: (vp * len(xs))(*[vp(x) for x in xs])  # noqa: E731
        ops._check(self.lib.rdp_alltoall(vp(send.data_ptr()), arr(self.sym.bases), arr(self._flag_ptrs), self.rank,
                                         self.world, self.chunk, self.stride, vp(self.epoch.data_ptr()),
                                         vp(self.counters.data_ptr()), vp(torch.cuda.current_stream().cuda_stream)))
        self._calls += 1
        return self._recv[self._calls & 1]

    def close(self) -> None:
        self.sym.close()
