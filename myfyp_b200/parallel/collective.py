"""Fused compute + collective kernels across the GPUs of one box (no NCCL on these paths).

``TwoShotFedAvg``: sample-weighted average of one slab per rank, as ONE kernel per GPU that reduce-scatters by coordinate
shard and all-gathers the reduced shards with peer loads, synchronised by flags in peer-mapped memory (kernel A2,
``csrc/collective.cu``).  Every rank owns a *symmetric region* ``[flag block | input slab | output slab]`` allocated with
``cudaMalloc`` through the native library and exported once over CUDA IPC.

Two ways to run it:

* ``TwoShotFedAvg(n, rank, world, device)`` under ``torchrun`` - one rank per GPU; handles travel through
  ``torch.distributed.all_gather_object`` at construction (plumbing), nothing but the kernel runs per call;
* ``TwoShotFedAvg.emulated(n, world, device)`` - all ranks in one launch on one GPU (every pointer local).  Kernels that wait
  on each other must be co-resident, which separate launches on one GPU do not guarantee; this is how the kernel is tested
  on a single-GPU box.

The NCCL path (``nccl_fedavg``) is the *baseline* the sweep benchmark compares against.
"""

from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence

import torch

from myfyp_b200.ops import native
from myfyp_b200.parallel.ipc import _wrap

MAX_RANKS = 16


class _Args(ctypes.Structure):
    _fields_ = [
        ("srcs", ctypes.c_void_p * MAX_RANKS), ("outs", ctypes.c_void_p * MAX_RANKS), ("flags", ctypes.c_void_p * MAX_RANKS),
        ("n", ctypes.c_int64), ("world", ctypes.c_int), ("rank", ctypes.c_int),
    ]


def _region_bytes(n: int, flag_bytes: int) -> int:
    return _pad(flag_bytes) + 2 * 4 * n


def _pad(b: int) -> int:
    return (b + 255) // 256 * 256


class TwoShotFedAvg:
    def __init__(self, n: int, rank: int, world: int, device: torch.device, _emulated: bool = False) -> None:
        if n % 4:
            raise ValueError("n must be a multiple of 4 (slabs are padded to 64 elements)")
        if world > MAX_RANKS:
            raise ValueError(f"at most {MAX_RANKS} ranks")
        self.n, self.rank, self.world, self.device = n, rank, world, device
        self.lib = native.load()
        self.lib.b2_arena_alloc.restype = ctypes.c_void_p
        self.lib.b2_ipc_open.restype = ctypes.c_void_p
        self.lib.b2_twoshot_flag_bytes.restype = ctypes.c_int64
        self.flag_bytes = int(self.lib.b2_twoshot_flag_bytes())
        self.epoch = 0
        self._emulated = _emulated
        self._bases: List[int] = []
        self._opened: List[int] = []
        nbytes = _region_bytes(n, self.flag_bytes)
        with torch.cuda.device(device):
            if _emulated:
                self._bases = [int(self.lib.b2_arena_alloc(ctypes.c_int64(nbytes)) or 0) for _ in range(world)]
            else:
                import torch.distributed as dist

                mine = int(self.lib.b2_arena_alloc(ctypes.c_int64(nbytes)) or 0)
                if not mine:
                    native.check(1, "b2_arena_alloc")
                buf = ctypes.create_string_buffer(64)
                native.check(self.lib.b2_ipc_export(ctypes.c_void_p(mine), buf), "b2_ipc_export")
                handles: List[Optional[bytes]] = [None] * world
                dist.all_gather_object(handles, bytes(buf.raw))
                for g, h in enumerate(handles):
                    if g == rank:
                        self._bases.append(mine)
                    else:
                        p = int(self.lib.b2_ipc_open(ctypes.c_char_p(h)) or 0)
                        if not p:
                            native.check(1, f"b2_ipc_open(rank {g})")
                        self._bases.append(p)
                        self._opened.append(p)
        if any(b == 0 for b in self._bases):
            native.check(1, "b2_arena_alloc")
        off_src, off_out = _pad(self.flag_bytes), _pad(self.flag_bytes) + 4 * n
        self.srcs = [_wrap(b + off_src, n, device) for b in self._bases]
        self.outs = [_wrap(b + off_out, n, device) for b in self._bases]
        self._off = (off_src, off_out)
        # the address table never changes: upload it once; weights and epoch travel as kernel arguments
        ranks = range(world) if _emulated else [rank]
        blob = b"".join(bytes(self._fill(r)) for r in ranks)
        self._args_dev = torch.frombuffer(bytearray(blob), dtype=torch.uint8).to(device)
        self._nranks = len(list(ranks))
        torch.cuda.synchronize(device)

    @classmethod
    def emulated(cls, n: int, world: int, device: torch.device) -> "TwoShotFedAvg":
        return cls(n, 0, world, device, _emulated=True)

    @property
    def src(self) -> torch.Tensor:
        """This rank's input slab (write the model here, or train directly in it)."""
        return self.srcs[self.rank]

    @property
    def out(self) -> torch.Tensor:
        return self.outs[self.rank]

    def _fill(self, rank: int) -> _Args:
        a = _Args()
        for g in range(self.world):
            a.srcs[g] = self._bases[g] + self._off[0]
            a.outs[g] = self._bases[g] + self._off[1]
            a.flags[g] = self._bases[g]
        a.n, a.world, a.rank = self.n, self.world, rank
        return a

    def __call__(self, weights: Sequence[float], max_ctas: int = 0) -> torch.Tensor:
        """Launch the fused kernel on the current stream; returns this rank's output slab (all ranks' in emulated mode: ``outs``)."""
        self.epoch += 1
        with native.device_guard(self.device):
            native.check(
                self.lib.b2_fedavg_twoshot(
                    native.ptr(self._args_dev), self._nranks, ctypes.c_int64(self.n), native.float_array(weights), self.world, ctypes.c_uint(self.epoch), max_ctas,
                    native.stream_ptr(self.device),
                ),
                "b2_fedavg_twoshot",
            )
        native.count_launch()
        return self.out

    def close(self) -> None:
        for p in self._opened:
            self.lib.b2_ipc_close(ctypes.c_void_p(p))
        self._opened.clear()


def nccl_fedavg(x: torch.Tensor, n_k: float, total: float) -> torch.Tensor:
    """BASELINE: pre-weight, NCCL all-reduce, done (what a torch.distributed FedAvg would do)."""
    import torch.distributed as dist

    y = x * (n_k / total)
    dist.all_reduce(y)
    return y


# ====================================================================================================== NVLS (in-switch reduction)
def nvls_available() -> bool:
    """True when the current GPU reports NVSwitch multicast support (``tools/nvls_probe.py`` goes further and creates an object).
    Touches no other GPU: a rank must not grow contexts on its peers' devices just to ask."""
    if not torch.cuda.is_available() or not native.available():
        return False
    return native.load().b2_multicast_supported() == 1


class NvlsFedAvg:
    """Sum (then scale) of one slab per rank through NVSwitch multicast: ``multimem.ld_reduce`` lets the switch add the ranks'
    values of an address in one load, ``multimem.st`` writes the result into every rank's output (kernel in ``csrc/nvls.cu``).
    Callers put their PRE-WEIGHTED model (``n_g / N * theta_g``, or a rank-local weighted partial average) into ``src``.

    * ``NvlsFedAvg(n, rank, world, device)`` under ``torchrun``: rank 0 creates the multicast object and hands its file
      descriptor to the other ranks over a unix socket (``SCM_RIGHTS``); every rank adds its device, then binds its buffer;
    * ``NvlsFedAvg.in_process(n, devices)``: one communicator per GPU of a single process (tests, the one-process deployment).

    No reference counterpart (p2pfl averages NumPy arrays on the host, fedavg.py:61-69); NCCL's own NVLS algorithm is the competitor.
    """

    def __init__(self, n: int, rank: int, world: int, device: torch.device, _fd: Optional[int] = None, _defer_bind: bool = False) -> None:
        if n % 4:
            raise ValueError("n must be a multiple of 4")
        self.n, self.rank, self.world, self.device = n, rank, world, device
        lib = self.lib = native.load()
        lib.b2_nvls_granularity.restype = ctypes.c_int64
        lib.b2_nvls_join.restype = ctypes.c_void_p
        lib.b2_nvls_join.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int64]
        lib.b2_nvls_local_ptr.restype = ctypes.c_void_p
        lib.b2_nvls_local_ptr.argtypes = [ctypes.c_void_p]
        lib.b2_nvls_flag_bytes.restype = ctypes.c_int64
        for f in (lib.b2_nvls_bind, lib.b2_nvls_destroy):
            f.argtypes = [ctypes.c_void_p]
        lib.b2_nvls_allreduce.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_uint, ctypes.c_float, ctypes.c_int, ctypes.c_void_p]
        lib.b2_nvls_last_timeout.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        gran = int(lib.b2_nvls_granularity(world))
        if gran <= 0:
            native.check(1, "b2_nvls_granularity")
        flag = int(lib.b2_nvls_flag_bytes())
        self.off_in, self.off_out = flag, flag + 4 * n
        self.bytes = (flag + 8 * n + gran - 1) // gran * gran
        self.epoch = 0
        fd = _fd
        if fd is None:
            fd = self._exchange_fd()
        with torch.cuda.device(device):
            self._h = lib.b2_nvls_join(fd, world, rank, ctypes.c_int64(self.bytes))
        if not self._h:
            native.check(1, "b2_nvls_join")
        if _fd is None:
            import os as _os

            _os.close(fd)
        if not _defer_bind:
            if _fd is None:
                import torch.distributed as dist

                dist.barrier()  # every device is in the team before anyone binds
            self._bind()
            if _fd is None:
                dist.barrier()  # ... and everybody has bound before anyone touches the multicast address (binding can take seconds)

    def _bind(self) -> None:
        with torch.cuda.device(self.device):
            native.check(self.lib.b2_nvls_bind(self._h), "b2_nvls_bind")
        base = int(self.lib.b2_nvls_local_ptr(self._h))
        self.src = _wrap(base + self.off_in, self.n, self.device)
        self.out = _wrap(base + self.off_out, self.n, self.device)

    def _exchange_fd(self) -> int:
        """Rank 0 creates the multicast object; its descriptor travels to the other ranks as ancillary data of a unix socket."""
        import os
        import socket

        import torch.distributed as dist

        path = f"/tmp/b2-nvls-{os.environ.get('MASTER_PORT', '0')}-{os.getpid() if self.world == 1 else 'job'}.sock"
        if self.rank == 0:
            fd = ctypes.c_int(-1)
            with torch.cuda.device(self.device):
                native.check(self.lib.b2_nvls_root_create(self.world, ctypes.c_int64(self.bytes), ctypes.byref(fd)), "b2_nvls_root_create")
            try:
                os.unlink(path)
            except OSError:
                pass
            srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            srv.bind(path)
            srv.listen(self.world)
            dist.barrier()
            for _ in range(self.world - 1):
                conn, _ = srv.accept()
                socket.send_fds(conn, [b"mc"], [fd.value])
                conn.close()
            srv.close()
            os.unlink(path)
            return fd.value
        dist.barrier()
        cli = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        cli.connect(path)
        _, fds, _, _ = socket.recv_fds(cli, 16, 1)
        cli.close()
        return fds[0]

    @classmethod
    def in_process(cls, n: int, devices: Sequence[torch.device]) -> List["NvlsFedAvg"]:
        import os

        lib = native.load()
        lib.b2_nvls_granularity.restype = ctypes.c_int64
        lib.b2_nvls_flag_bytes.restype = ctypes.c_int64
        world = len(devices)
        gran = int(lib.b2_nvls_granularity(world))
        if gran <= 0:
            native.check(1, "b2_nvls_granularity")
        nbytes = (int(lib.b2_nvls_flag_bytes()) + 8 * n + gran - 1) // gran * gran
        fd = ctypes.c_int(-1)
        with torch.cuda.device(devices[0]):
            native.check(lib.b2_nvls_root_create(world, ctypes.c_int64(nbytes), ctypes.byref(fd)), "b2_nvls_root_create")
        comms = [cls(n, r, world, d, _fd=fd.value, _defer_bind=True) for r, d in enumerate(devices)]
        os.close(fd.value)
        for c in comms:
            c._bind()
        return comms

    def __call__(self, scale: float = 1.0, max_ctas: int = 0, epoch: Optional[int] = None) -> torch.Tensor:
        """Launch on the current stream of this rank's device.  Every rank calls with the same, growing ``epoch`` (default: a
        call counter); an epoch that some rank never joins times out on the others (:meth:`timed_out`) and harms no later one."""
        self.epoch = self.epoch + 1 if epoch is None else int(epoch)
        with native.device_guard(self.device):
            native.check(
                self.lib.b2_nvls_allreduce(self._h, self.off_in, self.off_out, self.n, self.epoch, ctypes.c_float(scale), max_ctas, native.stream_ptr(self.device)),
                "b2_nvls_allreduce",
            )
        native.count_launch()
        return self.out

    def timed_out(self) -> bool:
        """After the stream has been synchronised: did the last call give up waiting for a rank?"""
        e = ctypes.c_uint(0)
        with native.device_guard(self.device):
            native.check(self.lib.b2_nvls_last_timeout(self._h, ctypes.byref(e)), "b2_nvls_last_timeout")
        return e.value == self.epoch and self.epoch > 0

    def close(self) -> None:
        if getattr(self, "_h", None):
            self.src = self.out = None
            self.lib.b2_nvls_destroy(self._h)
            self._h = None
