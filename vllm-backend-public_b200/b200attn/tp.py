"""Tensor-parallel runtime: one process per GPU, `--tensor-parallel-size N` of the launch contract
(VLLMConfig.tensor_parallel_size, reference benchmark-vllm/models.py:11 -> vllm_templates.py:55,226).

Three pieces, all single-node (NVLink / NVSwitch box):

* `TPGroup`      rank / size over an initialised torch.distributed world, the peer-mapped comm regions of
                 csrc/comm.cu (CUDA IPC handles exchanged once at start-up) and the two fused exchange ops the
                 model calls: `allreduce_rms_norm` (one-shot NVLink all-reduce + residual add + RMSNorm) and
                 `allgather_cols` (vocab-parallel logits).  NCCL is used only for start-up handshakes and for
                 the bandwidth-bound all-reduce of large prompt chunks.
* `StepChannel`  how rank 0 (the only rank with an HTTP front end and a request queue) tells the other ranks what
                 was admitted or aborted before each step: a single-slot mailbox in /dev/shm with a sequence
                 counter, polled by the workers (a few microseconds per step; no socket, no collective).  Every
                 rank runs the same deterministic C++ scheduler on the same inputs and samples from the same
                 gathered logits, so plans, CUDA-graph choices and token ids agree without being sent.
* `launch_workers` / `worker_main`   the server starts ranks 1..N-1 as child processes of rank 0.
"""
from __future__ import annotations

import ctypes as C
import json
import os
import pickle
import subprocess
import sys
import time
import uuid
from typing import List, Optional

import numpy as np
import torch

from . import _lib


class _RawCuda:
    """Zero-copy view of raw device memory for torch.as_tensor (__cuda_array_interface__ v3)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
                                         "strides": None}


def _device_view(ptr: int, nbytes: int, dev: torch.device) -> torch.Tensor:
    return torch.as_tensor(_RawCuda(ptr, nbytes), device=dev)


# ------------------------------------------------------------------------------------------------------ channel
class StepChannel:
    """Single-producer (rank 0) mailbox in shared memory, read by ranks 1..world-1.

    Layout (uint64 words): [0] messages published, [1] bytes of the current part, [2] 1 = more parts follow,
    [8 * r] parts consumed by rank r (one cache line per reader); payload from byte 4096.  The producer writes the
    payload, then bumps word 0; it reuses the slot only after every reader has acknowledged the previous part, which
    they do at the START of the step it describes -- so the producer never waits in steady state.  x86 keeps the
    store order (payload before counter); numpy does plain stores."""

    HEADER = 4096

    def __init__(self, path: str, world: int, rank: int, create: bool, capacity: int = 4 << 20):
        self.path, self.world, self.rank, self.capacity = path, world, rank, capacity
        size = self.HEADER + capacity
        if create:
            with open(path, "wb") as f:
                f.truncate(size)
        self._mm = np.memmap(path, dtype=np.uint8, mode="r+", shape=(size,))
        self._w = self._mm[: self.HEADER].view(np.uint64)
        self._payload = self._mm[self.HEADER:]
        self._seen = 0          # reader: parts consumed; producer: parts published

    def unlink(self) -> None:
        try:
            os.unlink(self.path)
        except OSError:
            pass

    @staticmethod
    def _wait(cond, timeout: float, what: str) -> None:
        t0 = time.perf_counter()
        while not cond():
            dt = time.perf_counter() - t0
            if dt > timeout:
                raise TimeoutError(f"tensor-parallel step channel: {what} (waited {timeout:.0f} s)")
            if dt > 2e-3:
                time.sleep(2e-4)

    def send(self, payload: bytes, timeout: float = 120.0) -> None:
        assert self.rank == 0
        view = memoryview(payload)
        off, n = 0, len(view)
        while True:
            part = min(self.capacity, n - off)
            seq = self._seen
            self._wait(lambda: all(int(self._w[8 * r]) == seq for r in range(1, self.world)), timeout,
                       "a worker rank stopped acknowledging steps")
            if part:
                self._payload[:part] = np.frombuffer(view[off:off + part], dtype=np.uint8)
            off += part
            self._w[1] = part
            self._w[2] = 1 if off < n else 0
            self._seen = seq + 1
            self._w[0] = self._seen
            if off >= n:
                return

    def recv(self, timeout: float = 1e9) -> bytes:
        assert self.rank > 0
        parts = []
        while True:
            want = self._seen + 1
            self._wait(lambda: int(self._w[0]) >= want, timeout, "rank 0 stopped publishing steps")
            n = int(self._w[1])
            more = int(self._w[2])
            parts.append(self._payload[:n].tobytes())
            self._seen = want
            self._w[8 * self.rank] = want
            if not more:
                return parts[0] if len(parts) == 1 else b"".join(parts)


# ------------------------------------------------------------------------------------------------------ group
class TPGroup:
    """This rank's handle on the tensor-parallel world.  torch.distributed must already be initialised with one
    rank per GPU on this node (torchrun, or `launch_workers`); `device` is this rank's GPU."""

    def __init__(self, device: torch.device, group=None):
        import torch.distributed as dist
        if not dist.is_initialized():
            raise RuntimeError("TPGroup needs an initialised torch.distributed world (one rank per GPU)")
        self.dist, self.group = dist, group
        self.rank, self.size = dist.get_rank(group), dist.get_world_size(group)
        if self.size not in (1, 2, 4, 8):
            raise ValueError(f"tensor_parallel_size {self.size}: the NVLink exchange kernels are built for 1, 2, 4, 8")
        self.dev = device
        self._lib = _lib.load()
        self._comm = None
        self._own_ptr = None
        self._peer_ptrs: List[int] = []
        self.max_rows = 0
        self.inputs: List[torch.Tensor] = []
        self.gathered: Optional[torch.Tensor] = None
        # step channel: rank 0 creates the mailbox, everyone maps it, then the name is removed (the mapping stays)
        path = [f"/dev/shm/b200attn-{os.getpid()}-{uuid.uuid4().hex[:8]}" if self.rank == 0 else None]
        dist.broadcast_object_list(path, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        if self.rank == 0:
            self.channel = StepChannel(path[0], self.size, 0, create=True)
        dist.barrier(group=group)
        if self.rank != 0:
            self.channel = StepChannel(path[0], self.size, self.rank, create=False)
        dist.barrier(group=group)
        if self.rank == 0:
            self.channel.unlink()

    # ---- collectives on host values (start-up only)
    def min_int(self, v: int) -> int:
        t = torch.tensor([int(v)], dtype=torch.int64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MIN, group=self.group)
        return int(t.item())

    def barrier(self) -> None:
        self.dist.barrier(group=self.group)
        torch.cuda.synchronize(self.dev)

    # ---- peer-mapped comm regions
    def setup_comm(self, max_rows: int, hidden: int, gather_rows: int, gather_cols: int) -> None:
        """Allocate this rank's region, exchange CUDA IPC handles, map the peers' regions, build the b2_comm handle.
        max_rows x hidden: largest partial sum on the one-shot path; gather_rows x gather_cols: the logits buffer."""
        lib = self._lib
        torch.cuda.set_device(self.dev)
        if self._comm:                               # a second engine on the same group: release the old regions first
            self.close()
        nbytes = int(lib.b2_comm_region_bytes(self.size, max_rows, hidden, gather_rows, gather_cols))
        if nbytes == 0:
            raise ValueError("comm region: bad sizes")
        ptr = C.c_void_p()
        handle = C.create_string_buffer(64)
        _lib.check(lib.b2_comm_alloc(nbytes, C.byref(ptr), handle))
        self._own_ptr = ptr.value
        handles: List[Optional[bytes]] = [None] * self.size
        self.dist.all_gather_object(handles, (os.getpid(), handle.raw), group=self.group)
        ptrs = []
        for r, (pid, h) in enumerate(handles):
            if r == self.rank:
                ptrs.append(self._own_ptr)
            elif pid == os.getpid():
                raise RuntimeError("tensor-parallel ranks must be separate processes (CUDA IPC cannot map a handle "
                                   "into the process that exported it)")
            else:
                p = C.c_void_p()
                _lib.check(lib.b2_comm_open(C.create_string_buffer(h, 64), C.byref(p)))
                ptrs.append(p.value)
                self._peer_ptrs.append(p.value)
        arr = (C.c_void_p * self.size)(*ptrs)
        comm = lib.b2_comm_create(self.rank, self.size, arr, nbytes, max_rows, hidden, gather_rows, gather_cols)
        if not comm:
            raise _lib.B2Error(_lib.B2_EINVAL, lib.b2_last_error().decode("utf-8", "replace"))
        self._comm = comm
        self._bind_views(max_rows, hidden, gather_rows, gather_cols)
        self.barrier()

    def _bind_views(self, max_rows: int, hidden: int, gather_rows: int, gather_cols: int) -> None:
        lib = self._lib
        self.max_rows, self.hidden, self.gather_rows, self.gather_cols = max_rows, hidden, gather_rows, gather_cols
        self.inputs = []
        for slot in (0, 1):
            p = lib.b2_comm_input(self._comm, slot)
            self.inputs.append(_device_view(p, max_rows * hidden * 2, self.dev).view(torch.bfloat16).view(max_rows, hidden))
        self.gathered = None
        if gather_rows and gather_cols:
            p = lib.b2_comm_gathered(self._comm)
            self.gathered = _device_view(p, gather_rows * gather_cols * 2, self.dev).view(torch.bfloat16) \
                .view(gather_rows, gather_cols)

    # ---- the two exchange ops of a decoder layer
    def allreduce_rms_norm(self, slot: int, T: int, weight: Optional[torch.Tensor], residual: Optional[torch.Tensor],
                           eps: float, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """out = RMSNorm(residual += sum over ranks of inputs[slot][:T]) * weight (b2_comm_allreduce_rms_norm).
        The caller's GEMM must have written inputs[slot][:T] on the current stream."""
        if out is None:
            out = torch.empty(T, self.hidden, dtype=torch.bfloat16, device=self.dev)
        if out.dtype != torch.bfloat16 or out.shape != (T, self.hidden) or out.stride(1) != 1:
            raise ValueError(f"out must be bf16 [{T}, {self.hidden}]")
        rs = 0
        if residual is not None:
            if residual.dtype != torch.bfloat16 or residual.shape != (T, self.hidden) or residual.stride(1) != 1:
                raise ValueError(f"residual must be bf16 [{T}, {self.hidden}]")
            rs = residual.stride(0) if T > 1 else self.hidden
        if weight is not None and (weight.dtype != torch.bfloat16 or weight.shape != (self.hidden,)
                                   or not weight.is_contiguous()):
            raise ValueError(f"weight must be a contiguous bf16 [{self.hidden}]")
        _lib.check(self._lib.b2_comm_allreduce_rms_norm(
            self._comm, slot, out.data_ptr(), None if residual is None else residual.data_ptr(),
            None if weight is None else weight.data_ptr(), T, out.stride(0) if T > 1 else self.hidden, rs, float(eps),
            torch.cuda.current_stream(self.dev).cuda_stream))
        return out

    PUSH_MAX_ROWS = 256          # csrc/comm.cu: every CTA of a push exchange must be resident

    def push_allreduce_rms_norm(self, part: torch.Tensor, weight: Optional[torch.Tensor],
                                residual: Optional[torch.Tensor], eps: float,
                                out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Same result as allreduce_rms_norm, by the push protocol (b2_comm_push_allreduce_rms_norm): `part` is this
        rank's partial sum, bf16 [T, hidden] in ordinary device memory, T <= PUSH_MAX_ROWS."""
        T = part.shape[0]
        if part.dtype != torch.bfloat16 or part.shape != (T, self.hidden) or part.stride(1) != 1:
            raise ValueError(f"part must be bf16 [T, {self.hidden}]")
        if out is None:
            out = torch.empty(T, self.hidden, dtype=torch.bfloat16, device=self.dev)
        if out.dtype != torch.bfloat16 or out.shape != (T, self.hidden) or out.stride(1) != 1:
            raise ValueError(f"out must be bf16 [{T}, {self.hidden}]")
        rs = 0
        if residual is not None:
            if residual.dtype != torch.bfloat16 or residual.shape != (T, self.hidden) or residual.stride(1) != 1:
                raise ValueError(f"residual must be bf16 [{T}, {self.hidden}]")
            rs = residual.stride(0) if T > 1 else self.hidden
        if weight is not None and (weight.dtype != torch.bfloat16 or weight.shape != (self.hidden,)
                                   or not weight.is_contiguous()):
            raise ValueError(f"weight must be a contiguous bf16 [{self.hidden}]")
        _lib.check(self._lib.b2_comm_push_allreduce_rms_norm(
            self._comm, part.data_ptr(), part.stride(0) if T > 1 else self.hidden, out.data_ptr(),
            None if residual is None else residual.data_ptr(), None if weight is None else weight.data_ptr(), T,
            out.stride(0) if T > 1 else self.hidden, rs, float(eps), torch.cuda.current_stream(self.dev).cuda_stream))
        return out

    def allgather_cols(self, local: torch.Tensor) -> torch.Tensor:
        """local bf16 [T, cols] (this rank's column block) -> view [T, size * cols] of the gathered buffer."""
        T, cols = local.shape
        if local.dtype != torch.bfloat16 or local.stride(1) != 1 or cols * self.size != self.gather_cols:
            raise ValueError(f"local must be bf16 [T, {self.gather_cols // self.size}]")
        _lib.check(self._lib.b2_comm_allgather_cols(self._comm, local.data_ptr(), local.stride(0) if T > 1 else cols, T,
                                                    cols, torch.cuda.current_stream(self.dev).cuda_stream))
        return self.gathered[:T]

    def all_reduce_nccl(self, t: torch.Tensor) -> torch.Tensor:
        """Bandwidth-bound partial sums (prompt chunks of thousands of rows): NCCL over NVLink / NVLS."""
        self.dist.all_reduce(t, group=self.group)
        return t

    def close(self) -> None:
        lib = self._lib
        if self._comm:
            torch.cuda.synchronize(self.dev)
            self.inputs, self.gathered = [], None
            lib.b2_comm_destroy(self._comm)
            self._comm = None
            for p in self._peer_ptrs:
                lib.b2_comm_close(p)
            self._peer_ptrs = []
            try:
                self.barrier()                       # nobody frees a region a peer still maps and may touch
            except Exception:
                pass
            if self._own_ptr:
                lib.b2_comm_free(self._own_ptr)
                self._own_ptr = None


# ------------------------------------------------------------------------------------------------------ messages
MSG_STEP, MSG_SHUTDOWN, MSG_CAPTURE = 0, 1, 2


def encode_step(incoming, aborts, cmd: int = MSG_STEP) -> bytes:
    """incoming: engine Request objects just admitted on rank 0.  An ordinary step with nothing new is b''."""
    if cmd == MSG_STEP and not incoming and not aborts:
        return b""
    reqs = [(r.id, r.prompt_ids, r.params, r.seed) for r in incoming]
    return pickle.dumps((cmd, reqs, list(aborts)), protocol=pickle.HIGHEST_PROTOCOL)


def decode_step(payload: bytes):
    if not payload:
        return MSG_STEP, [], []
    return pickle.loads(payload)


# ------------------------------------------------------------------------------------------------------ processes
def init_distributed(rank: int, world: int, device: torch.device, master_addr: str = "127.0.0.1",
                     master_port: Optional[int] = None) -> None:
    import torch.distributed as dist
    if dist.is_initialized():
        return
    os.environ.setdefault("MASTER_ADDR", master_addr)
    if master_port is not None:
        os.environ["MASTER_PORT"] = str(master_port)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)


def free_port() -> int:
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _exit_when_orphaned(launcher_pid: int) -> None:
    """Watcher thread of a worker rank: leave at once when the launcher (rank 0) is gone, whatever the main thread is
    doing.  A rank without rank 0 can only wait inside a collective with its GPU spinning; the driver releases the
    device with the process.  (Polling, not PR_SET_PDEATHSIG: that signal follows the launching THREAD, and rank 0
    starts its workers from the engine thread, which ends before the orderly shutdown does.)"""
    while True:
        time.sleep(1.0)
        if os.getppid() != launcher_pid:
            os._exit(1)


def launch_workers(world: int, spec: dict, master_port: int, log_dir: Optional[str] = None) -> List[subprocess.Popen]:
    """Start ranks 1..world-1 (`python -m b200attn.tp`), each bound to GPU `rank` of this node, with the engine spec
    rank 0 was started with.  The caller is rank 0 and must then call init_distributed(0, world, ...)."""
    procs = []
    launcher = os.getpid()
    pkg_parent = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for r in range(1, world):
        env = dict(os.environ)
        env.update({"RANK": str(r), "LOCAL_RANK": str(r), "WORLD_SIZE": str(world), "MASTER_ADDR": "127.0.0.1",
                    "MASTER_PORT": str(master_port), "B200ATTN_TP_SPEC": json.dumps(spec), "OMP_NUM_THREADS": "1",
                    "B200ATTN_TP_LAUNCHER_PID": str(launcher),
                    "PYTHONPATH": pkg_parent + os.pathsep + env.get("PYTHONPATH", "")})
        out = None
        if log_dir:
            out = open(os.path.join(log_dir, f"tp_worker_{r}.log"), "w")
        procs.append(subprocess.Popen([sys.executable, "-m", "b200attn.tp"], env=env, stdout=out,
                                      stderr=subprocess.STDOUT if out else None))
    return procs


def build_engine_from_spec(spec: dict, device: torch.device, group: "TPGroup"):
    from .engine import LLMEngine
    from .model import resolve_config
    from pathlib import Path
    cfg = resolve_config(spec["model"])
    wdir = Path(spec["model"]) if os.path.isdir(spec["model"]) else None
    return LLMEngine(cfg, device, tp=group, weights_dir=wdir, **spec.get("engine", {}))


def worker_main() -> int:
    """Entry point of ranks 1..N-1: build the same engine as rank 0 and mirror its steps until told to stop."""
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    spec = json.loads(os.environ["B200ATTN_TP_SPEC"])
    launcher = int(os.environ.get("B200ATTN_TP_LAUNCHER_PID", os.getppid()))
    if os.getppid() != launcher:
        return 1
    import threading
    threading.Thread(target=_exit_when_orphaned, args=(launcher,), name="b200attn-orphan-watch", daemon=True).start()
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    init_distributed(rank, world, dev)
    group = TPGroup(dev)
    eng = build_engine_from_spec(spec, dev, group)
    orderly = False
    try:
        orderly = eng.worker_loop()
    finally:
        if not orderly:
            # rank 0 is gone (or this rank failed): eng.close() holds a barrier nobody will match, and the process
            # group's destructor would wait on it too -- leave at once, the driver releases the GPU with the process
            sys.stdout.flush()
            sys.stderr.flush()
            os._exit(1)
    eng.close()
    import torch.distributed as dist
    if dist.is_initialized():
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(worker_main())
