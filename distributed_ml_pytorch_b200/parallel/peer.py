"""NVLink peer-memory group: symmetric buffers mapped into every rank with CUDA IPC, and the fused
push / pull / synchronous-update kernels that run on them.

This is the B200-native replacement for the reference's data plane (gloo ``isend``/``recv`` of CPU tensors to
a single server rank — example/main.py:151,165; asgd/optim/Asynchronous.py:34,49,59).  ``torch.distributed``
remains the *control* plane only: it carries the 64-byte IPC handles at set-up and the final barrier.

One process per GPU.  ``PeerGroup.symmetric(nbytes)`` allocates the same buffer on every rank and returns a
:class:`SymBuffer` holding the local tensor plus every peer's mapped pointer.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch
import torch.distributed as dist

from .. import _native

c_vp, c_int, c_ll, c_float, c_sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_size_t
_PP = ctypes.POINTER(ctypes.c_void_p)

_native.register("dmlb_peer_alloc", [c_sz, _PP])
_native.register("dmlb_peer_free", [c_vp])
_native.register("dmlb_ipc_get_handle", [c_vp, c_vp])
_native.register("dmlb_ipc_open_handle", [c_vp, _PP])
_native.register("dmlb_ipc_close", [c_vp])
_native.register("dmlb_enable_peer_access", [c_int])
_native.register("dmlb_comm_chunk_elems", [])
_native.register("dmlb_comm_flag_words", [])
_native.register("dmlb_push_add", [c_vp, _PP, c_int, c_ll, c_ll, c_vp])
_native.register("dmlb_pull", [c_vp, c_vp, _PP, _PP, c_vp, c_int, c_ll, c_ll, c_vp])
_native.register("dmlb_push_delta", [c_vp, c_vp, _PP, c_int, c_ll, c_ll, c_vp])
_native.register("dmlb_push_mailbox_delta", [c_vp, c_vp, c_vp, c_vp, _PP, _PP, c_int, c_int, c_ll, c_vp])
_native.register("dmlb_push_mailbox", [c_vp, c_vp, c_vp, _PP, _PP, c_int, c_int, c_ll, c_vp])
_native.register("dmlb_owner_apply", [c_vp, c_vp, c_vp, c_vp, c_vp, _PP, c_vp, c_int, c_int, c_ll, c_float, c_int, c_float, c_float, c_vp])
_native.register("dmlb_sync_rs_update_ag", [_PP, _PP, _PP, c_vp, c_int, c_int, c_ll, c_vp, c_float, c_float, c_float, c_vp])
_native.register("dmlb_sync_nvls_range", [c_vp, c_vp, c_vp, _PP, c_vp, c_int, c_int, c_ll, c_ll, c_vp, c_float, c_float, c_float, c_int, c_vp])
_native.register("dmlb_sync_rs_update_ag_range", [_PP, _PP, _PP, c_vp, c_int, c_int, c_ll, c_ll, c_vp, c_float, c_float, c_float, c_int, c_vp])


class _RawCudaBuffer:
    """Wraps a raw device pointer so ``torch.as_tensor`` can view it without copying."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None}


def _tensor_from_ptr(ptr: int, nbytes: int, device: torch.device) -> torch.Tensor:
    return torch.as_tensor(_RawCudaBuffer(ptr, nbytes), device=device)


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


class SymBuffer:
    """One symmetric allocation: ``local`` (uint8 tensor over this rank's memory) and ``ptrs[r]`` = address of
    rank r's copy as seen from this process (``ptrs[rank]`` is the local one)."""

    def __init__(self, local: torch.Tensor, ptrs: List[int], rank: int, owned_ptr: int, opened: List[int]):
        self.local = local
        self.ptrs = ptrs
        self.rank = rank
        self._owned_ptr = owned_ptr
        self._opened = opened
        self.table = (ctypes.c_void_p * len(ptrs))(*ptrs)

    def view(self, dtype: torch.dtype, numel: Optional[int] = None, offset_bytes: int = 0) -> torch.Tensor:
        t = self.local[offset_bytes:].view(dtype)
        return t if numel is None else t[:numel]

    def offset_table(self, offset_bytes: int):
        return (ctypes.c_void_p * len(self.ptrs))(*[p + offset_bytes for p in self.ptrs])


class PeerGroup:
    """All ranks of ``group`` (default: the world group), each driving one GPU of the same NVSwitch domain."""

    def __init__(self, group=None, device: Optional[torch.device] = None):
        self.group = group
        if dist.is_available() and dist.is_initialized():
            self.rank = dist.get_rank(group)
            self.world = dist.get_world_size(group)
        else:
            self.rank, self.world = 0, 1
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        if self.world > 16:
            raise ValueError("PeerGroup supports up to 16 GPUs (one NVSwitch domain)")
        self._buffers: List[SymBuffer] = []
        _native.lib()

    # ------------------------------------------------------------------------------------------
    def symmetric(self, nbytes: int) -> SymBuffer:
        nbytes = (int(nbytes) + 255) // 256 * 256
        with torch.cuda.device(self.device):
            out = ctypes.c_void_p()
            _native.call("dmlb_peer_alloc", nbytes, ctypes.byref(out))
            local_ptr = int(out.value)
            ptrs = [0] * self.world
            opened: List[int] = []
            if self.world > 1:
                handle = ctypes.create_string_buffer(64)
                _native.call("dmlb_ipc_get_handle", local_ptr, handle)
                gathered: List[Optional[bytes]] = [None] * self.world
                dist.all_gather_object(gathered, bytes(handle.raw), group=self.group)
                for r, h in enumerate(gathered):
                    if r == self.rank:
                        ptrs[r] = local_ptr
                        continue
                    p = ctypes.c_void_p()
                    _native.call("dmlb_ipc_open_handle", ctypes.create_string_buffer(h, 64), ctypes.byref(p))
                    ptrs[r] = int(p.value)
                    opened.append(int(p.value))
            else:
                ptrs[0] = local_ptr
            local = _tensor_from_ptr(local_ptr, nbytes, self.device)
        buf = SymBuffer(local, ptrs, self.rank, local_ptr, opened)
        self._buffers.append(buf)
        return buf

    def barrier(self) -> None:
        torch.cuda.synchronize(self.device)
        if self.world > 1:
            dist.barrier(group=self.group)

    def close(self) -> None:
        """Unmap peers and free local memory (call after a barrier; tensors viewing the buffers become invalid)."""
        torch.cuda.synchronize(self.device)
        for b in self._buffers:
            for p in b._opened:
                _native.call("dmlb_ipc_close", p)
            b._opened = []
        if self.world > 1:
            dist.barrier(group=self.group)
        for b in self._buffers:
            if b._owned_ptr:
                _native.call("dmlb_peer_free", b._owned_ptr)
                b._owned_ptr = 0
        self._buffers = []


# ------------------------------------------------------------------------------------------------
# kernel wrappers
# ------------------------------------------------------------------------------------------------
def chunk_elems() -> int:
    return int(_native.lib().dmlb_comm_chunk_elems())


def flag_words() -> int:
    return int(_native.lib().dmlb_comm_flag_words())


_native.register("dmlb_set_comm_max_blocks", [c_int], restype=None)


def set_comm_max_blocks(n: int) -> None:
    """Grid cap of the push / pull kernels launched next (0 = default, two CTAs per SM)."""
    _native.lib().dmlb_set_comm_max_blocks(int(n))


def set_comm_mover(mover: int) -> None:
    """Mover of the push / pull kernels launched next: 0 = chosen by size, 1 = per-thread ``ld`` / ``red.add``, 2 = TMA bulk copies
    (``cp.async.bulk`` through shared memory, ``cp.reduce.async.bulk.add.f32`` into the owner)."""
    _native.lib().dmlb_set_comm_mover(int(mover))


def push_add(acc: torch.Tensor, master_table, world: int, shard_size: int, n: int) -> None:
    _native.call("dmlb_push_add", acc.data_ptr(), master_table, world, shard_size, n, _stream())


def pull(params: torch.Tensor, master_table, world: int, shard_size: int, n: int, version_table=None, pulled_ver: Optional[torch.Tensor] = None,
         shadow: Optional[torch.Tensor] = None) -> None:
    _native.call("dmlb_pull", params.data_ptr(), shadow.data_ptr() if shadow is not None else None, master_table, version_table, pulled_ver.data_ptr() if pulled_ver is not None else None, world,
                 shard_size, n, _stream())


def push_delta(live: torch.Tensor, snapshot: torch.Tensor, master_table, world: int, shard_size: int, n: int) -> None:
    """``master[owner] += live - snapshot; snapshot = live`` (``dmlb_push_delta``): the accumulator-free push."""
    _native.call("dmlb_push_delta", live.data_ptr(), snapshot.data_ptr(), master_table, world, shard_size, n, _stream())


def push_mailbox_delta(live, snapshot, busy, pulled_ver, mailbox_table, full_table, rank, world, shard_size) -> None:
    _native.call("dmlb_push_mailbox_delta", live.data_ptr(), snapshot.data_ptr(), busy.data_ptr(), pulled_ver.data_ptr(), mailbox_table,
                 full_table, rank, world, shard_size, _stream())


def push_mailbox(acc, busy, pulled_ver, mailbox_table, full_table, rank, world, shard_size) -> None:
    _native.call("dmlb_push_mailbox", acc.data_ptr(), busy.data_ptr(), pulled_ver.data_ptr(), mailbox_table, full_table, rank, world, shard_size,
                 _stream())


def owner_apply(master, mom, version, full, mailbox, busy_table, stats, rank, world, shard_size, mu, stale_mode, stale_param, gain=1.0) -> None:
    _native.call("dmlb_owner_apply", master.data_ptr(), mom.data_ptr() if mom is not None else None, version.data_ptr(), full.data_ptr(),
                 mailbox.data_ptr(), busy_table, stats.data_ptr() if stats is not None else None, rank, world, shard_size, float(mu),
                 int(stale_mode), float(stale_param), float(gain), _stream())


def sync_rs_update_ag(grad_table, param_table, flag_table, mom, rank, world, shard_size, lr_dev, mu, wd, grad_scale) -> None:
    _native.call("dmlb_sync_rs_update_ag", grad_table, param_table, flag_table, mom.data_ptr() if mom is not None else None, rank, world,
                 shard_size, lr_dev.data_ptr(), float(mu), float(wd), float(grad_scale), _stream())


def sync_rs_update_ag_range(grad_table, param_table, flag_table, mom_ptr: int, rank, world, base, n, lr_dev, mu, wd, grad_scale, max_blocks=0) -> None:
    _native.call("dmlb_sync_rs_update_ag_range", grad_table, param_table, flag_table, mom_ptr or None, rank, world, base, n, lr_dev.data_ptr(),
                 float(mu), float(wd), float(grad_scale), int(max_blocks), _stream())


def sync_nvls_range(mc_grads: int, mc_params: int, params_local: torch.Tensor, flag_table, mom_ptr: int, rank, world, base, n, lr_dev, mu, wd,
                    grad_scale, max_blocks=0) -> None:
    _native.call("dmlb_sync_nvls_range", mc_grads, mc_params, params_local.data_ptr(), flag_table, mom_ptr or None, rank, world, base, n,
                 lr_dev.data_ptr(), float(mu), float(wd), float(grad_scale), int(max_blocks), _stream())


class SymmMemBuffer:
    """A symmetric buffer obtained from ``torch.distributed._symmetric_memory`` (used for plumbing only): besides the peer
    pointers it exposes ``mc_ptr``, the NVLS multicast address of the buffer (0 when the fabric offers no multicast)."""

    def __init__(self, nbytes: int, device, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        group = group if group is not None else dist.group.WORLD
        self.local = symm_mem.empty((int(nbytes) + 255) // 256 * 256, dtype=torch.uint8, device=device)
        self.local.zero_()
        self.handle = symm_mem.rendezvous(self.local, group=group)
        self.ptrs = [int(p) for p in self.handle.buffer_ptrs]
        self.mc_ptr = int(getattr(self.handle, "multicast_ptr", 0) or 0)
        self.table = (ctypes.c_void_p * len(self.ptrs))(*self.ptrs)

    def view(self, dtype: torch.dtype, numel: Optional[int] = None, offset_bytes: int = 0) -> torch.Tensor:
        t = self.local[offset_bytes:].view(dtype)
        return t if numel is None else t[:numel]
