"""Peer-visible device memory and the cross-GPU barrier of the data-parallel train step (csrc/peer.cu, include/foc_b200.h
`foc_peer_*`).  One process per GPU on one NVSwitch node: every rank cudaMalloc's an arena, the 64-byte IPC handles
travel through torch.distributed (any backend -- they are host bytes), every rank opens its peers' arenas and from then
on kernels load / store peer HBM directly.  torch.distributed is plumbing here; the data path has no collective."""
import ctypes

import torch
import torch.distributed as dist

from ._lib import check, lib, stream

MAX_PEERS = 8
PAD_BYTES = 256                       # signal pad at the start of every arena (24 uint32 used)

_TYPESTR = {torch.float16: "<f2", torch.float32: "<f4", torch.int32: "<i4", torch.uint8: "|u1"}


class FocPeerGroup(ctypes.Structure):
    """include/foc_b200.h: struct FocPeerGroup"""
    _fields_ = [("world", ctypes.c_int32), ("rank", ctypes.c_int32), ("flags", ctypes.c_void_p * MAX_PEERS), ("state", ctypes.c_void_p)]


class _Raw:
    """A span of raw device memory for torch.as_tensor (CUDA array interface v2); keeps its owner alive."""

    def __init__(self, owner, ptr, numel, dtype):
        self._owner = owner
        self.__cuda_array_interface__ = {"shape": (numel,), "typestr": _TYPESTR[dtype], "data": (ptr, False), "version": 2, "strides": None}


class PeerArena:
    """`nbytes` of zero-filled peer-visible memory per rank (+ the signal pad).  base[r] = device address of rank r's arena
    as mapped into THIS process (base[rank] is the local allocation)."""

    def __init__(self, nbytes, group=None, device=None):
        if not dist.is_initialized():
            raise RuntimeError("PeerArena needs an initialised torch.distributed process group (handles are exchanged through it)")
        self.group, self.world, self.rank = group, dist.get_world_size(group), dist.get_rank(group)
        if self.world > MAX_PEERS:
            raise RuntimeError(f"PeerArena: at most {MAX_PEERS} ranks (one NVSwitch node)")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.nbytes = PAD_BYTES + ((int(nbytes) + 255) // 256) * 256
        l = lib()
        with torch.cuda.device(self.device):
            p = ctypes.c_void_p()
            check(l.foc_peer_alloc(self.nbytes, ctypes.byref(p)), "peer_alloc")
            self._local = p.value
            handle = (ctypes.c_uint8 * 64)()
            check(l.foc_peer_export(self._local, handle), "peer_export")
            handles = [None] * self.world
            dist.all_gather_object(handles, bytes(handle), group=group)
            self.base, self._opened = [0] * self.world, []
            for r, h in enumerate(handles):
                if r == self.rank:
                    self.base[r] = self._local
                    continue
                q = ctypes.c_void_p()
                check(l.foc_peer_open((ctypes.c_uint8 * 64).from_buffer_copy(h), ctypes.byref(q)), f"peer_open(rank {r})")
                self.base[r] = q.value
                self._opened.append(q.value)
        self.state = torch.zeros(4, dtype=torch.int32, device=self.device)     # epoch, non-finite flag, status
        self._struct = FocPeerGroup(self.world, self.rank, (ctypes.c_void_p * MAX_PEERS)(*self.base), self.state.data_ptr())
        dist.barrier(group=group)                                              # every arena is mapped everywhere before first use

    def tensor(self, offset, numel, dtype, rank=None):
        """View of `numel` elements at byte `offset` of the data area of rank `rank`'s arena (default: the local one)."""
        r = self.rank if rank is None else rank
        if offset % 16 or offset + numel * torch.empty((), dtype=dtype).element_size() > self.nbytes - PAD_BYTES:
            raise ValueError("PeerArena.tensor: span outside the arena or not 16-byte aligned")
        return torch.as_tensor(_Raw(self, self.base[r] + PAD_BYTES + offset, numel, dtype), device=self.device)

    def address(self, rank, offset=0):
        return self.base[rank] + PAD_BYTES + offset

    def barrier(self, found_inf=None):
        """Device-side meeting of all ranks on the current stream (no host sync; graph-capturable)."""
        check(lib().foc_peer_barrier(ctypes.byref(self._struct), None if found_inf is None else found_inf.data_ptr(), stream()), "peer_barrier")

    def timed_out(self):
        """True when a barrier gave up waiting (a peer died or fell > 2 s behind); host sync."""
        return bool(self.state[2].item())

    def close(self):
        l = lib()
        torch.cuda.synchronize(self.device)
        for q in self._opened:
            l.foc_peer_close(q)
        self._opened = []
        if self._local:
            l.foc_peer_free(self._local)
            self._local = 0


class FocFrameOut(ctypes.Structure):
    """include/foc_b200.h: struct FocFrameOut"""
    _fields_ = [("world", ctypes.c_int32), ("rank", ctypes.c_int32), ("n_total", ctypes.c_uint32), ("frame", ctypes.c_void_p * MAX_PEERS)]


def render_finish(image, depth, weights_sum, nears, fars, bg, depth_from_first_sample, out_struct):
    """csrc/peer.cu k_render_finish: background mix + depth normalisation + packing into [r, g, b, depth, weights_sum] rows of
    the frame buffer(s) `out_struct` names, one launch."""
    N = weights_sum.shape[0]
    for t in (image, depth, weights_sum, nears, fars):
        if t.dtype != torch.float32 or not t.is_contiguous() or not t.is_cuda:
            raise RuntimeError("render_finish: contiguous float32 CUDA tensors expected")
    check(lib().foc_render_finish(image.data_ptr(), depth.data_ptr(), weights_sum.data_ptr(), nears.data_ptr(), fars.data_ptr(), N, float(bg),
                                  int(bool(depth_from_first_sample)), ctypes.byref(out_struct), stream()), "render_finish")


def local_frame(n_rays, device):
    """Single-GPU target of render_finish: ([n_rays, 5] buffer, FocFrameOut naming it)."""
    buf = torch.empty(n_rays, 5, dtype=torch.float32, device=device)
    fr = (ctypes.c_void_p * MAX_PEERS)()
    fr[0] = buf.data_ptr()
    return buf, FocFrameOut(1, 0, n_rays, fr)


class FrameExchange:
    """An n_rays-ray frame rendered by all ranks of the group: rays dealt in 128-ray blocks round-robin (dist.cyclic_tile),
    every rank's finished rows stored straight into every rank's frame buffer over peer memory by the render epilogue, one
    device-side barrier per frame -- no collective, no staging copy.  Two frame buffers alternate so that a fast rank's
    next frame never lands in the buffer a slow rank is still reading."""

    def __init__(self, n_rays, group=None, device=None):
        self.n_rays = int(n_rays)
        self._bytes = -(-self.n_rays * 20 // 256) * 256
        self.arena = PeerArena(2 * self._bytes, group=group, device=device)
        a = self.arena
        self.world, self.rank = a.world, a.rank
        self._frames = [a.tensor(k * self._bytes, self.n_rays * 5, torch.float32).view(self.n_rays, 5) for k in range(2)]
        self._structs = [FocFrameOut(a.world, a.rank, self.n_rays, (ctypes.c_void_p * MAX_PEERS)(*[a.address(r, k * self._bytes) for r in range(a.world)]))
                         for k in range(2)]
        self._k = 0

    def my_rays(self, device=None):
        """Indices (int64) of the frame's rays this rank renders."""
        from .dist import cyclic_tile
        return cyclic_tile(self.n_rays, self.rank, self.world, device=device)

    def finish(self, image, depth, weights_sum, nears, fars, bg, depth_from_first_sample):
        """-> the complete [n_rays, 5] frame on this rank (valid once the stream reaches this point)."""
        k = self._k
        self._k ^= 1
        render_finish(image, depth, weights_sum, nears, fars, bg, depth_from_first_sample, self._structs[k])
        self.arena.barrier()
        return self._frames[k]

    def close(self):
        self._frames = []
        self.arena.close()
