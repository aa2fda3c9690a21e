"""Multi-GPU plumbing for the ORB front end: frames and keyframe-pair jobs are independent, so they
are dealt round-robin to the ranks (frame f -> rank f mod G) with NO collective on the data path; the only exchange
is one final gather of what each rank found to one root rank.

Two forms.  `Comm` (below) is the product path: the C ABI's orbx_comm_* / orbx_gather_* (csrc/k_gather.cu) -- dense records,
peer window over NVLink, its own stream.  The torch.distributed helpers at the top (gather_frames: an all_gather of padded
records; global_order) are the portable form the world-size-2 gloo tests run on CPU.
Padded record layout per frame (cap = ORBextractor.max_keypoints()): n (int32), cap x 28 B keypoints, cap x 32 B descriptors.
"""
import torch
import torch.distributed as dist


def shard_indices(n_items, rank, world):
    """Global indices owned by `rank`: rank, rank + world, ... (static round-robin)."""
    return list(range(rank, n_items, world))


def local_count(n_items, rank, world):
    return (n_items - rank + world - 1) // world if rank < n_items else 0


def gather_frames(kps, desc, n, world, group=None, out=None):
    """kps (B, cap, 28) u8, desc (B, cap, 32) u8, n (B,) i32 on this rank -> the same three tensors for all ranks,
    rank-major: (world * B, ...).  B must be equal on all ranks (pad the last shard).  `out` = preallocated tensors."""
    if world == 1:
        return kps, desc, n
    if out is None:
        out = [torch.empty((world * kps.shape[0],) + tuple(kps.shape[1:]), dtype=kps.dtype, device=kps.device),
               torch.empty((world * desc.shape[0],) + tuple(desc.shape[1:]), dtype=desc.dtype, device=desc.device),
               torch.empty(world * n.shape[0], dtype=n.dtype, device=n.device)]
    dist.all_gather_into_tensor(out[0], kps.contiguous(), group=group)
    dist.all_gather_into_tensor(out[1], desc.contiguous(), group=group)
    dist.all_gather_into_tensor(out[2], n.contiguous(), group=group)
    return out


def global_order(gathered, n_items, world):
    """Undo the round-robin deal: returns index tensor `src` with result_of_frame[f] = gathered[src[f]]
    (rank-major gathered layout, B = ceil(n_items / world) slots per rank)."""
    B = (n_items + world - 1) // world
    f = torch.arange(n_items)
    return (f % world) * B + (f // world)


# ---------------------------------------------------------------------------------------------------------------------------
# The product path: the final gather behind the C ABI (include/orbx.h: orbx_comm_*, orbx_gather_frames, orbx_gather_matches;
# csrc/k_gather.cu).  Compact records only, to ONE root rank, on the communicator's own stream.  torch.distributed is used for
# nothing but handing NCCL's 128-byte unique id to the other ranks; any other transport does as well.
import ctypes
import numpy as np


class Comm:
    """orbx_comm: one per rank, tied to an ORBextractor's context (device + stream)."""

    def __init__(self, ex, world=1, rank=0, unique_id=None):
        from .binding import lib
        L = lib()
        vp, ci = ctypes.c_void_p, ctypes.c_int
        L.orbx_comm_unique_id.argtypes = [vp]
        L.orbx_comm_create.argtypes = [ctypes.POINTER(vp), vp, ci, ci, vp]
        L.orbx_comm_destroy.argtypes = [vp]
        L.orbx_comm_destroy.restype = None
        L.orbx_comm_last_error.argtypes = [vp]
        L.orbx_comm_last_error.restype = ctypes.c_char_p
        for f in (L.orbx_comm_mark, L.orbx_comm_wait):
            f.argtypes = [vp, vp]
        L.orbx_comm_sync.argtypes = [vp]
        L.orbx_comm_window.argtypes = [vp, ctypes.c_size_t, ctypes.POINTER(vp)]
        L.orbx_comm_stats.argtypes = [vp, ctypes.POINTER(ci), ctypes.POINTER(ci)]
        L.orbx_gather_frames.argtypes = [vp, vp, ci, vp, vp, vp, ci, ci, ci, vp, vp, vp, ctypes.c_size_t, ctypes.POINTER(ctypes.c_longlong)]
        L.orbx_gather_frames_dense.argtypes = L.orbx_gather_frames.argtypes
        L.orbx_gather_matches.argtypes = [vp, vp, ci, vp, vp, ci, ci, ci, vp, vp, ctypes.c_size_t, ctypes.POINTER(ctypes.c_longlong)]
        self._L, self._ex, self.world, self.rank = L, ex, world, rank
        self._h = vp()
        idp = None if unique_id is None else (ctypes.c_uint8 * 128).from_buffer_copy(bytes(unique_id))
        rc = L.orbx_comm_create(ctypes.byref(self._h), ex._h, world, rank, idp)
        if rc:
            raise RuntimeError(f"orbx_comm_create: {L.orbx_strerror(rc).decode()} ({rc})")

    @staticmethod
    def unique_id():
        from .binding import lib
        L = lib()
        L.orbx_comm_unique_id.argtypes = [ctypes.c_void_p]
        buf = (ctypes.c_uint8 * 128)()
        rc = L.orbx_comm_unique_id(buf)
        if rc:
            raise RuntimeError(f"orbx_comm_unique_id: {L.orbx_strerror(rc).decode()} ({rc})")
        return bytes(buf)

    @classmethod
    def from_torch(cls, ex):
        """Inside an initialised torch.distributed job: rank 0 makes the id, broadcast_object_list hands it round."""
        world, rank = dist.get_world_size(), dist.get_rank()
        box = [cls.unique_id() if rank == 0 else None]
        if world > 1:
            dist.broadcast_object_list(box, src=0)
        return cls(ex, world, rank, box[0] if world > 1 else None)

    def _chk(self, rc, what):
        if rc < 0:
            raise RuntimeError(f"{what}: {self._L.orbx_strerror(rc).decode()} ({rc}) {self._L.orbx_comm_last_error(self._h).decode()}")
        return rc

    def mark(self):
        """'The results enqueued on the extractor so far are complete' -> ticket for a later gather."""
        return self._chk(self._L.orbx_comm_mark(self._h, self._ex._h), "orbx_comm_mark")

    def wait(self):
        self._chk(self._L.orbx_comm_wait(self._h, self._ex._h), "orbx_comm_wait")

    def sync(self):
        self._chk(self._L.orbx_comm_sync(self._h), "orbx_comm_sync")

    def window(self, nbytes):
        """Collective (orbx_comm_window): this rank's peer window of `nbytes` as a torch uint8 CUDA tensor (None for 0 bytes), or
        None on every rank when the windows could not be set up (the gathers then go through ncclSend/ncclRecv).  Root-side
        gather buffers carved out of it (`carve`) are written by the other ranks' compaction kernels directly."""
        import torch
        base = ctypes.c_void_p()
        rc = self._L.orbx_comm_window(self._h, int(nbytes), ctypes.byref(base))
        if rc == -5:                                  # ORBX_E_UNSUPPORTED, on every rank alike
            return None
        self._chk(rc, "orbx_comm_window")
        if not base.value:
            return None

        class _Mem:                                   # the communicator owns the memory; the tensor is a view of it
            pass
        m = _Mem()
        m.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (base.value, False), "version": 3}
        m._owner = self
        return torch.as_tensor(m, device=torch.device("cuda", self._ex.device))

    @staticmethod
    def carve(window, *shapes_dtypes):
        """Views into a window tensor: carve(win, ((R, 28), torch.uint8), ((R, 32), torch.uint8)) -> tensors at 256-byte steps."""
        out, off = [], 0
        for shape, dtype in shapes_dtypes:
            import math
            nb = math.prod(shape) * dtype.itemsize
            out.append(window[off:off + nb].view(dtype).view(*shape))
            off += (nb + 255) & ~255
        return out

    @staticmethod
    def carve_bytes(*shapes_dtypes):
        import math
        return sum(((math.prod(s) * d.itemsize + 255) & ~255) for s, d in shapes_dtypes)

    def stats(self):
        """(gathers that went through the root's window, gathers staged through ncclSend/ncclRecv)"""
        a, b = ctypes.c_int(), ctypes.c_int()
        self._chk(self._L.orbx_comm_stats(self._h, ctypes.byref(a), ctypes.byref(b)), "orbx_comm_stats")
        return a.value, b.value

    def gather_frames(self, kps, desc, n, out, root=0, ticket=-1, dense=False):
        """kps (B, cap, 28) u8, desc (B, cap, 32) u8, n (B,) i32: torch CUDA tensors of this rank.  out = (all_n (world * B,) i32,
        all_kps (R, 28) u8, all_desc (R, 32) u8) CUDA tensors, used at the root.  Returns the total number of records.
        dense=True: kps / desc hold dense records already (extract_batch_device(dense=True)); no compaction pass."""
        P = lambda t: ctypes.c_void_p(t.data_ptr())
        total = ctypes.c_longlong()
        self._chk((self._L.orbx_gather_frames_dense if dense else self._L.orbx_gather_frames)(self._h, self._ex._h, ticket, P(kps), P(desc), P(n), kps.shape[0], kps.shape[1], root,
                                             P(out[0]), P(out[1]), P(out[2]), out[1].shape[0], ctypes.byref(total)), "orbx_gather_frames")
        return total.value

    def gather_matches(self, match, rows, out, root=0, ticket=-1):
        """match (pairs, cap) i32, rows (pairs,) i32 CUDA tensors; out = (all_n (world * pairs,) i32, all_pairs (R, 2) i32)."""
        P = lambda t: ctypes.c_void_p(t.data_ptr())
        total = ctypes.c_longlong()
        self._chk(self._L.orbx_gather_matches(self._h, self._ex._h, ticket, P(match), P(rows), match.shape[0], match.shape[1], root,
                                              P(out[0]), P(out[1]), out[1].shape[0], ctypes.byref(total)), "orbx_gather_matches")
        return total.value

    def close(self):
        if self._h:
            self._L.orbx_comm_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
