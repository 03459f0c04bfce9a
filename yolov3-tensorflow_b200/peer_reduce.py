"""All-reduce of the loss rows over the GPUs of one NVLink domain as ONE kernel over peer-mapped memory
(csrc/peer_reduce.cu, ``yv3_peer_allreduce``): an ordinary launch, so it can sit inside the captured step next to the
fold that produces the row -- which the NCCL call could not (round 1 kept it outside the CUDA graph, one call per
16-step replay).  torch supplies the plumbing only: a symmetric-memory allocation whose peer mappings come from
``torch.distributed._symmetric_memory.rendezvous`` (one process per GPU, same node).

``PeerAllReduce.create(group, cap)`` returns None when the peers cannot be mapped (single rank, another node, a torch
without symmetric memory); callers then keep the NCCL all-reduce.
"""
import ctypes

import torch

from . import _lib


class PeerAllReduce(object):
    def __init__(self, mailbox, ptrs, world, rank, cap, handle):
        self._mailbox, self._handle = mailbox, handle          # keep the mapping alive
        self.world, self.rank, self.cap = world, rank, cap
        self._ptrs = (ctypes.c_void_p * world)(*ptrs)
        self._lib = _lib.lib()

    @classmethod
    def create(cls, group=None, cap=512, device=None):
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()):
            return None
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        if world < 2 or world > 16:
            return None
        lib = _lib.lib()
        size = ctypes.c_size_t(0)
        _lib.check(lib.yv3_peer_mailbox_bytes(world, cap, ctypes.byref(size)), "yv3_peer_mailbox_bytes")
        device = torch.device("cuda", torch.cuda.current_device()) if device is None else device
        ok = torch.ones(1, dtype=torch.int32, device=device)
        mailbox = handle = ptrs = None
        try:
            import torch.distributed._symmetric_memory as symm
            mailbox = symm.empty(size.value, dtype=torch.uint8, device=device)
            mailbox.zero_()
            torch.cuda.synchronize(device)
            handle = symm.rendezvous(mailbox, group=(dist.group.WORLD if group is None else group))
            ptrs = [int(p) for p in handle.buffer_ptrs]
            if len(ptrs) != world or any(p == 0 for p in ptrs):
                raise RuntimeError("incomplete peer table")
        except Exception:
            ok.zero_()
        # every rank or none: a rank that failed must not leave the others polling
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
        if int(ok) == 0:
            return None
        self = cls(mailbox, ptrs, world, rank, cap, handle)
        dist.barrier(group=group)                               # every mailbox is zeroed before anyone pushes
        return self

    def __call__(self, rows, out=None):
        """out = sum over ranks of rows (fp32, contiguous, at most ``cap`` elements); in place by default.
        Stream-ordered on the current stream; every rank must make the same sequence of calls."""
        out = rows if out is None else out
        count = rows.numel()
        if rows.dtype != torch.float32 or not rows.is_contiguous() or count > self.cap or out.numel() != count:
            raise ValueError("rows must be a contiguous fp32 tensor of at most %d elements" % self.cap)
        with torch.cuda.device(rows.device):
            rc = self._lib.yv3_peer_allreduce(self._ptrs, self.world, self.rank, self.cap, _lib.ptr(rows), count,
                                              _lib.ptr(out), _lib.current_stream_ptr(rows.device))
        _lib.check(rc, "yv3_peer_allreduce")
        return out

    def check_status(self):
        rc = self._lib.yv3_peer_status(_lib.ptr(self._mailbox), _lib.current_stream_ptr(self._mailbox.device))
        if rc == 1:
            raise RuntimeError("peer all-reduce: a rank did not arrive within the poll bound")
        _lib.check(rc, "yv3_peer_status")
