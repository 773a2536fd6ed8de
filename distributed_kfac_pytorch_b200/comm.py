"""Communication backend for the K-FAC host logic.

Mirrors the public surface of the reference's ``kfac/comm.py`` (module-global
``backend``, ``init_comm_backend()``, ``CommGroup``, ``Ops``; file:line cites are
into /root/reference) so example scripts that call ``kfac.comm.backend.allreduce``
keep working.  ``torch.distributed`` (NCCL on the B200 box, gloo in CPU tests) is
the transport; Horovod is out of scope.
"""
import enum
import os

import torch.distributed as dist

# the current backend object; set once by init_comm_backend()   (kfac/comm.py:14-21)
backend = None


def init_comm_backend():
    global backend
    if backend is None:
        backend = TorchBackend() if dist.is_available() and dist.is_initialized() else CommBackend()


def reset_comm_backend():
    """Forget the cached backend (tests create/destroy process groups)."""
    global backend
    backend = None


class Ops(enum.Enum):
    Average = 'average'
    Sum = 'sum'


class CommGroup(object):
    """A set of ranks, with a c10d sub-group when it is a strict subset of the world
    (kfac/comm.py:53-64).  Must be constructed in the same order on every rank."""

    def __init__(self, ranks):
        self.ranks = list(ranks)
        self.group = None
        if (dist.is_available() and dist.is_initialized()
                and 1 < self.size < dist.get_world_size()):
            self.group = dist.new_group(self.ranks)

    @property
    def size(self):
        return len(self.ranks)


class CommBackend(object):
    """World-size-1 backend: every collective is a no-op (kfac/comm.py:66-141)."""
    Average = Ops.Average
    Sum = Ops.Sum

    def size(self):
        return 1

    def local_rank(self):
        return 0

    def rank(self):
        return 0

    def allreduce(self, tensor, op=Ops.Average, group=None, async_op=True):
        return None

    def broadcast(self, tensor, src, group=None, async_op=True):
        return None

    def reduce(self, tensor, dst, op=Ops.Average, async_op=True):
        return None

    def barrier(self):
        return None

    def sync(self, handles):
        return None

    def wait(self, handle):
        return None


class TorchBackend(CommBackend):
    """torch.distributed transport (kfac/comm.py:194-275).

    "Average" is SUM followed by an in-place divide by the WORLD size at sync time,
    exactly like the reference (comm.py:222,261), also for sub-groups."""

    def size(self):
        return dist.get_world_size()

    def local_rank(self):
        # the reference returns the raw env string (comm.py:198-203)
        try:
            return os.environ['LOCAL_RANK']
        except KeyError:
            raise RuntimeError('LOCAL_RANK must be set in the environment'
                               'when using torch.distributed')

    def rank(self):
        return dist.get_rank()

    @staticmethod
    def _group_kwargs(group):
        if group is None or group.group is None:
            return {}
        return {'group': group.group}

    def allreduce(self, tensor, op=Ops.Average, group=None, async_op=True):
        if group is not None and group.size <= 1:
            return None
        handle = dist.all_reduce(tensor, async_op=async_op, **self._group_kwargs(group))
        if not async_op:
            if op == Ops.Average:
                tensor /= self.size()
            return None
        return (handle, tensor) if op == Ops.Average else handle

    def broadcast(self, tensor, src, group=None, async_op=True):
        if group is not None and group.size <= 1:
            return None
        return dist.broadcast(tensor, src=src, async_op=async_op, **self._group_kwargs(group))

    def reduce(self, tensor, dst, op=Ops.Average, async_op=True):
        handle = dist.reduce(tensor, dst=dst, async_op=async_op)
        if not async_op:
            if op == Ops.Average:
                tensor /= self.size()
            return None
        return (handle, tensor) if op == Ops.Average else handle

    def barrier(self):
        dist.barrier()

    def sync(self, handles):
        if not isinstance(handles, list):
            handles = [handles]
        for h in handles:
            if isinstance(h, tuple):
                work, tensor = h
                self.wait(work)
                tensor /= self.size()
            else:
                self.wait(h)

    def wait(self, handle):
        if handle is not None:
            handle.wait()
