"""Work placement and tracing helpers.

Same public names and results as the reference's ``kfac/utils.py`` (file:line cites
are into /root/reference).  The placement functions are integer work and are
bit-exact with the reference (tests/test_host_logic.py checks them against
tests/golden/allocator.json and the reference's known-answer tables).
"""
import time

from . import comm

_FUNC_TRACES = {}


def clear_trace():
    # the reference's version rebinds a local and never clears (kfac/utils.py:11-12)
    _FUNC_TRACES.clear()


def get_trace(max_history=None):
    out = {}
    for fname, times in _FUNC_TRACES.items():
        if max_history is not None and len(times) > max_history:
            times = times[-max_history:]
        out[fname] = sum(times) / len(times)
    return out


def print_trace(max_history=None):
    for fname, t in get_trace(max_history).items():
        print('{}: {}'.format(fname, t))


def trace(sync=False):
    """Wall-clock a function, optionally between barriers (kfac/utils.py:39-56)."""
    def decorator(func):
        def func_timer(*args, **kwargs):
            if sync:
                comm.backend.barrier()
            t = time.time()
            out = func(*args, **kwargs)
            if sync:
                comm.backend.barrier()
            _FUNC_TRACES.setdefault(func.__name__, []).append(time.time() - t)
            return out
        return func_timer
    return decorator


def partition_grad_ranks(size, grad_workers):
    """Gradient broadcast groups: ranks congruent mod grad_workers (kfac/utils.py:150-153)."""
    return [list(range(first, size, grad_workers)) for first in range(grad_workers)]


def partition_inv_ranks(size, grad_workers):
    """Inverse broadcast groups: consecutive blocks of grad_workers ranks (kfac/utils.py:156-159)."""
    return [list(range(lo, min(lo + grad_workers, size))) for lo in range(0, size, grad_workers)]


def load_balance(n_workers, work):
    """Greedy longest-processing-time placement (kfac/utils.py:169-196).

    Items are visited in stable descending-cost order; each goes to the worker with
    the smallest load so far, the lowest index winning ties."""
    if not n_workers > 0:
        raise ValueError('n_workers must be > 0')
    if len(work) == 0:
        raise ValueError('work cannot be an empty list')
    load = [0] * n_workers
    placed = [0] * len(work)
    for idx in sorted(range(len(work)), key=lambda i: work[i], reverse=True):
        target = load.index(min(load))
        load[target] += work[idx]
        placed[idx] = target
    return placed


class WorkerAllocator(object):
    """Splits the world into inverse groups and gradient groups (kfac/utils.py:59-147).

    ``grad_workers = max(1, round(size * fraction))`` ranks hold a layer's inverses
    and precondition it; each of them then feeds one gradient broadcast group."""

    def __init__(self, size, compute_grad_fraction):
        grad_workers = max(1, round(size * compute_grad_fraction))
        if size % grad_workers != 0:
            raise ValueError('compute_grad_fraction must produce equally '
                             'size groups')
        self.size = size
        self.compute_grad_fraction = compute_grad_fraction
        self.bcast_grad_ranks = partition_grad_ranks(size, grad_workers)
        self.bcast_inv_ranks = partition_inv_ranks(size, grad_workers)
        # group creation order is part of the contract: inverse groups first
        self.bcast_inv_groups = [self.make_group(r) for r in self.bcast_inv_ranks]
        self.bcast_grad_groups = [self.make_group(r) for r in self.bcast_grad_ranks]

    @property
    def grad_groups(self):
        return len(self.bcast_grad_groups)

    @property
    def inv_groups(self):
        return len(self.bcast_inv_groups)

    def make_group(self, ranks):
        return comm.CommGroup(ranks)

    @staticmethod
    def _index_of(rank, partition):
        for i, members in enumerate(partition):
            if rank in members:
                return i
        raise ValueError('rank {} is in no group'.format(rank))

    def get_grad_ranks(self, rank):
        return self.bcast_grad_ranks[self._index_of(rank, self.bcast_grad_ranks)]

    def get_inv_ranks(self, rank):
        return self.bcast_inv_ranks[self._index_of(rank, self.bcast_inv_ranks)]

    def get_inv_group(self, rank):
        return self.bcast_inv_groups[self._index_of(rank, self.bcast_inv_ranks)]

    def get_grad_group(self, rank):
        return self.bcast_grad_groups[self._index_of(rank, self.bcast_grad_ranks)]

    def get_grad_groups(self, src_ranks):
        """For every rank: (source rank inside its gradient group, that group)."""
        sources = set(src_ranks)
        pairs = []
        for rank in range(self.size):
            members = self.get_grad_ranks(rank)
            src = [r for r in members if r in sources][0]
            pairs.append((src, self.get_grad_group(rank)))
        return pairs


def try_contiguous(x):
    return x if x.is_contiguous() else x.contiguous()


def get_block_boundary(index, block_count, shape):
    """Start/end indices of block ``index`` when cutting ``shape`` into
    ``block_count`` diagonal blocks (kfac/utils.py:199-212; unused by the hot path)."""
    if index >= block_count:
        raise ValueError("Index ({}) greater than number of requested blocks "
                         "({})".format(index, block_count))
    if block_count > min(shape):
        raise ValueError("Requested blocks ({}) greater than minimum possible "
                         "blocks for shape {}".format(block_count, shape))
    step = [extent // block_count for extent in shape]
    start = [s * index for s in step]
    last = index + 1 == block_count
    end = [extent if last else s * (index + 1) for s, extent in zip(step, shape)]
    return start, end
