"""Row-group sharding of one FLS table over the GPUs of a box.

Row groups are independent units (own segments, bases, dictionaries, FSST tables), so rank r of N decodes the
contiguous range [floor(r*R/N), floor((r+1)*R/N)) with its own handle, streams and buffers; output order is the
range order.  There is no collective on the data path; `torch.distributed` is only used by callers for barriers
and for summing counters."""


def rowgroup_range(rank: int, world: int, n_rowgroups: int):
    """Half-open row-group range of `rank`; ranges of ranks 0..world-1 tile [0, n_rowgroups) in order."""
    if not (0 <= rank < world):
        raise ValueError("rank %d outside world of %d" % (rank, world))
    return rank * n_rowgroups // world, (rank + 1) * n_rowgroups // world


def all_ranges(world: int, n_rowgroups: int):
    return [rowgroup_range(r, world, n_rowgroups) for r in range(world)]
