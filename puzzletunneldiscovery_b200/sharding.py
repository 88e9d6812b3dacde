"""Index-range sharding of sample batches over ranks / GPUs and the host-side gather.

The reference distributes by contiguous task chunks: `partt.get_task_chunk(shape, total_chunks, task_id)`
(/root/reference/src/GP/pipeline/partt.py:10-35) hands process `task_id` the flat indices
[int(k*avg), int((k+1)*avg)) with avg = n / total_chunks, the last chunk taking the remainder, and the
submitter concatenates the per-process result files.  Here one process per GPU does the same over
`torch.distributed`: every rank evaluates its own range with its own replica of the geometry and the label
ranges are all-gathered.  There is no collective on the per-sample path -- the gather moves finished results.

Host logic only (numpy + torch.distributed); the collision work is whatever `evaluate` the caller passes,
i.e. the engine's batched queries.
"""
import numpy as np


def task_chunk(n, total_chunks, index):
    """[lo, hi) of chunk `index`: partt.chunk_it's boundaries for a flat task list of length n."""
    if total_chunks <= 0 or not (0 <= index < total_chunks):
        raise ValueError("bad chunk index")
    avg = n / float(total_chunks)
    lo = int(index * avg)
    hi = n if index == total_chunks - 1 else int((index + 1) * avg)
    return min(lo, n), max(min(hi, n), min(lo, n))


def get_task_chunk(shape, total_chunks, index):
    """partt.get_task_chunk (/root/reference/src/GP/pipeline/partt.py:10-35), literally: the index tuples of
    numpy.ndindex(shape) cut by chunk_it -- boundaries from the RUNNING float sum `last += avg` (not k * avg), the
    last chunk taking what is left."""
    grand = [e for e in np.ndindex(tuple(int(x) for x in shape))]
    avg = len(grand) / float(total_chunks)
    out = []
    last = 0.0
    while last < len(grand) and len(out) < total_chunks:
        out.append(grand[int(last):int(last + avg)])
        last += avg
    out[-1] += grand[int(last):]
    if len(out) != total_chunks:
        raise ValueError("len(out) (%d) != %d" % (len(out), total_chunks))
    return out[index]


def engine_shard(n, parts, i):
    """The range device slot i gets inside one handle (csrc/engine.cu: shard)."""
    return n * i // parts, n * (i + 1) // parts


def balanced_ranges(cost, parts):
    """Contiguous ranges with (nearly) equal total cost: segments are balanced by their expected number of
    steps ceil(dist / verify_delta), known before launch, rather than by count."""
    cost = np.asarray(cost, dtype=np.float64)
    n = len(cost)
    if n == 0:
        return [(0, 0)] * parts
    c = np.concatenate([[0.0], np.cumsum(cost)])
    cuts = [0]
    for k in range(1, parts):
        cuts.append(int(np.searchsorted(c, c[-1] * k / parts, side="left")))
    cuts.append(n)
    cuts = np.maximum.accumulate(np.clip(cuts, 0, n))
    return [(int(cuts[k]), int(cuts[k + 1])) for k in range(parts)]


def expected_steps(dist, verify_delta):
    """Upper bound of the validity checks of a segment walk (unit_world.cc:445-470)."""
    d = np.asarray(dist, dtype=np.float64)
    return np.where(verify_delta >= d, 0.0, np.ceil(d / verify_delta))


def sharded_map(evaluate, arrays, out_dtype, group=None, cost=None):
    """Run `evaluate(*[a[lo:hi] for a in arrays])` on this rank's range and return the full result on every
    rank.  Ranges are count-balanced (task_chunk) or cost-balanced when `cost` is given.  Works with any
    torch.distributed backend (gloo on CPU tensors, nccl on CUDA tensors); single process when
    torch.distributed is not initialised."""
    import torch
    import torch.distributed as dist
    n = len(arrays[0])
    if not (dist.is_available() and dist.is_initialized()):
        return np.asarray(evaluate(*arrays), dtype=out_dtype)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    ranges = balanced_ranges(cost, world) if cost is not None else [task_chunk(n, world, r) for r in range(world)]
    lo, hi = ranges[rank]
    local = np.asarray(evaluate(*[a[lo:hi] for a in arrays]), dtype=out_dtype) if hi > lo else None
    if local is not None and len(local) != hi - lo:
        raise ValueError("evaluate returned %d rows for a range of %d" % (len(local), hi - lo))
    # A rank whose range is empty (n < world, or a cost-balanced cut) has nothing to infer the row shape from:
    # the ranks agree on it first, so that every rank pads to the same tensor shape.
    shapes = [None] * world
    dist.all_gather_object(shapes, None if local is None else tuple(local.shape[1:]), group=group)
    known = [t for t in shapes if t is not None]
    if len(set(known)) > 1:
        raise ValueError("ranks disagree on the row shape of the result: %r" % (sorted(set(known)),))
    trailing = known[0] if known else ()
    if local is None:
        local = np.zeros((0,) + trailing, out_dtype)
    longest = max(h - l for l, h in ranges)
    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    pad = np.zeros((longest,) + trailing, dtype=out_dtype)
    pad[: hi - lo] = local.reshape((hi - lo,) + trailing)
    mine = torch.from_numpy(pad).to(dev)
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine, group=group)
    out = np.empty((n,) + trailing, dtype=out_dtype)
    for r, (l, h) in enumerate(ranges):
        out[l:h] = parts[r][: h - l].cpu().numpy()
    return out


def max_over_ranks(value, group=None):
    """The timing rule of bench.py: a multi-GPU time is the maximum over ranks."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return float(value)
    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    t = torch.tensor([float(value)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())
