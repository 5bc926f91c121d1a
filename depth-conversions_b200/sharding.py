"""Frame sharding shared by bench.py and the tests: the path shards by frame
index with no data-path collective. `depth2x --devices N`
uses the same contiguous-block rule in C++ (host/apps/depth2x/pipeline.cpp
run_batch)."""


def contiguous_blocks(start, end, parts):
    """Inclusive index range [start, end] cut into `parts` contiguous blocks
    (the first `total % parts` ones one index longer, so that no GPU idles: 9
    frames on 8 GPUs are 2+1+1+1+1+1+1+1): a list of (first, last). Reversed bounds are swapped, as
    parallel_processing.h:31-34 does."""
    if start > end:
        start, end = end, start
    total = end - start + 1
    parts = max(1, min(parts, total))
    base, extra = divmod(total, parts)
    out, first = [], start
    for g in range(parts):
        n = base + (1 if g < extra else 0)
        out.append((first, first + n - 1))
        first += n
    return out


def rank_frames(rank, world, frames_per_rank):
    """bench.py (weak scaling): rank r owns synthetic frames
    [r * F, (r + 1) * F) of every step."""
    return range(rank * frames_per_rank, (rank + 1) * frames_per_rank)
