"""Frame sharding shared by bench.py and the tests: the path shards by frame
index with no data-path collective. `depth2x --devices N`
uses the same contiguous-block rule in C++ (host/apps/depth2x/pipeline.cpp
run_batch)."""


def contiguous_blocks(start, end, parts):
    """Inclusive index range [start, end] cut into `parts` contiguous blocks
    (the first ones one index longer when it does not divide): a list of
    (first, last) with empty blocks dropped. Reversed bounds are swapped, as
    parallel_processing.h:31-34 does."""
    if start > end:
        start, end = end, start
    total = end - start + 1
    parts = max(1, min(parts, total))
    per = (total + parts - 1) // parts
    out = []
    for g in range(parts):
        first = start + g * per
        last = min(end, first + per - 1)
        if first <= last:
            out.append((first, last))
    return out


def rank_frames(rank, world, frames_per_rank):
    """bench.py (weak scaling): rank r owns synthetic frames
    [r * F, (r + 1) * F) of every step."""
    return range(rank * frames_per_rank, (rank + 1) * frames_per_rank)
