"""Train-set partition used by every multi-GPU path: rank g of G
holds train rows [g * ceil(xrows / G), min((g + 1) * ceil(xrows / G), xrows)),
query rows are replicated, per-rank top-2 keys are all-gathered and merged.
host_ctx.cu (single process, SPECTAVI_GPUS > 1) uses the same rule."""


def shard_range(xrows, world, rank):
    """(first row, number of rows) of rank's train shard; empty shards are legal."""
    per = (xrows + world - 1) // world
    x0 = min(rank * per, xrows)
    return x0, max(0, min(per, xrows - x0))
