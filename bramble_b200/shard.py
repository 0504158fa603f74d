"""Multi-GPU sharding of the projection path: bundles are independent (all coupling -- pairing, NH/HI,
primary choice -- is inside one read-name group and groups never span bundles, reference
src/bramble.cpp:362,396-398), so bundle k goes to rank k mod world_size, the annotation index is
replicated on every GPU, and the host merges the per-bundle row blocks back in bundle order.
No collective is needed on the data path; `gather_in_order` is host-side plumbing only."""


def bundles_of_rank(n_bundles, world_size, rank):
    return list(range(rank, n_bundles, world_size))


def owner_of(bundle, world_size):
    return bundle % world_size


def gather_in_order(local_results, n_bundles, world_size, rank, dist=None):
    """local_results: {bundle_index: rows}.  Returns the full list in bundle order on rank 0 (None elsewhere).
    `dist` is torch.distributed (gloo or nccl object collectives); None means single process."""
    if dist is None or world_size == 1:
        return [local_results[k] for k in range(n_bundles)]
    gathered = [None] * world_size
    dist.all_gather_object(gathered, local_results)
    if rank != 0:
        return None
    merged = {}
    for part in gathered:
        merged.update(part)
    assert sorted(merged) == list(range(n_bundles)), "a bundle was lost or duplicated"
    return [merged[k] for k in range(n_bundles)]
