"""Frame sharding across the GPUs of one box: frames and stereo pairs are independent, so
frame i goes to rank i mod world and there is NO collective on the hot path.  The only communication is the optional
gather of per-frame results to rank 0 (host objects over torch.distributed; gloo on CPU, nccl-launched jobs use the
same code through the default group's object collectives)."""


def frames_for_rank(n_frames, rank, world):
    """Indices of the frames rank `rank` extracts (round-robin keeps the load even when frames differ in content)."""
    return list(range(rank, n_frames, world))


def extract_sharded(frames, extractor, rank, world, batch=64):
    """Run `extractor.extract_batch` on this rank's share of `frames`; returns {frame index: (kps, desc)}."""
    mine = frames_for_rank(len(frames), rank, world)
    out = {}
    for s in range(0, len(mine), batch):
        idx = mine[s:s + batch]
        for i, r in zip(idx, extractor.extract_batch([frames[i] for i in idx])):
            out[i] = r
    return out


def gather_results(local, n_frames, dist=None, dst=0):
    """Collect the per-rank dictionaries on rank `dst` and return the results in frame order (None on other ranks).
    Output bytes are identical for every world size: ordering is by frame index, never by arrival."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return [local[i] for i in range(n_frames)]
    parts = [None] * dist.get_world_size() if dist.get_rank() == dst else None
    dist.gather_object(local, parts, dst=dst)
    if dist.get_rank() != dst:
        return None
    merged = {}
    for p in parts:
        merged.update(p)
    return [merged[i] for i in range(n_frames)]


# ---- large brute-force match (10 000 x 10 000, BASELINE.json configs[3]): query ROWS are split into `world` contiguous
# blocks, the train set (320 KB) is replicated to every GPU, and the per-block (best index, best distance, second distance)
# arrays are concatenated in block order.  Still no collective; column sharding, which would need a (best, second) merge,
# is not used.  The best index of a row does not depend on which rows share its launch, so the result is the unsharded one.

def query_block(nq, rank, world):
    """[lo, hi) of the query rows of `rank`: contiguous blocks whose sizes differ by at most one row."""
    base, rem = divmod(nq, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def match_rows_for_rank(q, t, matcher, rank, world):
    """This rank's block of `matcher.match_bruteforce(q, t)` -> (lo, hi, best_idx, best_dist, second_dist)."""
    lo, hi = query_block(len(q), rank, world)
    bi, bd, sd = matcher.match_bruteforce(q[lo:hi], t)
    return lo, hi, bi, bd, sd


def match_sharded(q, t, matchers):
    """One process, several GPUs: matchers[g] (one per device) scans block g on its own host thread; returns the three
    arrays of the unsharded call.  ctypes releases the GIL inside the C call, so the blocks run concurrently."""
    import threading
    import numpy as np
    world = len(matchers)
    parts, errs = [None] * world, []

    def work(g):
        try:
            parts[g] = match_rows_for_rank(q, t, matchers[g], g, world)
        except Exception as e:        # re-raised on the caller's thread
            errs.append(e)

    th = [threading.Thread(target=work, args=(g,)) for g in range(world)]
    for x in th:
        x.start()
    for x in th:
        x.join()
    if errs:
        raise errs[0]
    return tuple(np.concatenate([p[k] for p in parts]) for k in (2, 3, 4))


def gather_match(local, nq, dist=None, dst=0):
    """torchrun form: every rank passes its match_rows_for_rank result; rank `dst` gets the three arrays in row order."""
    import numpy as np
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return local[2], local[3], local[4]
    parts = [None] * dist.get_world_size() if dist.get_rank() == dst else None
    dist.gather_object(local, parts, dst=dst)
    if dist.get_rank() != dst:
        return None
    parts.sort(key=lambda p: p[0])
    assert parts[0][0] == 0 and parts[-1][1] == nq and all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
    return tuple(np.concatenate([p[k] for p in parts]) for k in (2, 3, 4))
