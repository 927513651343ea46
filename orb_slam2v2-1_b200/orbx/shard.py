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
