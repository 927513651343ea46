"""Frame sharding across ranks: frame i -> rank i mod world, no collective on the hot path,
results gathered in frame order so the output bytes do not depend on the world size.

CPU leg: world_size-2 gloo job; the extractor behind the sharding logic is the CPU oracle (checker standing in for the
device, host logic only).  GPU leg: the same job with the CUDA extractor, both ranks on cuda:0."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class _OracleBatch:
    def __init__(self, nf):
        import oracle_py as O
        self.e = O.OracleExtractor(nf, 1.2, 8, 20, 7)

    def extract_batch(self, images):
        return [self.e(im) for im in images]


def _worker(rank, world, port, use_gpu, n_frames, w, h, nf, q):
    sys.path.insert(0, os.path.join(ROOT, "orb_slam2v2-1_b200"))
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch.distributed as dist
    from orbx import shard, synth
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        frames = [synth.rects(w, h, 9000 + i) for i in range(n_frames)]
        if use_gpu:
            import orbx
            ex = orbx.ORBextractor(nf, 1.2, 8, 20, 7, max_batch=4)
        else:
            ex = _OracleBatch(nf)
        local = shard.extract_sharded(frames, ex, rank, world, batch=4)
        assert sorted(local) == shard.frames_for_rank(n_frames, rank, world)
        res = shard.gather_results(local, n_frames, dist)
        if rank == 0:
            q.put([(k.tobytes(), d.tobytes()) for k, d in res])
        else:
            assert res is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


def _run(world, use_gpu, n_frames=5, w=320, h=240, nf=300):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, use_gpu, n_frames, w, h, nf, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    return out


def _oracle_serial(n_frames=5, w=320, h=240, nf=300):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    from orbx import synth
    e = _OracleBatch(nf)
    return [(k.tobytes(), d.tobytes()) for k, d in e.extract_batch([synth.rects(w, h, 9000 + i) for i in range(n_frames)])]


def test_round_robin_partition():
    from orbx import shard
    for n in (0, 1, 7, 64):
        for world in (1, 2, 4, 8):
            parts = [shard.frames_for_rank(n, r, world) for r in range(world)]
            assert sorted(sum(parts, [])) == list(range(n))
            assert max(map(len, parts)) - min(map(len, parts)) <= 1


def test_gloo_world2_output_independent_of_world_size():
    assert _run(2, use_gpu=False) == _oracle_serial()


@pytest.mark.gpu
def test_gpu_world2_matches_oracle():
    assert _run(2, use_gpu=True) == _oracle_serial()


# ---- query-row sharding of the large brute-force match

class _OracleMatcher:
    def match_bruteforce(self, q, t):
        import oracle_py as O
        return O.match_bruteforce(q, t)


def _descriptor_sets(nq, nt, seed=5):
    rng = np.random.default_rng(seed)
    base = rng.integers(0, 256, (16, 32)).astype(np.uint8)          # clustered: equal distances and equal best / second occur
    def cluster(n):
        d = base[rng.integers(0, 16, n)].copy()
        d[np.arange(n), rng.integers(0, 32, n)] ^= (1 << rng.integers(0, 8, n)).astype(np.uint8)
        return d
    return cluster(nq), cluster(nt)


def test_query_blocks_are_contiguous_and_even():
    from orbx import shard
    for nq in (0, 1, 7, 64, 10000):
        for world in (1, 2, 3, 4, 8):
            blocks = [shard.query_block(nq, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == nq
            assert all(a[1] == b[0] for a, b in zip(blocks, blocks[1:]))
            sizes = [hi - lo for lo, hi in blocks]
            assert max(sizes) - min(sizes) <= 1


def test_match_sharded_equals_unsharded_on_the_oracle():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    from orbx import shard
    q, t = _descriptor_sets(301, 257)
    whole = _OracleMatcher().match_bruteforce(q, t)
    for world in (1, 2, 3, 8):
        got = shard.match_sharded(q, t, [_OracleMatcher() for _ in range(world)])
        assert all(np.array_equal(a, b) for a, b in zip(got, whole)), world


def _match_worker(rank, world, port, nq, nt, out):
    sys.path.insert(0, os.path.join(ROOT, "orb_slam2v2-1_b200"))
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch.distributed as dist
    from orbx import shard
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        q, t = _descriptor_sets(nq, nt)
        res = shard.gather_match(shard.match_rows_for_rank(q, t, _OracleMatcher(), rank, world), nq, dist)
        if rank == 0:
            out.put([a.tobytes() for a in res])
        else:
            assert res is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_gloo_world2_match_rows_gathered_in_row_order():
    import torch.multiprocessing as mp
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_match_worker, args=(r, 2, port, 203, 190, out)) for r in range(2)]
    for p in procs:
        p.start()
    got = out.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    q, t = _descriptor_sets(203, 190)
    assert got == [a.tobytes() for a in _OracleMatcher().match_bruteforce(q, t)]


@pytest.mark.gpu
def test_gpu_match_sharded_over_every_device_equals_one_device():
    import torch
    import orbx
    from orbx import shard
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    q, t = _descriptor_sets(10000, 10000)
    ndev = torch.cuda.device_count()
    devices = list(range(ndev)) if ndev > 1 else [0, 0, 0]           # one GPU: three contexts (streams) on it
    whole = orbx.ORBmatcher(0.9, True, device=0).match_bruteforce(q, t)
    got = shard.match_sharded(q, t, [orbx.ORBmatcher(0.9, True, device=d) for d in devices])
    assert all(np.array_equal(a, b) for a, b in zip(got, whole))
    small = _OracleMatcher().match_bruteforce(q[:400], t)
    assert all(np.array_equal(a[:400], b) for a, b in zip(got, small))
