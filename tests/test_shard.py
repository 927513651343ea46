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
