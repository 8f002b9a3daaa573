"""Host-side multi-GPU logic on CPU: shard arithmetic and the torch.distributed handle
exchange, run with the gloo backend at world_size 2 (no GPU needed)."""
import os
import socket
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_problem_shard_partitions_exactly():
    from linear_programming_b200.sharded import problem_shard
    for total in (0, 1, 7, 64, 65536, 65537):
        for world in (1, 2, 3, 8):
            spans = [problem_shard(total, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            for (a, b), (c, d) in zip(spans, spans[1:]):
                assert b == c and a <= b
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        problem_shard(5, 2, 2)


def test_column_shard_covers_all_columns():
    from linear_programming_b200 import build
    from linear_programming_b200.sharded import column_shard
    build.build()
    for m in (1, 3, 15, 16, 17, 100, 32767):
        for world in (1, 2, 3, 8):
            got = [column_shard(m, world, r) for r in range(world)]
            width = got[0][1] if world > 1 and got[0][1] < m else None
            cover = 0
            for lo, cnt in got:
                assert cnt >= 0
                if cnt:
                    assert lo == cover  # contiguous, in rank order
                    cover += cnt
            assert cover == m
            # positions are dealt in blocks of a multiple of 16 (TMA/vector alignment)
            if width is not None:
                assert width % 16 == 0


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        sys.path.insert(0, ROOT)
        import torch.distributed as dist
        from linear_programming_b200.sharded import gather_handles, problem_shard
        dist.init_process_group("gloo", rank=rank, world_size=world)
        handle = bytes([(rank * 37 + i) & 0xff for i in range(64)])
        allh = gather_handles(handle)
        ok = len(allh) == 64 * world
        for r in range(world):
            ok = ok and allh[64 * r:64 * (r + 1)] == bytes([(r * 37 + i) & 0xff for i in range(64)])
        # batch sharding: the ranks' blocks tile the batch exactly once (checked collectively)
        lo, hi = problem_shard(1001, world, rank)
        spans = [None] * world
        dist.all_gather_object(spans, (lo, hi))
        ok = ok and spans[0][0] == 0 and spans[-1][1] == 1001 and all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, bool(ok)))
    except Exception as e:  # pragma: no cover
        q.put((rank, repr(e)))


def test_gloo_world_size_2_handle_exchange_and_sharding():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == {0: True, 1: True}, res
