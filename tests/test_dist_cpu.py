"""CPU: the N>1 host logic (batch sharding + fusion-gradient allreduce) on gloo, world_size 2."""
import os

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from frn_b200.dist import GradBucket, allreduce_fusion_grads, shard_batch


def test_shard_batch_covers_everything():
    for gb in (1, 7, 8, 64):
        for world in (1, 2, 4, 8):
            spans = [shard_batch(gb, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == gb
            for (a, b), (c, d) in zip(spans, spans[1:]):
                assert b == c
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import frn_b200
    torch.manual_seed(0)
    m = frn_b200.REFusion(16, 16)                      # same weights on every rank
    params = frn_b200.fusion_parameters(m)
    for i, p in enumerate(params):
        p.grad = torch.full_like(p, float(rank + 1) * (i + 1))
    bucket = GradBucket(params)
    bucket.allreduce(average=True)
    ok = all(torch.allclose(p.grad, torch.full_like(p, 1.5 * (i + 1))) for i, p in enumerate(params))
    # second step through the packed path
    for i, p in enumerate(params):
        p.grad.fill_(float(rank) + i)
    allreduce_fusion_grads(params, average=False)
    ok = ok and all(torch.allclose(p.grad, torch.full_like(p, 1.0 + 2 * i)) for i, p in enumerate(params))
    ok = ok and all(p.grad.data_ptr() >= bucket.flat.data_ptr() for p in params)
    q.put((rank, bool(ok)))
    dist.destroy_process_group()


def test_gradient_allreduce_gloo_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(60)
    assert res == {0: True, 1: True}
