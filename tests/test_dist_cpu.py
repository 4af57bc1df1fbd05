"""CPU: the N>1 host logic (batch sharding + fusion-gradient allreduce) on gloo, world_size 2."""
import os

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from frn_b200.dist import GradBucket, LevelBuckets, allreduce_fusion_grads, shard_batch


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
    # third step the way the reference loop runs (train_dsec_det_fast.py:237): optimizer.zero_grad() with the
    # torch >= 2.0 default set_to_none=True detaches every .grad from the bucket; a backward then allocates fresh
    # gradients outside it.  allreduce() must reduce THOSE, not the stale bucket, and re-alias them.
    opt = torch.optim.SGD(params, lr=0.1)
    opt.zero_grad()
    ok = ok and all(p.grad is None for p in params)
    loss = sum((p * float(rank + 1) * (i + 1)).sum() for i, p in enumerate(params))
    loss.backward()                                     # dL/dp = (rank + 1) * (i + 1), in fresh tensors
    ok = ok and all(p.grad.data_ptr() != v.data_ptr() for p, v in zip(params, bucket.views))
    bucket.allreduce(average=True)
    ok = ok and all(torch.allclose(p.grad, torch.full_like(p, 1.5 * (i + 1))) for i, p in enumerate(params))
    ok = ok and all(p.grad.data_ptr() == v.data_ptr() for p, v in zip(params, bucket.views))
    # a parameter that received no gradient this step contributes zeros, not last step's values
    opt.zero_grad()
    params[0].grad = torch.full_like(params[0], float(rank + 1))
    bucket.allreduce(average=False)
    ok = ok and torch.allclose(params[0].grad, torch.full_like(params[0], 3.0))
    ok = ok and all(float(p.grad.abs().max()) == 0.0 for p in params[1:])
    # per-level buckets issued asynchronously in a fixed level order (bench.py / SURVEY.md 8.2 e)
    mods = [frn_b200.REFusion(8, 8), frn_b200.REFusion(16, 16)]
    lb = LevelBuckets(mods)
    for k, mod in enumerate(mods):
        for i, p in enumerate(mod.parameters()):
            p.grad.fill_(float(rank + 1) * (k + 1))
        lb.allreduce_level(k, average=True)
    lb.wait()
    ok = ok and all(torch.allclose(p.grad, torch.full_like(p, 1.5 * (k + 1))) for k, mod in enumerate(mods) for p in mod.parameters())
    ok = ok and lb.numel == sum(p.numel() for mod in mods for p in mod.parameters())
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
