"""GPU: the library is re-entrant (include/frn_b200.h conventions; the reference is driven by nn.DataParallel,
train_dsec.py:113, i.e. one Python thread per replica, and by the autograd engine's own thread in backward):
two host threads run REFusion fwd+bwd concurrently on their own CUDA streams and must reproduce, bit for bit,
what each computes alone."""
import threading

import pytest
import torch

import frn_b200

pytestmark = pytest.mark.gpu


def _job(mod, a, e, go, stream, out, key, reps):
    with torch.cuda.stream(stream):
        for _ in range(reps):
            x, y = a.clone().requires_grad_(True), e.clone().requires_grad_(True)
            mod.zero_grad()
            o = mod(x, y)
            o.backward(go)
        stream.synchronize()
        out[key] = (o.detach().clone(), x.grad.clone(), y.grad.clone(),
                    {k: p.grad.clone() for k, p in mod.named_parameters()})


def test_two_threads_two_streams_bitwise():
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    jobs = []
    for c, h, w in ((256, 24, 32), (512, 12, 16)):           # streaming attention (d = 32 / 64) on both threads
        mod = frn_b200.REFusion(c, c).to(dev)
        g = torch.Generator().manual_seed(c)
        a = torch.randn(2, c, h, w, generator=g).relu().to(dev)
        e = torch.randn(2, c, h, w, generator=g).relu().to(dev)
        go = torch.randn(2, 2 * c, h, w, generator=g).to(dev)
        jobs.append((mod, a, e, go))
    torch.cuda.synchronize()
    alone, together = {}, {}
    for i, (mod, a, e, go) in enumerate(jobs):
        _job(mod, a, e, go, torch.cuda.Stream(dev), alone, i, 1)
    threads = [threading.Thread(target=_job, args=(*job, torch.cuda.Stream(dev), together, i, 8))
               for i, job in enumerate(jobs)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    torch.cuda.synchronize()
    for i in alone:
        o1, da1, de1, g1 = alone[i]
        o2, da2, de2, g2 = together[i]
        assert torch.equal(o1, o2) and torch.equal(da1, da2) and torch.equal(de1, de2)
        for k in g1:
            assert torch.equal(g1[k], g2[k]), k
