"""GPU: the C ABI only enqueues work (no allocation, no synchronisation, side streams forked from and joined into the
caller's stream), so a whole fwd+bwd step can be captured into ONE CUDA graph and replayed (VERDICT r1 item 5;
model.py:305-308 is the loop a user would capture).  The replay must reproduce the eagerly launched step bit for bit,
for the fp32 training path with the levels on their own streams and for the bf16 inference path."""
import pytest
import torch

import frn_b200

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _step(mods, inputs, douts, streams):
    main = torch.cuda.current_stream()
    outs = []
    for s, m, (a, e), go in zip(streams, mods, inputs, douts):
        s.wait_stream(main)
        with torch.cuda.stream(s):
            a.grad = None; e.grad = None
            out = m(a, e)
            out.backward(go)
            outs.append(out.detach())
    for s in streams:
        main.wait_stream(s)
    return outs


def test_training_step_replays_from_one_cuda_graph():
    torch.manual_seed(0)
    shapes = [(2, 256, 24, 32), (2, 512, 12, 16), (2, 1024, 6, 8)]        # d = 32, 64 (streaming kernels), 128 (GEMM path)
    mods = [frn_b200.REFusion(c, c).to(DEV) for _, c, _, _ in shapes]
    inputs = [(torch.randn(s, device=DEV).relu().requires_grad_(True), torch.randn(s, device=DEV).relu().requires_grad_(True))
              for s in shapes]
    douts = [torch.randn(b, 2 * c, h, w, device=DEV) for b, c, h, w in shapes]
    streams = [torch.cuda.Stream() for _ in shapes]
    for m in mods:
        for p in m.parameters():
            p.grad = torch.zeros_like(p)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):                       # warm-up: lazily created side streams, smem attributes, allocator
        for _ in range(3):
            _step(mods, inputs, douts, streams)
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()

    def zero_param_grads():                 # parameter gradients ACCUMULATE (autograd adds into existing .grad)
        for m in mods:
            for p in m.parameters():
                p.grad.zero_()

    zero_param_grads()
    eager_out = [o.clone() for o in _step(mods, inputs, douts, streams)]
    torch.cuda.synchronize()
    eager_da = [a.grad.clone() for a, _ in inputs]
    eager_gw = [m.conv0_rgb.weight.grad.clone() for m in mods]
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        outs = _step(mods, inputs, douts, streams)
    for _ in range(2):
        for a, e in inputs:
            a.grad.zero_(); e.grad.zero_()
        zero_param_grads()
        graph.replay()
    torch.cuda.synchronize()
    for o, ref in zip(outs, eager_out):
        assert torch.equal(o, ref)
    for (a, _), ref in zip(inputs, eager_da):
        assert torch.equal(a.grad, ref)
    for m, ref in zip(mods, eager_gw):
        assert torch.equal(m.conv0_rgb.weight.grad, ref)


def test_bf16_inference_replays_from_one_cuda_graph():
    torch.manual_seed(1)
    m = frn_b200.REFusion(256, 256).to(DEV).eval()
    a = torch.randn(3, 256, 17, 22, device=DEV).relu().bfloat16()          # DDD17-like odd N: padded maps
    e = torch.randn(3, 256, 17, 22, device=DEV).relu().bfloat16()
    with torch.no_grad():
        for _ in range(2):
            ref = m(a, e)
        torch.cuda.synchronize()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            m(a, e)
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            out = m(a, e)
        graph.replay()
        torch.cuda.synchronize()
    assert out.dtype == torch.bfloat16 and torch.equal(out, ref)
