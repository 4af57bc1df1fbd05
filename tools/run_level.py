"""Run REFusion fwd+bwd at one DSEC / DDD17 level shape a few times (the command ncu / compute-sanitizer wrap).

    python tools/run_level.py [level 0-3 | all] [batch] [iters] [dsec|ddd17] [fwd|fwdbwd] [f32|bf16]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import frn_b200

LEVELS = {"dsec": [(256, 120, 160), (512, 60, 80), (1024, 30, 40), (2048, 15, 20)],
          "ddd17": [(256, 65, 87), (512, 33, 44), (1024, 17, 22), (2048, 9, 11)]}
lvl = sys.argv[1] if len(sys.argv) > 1 else "0"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 8
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 1
shapes = LEVELS[sys.argv[4] if len(sys.argv) > 4 else "dsec"]
shapes = shapes if lvl == "all" else [shapes[int(lvl)]]
mode = sys.argv[5] if len(sys.argv) > 5 else "fwdbwd"
dtype = torch.bfloat16 if (len(sys.argv) > 6 and sys.argv[6] == "bf16") else torch.float32
dev = torch.device("cuda:0")
torch.manual_seed(0)
for c, h, w in shapes:
    m = frn_b200.REFusion(c, c).to(dev)
    a = torch.randn(B, c, h, w, device=dev).relu().to(dtype).requires_grad_(mode == "fwdbwd")
    e = torch.randn(B, c, h, w, device=dev).relu().to(dtype).requires_grad_(mode == "fwdbwd")
    go = torch.randn(B, 2 * c, h, w, device=dev).to(dtype)
    for _ in range(iters):
        if mode == "fwdbwd":
            m(a, e).backward(go)
        else:
            with torch.no_grad():
                m(a, e)
    torch.cuda.synchronize()
    print(f"ok: C={c} {h}x{w} B={B} x{iters} {mode} {dtype}")
