"""Stand-in for fvcore.nn.weight_init (fvcore is not installed in this image).

TEST INFRASTRUCTURE ONLY.  The reference imports `fvcore.nn.weight_init` at
retinanet/Attention.py:5 and uses one function from it, at Attention.py:16-17.
fvcore's published `c2_xavier_fill` is Caffe2's XavierFill: kaiming_uniform_
with a=1 on the weight and a zero bias.  Only oracle/make_golden.py puts this
directory on sys.path, and only to import the unmodified reference module.
"""
import torch.nn as nn


def c2_xavier_fill(module):
    nn.init.kaiming_uniform_(module.weight, a=1)
    if module.bias is not None:
        nn.init.constant_(module.bias, 0)
