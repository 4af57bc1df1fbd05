"""Install the B200 modules into an importable copy of the reference package.

retinanet/model.py:10 does `from retinanet.fusion import REFusion` (and builds
`self.fus` at model.py:247-252), so the name has to be rebound in
`retinanet.model` as well as in `retinanet.fusion` / `Attention` / `AdaIN`
before `model.resnet50(...)` is called.  model.py itself is not touched.
"""
from __future__ import annotations

import importlib
import sys


def patch_reference(package: str = "retinanet"):
    """Rebind REFusion / CrossAttentionBlock / AdaIN_block (and Anchors / FocalLoss, row f2) inside `package`.  Returns the
    list of (module, attribute) pairs that were rebound."""
    from . import detection, modules

    done = []
    targets = {
        f"{package}.AdaIN": ("AdaIN_block",),
        f"{package}.Attention": ("CrossAttentionBlock", "AdaIN_block"),
        f"{package}.fusion": ("REFusion", "CrossAttentionBlock", "AdaIN_block"),
        f"{package}.model": ("REFusion", "Anchors"),          # model.py:7,221: `self.anchors = Anchors()`
        f"{package}.anchors": ("Anchors",),
        f"{package}.losses": ("FocalLoss",),                  # model.py:227: `losses.FocalLoss()`
    }
    for mod_name, attrs in targets.items():
        mod = sys.modules.get(mod_name)
        if mod is None:
            try:
                mod = importlib.import_module(mod_name)
            except Exception:  # the reference's optional deps (skimage, fvcore ...) may be absent
                continue
        for attr in attrs:
            if hasattr(mod, attr):
                setattr(mod, attr, getattr(modules, attr, None) or getattr(detection, attr))
                done.append((mod_name, attr))
    return done
