"""frn_b200 -- B200-native (sm_100a) implementation of FRN's cross-modal fusion block.

Drop-in replacements for retinanet/fusion.py, retinanet/Attention.py and
retinanet/AdaIN.py of JinkaiLiu/FRN; see DESIGN.md and INTEGRATION.md.
"""
from .modules import (AdaIN_block, Conv1x1, CrossAttentionBlock, REFusion, REFusionLateral, c2_xavier_fill,
                      convert_fpn_laterals, fusion_parameters)
from .detection import Anchors, FocalLoss, decode_boxes, detect_postprocess, nms
from .events import normalize_event_data, preprocess_events, voxelize_events
from .patch import patch_reference
from .dist import allreduce_fusion_grads, shard_batch

__all__ = ["AdaIN_block", "CrossAttentionBlock", "REFusion", "c2_xavier_fill", "fusion_parameters", "Conv1x1", "REFusionLateral", "convert_fpn_laterals",
           "Anchors", "FocalLoss", "detect_postprocess", "decode_boxes", "nms", "voxelize_events", "preprocess_events", "normalize_event_data", "patch_reference", "allreduce_fusion_grads", "shard_batch"]
__version__ = "0.1.0"
