"""EvalWrapper -- the caller of the voting layer in the reference, with the producer side fused.

Mirrors the forward signature of
    tools/demo.py:46-55               EvalWrapper.forward(seg_pred, vertex_pred, use_argmax=True)      -> v3(512, 0.99)
    tools/train_linemod.py:116-125    EvalWrapper.forward(mask_pred, vertex_pred, use_argmax, use_uncertainty)
    lib/test_linemod_pbr.py:102-114   EvalWrapper.forward(seg_pred, vertex_pred, use_argmax, use_uncertainty)
The three reference call sites hard-code different layer arguments; `EvalWrapper.preset(name)` reproduces each one:
    "demo"              v3(512, inlier_thresh=0.99)                      (max_num = the layer's default 30000)
    "train_linemod"     v3 / v5(128, inlier_thresh=0.99, max_num=100)    tools/train_linemod.py:122-125
    "test_linemod_pbr"  v3 / v5(128, inlier_thresh=0.99, max_num=100)    lib/test_linemod_pbr.py:108-113
train_linemod's wrapper ignores `use_argmax` and is handed an already arg-maxed [b,h,w] mask (tools/train_linemod.py:254):
a 3-D input is therefore always taken as the mask, whatever `use_argmax` says; only a 4-D input is treated as logits.
In the reference `mask = torch.argmax(seg_pred, 1)` writes an int64 [b,h,w] tensor (2.46 MB per 480x640 image) that the
voting layer reads back at once.  Here the logits go straight to the compaction pass (PVV_MASK_LOGITS_F32), which takes
the argmax per pixel while it builds the foreground bitmask: the mask never exists in memory.  vertex_pred is consumed
as the permuted view of the network's [b,2*vn,h,w] output, exactly as the reference builds it.
"""
from __future__ import annotations

import torch
from torch import nn

from .ransac_voting_gpu import ransac_voting_layer_v3, ransac_voting_layer_v5


class EvalWrapper(nn.Module):
    PRESETS = {"demo": dict(round_hyp_num=512, inlier_thresh=0.99),
               "train_linemod": dict(round_hyp_num=128, inlier_thresh=0.99, max_num=100),
               "test_linemod_pbr": dict(round_hyp_num=128, inlier_thresh=0.99, max_num=100)}

    def __init__(self, round_hyp_num=512, inlier_thresh=0.99, **layer_kwargs):
        super().__init__()
        self.round_hyp_num, self.inlier_thresh, self.layer_kwargs = round_hyp_num, inlier_thresh, layer_kwargs

    @classmethod
    def preset(cls, name, **layer_kwargs):
        """The layer arguments one of the reference's call sites hard-codes (see the module docstring)."""
        return cls(**{**cls.PRESETS[name], **layer_kwargs})

    @torch.no_grad()
    def forward(self, seg_pred, vertex_pred, use_argmax=True, use_uncertainty=False):
        vertex_pred = vertex_pred.permute(0, 2, 3, 1)
        b, h, w, vn_2 = vertex_pred.shape
        vertex_pred = vertex_pred.view(b, h, w, vn_2 // 2, 2)
        # 4-D: logits [b,C,h,w], the argmax is fused into the compaction pass; 3-D: already the mask [b,h,w] (what
        # train_linemod's wrapper receives with use_argmax left at its default)
        if seg_pred.dim() == 4 and not use_argmax:
            raise ValueError("use_argmax=False expects a [b,h,w] mask, got logits [b,C,h,w]")
        mask = seg_pred
        layer = ransac_voting_layer_v5 if use_uncertainty else ransac_voting_layer_v3
        return layer(mask, vertex_pred, self.round_hyp_num, inlier_thresh=self.inlier_thresh, **self.layer_kwargs)
