"""EvalWrapper -- the caller of the voting layer in the reference, with the producer side fused.

Mirrors (same forward signature, same outputs)
    tools/demo.py:46-55               EvalWrapper.forward(seg_pred, vertex_pred, use_argmax=True)      -> v3(512, 0.99)
    tools/train_linemod.py:116-125    EvalWrapper.forward(mask_pred, vertex_pred, use_argmax, use_uncertainty)
    lib/test_linemod_pbr.py:102-114   EvalWrapper.forward(seg_pred, vertex_pred, use_argmax, use_uncertainty)
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
    def __init__(self, round_hyp_num=512, inlier_thresh=0.99, **layer_kwargs):
        super().__init__()
        self.round_hyp_num, self.inlier_thresh, self.layer_kwargs = round_hyp_num, inlier_thresh, layer_kwargs

    @torch.no_grad()
    def forward(self, seg_pred, vertex_pred, use_argmax=True, use_uncertainty=False):
        vertex_pred = vertex_pred.permute(0, 2, 3, 1)
        b, h, w, vn_2 = vertex_pred.shape
        vertex_pred = vertex_pred.view(b, h, w, vn_2 // 2, 2)
        # use_argmax: seg_pred holds logits [b,C,h,w] (fused argmax); otherwise it already is the mask [b,h,w]
        if use_argmax and seg_pred.dim() != 4:
            raise ValueError("use_argmax=True expects seg_pred logits [b,C,h,w]")
        mask = seg_pred
        layer = ransac_voting_layer_v5 if use_uncertainty else ransac_voting_layer_v3
        return layer(mask, vertex_pred, self.round_hyp_num, inlier_thresh=self.inlier_thresh, **self.layer_kwargs)
