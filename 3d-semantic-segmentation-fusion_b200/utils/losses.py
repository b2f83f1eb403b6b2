"""Drop-in for the reference's utils/losses.py (CELSLoss, utils/losses.py:7-70) on the CUDA kernels
of libtdssf.so (csrc/cels.cu): compaction, softmax, weighted CE, per-class segmented radix sort,
Jaccard gradient, dot product and the closed-form backward."""
from __future__ import annotations

import torch
import torch.nn as nn

from tdssf import ops


class CELSLoss(nn.Module):
    """Cross-Entropy + lambda * Lovasz-Softmax for point-level segmentation.

    weight: class weights of the Cross-Entropy term (16 floats); ignore_index: noise label (0);
    lamda_lovasz: weight of the Lovasz term (spelling follows the reference).
    """

    def __init__(self, weight=None, ignore_index=0, lamda_lovasz=1.0):
        super().__init__()
        self.ce_loss = nn.CrossEntropyLoss(weight=weight)  # kept for state_dict / attribute compatibility
        self.ignore_index = ignore_index
        self.lamda_lovasz = lamda_lovasz

    def forward(self, pred_scores, gt_labels, mask=None):
        """pred_scores (B,P,C) logits, gt_labels (B,P) in 0..C, mask (B,P') optional.
        Returns (total_loss, ce_loss, lovasz_loss, predictions (N,) in 1..C, gt_labels_valid (N,))."""
        B, P, C = pred_scores.shape
        if C != 16:
            raise ValueError(f"tdssf CELS kernels are built for 16 classes, got {C}")
        if mask is not None:
            # length alignment and its warnings: utils/losses.py:32-43
            if mask.shape[1] > P:
                print(f"[CELSLoss] Warning: mask longer ({mask.shape[1]}) than pred_scores ({P}), truncating mask.")
                mask = mask[:, :P]
            elif mask.shape[1] < P:
                print(f"[CELSLoss] Warning: pred_scores longer ({P}) than mask ({mask.shape[1]}), truncating preds.")
                pred_scores = pred_scores[:, :mask.shape[1]]
                gt_labels = gt_labels[:, :mask.shape[1]]
                P = mask.shape[1]
            mask = mask.to(pred_scores.device)
        gt_labels = gt_labels.to(pred_scores.device)
        weight = self.ce_loss.weight
        if weight is None:
            weight = torch.ones(C, dtype=torch.float32, device=pred_scores.device)
        total, ce, lovasz, preds, gt = ops.cels(pred_scores, gt_labels, mask, weight, self.lamda_lovasz,
                                                self.ignore_index)
        if preds.numel() == 0:
            # the reference's empty-selection behaviour: lovasz_softmax_flat returns probas*0
            # (an empty (0,C) tensor), CE over nothing is NaN, and their sum broadcasts to an empty tensor
            lovasz = pred_scores.new_zeros((0, C))
            ce = pred_scores.new_tensor(float("nan"))
            total = ce + self.lamda_lovasz * lovasz
        return total, ce, lovasz, preds, gt
