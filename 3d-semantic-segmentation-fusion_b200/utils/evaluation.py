"""Drop-in for the reference's utils/evaluation.py: evaluate, compute_confusion_matrix, compute_iou,
per_class_accuracy, overall_accuracy, precision_recall_f1, measure_efficiency
(utils/evaluation.py:5-81).  The confusion matrix is the int64 histogram kernel of libtdssf.so
(csrc/confusion.cu); every derived metric is computed from its three marginals."""
import time

import torch

from tdssf import ops


def _marginals(cm: torch.Tensor):
    """diag = correct per class, truth = row sums, predicted = column sums, all as fp32."""
    return torch.diag(cm).float(), cm.sum(dim=1).float(), cm.sum(dim=0).float()


def compute_confusion_matrix(preds, labels, num_classes, out=None):
    """preds, labels in 1..C on any device (host tensors are uploaded once — the reference's loops hand
    CPU tensors here, utils/test.py:82-87).  Returns the int64 (C,C) matrix, rows = truth, on the inputs'
    device.  `out` (CUDA int64 (C,C)) accumulates across calls without leaving the device.

    Values outside 1..C (torch.bincount would raise on a negative index and grow the matrix on a large one,
    utils/evaluation.py:38-40) raise RuntimeError.  The check never adds a device synchronisation to a loop:
    host inputs are checked on the host before the upload; device inputs without `out` are checked through the one
    scalar the caller is about to read anyway (the matrix total); with `out` the kernel drops such pairs and the
    caller can compare `out.sum()` with its own point count after the last batch."""
    if not torch.cuda.is_available():
        raise RuntimeError("tdssf confusion matrix runs on a CUDA device only (no CPU fallback)")
    home = preds.device
    dev = out.device if out is not None else (home if preds.is_cuda else torch.device("cuda"))
    n = preds.numel()
    if n and not preds.is_cuda and not labels.is_cuda:
        lo = min(int(preds.min()), int(labels.min()))
        hi = max(int(preds.max()), int(labels.max()))
        if lo < 1 or hi > num_classes:
            raise RuntimeError(f"confusion matrix: values must lie in 1..{num_classes} (got {lo}..{hi})")
    cm = ops.confusion(preds.to(dev), labels.to(dev), num_classes, out=out)
    if out is not None:
        return cm
    if n and (preds.is_cuda or labels.is_cuda) and int(cm.sum()) != n:
        raise RuntimeError(f"confusion matrix: {n - int(cm.sum())} of {n} pairs lie outside 1..{num_classes}")
    return cm.to(home)


def compute_iou(conf_matrix):
    # the union is formed in int64 like the reference (utils/evaluation.py:46-49): exact beyond 2^24 points
    hit = torch.diag(conf_matrix)
    union = conf_matrix.sum(dim=1) + conf_matrix.sum(dim=0) - hit
    iou = hit.float() / union.float().clamp(min=1)
    return iou, iou.mean()


def per_class_accuracy(conf_matrix):
    hit, truth, _ = _marginals(conf_matrix)
    acc = hit / truth.clamp(min=1)
    return acc, acc.mean()


def overall_accuracy(conf_matrix):
    # integer sums first, one conversion each (utils/evaluation.py:59-62)
    return torch.diag(conf_matrix).sum().float() / conf_matrix.sum().float().clamp(min=1)


def precision_recall_f1(conf_matrix):
    hit, truth, predicted = _marginals(conf_matrix)
    # the reference forms fp = column sum - tp and fn = row sum - tp in fp32 and adds tp back (utils/evaluation.py:66-71);
    # the same two roundings are kept so that the macro averages agree bit for bit
    precision = (hit / (hit + (predicted - hit)).clamp(min=1)).mean()
    recall = (hit / (hit + (truth - hit)).clamp(min=1)).mean()
    f1 = 2 * precision * recall / (precision + recall).clamp(min=1e-6)
    return precision, recall, f1


def evaluate(all_preds, all_labels, num_classes, total_loss, total_correct, total_points, iteration):
    """Prints the reference's report and returns its dict (utils/evaluation.py:5-32)."""
    cm = compute_confusion_matrix(all_preds, all_labels, num_classes)
    iou, miou = compute_iou(cm)
    acc, macc = per_class_accuracy(cm)
    precision, recall, f1 = precision_recall_f1(cm)
    fmt = lambda t: [f"{v:.4f}" for v in t.tolist()]
    report = [
        "\n====== EVALUATION METRICS ======",
        f"Loss: {total_loss/iteration:.4f}, Overall Acc: {total_correct/total_points:.4f}",
        f"Per-Class Acc: {fmt(acc)}, Mean Per-Class Acc: {macc.item():.4f}",
        f"Per-Class IoU: {fmt(iou)}, Mean IoU: {miou.item():.4f}",
        f"Precision: {precision.item():.4f}, Recall: {recall.item():.4f}, F1: {f1.item():.4f}",
        "=================================\n",
    ]
    print("\n".join(report))
    return {
        "loss": total_loss / iteration,
        "overall_acc": total_correct / total_points,
        "iou_per_class": iou.tolist(),
        "mean_iou": miou.item(),
        "mean_per_class_acc": macc.item(),
        "precision": precision.item(),
        "recall": recall.item(),
        "f1": f1.item(),
    }


def measure_efficiency(func, *args, **kwargs):
    """Wall time and allocated device memory of one call (utils/evaluation.py:76-81)."""
    t0 = time.time()
    result = func(*args, **kwargs)
    stats = {"runtime_sec": time.time() - t0,
             "mem_alloc_MB": torch.cuda.memory_allocated() / (1024 ** 2) if torch.cuda.is_available() else 0}
    return result, stats
