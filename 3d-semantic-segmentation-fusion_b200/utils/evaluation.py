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
    device.  `out` (CUDA int64 (C,C)) accumulates across calls without leaving the device."""
    if not torch.cuda.is_available():
        raise RuntimeError("tdssf confusion matrix runs on a CUDA device only (no CPU fallback)")
    home = preds.device
    dev = out.device if out is not None else (home if preds.is_cuda else torch.device("cuda"))
    if preds.numel():
        lo = min(int(preds.min()), int(labels.min()))
        hi = max(int(preds.max()), int(labels.max()))
        if lo < 1 or hi > num_classes:  # torch.bincount would raise / grow on these
            raise RuntimeError(f"confusion matrix: values must lie in 1..{num_classes} (got {lo}..{hi})")
    cm = ops.confusion(preds.to(dev), labels.to(dev), num_classes, out=out)
    return cm if out is not None else cm.to(home)


def compute_iou(conf_matrix):
    hit, truth, predicted = _marginals(conf_matrix)
    iou = hit / (truth + predicted - hit).clamp(min=1)
    return iou, iou.mean()


def per_class_accuracy(conf_matrix):
    hit, truth, _ = _marginals(conf_matrix)
    acc = hit / truth.clamp(min=1)
    return acc, acc.mean()


def overall_accuracy(conf_matrix):
    hit, truth, _ = _marginals(conf_matrix)
    return hit.sum() / truth.sum().clamp(min=1)


def precision_recall_f1(conf_matrix):
    hit, truth, predicted = _marginals(conf_matrix)
    # tp + fp = column sum, tp + fn = row sum
    precision = (hit / predicted.clamp(min=1)).mean()
    recall = (hit / truth.clamp(min=1)).mean()
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
