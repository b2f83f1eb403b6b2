"""Drop-in for the evaluation loops of the reference's utils/test.py (`test_model`, utils/test.py:12-91, and
`test_sample`, utils/test.py:93-173), SURVEY f3.  The reference's own utils/test.py also runs unchanged over the other
drop-in modules (tests/test_reference_loops.py); this module is the variant that keeps the statistics on the device.

Same signature, same printed report and returned dict (through utils.evaluation).  What changes is where the
statistics live: the reference appends every batch's predictions and labels to host lists (a D2H copy + sync per
batch, O(dataset) host memory, utils/test.py:69-70) and builds the confusion matrix at the end; here one int64 (C,C)
matrix stays on the device and every batch's `tdssf_confusion` launch accumulates into it, the loss and the correct /
point counters are device scalars, and the host reads them once after the last batch.

With torch.distributed initialised (one process per GPU, every rank iterating its own shard of the test set) the
matrix and the three counters are all-reduced once at the end (tdssf.parallel), so every rank returns the metrics of
the whole set; the integer matrix makes the result independent of the number of ranks.
"""
import torch
from tqdm import tqdm

from tdssf import ops, parallel
from utils.evaluation import compute_iou, per_class_accuracy, precision_recall_f1


def _report(cm, total_loss, total_correct, total_points, iterations):
    """The report and dict of utils/evaluation.py:5-32 from an already accumulated confusion matrix."""
    iou, miou = compute_iou(cm)
    acc, macc = per_class_accuracy(cm)
    precision, recall, f1 = precision_recall_f1(cm)
    fmt = lambda t: [f"{v:.4f}" for v in t.tolist()]
    print("\n".join([
        "\n====== EVALUATION METRICS ======",
        f"Loss: {total_loss/iterations:.4f}, Overall Acc: {total_correct/total_points:.4f}",
        f"Per-Class Acc: {fmt(acc)}, Mean Per-Class Acc: {macc.item():.4f}",
        f"Per-Class IoU: {fmt(iou)}, Mean IoU: {miou.item():.4f}",
        f"Precision: {precision.item():.4f}, Recall: {recall.item():.4f}, F1: {f1.item():.4f}",
        "=================================\n",
    ]))
    return {
        "loss": total_loss / iterations,
        "overall_acc": total_correct / total_points,
        "iou_per_class": iou.tolist(),
        "mean_iou": miou.item(),
        "mean_per_class_acc": macc.item(),
        "precision": precision.item(),
        "recall": recall.item(),
        "f1": f1.item(),
    }


@torch.no_grad()
def test_model(dataloaders, image_encoder, pcd_encoder, model, criterion, device):
    """Evaluate the model on dataloaders['test'] (this rank's shard of it under torch.distributed)."""
    model.eval()
    device = torch.device(device)
    cm = None
    stats = torch.zeros(3, dtype=torch.float64, device=device)   # sum of batch losses, correct points, points
    iterations = 0

    print("🔍 Running test evaluation...")
    for batch in tqdm(dataloaders['test'], desc="Testing", leave=False):
        images, image_sizes, lidar_points, labels, mask, cam_intrinsics, lidar2cam_extrinsics = batch
        images = images.to(device)              # (B, 6, C, H, W)
        image_sizes = image_sizes.to(device)    # (B, 2) (H, W)
        lidar_points = lidar_points.to(device)  # (B, max_P, 4)
        labels = labels.to(device)              # (B, max_P)
        mask = mask.to(device)                  # (B, max_P)

        views = [image_encoder(images[:, v])["patch_features"] for v in range(images.shape[1])]
        patch_tokens = torch.stack(views, dim=1)                                   # (B, 6, M, patch_dim)
        voxel_features, voxel_raw, voxel_coords, voxel_mask = pcd_encoder(lidar_points)
        outputs = model(patch_tokens, voxel_features, voxel_raw, voxel_coords, image_sizes, cam_intrinsics,
                        lidar2cam_extrinsics)                                      # (B, P, num_classes)
        combined_loss, ce_loss, lovasz_loss, predictions, gt_labels_valid = criterion(outputs, labels, mask=mask)

        num_classes = outputs.shape[-1]
        if cm is None:
            cm = torch.zeros((num_classes, num_classes), dtype=torch.int64, device=device)
        if predictions.numel():
            ops.confusion(predictions, gt_labels_valid, num_classes, out=cm)       # accumulates on the device
        stats[0] += combined_loss.detach().double()
        stats[1] += (predictions == gt_labels_valid).sum()
        stats[2] += gt_labels_valid.numel()
        iterations += 1

    if cm is None:
        raise RuntimeError("test_model: the test loader produced no batches")
    if parallel.world()[1] > 1:
        parallel.allreduce_confusion(cm)
        torch.distributed.all_reduce(stats)
        iterations = int(parallel.sum_over_ranks(iterations, device))
    total_loss, total_correct, total_points = stats.tolist()                       # the one host read
    return _report(cm, total_loss, total_correct, total_points, iterations)


@torch.no_grad()
def test_sample(config, dataloaders, image_encoder, pcd_encoder, model, criterion, device):
    """One batch of dataloaders['test'] through the model, metrics + the scene arrays (utils/test.py:93-173).
    The projection plot is drawn by the reference's utils.plot when that module (matplotlib / plotly / cv2) is
    importable; the returned dict is the same either way."""
    model.eval()
    device = torch.device(device)
    batch = next(iter(dataloaders["test"]))
    images, image_sizes, lidar_points, labels, mask, cam_intrinsics, lidar2cam_extrinsics = batch
    images = images.to(device)
    image_sizes = image_sizes.to(device)
    lidar_points = lidar_points.to(device)
    labels = labels.to(device)
    mask = mask.to(device)
    cam_intrinsics = [c.to(device) for c in cam_intrinsics]
    lidar2cam_extrinsics = [c.to(device) for c in lidar2cam_extrinsics]

    views = [image_encoder(images[:, v])["patch_features"] for v in range(images.shape[1])]
    patch_tokens = torch.stack(views, dim=1)
    voxel_features, voxel_raw, voxel_coords, voxel_mask = pcd_encoder(lidar_points)
    outputs = model(patch_tokens, voxel_features, voxel_raw, voxel_coords, image_sizes, cam_intrinsics,
                    lidar2cam_extrinsics)
    combined_loss, ce_loss, lovasz_loss, predictions, gt_labels_valid = criterion(outputs, labels, mask=mask)

    num_classes = outputs.shape[-1]
    cm = torch.zeros((num_classes, num_classes), dtype=torch.int64, device=device)
    if predictions.numel():
        ops.confusion(predictions, gt_labels_valid, num_classes, out=cm)
    total_correct = (predictions == gt_labels_valid).sum().double()
    total_points = gt_labels_valid.size(0)
    metrics = _report(cm, combined_loss, total_correct, total_points, 1)

    valid_pts = lidar_points[0][mask.bool()[0]]            # (N_valid, 4), first sample like the reference
    all_preds, all_labels = predictions.cpu(), gt_labels_valid.cpu()
    results = {**metrics, "points": valid_pts.cpu().numpy(), "predictions": all_preds, "labels": all_labels}
    try:
        from utils.plot import plot_images_with_point_cloud
    except ImportError:
        plot_images_with_point_cloud = None
    if plot_images_with_point_cloud is not None:
        plot_images_with_point_cloud(config=config, images=images, points=valid_pts, pred_labels=all_preds,
                                     gt_labels=all_labels, cam_intrinsics=cam_intrinsics,
                                     lidar2cam_extrinsics=lidar2cam_extrinsics,
                                     save_dir=config['test_params']['checkpoint_path'])
    return results
