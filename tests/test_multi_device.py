"""One process, two GPUs: the library keeps its per-device state (SM count, zero rows, debug words, kernel attributes, capture
stream) by device, and tdssf/ops.py makes the tensors' device current around every call - a model on cuda:1 in a process whose
current device is cuda:0 must give the same bits as the same model on cuda:0.  Skipped with fewer than two visible GPUs."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from tdssf import synth  # noqa: E402

W_KEYS = ("mlp.0.weight", "mlp.0.bias", "mlp.2.weight", "mlp.2.bias", "mlp.4.weight", "mlp.4.bias")


def test_second_device_without_set_device():
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs two visible GPUs")
    from tdssf import ops
    from utils.fusion_model import FeatureFusionModel
    from utils.losses import CELSLoss
    torch.cuda.set_device(0)
    sw = synth.make_sweeps(2, 5000, "dinov3", seed=71)
    w = synth.make_mlp_weights(71)

    class Pcd(torch.nn.Module):
        def devoxelize(self, s):
            return s, None

    def run(dev):
        dev = torch.device(dev)
        packed = ops.pack_weights(*[w[k].to(dev) for k in W_KEYS])
        logits, rec = ops.fusion_fwd(sw.points.to(dev), torch.stack(sw.K).to(dev), torch.stack(sw.Rt).to(dev), sw.image_sizes.float().to(dev),
                                     sw.encoder, sw.tokens.to(dev), sw.point_feat.to(dev), packed, want_rec=True)
        model = FeatureFusionModel(sw.encoder, Pcd(), 64, 384, 16, device=dev).to(dev)
        model.load_state_dict(w, strict=True)
        feat = sw.point_feat.to(dev).requires_grad_(True)
        crit = CELSLoss(weight=torch.tensor(synth.CLASS_WEIGHTS, device=dev))
        total, _, _, preds, gt = crit(model(sw.tokens.to(dev), feat, sw.points.to(dev), None, sw.image_sizes, sw.K, sw.Rt),
                                      sw.labels.to(dev), sw.mask.to(dev))
        total.backward()
        cm = ops.confusion(preds, gt, 16)
        assert logits.device == dev and cm.device == dev and feat.grad.device == dev
        return [t.detach().cpu() for t in (logits, rec, total, preds, cm, feat.grad, *[p.grad for p in model.mlp.parameters()])]

    a = run("cuda:0")
    assert torch.cuda.current_device() == 0
    b = run("cuda:1")          # current device stays 0: every op must launch where its tensors live
    assert torch.cuda.current_device() == 0
    for x, y in zip(a[:5], b[:5]):
        assert torch.equal(x, y)   # forward, records, loss, predictions, confusion: bit-identical across devices
    for x, y in zip(a[5:], b[5:]):  # gradients: the weight-gradient partials are summed in a fixed order -> identical too
        assert torch.equal(x, y)
