"""The repo's ResNet-FPN backbone (frozen batch norm folded into the convolutions) against torchvision's
resnet50(FrozenBatchNorm2d) + BackboneWithFPN with the same weights: same state-dict names, same outputs, same gradients."""
import pytest
import torch

from b200flow.models.resnet_fpn import ResNetFPN


@pytest.mark.parametrize("depth", [18, 50])
def test_folded_backbone_matches_torchvision(depth):
    torchvision = pytest.importorskip("torchvision")
    from torchvision.models.detection.backbone_utils import BackboneWithFPN
    from torchvision.ops import FrozenBatchNorm2d

    torch.manual_seed(0)
    mk = torchvision.models.resnet50 if depth == 50 else torchvision.models.resnet18
    chans = [256, 512, 1024, 2048] if depth == 50 else [64, 128, 256, 512]
    body = mk(weights=None, norm_layer=FrozenBatchNorm2d)
    ref = BackboneWithFPN(body, {"layer1": "0", "layer2": "1", "layer3": "2", "layer4": "3"}, chans, 64)
    with torch.no_grad():       # non-trivial frozen statistics
        for n, b in ref.named_buffers():
            if n.endswith("running_var"):
                b.uniform_(0.5, 2.0)
            elif n.endswith(("running_mean", "bias")):
                b.normal_(0, 0.2)
            elif n.endswith("weight"):
                b.uniform_(0.5, 1.5)
    ours = ResNetFPN(depth, 64)
    sd = {k: v for k, v in ref.state_dict().items() if not k.startswith("body.fc")}
    missing, unexpected = ours.load_state_dict(sd, strict=False)
    assert not missing and not unexpected, (missing, unexpected)
    frozen = [n for n, p in ours.named_parameters() if not p.requires_grad]
    assert frozen and all(n.startswith(("body.conv1", "body.layer1")) for n in frozen)
    x = torch.randn(2, 3, 96, 128)
    a, b = ref(x), ours(x)
    assert list(a) == list(b)
    for k in a:
        assert a[k].shape == b[k].shape
        assert torch.allclose(a[k], b[k], atol=1e-4 * float(a[k].abs().max()), rtol=1e-4), (k, (a[k] - b[k]).abs().max())
    # gradients of a trainable convolution and of an FPN convolution agree as well
    sum(v.square().mean() for v in a.values()).backward()
    sum(v.square().mean() for v in b.values()).backward()
    ga = dict(ref.named_parameters())
    gb = dict(ours.named_parameters())
    for name in ("body.layer3.0.conv2.weight", "fpn.inner_blocks.1.0.weight", "body.layer2.0.downsample.0.weight"):
        assert torch.allclose(ga[name].grad, gb[name].grad, atol=1e-4 * float(ga[name].grad.abs().max()), rtol=2e-3), name
    assert gb["body.layer1.0.conv1.weight"].grad is None
