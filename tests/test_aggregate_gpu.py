"""GPU: multi-level feature aggregation (network/fusion_depthnet.py:57-66) through the C ABI against
the golden fixture taken from the reference's FusedDepthNet and against the oracle on seeded inputs, including odd
level sizes (ResNet strides round up), a single level, and the hand-off into VFNet (channel-last, no copy)."""
import pytest
import torch
import torch.nn as nn

from tests import helpers as H
from oracle import vfdepth_oracle as orc
from vfdepth_b200 import aggregate

pytestmark = pytest.mark.gpu
TOL = 1e-4      # fp32; the 1x1 conv is applied before the resize instead of after it (both linear)


def _dev():
    if not torch.cuda.is_available():
        pytest.skip('needs a GPU')
    return torch.device('cuda:0')


def _block(cin, cout, dev):
    # the reference's conv2d() factory with kernel_size=1 (network/blocks.py:41-61)
    return nn.Sequential(nn.Conv2d(cin, cout, kernel_size=1, padding=0, bias=True, padding_mode='reflect'),
                         nn.Identity(), nn.LeakyReLU(0.1, inplace=True)).to(dev)


def test_golden_feature_aggregation():
    """forward and all gradients against the reference's own FusedDepthNet (fixture: tests/golden/make_golden.py)."""
    dev = _dev()
    g = H.load_npz('golden_feat_agg.npz')
    levels = [g['feat_%d' % i].to(dev).requires_grad_(True) for i in range(3)]
    blk = _block(896, 64, dev)
    with torch.no_grad():
        blk[0].weight.copy_(g['weight'].to(dev))
        blk[0].bias.copy_(g['bias'].to(dev))
    out = aggregate.aggregate_features(levels, 0, blk, 1, 6)
    assert out.shape == (1, 6, 64, 8, 12)
    ref = g['agg'].to(dev).view(1, 6, 64, 8, 12)
    assert H.rel_err(out, ref) < TOL
    grads = torch.autograd.grad((out * g['g_up'].to(dev).view_as(out)).sum(), levels + [blk[0].weight, blk[0].bias])
    names = ['d_feat_0', 'd_feat_1', 'd_feat_2', 'd_weight', 'd_bias']
    for n, a in zip(names, grads):
        assert H.rel_err(a, g[n].to(dev)) < TOL, (n, H.rel_err(a, g[n].to(dev)))


@pytest.mark.parametrize('sizes,chans,cout', [(((9, 15), (5, 8), (3, 4)), (128, 256, 512), 256),
                                              (((11, 20), (6, 10)), (24, 40), 32),
                                              (((7, 5),), (16,), 8),
                                              (((12, 16), (6, 8), (3, 4), (1, 1)), (8, 8, 8, 8), 16)])
def test_aggregation_matches_oracle(sizes, chans, cout):
    """odd sizes, two / one / four levels (a 1x1 coarsest level: scale 0), against the oracle on the same inputs;
    the result is deterministic (two runs bit-identical, gradients included)."""
    dev = _dev()
    gen = torch.Generator().manual_seed(sum(chans) + cout)
    n = 4
    levels_cpu = [torch.randn((n, c, h, w), generator=gen) for (h, w), c in zip(sizes, chans)]
    blk = _block(sum(chans), cout, dev)
    with torch.no_grad():
        blk[0].bias.normal_(0, 0.3)
    up = torch.randn((n, cout) + sizes[0], generator=gen)
    runs = []
    for _ in range(2):
        levels = [x.to(dev).requires_grad_(True) for x in levels_cpu]
        out = aggregate.aggregate_features(levels, 0, blk, 2, 2)
        grads = torch.autograd.grad((out * up.to(dev).view_as(out)).sum(), levels + [blk[0].weight, blk[0].bias])
        runs.append([out.detach()] + list(grads))
    for a, b in zip(*runs):
        assert torch.equal(a, b)
    lv = [x.clone().requires_grad_(True) for x in levels_cpu]
    w, b = blk[0].weight.detach().cpu().requires_grad_(True), blk[0].bias.detach().cpu().requires_grad_(True)
    ref = orc.aggregate_features(lv, 0, w, b)
    # LeakyReLU kink: compare gradients with the oracle's slopes pinned to the kernel's where they differ near zero
    out = runs[0][0].cpu().reshape(ref.shape)
    assert H.rel_err(out, ref) < TOL
    flip = (out > 0) != (ref > 0)
    assert int(flip.sum()) <= max(1, int(1e-4 * out.numel()))
    pre = torch.where(ref > 0, ref, ref / 0.1)
    pinned = pre * torch.where(out > 0, 1.0, 0.1)
    rg = torch.autograd.grad((pinned * up).sum(), lv + [w, b])
    for i, (a, r) in enumerate(zip(runs[0][1:], rg)):
        assert H.rel_err(a.cpu(), r) < TOL, (i, H.rel_err(a.cpu(), r))


def test_aggregation_feeds_vfnet_without_copy():
    """the output is a [B,cams,C,h,w] view of channel-last storage: VFNet's layout conversion is a no-op on it."""
    dev = _dev()
    from vfdepth_b200.vfnet import VFNet
    gen = torch.Generator().manual_seed(3)
    levels = [torch.randn((12, 32, 12, 20), generator=gen).to(dev), torch.randn((12, 64, 6, 10), generator=gen).to(dev)]
    blk = _block(96, 64, dev)
    out = aggregate.aggregate_features(levels, 0, blk, 2, 6)
    cl = VFNet._channel_last(out)
    assert cl.data_ptr() == out.data_ptr() and cl.is_contiguous()


def test_aggregation_rejects_bad_arguments():
    dev = _dev()
    blk = _block(96, 64, dev)
    levels = [torch.randn((12, 32, 12, 20), device=dev), torch.randn((12, 32, 6, 10), device=dev)]
    with pytest.raises(ValueError):
        aggregate.aggregate_features(levels, 0, blk, 2, 6)          # channel count != conv1x1 input
    levels[1] = torch.randn((12, 64, 6, 10), device=dev)
    with pytest.raises(ValueError):
        aggregate.aggregate_features(levels, 0, blk, 2, 5)          # batch * cams mismatch
