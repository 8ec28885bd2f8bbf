"""GPU: the sub-method surface of the view synthesis -- Projection.backproject / reproject / forward
(models/geometry/geometry_util.py:33-82) and ViewRendering.get_virtual_image (view_rendering.py:61-82) -- through the C
ABI against the oracle (which is pinned bit for bit to the reference, tests/test_oracle_golden.py): coordinates and
masks bit-exact, samples and gradients within 1e-4."""
import pytest
import torch

from oracle import vfdepth_oracle as orc
from tests import helpers as H
from vfdepth_b200 import config as vcfg
from vfdepth_b200.rig import make_rig

pytestmark = pytest.mark.gpu
TOL = 1e-4


def _dev():
    if not torch.cuda.is_available():
        pytest.skip('needs a GPU')
    return torch.device('cuda:0')


def _pair(dataset, seed, cam=0, src=None, frame=-1):
    """one (source, target) pair of a seeded rig: temporal (src None: same camera, frame f) or spatial (neighbour)."""
    cfg = vcfg.default_config(dataset, batch_size=2, training__height=96, training__width=160,
                              training__focal_length_scale=75)
    inp, ex = make_rig(cfg, seed=seed, with_feats=False)
    if src == 'nbr':
        src = cfg['data']['rel_cam_list'][cam][0]
    s = cam if src is None else src
    T = ex['cam_T_cam'][frame][:, cam]
    if src is not None:
        T = inp['extrinsics_inv'][:, src] @ inp['extrinsics'][:, cam]
        frame = 0
    return dict(src_img=inp[('color', frame, 0)][:, s].contiguous(), src_mask=inp['mask'][:, s].contiguous(),
                depth=ex['depth'][cam].contiguous(), invK=inp[('inv_K', 0)][:, cam].contiguous(),
                K=inp[('K', 0)][:, s].contiguous(), T=T.contiguous())


@pytest.mark.parametrize('dataset,cam,src', [('ddad', 0, None), ('ddad', 2, 'nbr'), ('nuscenes', 1, 'nbr')])
def test_projection_grid_bit_exact_and_gradients(dataset, cam, src):
    from vfdepth_b200.projection import Projection
    dev = _dev()
    p = _pair(dataset, 11, cam, src)
    B, _, Hh, Ww = p['depth'].shape
    proj = Projection(B, Hh, Ww, dev)
    d = {k: v.to(dev) for k, v in p.items()}
    depth = d['depth'].clone().requires_grad_(True)
    T = d['T'].clone().requires_grad_(True)
    K = d['K'].clone().requires_grad_(True)
    grid = proj(depth, T, d['invK'], K)
    dr = p['depth'].clone().requires_grad_(True)
    Tr = p['T'].clone().requires_grad_(True)
    Kr = p['K'].clone().requires_grad_(True)
    ref = orc.reproject_grid(dr, Tr, p['invK'], Kr)
    assert torch.equal(grid.cpu(), ref)
    # the two halves on their own
    pts = proj.backproject(d['invK'], d['depth'])
    homo = orc.homo_points(B, Hh, Ww)
    ref_pts = torch.cat([p['depth'].view(B, 1, -1) * torch.matmul(p['invK'][:, :3, :3], homo), homo[:, 2:3]], 1)
    assert torch.equal(pts.cpu(), ref_pts)
    assert torch.equal(proj.reproject(d['K'], pts, d['T']).cpu(), ref)
    gen = torch.Generator().manual_seed(1)
    up = torch.randn(ref.shape, generator=gen)
    got = torch.autograd.grad((grid * up.to(dev)).sum(), [depth, T, K])
    want = torch.autograd.grad((ref * up).sum(), [dr, Tr, Kr])
    for n, a, b in zip(('d_depth', 'd_T', 'd_K'), got, want):
        assert H.rel_err(a.cpu(), b) < TOL, (n, H.rel_err(a.cpu(), b))
    with pytest.raises(ValueError):
        proj(depth[:, :, :-1], T, d['invK'], K)


@pytest.mark.parametrize('dataset,cam,src', [('ddad', 0, None), ('ddad', 3, 'nbr'), ('nuscenes', 5, 'nbr')])
def test_get_virtual_image_matches_oracle(dataset, cam, src):
    from vfdepth_b200 import view
    dev = _dev()
    p = _pair(dataset, 23, cam, src)
    cfg = vcfg.default_config(dataset, batch_size=2, training__height=96, training__width=160)
    vr = view.ViewRendering(cfg, 0)
    d = {k: v.to(dev) for k, v in p.items()}
    depth = d['depth'].clone().requires_grad_(True)
    T = d['T'].clone().requires_grad_(True)
    img, mask = vr.get_virtual_image(d['src_img'], d['src_mask'], depth, d['invK'], d['K'], T)
    dr = p['depth'].clone().requires_grad_(True)
    Tr = p['T'].clone().requires_grad_(True)
    ref_img, ref_mask = orc.virtual_image(p['src_img'], p['src_mask'], dr, p['invK'], p['K'], Tr)
    assert torch.equal(mask.cpu(), ref_mask.float())
    assert float((img.cpu() - ref_img).abs().max()) < 1e-5
    assert 200 < float(mask.sum()) < mask.numel()   # the pair really overlaps (neighbouring cameras: a narrow strip)
    gen = torch.Generator().manual_seed(2)
    up = torch.randn(ref_img.shape, generator=gen)
    got = torch.autograd.grad((img * up.to(dev)).sum(), [depth, T])
    want = torch.autograd.grad((ref_img * up).sum(), [dr, Tr])
    # depth gradient: per pixel, compared on the scale of the whole map
    assert H.rel_err(got[0].cpu(), want[0]) < 1e-3
    assert H.rel_err(got[1].cpu(), want[1]) < 1e-3


def test_virtual_image_nan_and_behind_camera():
    """a NaN depth gives 2.0 / mask 0 (view_rendering.py:73-76); points behind the source camera are NOT rejected (the
    reference has no z > 0 test): whatever the projection yields is sampled."""
    from vfdepth_b200 import projection
    dev = _dev()
    p = _pair('ddad', 5, 0, None)
    depth = p['depth'].clone()
    depth[:, :, 10:12, 20:30] = float('nan')
    depth[:, :, 40:44, 50:60] = -3.0
    d = {k: v.to(dev) for k, v in p.items()}
    img, mask = projection.get_virtual_image(d['src_img'], d['src_mask'], depth.to(dev), d['invK'], d['K'], d['T'])
    ref_img, ref_mask = orc.virtual_image(p['src_img'], p['src_mask'], depth, p['invK'], p['K'], p['T'])
    assert torch.equal(mask.cpu(), ref_mask.float())
    assert float((img.cpu() - ref_img).abs().max()) < 1e-5
    assert float(img[:, :, 10:12, 20:30].min()) == 2.0 and float(mask[:, :, 10:12, 20:30].max()) == 0.0


@pytest.mark.parametrize('dataset,cam', [('ddad', 0), ('nuscenes', 4)])
def test_get_virtual_depth_matches_oracle(dataset, cam):
    """view_rendering.py:84-116 (forward-only sub-method): a neighbour's depth map seen from a slightly rotated copy of
    the camera, mask identical (up to 1e-4 of the pixels at the depth-range thresholds), depth within 1e-4."""
    from vfdepth_b200 import view
    from vfdepth_b200.rig import axis_angle_to_matrix
    dev = _dev()
    cfg = vcfg.default_config(dataset, batch_size=2, training__height=96, training__width=160,
                              training__focal_length_scale=75)
    inp, ex = make_rig(cfg, seed=31, with_feats=False)
    src = cfg['data']['rel_cam_list'][cam][0]
    tform = torch.eye(4).repeat(2, 1, 1)
    tform[:, :3, :3] = axis_angle_to_matrix(torch.tensor([[0.03, -0.02, 0.05], [-0.04, 0.01, 0.02]]))
    ext_aug = tform @ inp['extrinsics'][:, cam]
    T = torch.inverse(ext_aug) @ inp['extrinsics'][:, src]                 # source -> (novel) target
    t = cfg['training']
    args = (ex['depth'][src], inp['mask'][:, src], inp[('inv_K', 0)][:, src], inp[('K', 0)][:, src], ex['depth'][cam],
            inp[('inv_K', 0)][:, cam], inp[('K', 0)][:, cam], T)
    ref_d, ref_m = orc.virtual_depth(*[a.clone() for a in args], t['min_depth'], t['max_depth'])
    vr = view.ViewRendering(cfg, 0)
    got_d, got_m = vr.get_virtual_depth(*[a.to(dev) for a in args], t['min_depth'], t['max_depth'])
    assert got_d.shape == ref_d.shape and got_m.shape == ref_m.shape
    flips = got_m.cpu() != ref_m                         # the depth-range tests sit on a computed value
    assert float(flips.float().mean()) < 1e-4
    assert 200 < float(got_m.sum())
    assert H.rel_err(got_d.cpu(), ref_d) < TOL           # same tolerance as the fused form (tests/test_warp_gpu.py)


@pytest.mark.parametrize('empty', [False, True])
def test_get_norm_image_single_matches_oracle(empty):
    """view_rendering.py:39-59: aligned image and its gradient w.r.t. the warped image; with an empty overlap in ONE batch
    element the whole batch is returned unchanged (the reference's batch-wide early exit)."""
    from vfdepth_b200 import view
    dev = _dev()
    p = _pair('ddad', 41, 0, None)
    cfg = vcfg.default_config('ddad', batch_size=2, training__height=96, training__width=160)
    vr = view.ViewRendering(cfg, 0)
    ref_img = p['src_img'].flip(-1).contiguous() * 0.8 + 0.1       # a "target" image with different statistics
    ref_mask = p['src_mask']
    warp_img, warp_mask = orc.virtual_image(p['src_img'], p['src_mask'], p['depth'], p['invK'], p['K'], p['T'])
    warp_mask = warp_mask.float()
    if empty:
        warp_mask = warp_mask.clone()
        warp_mask[1] = 0
    wr = warp_img.clone().requires_grad_(True)
    ref = orc.norm_image_single(ref_img, ref_mask, wr, warp_mask)
    wg = warp_img.to(dev).requires_grad_(True)
    got = vr.get_norm_image_single(ref_img.to(dev), ref_mask.to(dev), wg, warp_mask.to(dev))
    if empty:
        assert torch.equal(got.cpu(), warp_img)
    assert H.rel_err(got.cpu(), ref.detach()) < TOL
    gen = torch.Generator().manual_seed(4)
    up = torch.randn(ref.shape, generator=gen)
    g_ref, = torch.autograd.grad((ref * up).sum(), wr)
    g_got, = torch.autograd.grad((got * up.to(dev)).sum(), wg)
    assert H.rel_err(g_got.cpu(), g_ref) < TOL


def test_photometric_free_functions_match_oracle():
    """loss_util.compute_ssim_loss / compute_photometric_loss (:43-78) on two images: values within 1e-5 absolute
    (fp32 noise floor of E[x^2] - mu^2), gradients w.r.t. pred within 1e-3 of the map's scale (the clamp and |.| kinks
    are decided on values that differ by an ulp)."""
    from vfdepth_b200 import loss_util
    dev = _dev()
    p = _pair('ddad', 61, 0, None)
    target = p['src_img']
    pred, _ = orc.virtual_image(p['src_img'], p['src_mask'], p['depth'], p['invK'], p['K'], p['T'])
    pr = pred.clone().requires_grad_(True)
    pg = pred.to(dev).requires_grad_(True)
    gen = torch.Generator().manual_seed(8)
    for fn_ref, fn_got, ch in ((orc.ssim_loss, loss_util.compute_ssim_loss, 3),
                               (orc.photometric_loss, loss_util.compute_photometric_loss, 1)):
        ref = fn_ref(pr, target)
        got = fn_got(pg, target.to(dev))
        assert got.shape == ref.shape == (2, ch, 96, 160)
        assert float((got.cpu() - ref).abs().max()) < 1e-5
        up = torch.randn(ref.shape, generator=gen)
        g_ref, = torch.autograd.grad((ref * up).sum(), pr)
        g_got, = torch.autograd.grad((got * up.to(dev)).sum(), pg)
        assert H.rel_err(g_got.cpu(), g_ref) < 1e-3


def test_loss_free_functions_match_reference_fixture():
    """loss_util.compute_auto_masks / compute_masked_loss / compute_edg_smooth_loss (:6-40) and the photometric pair on
    the fixture the reference itself produced (tests/golden/golden_loss_free.npz): the auto-mask bit-exact (ties and
    all), the masked-loss gradient bit-exact, scalars within 1e-6 relative (f64 sums against torch's fp32 tree),
    the smoothness gradient within 1e-5 of its scale (expf against the host's exp)."""
    from vfdepth_b200 import loss_util
    dev = _dev()
    g = {k: v.to(dev) for k, v in H.load_npz('golden_loss_free.npz').items()}
    assert torch.equal(loss_util.compute_auto_masks(g['reproj'], g['ident']), g['auto_mask'])
    assert torch.equal(loss_util.compute_auto_masks(g['reproj'], None), g['auto_mask_none'])
    # NaN is the smallest value of torch.argmin; the first one wins
    rp, idl = g['reproj'].clone(), g['ident'].clone()
    rp[0, 0, 0, :5] = float('nan')
    idl[1, 1, 3, :7] = float('nan')
    idl[0, 0, 0, 2] = float('nan')
    assert torch.equal(loss_util.compute_auto_masks(rp, idl).cpu(), orc.auto_masks(rp.cpu(), idl.cpu()))

    ll = g['photo'].clone().requires_grad_(True)
    ml = loss_util.compute_masked_loss(ll, g['mask'])
    assert ml.shape == g['masked'].shape
    assert abs(float(ml) - float(g['masked'])) < 1e-6 * abs(float(g['masked']))
    d_loss, = torch.autograd.grad(ml, ll)
    assert H.rel_err(d_loss.cpu(), g['d_loss'].cpu()) < 1e-6
    zero = loss_util.compute_masked_loss(g['photo'], torch.zeros_like(g['mask']))
    assert float(zero) == 0.0                                           # empty mask: 0 / 1e-8

    dd = g['disp'].clone().requires_grad_(True)
    sm = loss_util.compute_edg_smooth_loss(g['target'], dd)
    assert abs(float(sm) - float(g['smooth'])) < 1e-6 * abs(float(g['smooth']))
    d_disp, = torch.autograd.grad(sm * 3.0, dd)
    assert H.rel_err(d_disp.cpu(), 3.0 * g['d_disp'].cpu()) < 1e-5

    pr = g['pred'].clone().requires_grad_(True)
    photo = loss_util.compute_photometric_loss(pr, g['target'])
    assert float((photo - g['photo']).abs().max()) < 1e-5
    assert float((loss_util.compute_ssim_loss(pr, g['target']) - g['ssim']).abs().max()) < 1e-5
    with pytest.raises(ValueError):
        loss_util.compute_masked_loss(g['photo'].cpu(), g['mask'].cpu())


def test_loss_free_functions_full_size():
    """the three free functions at 8 x 384 x 640 against the oracle on the same seeded inputs."""
    from vfdepth_b200 import loss_util
    dev = _dev()
    gen = torch.Generator().manual_seed(77)
    B, Hh, Ww = 8, 384, 640
    rgb = torch.nn.functional.interpolate(torch.rand(B, 3, 48, 80, generator=gen), size=(Hh, Ww), mode='bilinear',
                                          align_corners=True)
    disp = torch.rand(B, 1, Hh, Ww, generator=gen).requires_grad_(True)
    rp, idl = torch.rand(B, 1, Hh, Ww, generator=gen), torch.rand(B, 2, Hh, Ww, generator=gen)
    mask = orc.auto_masks(rp, idl)
    assert torch.equal(loss_util.compute_auto_masks(rp.to(dev), idl.to(dev)).cpu(), mask)
    ref = orc.edg_smooth_loss_raw(rgb, disp) + orc.masked_loss(disp, mask)
    dg = disp.detach().to(dev).requires_grad_(True)
    got = loss_util.compute_edg_smooth_loss(rgb.to(dev), dg) + loss_util.compute_masked_loss(dg, mask.to(dev))
    assert abs(float(got) - float(ref)) < 1e-5 * abs(float(ref))
    g_ref, = torch.autograd.grad(ref, disp)
    g_got, = torch.autograd.grad(got, dg)
    assert H.rel_err(g_got.cpu(), g_ref) < 1e-5
