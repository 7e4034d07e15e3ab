"""GPU parity of configurations beyond the two golden ones, against the CPU oracle on the
same seeded weights and inputs: the 512-pixel generator branch (BASELINE configs[4] shape of
network: up_3a / up_3b / conv_img512, tanh of the summed to-RGB outputs) in training mode
with sync-batch norms, then in eval mode on the running statistics it just accumulated; and
multi-SPADE conditioning as the vid2vid generators use it (two conditioning maps, 1x1
projection without hidden layer, one of them optionally absent)."""
import pytest
import torch

from oracle import spade_oracle as so
from tests import helpers as H

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def _tiny_512_cfg():
    from imaginaire_b200.config import AttrDict
    gen = AttrDict(type='imaginaire.generators.spade', style_dims=32, num_filters=16, kernel_size=3,
                   weight_norm_type='spectral', global_adaptive_norm_type='sync_batch',
                   use_posenc_in_input_layer=False,
                   activation_norm_params=AttrDict(num_filters=16, kernel_size=3, separate_projection=True,
                                                   activation_norm_type='sync_batch'),
                   style_enc=AttrDict(num_filters=8, kernel_size=3))
    data = AttrDict(input_types=[AttrDict(images=AttrDict(num_channels=3)),
                                 AttrDict(seg_maps=AttrDict(num_channels=7, use_dont_care=True))],
                    input_image=['images'], input_labels=['seg_maps'],
                    train=AttrDict(augmentations=AttrDict(random_crop_h_w='512, 512')))
    return gen, data


def test_generator_512_branch_train_then_eval_vs_oracle():
    from imaginaire_b200.generators.spade import Generator
    from imaginaire_b200.utils.init_weight import weights_init
    gen_cfg, data_cfg = _tiny_512_cfg()
    torch.manual_seed(7)
    net = Generator(gen_cfg, data_cfg)
    net.apply(weights_init('xavier', 0.02))
    assert hasattr(net.spade_generator, 'up_3a') and net.spade_generator.base == 32
    sd0 = {k: v.clone() for k, v in net.state_dict().items()}
    label = H.onehot_label(71, 2, 8, 512, 512, blocky=8)
    images = torch.rand(2, 3, 512, 512, generator=torch.Generator().manual_seed(72)) * 2 - 1
    eps = torch.randn(2, 32, generator=torch.Generator().manual_seed(73))
    spec = so.generator_spec_from_cfg(gen_cfg, (512, 512))

    # ---- training mode: batch statistics, running-stat update, one power iteration
    ref = so.generator(sd0, {'label': label, 'images': images}, spec, True, eps)
    so.QUANT = lambda t: t.bfloat16().float()
    try:
        emu = so.generator(sd0, {'label': label, 'images': images}, spec, True, eps)
    finally:
        so.QUANT = None
    emu_mean = float((emu['fake_images'] - ref['fake_images']).abs().mean())
    emu_max = float((emu['fake_images'] - ref['fake_images']).abs().max())
    net = net.to(DEV).train()
    net.style_encoder.freeze_random = True
    net.style_encoder.eps = eps.to(DEV)
    with torch.no_grad():
        out = net({'label': label.to(DEV), 'images': images.to(DEV)})
    img = out['fake_images'].cpu()
    assert img.shape == (2, 3, 512, 512)
    diff = (img - ref['fake_images']).abs()
    print('512 branch train: max %.4f mean %.5f (bf16-operand oracle: max %.4f mean %.5f)' % (float(diff.max()), float(diff.mean()), emu_max, emu_mean))
    assert float(diff.mean()) <= max(2.5 * emu_mean, 2e-3) and float(diff.max()) <= max(3.0 * emu_max, 2e-2)
    sd1 = net.state_dict()
    for k, v in ref['state_dict'].items():
        if k.endswith(('running_mean', 'running_var')):
            assert H.rel_err(sd1[k].cpu(), v) < 3e-2, k
        if k.endswith(('weight_u', 'weight_v')) and 'spade_generator' in k and '.fc' not in k:
            assert H.rel_err(sd1[k].cpu(), v) < 1e-3, k

    # ---- eval mode on the oracle's updated state: running statistics, no power iteration
    sd_eval = {k: v.clone() for k, v in ref['state_dict'].items()}
    ref_eval = so.generator(sd_eval, {'label': label, 'images': images}, spec, False, eps)
    net.load_state_dict(sd_eval, strict=True)
    net.eval()
    with torch.no_grad():
        out_eval = net({'label': label.to(DEV), 'images': images.to(DEV)})
    diff = (out_eval['fake_images'].cpu() - ref_eval['fake_images']).abs()
    print('512 branch eval: max %.4f mean %.5f' % (float(diff.max()), float(diff.mean())))
    assert float(diff.mean()) <= max(2.5 * emu_mean, 2e-3) and float(diff.max()) <= max(3.0 * emu_max, 2e-2)


@pytest.mark.parametrize('second_absent', [False, True])
@pytest.mark.parametrize('grad', [False, True])
def test_multi_spade_conditioning_vs_oracle(second_absent, grad):
    from imaginaire_b200 import functional as NF, ops
    from imaginaire_b200.layers import SpatiallyAdaptiveNorm
    torch.manual_seed(9)
    m = SpatiallyAdaptiveNorm(64, [16, 24], num_filters=0, kernel_size=1, weight_norm_type='none',
                              separate_projection=False, activation_norm_type='instance')
    sd = {k: v.clone() for k, v in m.state_dict().items()}
    g = torch.Generator().manual_seed(10)
    x = torch.randn(2, 64, 12, 20, generator=g) * 1.3 + 0.2
    c1 = torch.randn(2, 16, 24, 40, generator=g)
    c2 = None if second_absent else torch.randn(2, 24, 12, 20, generator=g)
    p = dict(kernel_size=1, separate_projection=False, num_filters=0, norm_kind='instance')
    ref = so.leaky_relu(so.spatially_adaptive_norm(sd, '', x, [c1, c2], p, True))
    m = m.to(DEV).train()
    with torch.set_grad_enabled(grad):
        y = m.forward_fused(x.to(DEV).requires_grad_(grad), c1.to(DEV), None if c2 is None else c2.to(DEV),
                            act=ops.ACT_LRELU)
    y = NF.to_nchw(y).detach().cpu()
    assert H.rel_err(y, ref) < 3e-2
    assert float((y - ref).abs().mean() / ref.abs().mean()) < 8e-3
