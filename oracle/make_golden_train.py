"""Generates the fixtures that pin the TIMED training step against the reference's own trainer:

  tests/golden/train_step_c3mini.npz   one dis_update + gen_update of imaginaire.trainers.spade.Trainer
                                       (trainers/base.py:594-666, trainers/spade.py:128-187) on CPU fp32: every loss
                                       term, gradients, the weights after the Adam steps, the EMA copy
  tests/golden/discriminator_c3mini.npz  imaginaire.discriminators.spade.Discriminator.forward: all outputs and
                                       feature lists, real and fake (discriminators/spade.py:73-117,
                                       fpse.py:87-132, multires_patch.py:250-313)
  tests/golden/generator_c3mini.npz    the C3-style generator (sync_batch, separate 5x5 projections, spectral
                                       SPADE convs) forward in train mode (generators/spade.py:401-493)
  tests/golden/adaptive_norm_*.npz     AdaptiveNorm (layers/activation_norm.py:22-106) forward and gradients

TEST INFRASTRUCTURE ONLY.  Runs only in the build container: imports the reference from /root/reference
through oracle/ref_shims.py.  What the shims change, none of it arithmetic of the path: apex.amp is its
pass-through form (fp32, loss scale 1), apex FusedAdam -> torch.optim.Adam (`fused_opt: False`), `.cuda()` is a
no-op, and torchvision's VGG-19 is built with seeded random weights instead of the ImageNet download (the test
rebuilds the same weights from the same seed; their digest is in the fixture).

The configuration is configs/spade_c3_mini.yaml: BASELINE configs[2]'s architecture and trainer settings at 16
base filters and 256x256.  Networks and inputs are rebuilt from seeds by the test (digests stored here), so the
fixtures hold results only.

    python -m oracle.make_golden_train
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import ref_shims  # noqa: E402
from oracle.make_golden import onehot_label, save, sd_arrays, state_dict_digest  # noqa: E402

CONFIG = os.path.join(ROOT, 'configs', 'spade_c3_mini.yaml')
VGG_SEED = 1234
SEED = 0


def seeded_vgg19():
    """torchvision vgg19 with its default (kaiming) initialisation under a fixed seed; the global RNG stream is
    left untouched."""
    from torchvision.models.vgg import vgg19      # the builder itself (main() replaces torchvision.models.vgg19)
    with torch.random.fork_rng():
        torch.manual_seed(VGG_SEED)
        return vgg19(weights=None)


def synthetic_batch(n=2, h=256, w=256, classes=35):
    """Labels: blocky one-hot map (what real label maps look like); images: U(-1, 1)."""
    label = onehot_label(5, n, classes, h, w, blocky=8)
    g = torch.Generator().manual_seed(6)
    images = torch.rand(n, 3, h, w, generator=g) * 2 - 1
    return {'label': label, 'images': images}


def style_noise(seed, n, dims):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(n, dims, generator=g)


def fix_style_noise(net_G_module, eps):
    """The style encoder draws eps = randn_like(std) (generators/spade.py:558); hold it to a known tensor."""
    enc = net_G_module.style_encoder
    orig = torch.randn_like

    def forward(input_x):
        torch.randn_like = lambda t, *a, **k: eps.to(t)
        try:
            return type(enc).forward(enc, input_x)
        finally:
            torch.randn_like = orig
    enc.forward = forward


def grads(net, prefix=''):
    return {prefix + n: p.grad.detach().clone() for n, p in net.named_parameters() if p.grad is not None}


def sub(t, limit=20000):
    """Fixtures stay small: tensors above `limit` elements are stored as a strided sample of their flattened
    form (tests/helpers.py subsample() applies the same rule)."""
    flat = t.reshape(-1)
    return flat[::-(-flat.numel() // limit)] if flat.numel() > limit else t


def pick(names, want):
    """A spread of parameter names: every name matching one of `want` (substring), first match each."""
    out = []
    for w in want:
        for n in names:
            if w in n and n not in out:
                out.append(n)
                break
    return out


def main():
    ref_shims.install()
    import torchvision
    torchvision.models.vgg19 = lambda pretrained=False, **kw: seeded_vgg19()   # no network: seeded weights
    from imaginaire.config import Config as RefConfig                           # reference
    from imaginaire.utils.trainer import get_model_optimizer_and_scheduler, get_trainer   # reference

    # The fork normalises FOUR image channels (utils/misc.py:233-234) and then keeps the first three
    # (losses/perceptual.py:97): with 3-channel images it raises.  A dummy fourth channel goes in and is
    # dropped by that slice; the three real channels see the reference's own arithmetic.
    import imaginaire.losses.perceptual as ref_perceptual
    ref_norm = ref_perceptual.apply_imagenet_normalization
    ref_perceptual.apply_imagenet_normalization = \
        lambda x: ref_norm(torch.cat([x, torch.zeros_like(x[:, :1])], 1) if x.shape[1] == 3 else x)

    cfg = RefConfig(CONFIG)
    cfg.gen_opt.fused_opt = False        # apex FusedAdam is not installable: torch.optim.Adam (same update rule)
    cfg.dis_opt.fused_opt = False
    cfg.trainer.amp = 'O0'
    cfg.local_rank = 0
    cfg.logdir = '/tmp/imx_golden_logdir'

    def build():
        net_G, net_D, opt_G, opt_D, sch_G, sch_D = get_model_optimizer_and_scheduler(cfg, seed=SEED)
        trainer = get_trainer(cfg, net_G, net_D, opt_G, opt_D, sch_G, sch_D, object(), None)
        net_G.train()
        net_D.train()
        return trainer, net_G, net_D, opt_G

    trainer, net_G, net_D, opt_G = build()
    G = trainer.net_G_module
    D = net_D.module
    out = {'g_sha256': np.array(state_dict_digest(G.state_dict())),
           'd_sha256': np.array(state_dict_digest(D.state_dict())),
           'vgg_sha256': np.array(state_dict_digest(seeded_vgg19().features.state_dict())),
           'g_keys': np.array(list(G.state_dict().keys())), 'd_keys': np.array(list(D.state_dict().keys()))}
    data = synthetic_batch()
    n, sdims = data['label'].shape[0], cfg.gen.style_dims
    eps_d, eps_g = style_noise(11, n, sdims), style_noise(12, n, sdims)
    net_G.train()
    net_D.train()

    # ------------------------------------------------------------------ forward-only fixtures (separate copies)
    import copy
    with torch.no_grad():
        G2 = copy.deepcopy(G).train()
        fix_style_noise(G2, eps_d)
        g_out = G2(data)
        save('generator_c3mini.npz', fake_images=g_out['fake_images'].half(), mu=g_out['mu'], logvar=g_out['logvar'],
             eps=eps_d, state_dict_sha256=out['g_sha256'],
             state_dict_after_sha256=np.array(state_dict_digest(G2.state_dict())))
        D2 = copy.deepcopy(D).train()
        gfk = torch.Generator().manual_seed(7)
        fake = torch.rand(n, 3, 256, 256, generator=gfk) * 2 - 1
        d_out = D2(data, {'fake_images': fake})
        arrays = {}
        for kind in ('real', 'fake'):
            for i, o in enumerate(d_out[kind + '_outputs']):
                arrays['%s_outputs.%d' % (kind, i)] = o
            for i, feats in enumerate(d_out[kind + '_features']):
                for j, f in enumerate(feats):
                    # fixtures stay small: fp16, and maps above 32x32 as a stride-2 subsample
                    arrays['%s_features.%d.%d' % (kind, i, j)] = (f[:, :, ::2, ::2] if f.shape[-1] > 32 else f).half()
        after = D2.state_dict()
        save('discriminator_c3mini.npz', state_dict_sha256=out['d_sha256'],
             state_dict_after_sha256=np.array(state_dict_digest(after)),
             **{k: v for k, v in sd_arrays(after, 'sd_after.').items() if k.endswith(('weight_u', 'weight_v'))},
             **arrays)

    # ------------------------------------------------------------------ one training iteration (train.py:76-79)
    g_before = {k: v.detach().clone() for k, v in G.state_dict().items()}
    d_before = {k: v.detach().clone() for k, v in D.state_dict().items()}
    fix_style_noise(G, eps_d)
    trainer.dis_update(data)
    for k, v in trainer.dis_losses.items():
        out['dis_loss.' + k] = v.detach().reshape(())
    d_grads = grads(D)
    d_after = {k: v.detach().clone() for k, v in D.state_dict().items()}
    g_mid = {k: v.detach().clone() for k, v in G.state_dict().items()}      # buffers moved (u, v, running stats)

    fix_style_noise(G, eps_g)
    trainer.gen_update(data)
    for k, v in trainer.gen_losses.items():
        out['seq_gen_loss.' + k] = v.detach().reshape(())
    out['eps_dis'], out['eps_gen'] = eps_d, eps_g
    out['seq_ema_num_updates'] = net_G.module.num_updates_tracked.detach().clone()

    # ------------------------------------------------------------------ gen_update alone, from the initial state
    # After its first Adam step (beta1 = 0: every weight moves by +-lr, whatever the size of its gradient) the
    # discriminator is a discontinuous function of its gradients' signs, so the generator losses measured through it
    # in the sequence above are comparable only loosely between two arithmetics.  The generator update is therefore
    # also recorded against the INITIAL discriminator, where it can be held tightly.
    trainer, net_G, net_D, opt_G = build()
    G = trainer.net_G_module
    g_before = {k: v.detach().clone() for k, v in G.state_dict().items()}
    g_mid = g_before
    fix_style_noise(G, eps_g)
    trainer.gen_update(data)
    for k, v in trainer.gen_losses.items():
        out['gen_loss.' + k] = v.detach().reshape(())
    g_grads = grads(G)
    g_after = {k: v.detach().clone() for k, v in G.state_dict().items()}
    ema_after = {k: v.detach().clone() for k, v in net_G.module.averaged_model.state_dict().items()}

    # every parameter: gradient norm and norm of the update; a spread of tensors in full
    for tag, gr, before, after in (('d', d_grads, d_before, d_after), ('g', g_grads, g_before, g_after)):
        names = list(gr.keys())
        out[tag + '_grad_names'] = np.array(names)
        out[tag + '_grad_norms'] = np.array([float(gr[k].norm()) for k in names], np.float64)
        out[tag + '_update_norms'] = np.array([float((after[k] - before[k]).norm()) for k in names], np.float64)
    full_d = pick(list(d_grads), ['discriminators.0.layer0', 'discriminators.0.layer2', 'discriminators.1.layer4',
                                  'discriminators.0.layer5.0.layers.conv.bias', 'fpse_discriminator.enc1',
                                  'fpse_discriminator.lat3', 'fpse_discriminator.final2', 'fpse_discriminator.output',
                                  'fpse_discriminator.seg.layers.conv.weight', 'fpse_discriminator.embedding'])
    full_g = pick(list(g_grads), ['style_encoder.enc_layers.0', 'style_encoder.fc_mu.layers.linear.weight', 'fc_0',
                                  'head_0.layers.conv.weight_orig', 'cbn_head_0.layers.norm.fc', 'head_1.conv_block_0.layers.conv.weight_orig',
                                  'head_1.conv_block_0.layers.norm.gammas.0', 'head_1.conv_block_0.layers.norm.mlps.0',
                                  'up_0a.conv_block_s.layers.conv.weight_orig', 'up_0a.conv_block_s.layers.norm.betas.0',
                                  'up_1b.conv_block_1.layers.conv.bias', 'up_2a.conv_block_0.layers.norm.norm.weight',
                                  'up_2b.conv_block_1.layers.conv.weight_orig', 'conv_img256.layers.conv.weight_orig',
                                  'conv_img256.layers.conv.bias'])
    for tag, names, gr, before, after in (('d', full_d, d_grads, d_before, d_after), ('g', full_g, g_grads, g_before, g_after)):
        out[tag + '_full_names'] = np.array(names)
        for k in names:
            out['%s_grad.%s' % (tag, k)] = sub(gr[k])
            out['%s_after.%s' % (tag, k)] = sub(after[k])
            out['%s_before.%s' % (tag, k)] = sub(before[k])
    # buffers after the iteration: power-iteration vectors of two layers, running statistics of two norms
    for k in pick(list(g_after), ['head_0.layers.conv.weight_u', 'up_2b.conv_block_1.layers.conv.weight_v',
                                  'cbn_head_0.layers.norm.norm.running_mean', 'up_1a.conv_block_0.layers.norm.norm.running_var']):
        out['g_buffer_after.' + k] = g_after[k]
        out['g_buffer_mid.' + k] = g_mid[k]
    # the EMA copy (beta = 0 during warm-up: the current weights, W / sigma for spectral layers; model_average.py:87-115)
    for k in pick(list(ema_after), ['head_0.layers.conv.weight', 'up_2b.conv_block_1.layers.conv.weight',
                                    'conv_img256.layers.conv.bias', 'up_0a.conv_block_0.layers.norm.norm.running_mean']):
        out['ema_after.' + k] = sub(ema_after[k])
    out['ema_num_updates'] = net_G.module.num_updates_tracked.detach().clone()
    out['opt_g_step'] = np.array(float(next(iter(opt_G.state.values()))['step']))
    save('train_step_c3mini.npz', **out)
    for k in sorted(out):
        if 'loss' in k:
            print('%-32s %.6f' % (k, float(out[k])))

    # ------------------------------------------------------------------ AdaptiveNorm (activation_norm.py:22-106)
    from imaginaire.layers.activation_norm import AdaptiveNorm
    cases = [('a', dict(num_features=32, cond_dims=24, weight_norm_type='spectral', separate_projection=False,
                        activation_norm_type='sync_batch', activation_norm_params=types.SimpleNamespace(affine=False)), True),
             ('b', dict(num_features=40, cond_dims=16, weight_norm_type='', separate_projection=True,
                        activation_norm_type='instance', activation_norm_params=types.SimpleNamespace(affine=True)), True),
             ('c', dict(num_features=16, cond_dims=8, weight_norm_type='spectral', separate_projection=False,
                        activation_norm_type='sync_batch', activation_norm_params=types.SimpleNamespace(affine=True)), False)]
    for tag, kw, training in cases:
        torch.manual_seed(40)
        m = AdaptiveNorm(**kw)
        for p_name, p in m.named_parameters():
            if 'norm.' in p_name:
                with torch.no_grad():
                    p.uniform_(0.5, 1.5) if p_name.endswith('weight') else p.uniform_(-0.5, 0.5)
        m.train(training)
        if not training:
            m.norm.running_mean.uniform_(-0.3, 0.3)
            m.norm.running_var.uniform_(0.5, 1.5)
        sd_before = {k: v.clone() for k, v in m.state_dict().items()}
        x = (torch.randn(3, kw['num_features'], 10, 14) * 1.5 + 0.3).requires_grad_(True)
        y = torch.randn(3, kw['cond_dims']).requires_grad_(True)
        gout = torch.randn(3, kw['num_features'], 10, 14)
        o = m(x, y)
        o.backward(gout)
        save('adaptive_norm_%s.npz' % tag, x=x, y=y, gout=gout, out=o, gx=x.grad, gy=y.grad, training=np.array(training),
             separate_projection=np.array(kw['separate_projection']), weight_norm_type=np.array(kw['weight_norm_type']),
             norm_kind=np.array(kw['activation_norm_type']), affine=np.array(kw['activation_norm_params'].affine),
             **sd_arrays(sd_before, 'sd.'), **sd_arrays(m.state_dict(), 'sd_after.'),
             **{'grad.' + n: p.grad for n, p in m.named_parameters() if p.grad is not None})


if __name__ == '__main__':
    main()
