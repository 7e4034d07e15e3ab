"""oracle/train_oracle.py (CPU restatement of the SPADE discriminator, the loss terms and one G+D iteration) against the
REFERENCE's own trainer: tests/golden/{discriminator,train_step}_c3mini.npz, written by oracle/make_golden_train.py from
the imported reference in fp32.  Networks and inputs are rebuilt from the same seeds (sha256 digests in the fixtures
prove the bits); fp32 against fp32, so the tolerances are accumulation-order sized."""
import os

import numpy as np
import pytest
import torch

from oracle import train_oracle as to
from tests import helpers as H

CONFIG = 'spade_c3_mini.yaml'


def _cfg():
    from imaginaire_b200.config import Config
    return Config(os.path.join(H.ROOT, 'configs', CONFIG))


def _seeded_state_dicts(cfg):
    """The reference builds G, then D, then initialises G, then D under one seed (utils/trainer.py:93-106); module
    construction and initialisation are host-side torch, so the native package's classes give the same bits on CPU."""
    from imaginaire_b200.discriminators.spade import Discriminator
    from imaginaire_b200.generators.spade import Generator
    from imaginaire_b200.utils.init_weight import weights_init
    torch.manual_seed(0)
    net_G = Generator(cfg.gen, cfg.data)
    net_D = Discriminator(cfg.dis, cfg.data)
    fn = weights_init(cfg.trainer.init.type, cfg.trainer.init.gain)
    net_G.apply(fn)
    net_D.apply(fn)
    return net_G.state_dict(), net_D.state_dict()


def _rel_l2(a, b):
    a, b = torch.as_tensor(a).double().reshape(-1), torch.as_tensor(b).double().reshape(-1)
    return float((a - b).norm() / (b.norm() + 1e-30))


@pytest.fixture(scope='module')
def setup():
    cfg = _cfg()
    sd_g, sd_d = _seeded_state_dicts(cfg)
    d = H.load_golden('train_step_c3mini.npz')
    assert H.state_dict_digest(sd_g) == str(d['g_sha256'])
    assert H.state_dict_digest(sd_d) == str(d['d_sha256'])
    vgg = H.seeded_vgg19_features().state_dict()
    assert H.state_dict_digest(vgg) == str(d['vgg_sha256'])
    spec = to.trainer_spec_from_cfg(cfg, (256, 256))
    return cfg, sd_g, sd_d, vgg, spec, d


def test_discriminator_oracle_vs_reference(setup):
    """discriminators/spade.py:73-117, fpse.py:87-132, multires_patch.py:250-313: every output and feature map of the
    real and the fake pass, and the spectral-norm vectors after the call's two power iterations."""
    cfg, _, sd_d, _, spec, _ = setup
    d = H.load_golden('discriminator_c3mini.npz')
    assert H.state_dict_digest(sd_d) == str(d['state_dict_sha256'])
    sd = {k: v.clone() for k, v in sd_d.items()}
    g = torch.Generator().manual_seed(7)
    fake = torch.rand(2, 3, 256, 256, generator=g) * 2 - 1
    with torch.no_grad():
        out = to.discriminator(sd, H.c3mini_batch(), fake, spec['dis'], True)
    for kind in ('real', 'fake'):
        outs, feats = out[kind + '_outputs'], out[kind + '_features']
        assert len(outs) == 5 and len(feats) == 2
        for i, o in enumerate(outs):
            ref = torch.from_numpy(d['%s_outputs.%d' % (kind, i)]).float()
            assert o.shape == ref.shape
            assert _rel_l2(o, ref) < 1e-5, (kind, i)
        for i, fl in enumerate(feats):
            assert len(fl) == cfg.dis.num_layers + 1
            for j, f in enumerate(fl):
                ref = torch.from_numpy(d['%s_features.%d.%d' % (kind, i, j)]).float()      # stored in fp16
                f = f[:, :, ::2, ::2] if f.shape[-1] > 32 else f
                assert f.shape == ref.shape
                assert _rel_l2(f, ref) < 1e-3, (kind, i, j)
    for k, v in d.items():
        if k.startswith('sd_after.'):
            assert H.rel_err(sd[k[len('sd_after.'):]], v) < 1e-5, k


def _check_update(tag, d, grads, before, after, lr):
    names = [str(n) for n in d[tag + '_grad_names']]
    for n, rn in zip(names, d[tag + '_grad_norms']):
        if rn > 1e-6:
            # gradients of ~1e-5 (a bias in front of a batch norm: analytically zero) are fp32 rounding residue of
            # sums of O(1) terms; they agree to their own noise level only
            tol = 2e-3 if rn > 1e-3 else 0.3
            assert abs(float(grads[n].norm()) / rn - 1) < tol, (n, float(grads[n].norm()), rn)
    for n in [str(x) for x in d[tag + '_full_names']]:
        ref_grad = d['%s_grad.%s' % (tag, n)]
        if np.abs(ref_grad).max() < 1e-6:
            continue        # a bias in front of a batch norm: the true gradient is zero, both sides hold rounding residue
        # the generator's gradients pass through 13 batch-normalised blocks whose backward subtracts nearly equal sums:
        # two fp32 evaluation orders differ by ~3e-3 at the first layers
        assert _rel_l2(H.subsample(grads[n]), ref_grad) < (2e-3 if tag == 'd' else 1e-2), n
        # first Adam step with beta1 = 0: w - lr * g / (|g| + eps); where the gradient is well above fp32 noise the
        # element moved by -lr * sign(g) exactly as in the reference
        ref_delta = (d['%s_after.%s' % (tag, n)] - d['%s_before.%s' % (tag, n)]).reshape(-1)
        ours_delta = H.subsample(after[n] - before[n]).reshape(-1).numpy()
        strong = np.abs(ref_grad.reshape(-1)) > 2e-2 * np.abs(ref_grad).max()
        assert np.abs(ours_delta[strong] - ref_delta[strong]).max() < 1e-2 * lr, n
        moved = np.abs(ref_delta) > 0.5 * lr           # elsewhere a gradient of rounding-residue size may step the other way
        assert np.mean(np.sign(ours_delta[moved]) == np.sign(ref_delta[moved])) > 0.98, n
        assert np.abs(ours_delta).max() <= lr * 1.001


def test_discriminator_update_oracle_vs_reference_trainer(setup):
    """trainers/base.py:635-666 + trainers/spade.py:168-187: loss terms, gradients, the Adam step."""
    cfg, sd_g, sd_d, _, spec, d = setup
    losses, g_mid, d_after, grads = to.dis_update(sd_g, sd_d, H.c3mini_batch(), spec, torch.from_numpy(d['eps_dis']), {})
    for k in ('GAN/fake', 'GAN/true', 'GAN', 'total'):
        assert abs(losses[k] - float(d['dis_loss.' + k])) <= 1e-5 * max(1.0, abs(float(d['dis_loss.' + k]))), k
    _check_update('d', d, grads, sd_d, d_after, cfg.dis_opt.lr)


def test_generator_update_oracle_vs_reference_trainer(setup):
    """trainers/base.py:594-633 + trainers/spade.py:128-166 from the initial state: the four loss terms and their
    weighted total, gradients, the Adam step, running statistics / spectral-norm vectors, the moving average."""
    cfg, sd_g, sd_d, vgg, spec, d = setup
    losses, g_after, _, grads = to.gen_update(sd_g, sd_d, vgg, H.c3mini_batch(), spec, torch.from_numpy(d['eps_gen']), {})
    for k in ('GAN', 'FeatureMatching', 'GaussianKL', 'Perceptual', 'total'):
        ref = float(d['gen_loss.' + k])
        assert abs(losses[k] - ref) <= 2e-5 * max(1.0, abs(ref)), (k, losses[k], ref)
    _check_update('g', d, grads, sd_g, g_after, cfg.gen_opt.lr)
    for k, v in d.items():
        if k.startswith('g_buffer_after.'):
            assert H.rel_err(g_after[k[len('g_buffer_after.'):]], v) < 1e-4, k
    # moving average with beta = 0 (warm-up): the current value, W / (u^T W v) for spectrally normalised filters
    ema0 = {k[len('ema_after.'):]: None for k in d if k.startswith('ema_after.')}
    ema = to.model_average_update({k: torch.zeros_like(g_after.get(k, g_after.get(k + '_orig'))) for k in ema0}, g_after, 0.0)
    for k in ema0:
        ref = d['ema_after.' + k]
        if np.abs(ref).max() > 0:
            # the filters moved by +-lr with the sign of every element's gradient: the few elements whose gradient is
            # fp32 rounding residue may have stepped the other way (2 lr / sigma apart); all others agree closely
            diff = np.abs(H.subsample(ema[k]).numpy() - ref)
            assert np.mean(diff <= 1e-4 * np.abs(ref).max()) > 0.97, k
            assert _rel_l2(H.subsample(ema[k]), ref) < 5e-2, k


def test_full_iteration_oracle_vs_reference_trainer(setup):
    """train.py:76-79: dis_update then gen_update through the discriminator that has just stepped, then the moving
    average.  The discriminator terms are held as tightly as above; the generator terms that pass through the stepped
    discriminator are compared loosely (its first Adam step moves every weight by +-lr, so elements whose gradient is
    fp32 rounding residue may step the other way in two fp32 evaluation orders), the others tightly."""
    cfg, sd_g, sd_d, vgg, spec, d = setup
    state = {'opt_d': {}, 'opt_g': {}, 'ema_updates': 0, 'ema_start': cfg.trainer.model_average_start_iteration,
             'ema_beta': cfg.trainer.model_average_beta,
             'ema': {k[len('ema_after.'):]: torch.zeros(1) for k in d if k.startswith('ema_after.')}}
    dl, gl, g_after, d_after, state = to.train_iteration(sd_g, sd_d, vgg, H.c3mini_batch(), spec,
                                                         torch.from_numpy(d['eps_dis']), torch.from_numpy(d['eps_gen']), state)
    for k in ('GAN/fake', 'GAN/true', 'GAN', 'total'):
        assert abs(dl[k] - float(d['dis_loss.' + k])) <= 1e-5 * max(1.0, abs(float(d['dis_loss.' + k]))), k
    for k in ('GAN', 'FeatureMatching', 'GaussianKL', 'Perceptual', 'total'):
        ref = float(d['seq_gen_loss.' + k])
        tol = 1e-3 if k in ('GaussianKL', 'Perceptual') else 5e-2
        assert abs(gl[k] - ref) <= tol * max(abs(ref), 1e-3), (k, gl[k], ref)
    assert state['ema_updates'] == int(d['seq_ema_num_updates'])
    # the generator's buffers moved twice (dis_update's forward, gen_update's forward); its parameters once
    moved = [k for k in sd_g if to.is_parameter(k) and not torch.equal(sd_g[k], g_after[k])]
    assert len(moved) > 0.9 * sum(to.is_parameter(k) for k in sd_g)
    assert all(float((g_after[k] - sd_g[k]).abs().max()) <= cfg.gen_opt.lr * 1.001 for k in moved)
