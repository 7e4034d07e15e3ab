"""The timed training step against the REFERENCE's own trainer, discriminator, generator and AdaptiveNorm.

Fixtures: tests/golden/{train_step,discriminator,generator}_c3mini.npz and adaptive_norm_*.npz, produced by
oracle/make_golden_train.py from the imported reference (fp32, CPU) on configs/spade_c3_mini.yaml -- BASELINE
configs[2]'s architecture (sync batch norm, separate 5x5 SPADE projections, spectral norm everywhere, SPADE
discriminator = FPSE + 2 patch discriminators, hinge + feature matching + KL + VGG-19 perceptual loss, Adam, EMA) at
16 base filters and 256x256.  Networks, inputs and the (seeded random) VGG-19 are rebuilt here from the same seeds;
sha256 digests in the fixtures prove they are the same bits.

The native path computes in bf16 storage with fp32 accumulation, the reference in fp32: tolerances are bf16-sized and
written beside each assertion.
"""
import os

import numpy as np
import pytest
import torch

from tests import helpers as H

pytestmark = pytest.mark.gpu
CONFIG = 'spade_c3_mini.yaml'


def _dev():
    assert torch.cuda.is_available(), 'GPU tests need a CUDA device'
    return torch.device('cuda:0')


def _cfg():
    from imaginaire_b200.config import Config
    return Config(os.path.join(H.ROOT, 'configs', CONFIG))


def _rel_l2(a, b):
    a, b = torch.as_tensor(a).double().reshape(-1), torch.as_tensor(b).double().reshape(-1)
    return float((a - b).norm() / (b.norm() + 1e-30))


def _cos(a, b):
    a, b = torch.as_tensor(a).double().reshape(-1), torch.as_tensor(b).double().reshape(-1)
    return float((a * b).sum() / (a.norm() * b.norm() + 1e-30))


def _fix_eps(net_G_module, eps):
    enc = net_G_module.style_encoder
    enc.freeze_random = True
    enc.eps = eps.to(_dev())


def _build_trainer():
    """get_model_optimizer_and_scheduler + get_trainer exactly as train.py:45-60 calls them."""
    from imaginaire_b200.utils.trainer import get_model_optimizer_and_scheduler, get_trainer
    cfg = _cfg()
    net_G, net_D, opt_G, opt_D, sch_G, sch_D = get_model_optimizer_and_scheduler(cfg, seed=0)
    trainer = get_trainer(cfg, net_G, net_D, opt_G, opt_D, sch_G, sch_D, (), None)
    vgg = H.seeded_vgg19_features()
    trainer.criteria['Perceptual'].model.load_torchvision_state_dict(vgg.state_dict())
    return cfg, trainer, vgg


def test_seeded_networks_are_the_references_bits():
    """Same seed -> same construction order -> bit-identical initial weights as the reference trainer's networks
    (and the same VGG-19), so every comparison below starts from identical parameters."""
    d = H.load_golden('train_step_c3mini.npz')
    cfg, trainer, vgg = _build_trainer()
    G, D = trainer.net_G_module, trainer.net_D.module
    assert list(G.state_dict().keys()) == [str(k) for k in d['g_keys']]
    assert list(D.state_dict().keys()) == [str(k) for k in d['d_keys']]
    assert H.state_dict_digest(G.state_dict()) == str(d['g_sha256'])
    assert H.state_dict_digest(D.state_dict()) == str(d['d_sha256'])
    assert H.state_dict_digest(vgg.state_dict()) == str(d['vgg_sha256'])


def test_c3_style_generator_forward_vs_reference():
    """generators/spade.py:401-493 with sync_batch + separate 5x5 projections + spectral SPADE convs (the
    headline architecture), train mode, grad and no-grad paths."""
    from imaginaire_b200.generators.spade import Generator
    from imaginaire_b200.utils.init_weight import weights_init
    from oracle import spade_oracle as so
    d = H.load_golden('generator_c3mini.npz')
    cfg = _cfg()
    ref = torch.from_numpy(d['fake_images']).float()
    data = {k: v.to(_dev()) for k, v in H.c3mini_batch().items()}
    from imaginaire_b200.discriminators.spade import Discriminator
    emu = None
    for grad in (False, True):
        # the reference trainer builds G, then D, then initialises G (utils/trainer.py:93-106): same draws here
        torch.manual_seed(0)
        net = Generator(cfg.gen, cfg.data)
        Discriminator(cfg.dis, cfg.data)
        net.apply(weights_init(cfg.trainer.init.type, cfg.trainer.init.gain))
        assert H.state_dict_digest(net.state_dict()) == str(d['state_dict_sha256'])
        if emu is None:
            # what bf16 storage itself costs on this network: the fp32 oracle with every tensor the pipeline keeps in
            # HBM rounded to bf16 (profiles/r02_generator_error_trace.md has the same block by block)
            bf = (lambda t: t.bfloat16().float())
            so.QUANT, so.STORE = bf, bf
            try:
                emu = so.generator(net.state_dict(), H.c3mini_batch(), so.generator_spec_from_cfg(cfg.gen, (256, 256)),
                                   True, torch.from_numpy(d['eps']))['fake_images']
            finally:
                so.QUANT, so.STORE = None, None
            emu_diff = (emu - ref).abs()
            emu_max, emu_mean = float(emu_diff.max()), float(emu_diff.mean())
            print('bf16-storage oracle vs fp32 reference: max-abs %.4f mean-abs %.5f' % (emu_max, emu_mean))
        net = net.to(_dev()).train()
        _fix_eps(net, torch.from_numpy(d['eps']))
        with torch.set_grad_enabled(grad):
            out = net(dict(data))
        img = out['fake_images'].float().cpu()
        diff = (img - ref).abs()
        print('c3-style generator grad=%s: max-abs %.4f mean-abs %.5f (ref mean-abs %.4f)'
              % (grad, float(diff.max()), float(diff.mean()), float(ref.abs().mean())))
        assert H.rel_err(out['mu'].cpu(), d['mu']) < 3e-2 and H.rel_err(out['logvar'].cpu(), d['logvar']) < 3e-2
        # Stated bf16 tolerance after tanh: mean-abs within 1.1x, max-abs (one pixel out of 393 216) within 1.5x of what
        # bf16 storage alone costs on this randomly initialised 13-block network, and absolute caps.  north_star's
        # example figures (2e-2 / 2e-3) are below that floor: operand rounding alone gives 7.3e-2 / 6.8e-3.
        assert float(diff.mean()) <= 1.1 * emu_mean and float(diff.max()) <= 1.5 * emu_max
        assert float(diff.max()) <= 0.15 and float(diff.mean()) <= 1.2e-2


def test_discriminator_forward_vs_reference():
    """discriminators/spade.py:73-117, fpse.py:87-132, multires_patch.py:250-313: every output and every feature
    map, real and fake pass, plus the spectral-norm vectors after the two power iterations."""
    from imaginaire_b200.discriminators.spade import Discriminator
    from imaginaire_b200.utils.init_weight import weights_init
    d = H.load_golden('discriminator_c3mini.npz')
    cfg = _cfg()
    from imaginaire_b200.generators.spade import Generator
    # the reference builds G, then D, then initialises G, then D under one seed (utils/trainer.py:93-106): D's
    # weights depend on the draws G consumed before it
    torch.manual_seed(0)
    net_G = Generator(cfg.gen, cfg.data)
    net = Discriminator(cfg.dis, cfg.data)
    fn = weights_init(cfg.trainer.init.type, cfg.trainer.init.gain)
    net_G.apply(fn)
    net.apply(fn)
    assert H.state_dict_digest(net.state_dict()) == str(d['state_dict_sha256'])
    net = net.to(_dev()).train()
    data = {k: v.to(_dev()) for k, v in H.c3mini_batch().items()}
    g = torch.Generator().manual_seed(7)
    fake = (torch.rand(2, 3, 256, 256, generator=g) * 2 - 1).to(_dev())
    with torch.no_grad():
        out = net(data, {'fake_images': fake})
    worst_o, worst_f = 0.0, 0.0
    for kind in ('real', 'fake'):
        outs, feats = out[kind + '_outputs'], out[kind + '_features']
        assert len(outs) == 5 and len(feats) == 2          # 3 FPSE heads + 2 patch discriminators
        for i, o in enumerate(outs):
            ref = torch.from_numpy(d['%s_outputs.%d' % (kind, i)]).float()
            o = o.float().cpu()
            assert o.shape == ref.shape, (kind, i, o.shape, ref.shape)
            worst_o = max(worst_o, _rel_l2(o, ref))
        for i, fl in enumerate(feats):
            assert len(fl) == cfg.dis.num_layers + 1      # every layer's activation but the last (multires_patch.py:305-313)
            for j, f in enumerate(fl):
                ref = torch.from_numpy(d['%s_features.%d.%d' % (kind, i, j)]).float()
                f = f.float().cpu()
                f = f[:, :, ::2, ::2] if f.shape[-1] > 32 else f
                assert f.shape == ref.shape, (kind, i, j, f.shape, ref.shape)
                worst_f = max(worst_f, _rel_l2(f, ref))
    print('discriminator: worst relative L2 error outputs %.4f features %.4f' % (worst_o, worst_f))
    assert worst_o < 2e-2 and worst_f < 2e-2               # bf16 storage through <= 6 layers
    sd = net.state_dict()
    for k, v in d.items():
        if k.startswith('sd_after.'):
            assert H.rel_err(sd[k[len('sd_after.'):]].cpu(), v) < 2e-3, k


def test_one_training_iteration_vs_reference_trainer():
    """dis_update then gen_update (train.py:76-79; trainers/base.py:594-666, trainers/spade.py:128-187): every loss
    term, gradients of every parameter (norms) and of a spread of tensors (direction), the weights after the two
    Adam steps, the EMA copy, spectral-norm vectors and running statistics."""
    d = H.load_golden('train_step_c3mini.npz')
    cfg, trainer, _ = _build_trainer()
    G, D = trainer.net_G_module, trainer.net_D.module
    data = trainer.start_of_iteration(H.c3mini_batch(), current_iteration=1)

    # ---- discriminator update
    _fix_eps(G, torch.from_numpy(d['eps_dis']))
    d_before = {k: v.detach().clone() for k, v in D.state_dict().items()}
    trainer.dis_update(data)
    for k in ('GAN/fake', 'GAN/true', 'GAN', 'total'):
        ours, ref = float(trainer.dis_losses[k]), float(d['dis_loss.' + k])
        print('dis %-16s ours %.6f reference %.6f' % (k, ours, ref))
        assert abs(ours - ref) <= 2e-3 * max(1.0, abs(ref)), k
    _check_grads_and_update('d', d, dict(D.named_parameters()), d_before, D.state_dict(), lr=cfg.dis_opt.lr)

    # ---- generator update, measured through the discriminator that has just stepped
    _fix_eps(G, torch.from_numpy(d['eps_gen']))
    trainer.gen_update(data)
    for k in ('GAN', 'FeatureMatching', 'GaussianKL', 'Perceptual', 'total'):
        ours, ref = float(trainer.gen_losses[k]), float(d['seq_gen_loss.' + k])
        print('gen (after D step) %-16s ours %.6f reference %.6f' % (k, ours, ref))
        # First Adam step with beta1 = 0: every discriminator weight moved by +-lr whatever the size of its gradient,
        # so the stepped discriminator depends discontinuously on the signs of gradients that are bf16 noise in one
        # arithmetic and fp32 noise in the other (and the step, 4e-4, is the size of the freshly initialised weights
        # themselves).  Terms measured through it are reported, not asserted -- they also move from run to run of
        # the CUDA path itself (unordered atomics); the terms that do not involve it agree tightly.  test_generator_update_vs_reference_trainer holds the same update tightly
        # against the initial discriminator.
        if k in ('GaussianKL', 'Perceptual'):
            assert abs(ours - ref) <= 5e-2 * max(abs(ref), 1e-3) + 1e-4, k
        else:
            assert np.isfinite(ours), k
    assert int(trainer.net_G.module.num_updates_tracked) == int(d['seq_ema_num_updates'])


def test_generator_update_vs_reference_trainer():
    """gen_update from the initial state (trainers/base.py:594-633, trainers/spade.py:128-166): every loss term, the
    gradient of every generator parameter (norm) and of a spread of tensors (direction), the weights after the Adam
    step, the EMA copy (utils/model_average.py:87-115), spectral-norm vectors and running statistics."""
    d = H.load_golden('train_step_c3mini.npz')
    cfg, trainer, _ = _build_trainer()
    G = trainer.net_G_module
    data = trainer.start_of_iteration(H.c3mini_batch(), current_iteration=1)
    _fix_eps(G, torch.from_numpy(d['eps_gen']))
    g_before = {k: v.detach().clone() for k, v in G.state_dict().items()}
    trainer.gen_update(data)
    for k in ('GAN', 'FeatureMatching', 'GaussianKL', 'Perceptual', 'total'):
        ours, ref = float(trainer.gen_losses[k]), float(d['gen_loss.' + k])
        print('gen %-16s ours %.6f reference %.6f' % (k, ours, ref))
        assert abs(ours - ref) <= 2e-2 * abs(ref) + 2e-4, k
    _check_grads_and_update('g', d, dict(G.named_parameters()), g_before, G.state_dict(), lr=cfg.gen_opt.lr)
    sd = G.state_dict()
    for k, v in d.items():
        if k.startswith('g_buffer_after.'):
            name = k[len('g_buffer_after.'):]
            tol = 2e-2 if 'running' in name else 5e-3
            assert H.rel_err(sd[name].cpu(), v) < tol, name
    # The EMA copy (utils/model_average.py:87-115; beta = 0 during warm-up): exactly the reference's rule applied to
    # OUR weights after the step -- the current value, W / (u^T W v) for spectrally normalised filters -- and close to
    # the reference's copy (its weights moved by +-lr with the signs of ITS gradients, which differ where a gradient
    # is rounding noise: up to ~10 % of the elements of the earliest layers).
    ema = trainer.net_G.module.averaged_model.state_dict()
    for k, v in d.items():
        if k.startswith('ema_after.'):
            name = k[len('ema_after.'):]
            if name + '_orig' in sd:
                w, u, vv = sd[name + '_orig'], sd[name + '_u'], sd[name + '_v']
                expect = w / torch.sum(u * torch.mv(w.reshape(w.shape[0], -1), vv))
            else:
                expect = sd[name]
            assert _rel_l2(ema[name].float().cpu(), expect.float().cpu()) < 1e-3, name
            assert _rel_l2(H.subsample(ema[name].float().cpu()), v) < 0.15, name
    assert int(trainer.net_G.module.num_updates_tracked) == int(d['ema_num_updates'])
    assert float(next(iter(trainer.opt_G.state.values()))['step']) == float(d['opt_g_step'])


def test_graphed_iteration_equals_the_eager_iteration():
    """utils/cuda_graph.py: dis_update + gen_update captured once and replayed.  Same seeds and batches through an
    eager trainer and a graphed one, with learning rates of 1e-7 so that the two runs stay on the same trajectory
    (at the configured rates the first sign-like Adam steps make a GAN iteration chaotic, see above): every loss
    term of the three replayed iterations agrees with the eager ones, the weights MOVE in the replays and move the
    same way, Adam's device step counters and the EMA counters advance."""
    from imaginaire_b200.utils.cuda_graph import GraphedTrainerStep
    d = H.load_golden('train_step_c3mini.npz')
    batch = {k: v.to(_dev()) for k, v in H.c3mini_batch().items()}
    names = ['discriminators.0.layer2.0.layers.conv.weight_orig', 'fpse_discriminator.enc1.layers.conv.weight_orig']
    g_names = ['spade_generator.up_2b.conv_block_1.layers.conv.weight_orig', 'spade_generator.conv_img256.layers.conv.weight_orig']
    results = []
    for graphed in (False, True):
        cfg, trainer, _ = _build_trainer()
        for opt in (trainer.opt_G, trainer.opt_D):
            for group in opt.param_groups:
                group['lr'] = 1e-7
        trainer.net_G.train(); trainer.net_D.train()
        G, D = trainer.net_G_module, trainer.net_D.module
        _fix_eps(G, torch.from_numpy(d['eps_gen']))
        step = trainer.step
        trainer.step(dict(batch))                              # one eager iteration either way (lazy set-up)
        if graphed:
            step = GraphedTrainerStep(trainer, batch)
        rows, deltas = [], []
        for _ in range(3):
            before = {n: dict(D.named_parameters())[n].detach().clone() for n in names}
            before.update({n: dict(G.named_parameters())[n].detach().clone() for n in g_names})
            step(dict(batch))
            rows.append({k: float(v) for k, v in list(trainer.dis_losses.items()) + [('gen/' + k, v) for k, v in trainer.gen_losses.items()]})
            now = dict(D.named_parameters()); now.update(dict(G.named_parameters()))
            deltas.append({n: (now[n].detach() - before[n]).float().cpu() for n in before})
        results.append((rows, deltas, float(next(iter(trainer.opt_G.state.values()))['step']),
                        int(trainer.net_G.module.num_updates_tracked), trainer.net_G.module._steps))
        if graphed:
            assert step.launches_per_replay > 500
    (eager, e_delta, e_step, e_ema, e_host), (graph, g_delta, g_step, g_ema, g_host) = results
    assert (e_step, e_ema, e_host) == (g_step, g_ema, g_host) == (4.0, 4, 4)
    for it, (a, b) in enumerate(zip(eager, graph)):
        for k in a:
            print('iteration %d %-20s eager %.6f graph %.6f' % (it + 2, k, a[k], b[k]))
            assert abs(a[k] - b[k]) <= 1e-2 * max(abs(a[k]), 1e-3), (it, k)
    for it, (a, b) in enumerate(zip(e_delta, g_delta)):
        for n in a:
            assert float(b[n].abs().max()) > 0, n               # the replay did update the weights
            print('iteration %d update cosine %.4f %s' % (it + 2, _cos(a[n], b[n]), n))
            assert _cos(a[n], b[n]) > 0.9, (it, n)


def _check_grads_and_update(tag, d, params, before, after, lr):
    names = [str(n) for n in d[tag + '_grad_names']]
    ref_norms = d[tag + '_grad_norms']
    missing = [n for n in names if params[n].grad is None]
    assert not missing, 'no gradient for %s' % missing[:5]
    ratios = []
    for n, rn in zip(names, ref_norms):
        if rn > 1e-6:      # below that the reference's own gradient is rounding residue (e.g. the hinge terms' +1 / -1 on a bias)
            ratios.append(float(params[n].grad.float().norm()) / rn)
            if not 0.8 < ratios[-1] < 1.25:
                print('%s: gradient norm of %s: ours %.4e reference %.4e' % (tag, n, ratios[-1] * rn, rn))
    ratios = np.array(ratios)
    print('%s: gradient-norm ratio ours/reference over %d tensors: median %.4f min %.4f max %.4f'
          % (tag, len(ratios), np.median(ratios), ratios.min(), ratios.max()))
    assert abs(np.median(ratios) - 1) < 2e-2 and ratios.min() > 0.8 and ratios.max() < 1.25
    worst, bad = 1.0, []
    for n in [str(x) for x in d[tag + '_full_names']]:
        ours = H.subsample(params[n].grad.float().cpu())
        ref_grad = d['%s_grad.%s' % (tag, n)]
        if np.abs(ref_grad).max() < 1e-6:
            # a bias whose effect the following batch norm removes again: the true gradient is zero and both sides
            # hold rounding residue
            continue
        cos = _cos(ours, ref_grad)
        worst = min(worst, cos)
        print('%s: gradient cosine %.5f  %s' % (tag, cos, n))
        # bf16 storage of the activation gradients costs direction on the way back through the 13 blocks: the fp32
        # oracle with its stored tensors and their gradients rounded to bf16 loses the same amount
        # (profiles/r02_generator_grad_trace.md, CUDA vs emulation parameter by parameter: 0.97 at the head for a
        # linear loss; through the GAN / feature-matching / VGG losses 0.93-0.94, and 0.85-0.90 run to run for
        # 16-element bias vectors)
        floor = 0.98 if tag == 'd' else (0.9 if ref_grad.size > 64 else 0.7)
        bad += [] if cos > floor else [(n, cos)]
        # after the first Adam step (beta1 = 0): w - lr * g / (|g| + eps); elements whose gradient is far above the
        # bf16 noise moved by -lr * sign(g) exactly as in the reference
        ref_after, ref_before = d['%s_after.%s' % (tag, n)], d['%s_before.%s' % (tag, n)]
        ours_delta = H.subsample((after[n] - before[n]).float().cpu()).reshape(-1).numpy()
        ref_delta = (ref_after - ref_before).reshape(-1)
        g_ref = d['%s_grad.%s' % (tag, n)].reshape(-1)
        strong = np.abs(g_ref) > 0.5 * np.abs(g_ref).max()
        agree = np.mean(np.sign(ours_delta[strong]) == np.sign(ref_delta[strong]))
        # (generator: a few of the largest-gradient elements of a 16-256 element bias sit within the bf16 noise of zero
        # in one arithmetic or the other; 0.86-1.0 from run to run)
        bad += [] if agree > (0.99 if tag == 'd' else 0.8) else [(n, 'update sign agreement %.4f' % agree)]
        assert np.abs(ours_delta).max() <= lr * 1.001
    print('%s: worst gradient cosine over the fully stored tensors %.5f' % (tag, worst))
    assert not bad, bad


@pytest.mark.parametrize('tag', ['a', 'b', 'c'])
def test_adaptive_norm_vs_reference(tag):
    """layers/activation_norm.py:22-106: forward, gradients of x / y / parameters, running statistics; sync_batch and
    instance inner norms, joint and separate projections, spectral and plain linear layers."""
    from types import SimpleNamespace
    from imaginaire_b200 import functional as NF
    from imaginaire_b200.layers.activation_norm import AdaptiveNorm
    d = H.load_golden('adaptive_norm_%s.npz' % tag)
    x0, y0 = torch.from_numpy(d['x']), torch.from_numpy(d['y'])
    m = AdaptiveNorm(x0.shape[1], y0.shape[1], weight_norm_type=str(d['weight_norm_type']),
                     separate_projection=bool(d['separate_projection']), activation_norm_type=str(d['norm_kind']),
                     activation_norm_params=SimpleNamespace(affine=bool(d['affine'])))
    m.load_state_dict(H.split_sd(d, 'sd.'), strict=True)
    m = m.to(_dev()).train(bool(d['training']))
    x = x0.to(_dev()).requires_grad_(True)
    y = y0.to(_dev()).requires_grad_(True)
    out = NF.to_nchw(m(x, y), x0.shape[1])
    ref = torch.from_numpy(d['out'])
    assert H.rel_err(out.detach().cpu(), ref) < 1e-2                   # bf16 storage of x and of the output
    out.backward(torch.from_numpy(d['gout']).to(_dev()))
    assert _rel_l2(x.grad.cpu(), d['gx']) < 2e-2
    assert _rel_l2(y.grad.cpu(), d['gy']) < 2e-2
    for n, p in m.named_parameters():
        key = 'grad.' + n
        if key in d:
            assert p.grad is not None, n
            assert _rel_l2(p.grad.cpu(), d[key]) < 3e-2, n
    after = H.split_sd(d, 'sd_after.')
    sd = m.state_dict()
    for k in after:
        if k.endswith(('running_mean', 'running_var', 'weight_u', 'weight_v')):
            assert H.rel_err(sd[k].cpu(), after[k]) < 5e-3, k
    with torch.no_grad():                                              # the no-grad path (fused kernels)
        out2 = NF.to_nchw(m(x0.to(_dev()), y0.to(_dev())), x0.shape[1])
    if not bool(d['training']):
        assert H.rel_err(out2.cpu(), ref) < 1e-2
