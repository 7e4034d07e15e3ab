"""CPU oracle of one SPADE G+D training iteration -- TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

A functional, fp32, CPU restatement of what the reference's trainer computes around the generator of
``oracle/spade_oracle.py``: the SPADE discriminator (feature-pyramid semantics-embedding discriminator + multi-resolution
patch discriminators), the four loss terms, the two Adam steps and the generator's moving average.  Driven by
reference-format ``state_dict``s (same keys as ``imaginaire.discriminators.spade.Discriminator``,
``imaginaire.generators.spade.Generator`` and torchvision's ``vgg19().features``); every function cites the reference
lines it follows.  As in spade_oracle.py the convolution products are torch's CPU ``F.conv2d``; everything composed
around them is written out.

Pinned against the reference itself: ``oracle/make_golden_train.py`` runs the reference's own trainer (imported from
/root/reference in the build container) on ``configs/spade_c3_mini.yaml`` and stores discriminator outputs, loss terms,
gradients and updated weights under ``tests/golden/{discriminator,train_step}_c3mini.npz``;
``tests/test_train_oracle.py`` holds this file to those fixtures (fp32 against fp32: 1e-4 .. 1e-6 relative).

Users: tests/, and bench.py's CPU legs (``cpu_baseline`` / ``--impl reference``), which time ``train_iteration`` on the
host cores as the reference's CPU path for the benchmark's metric.  Nothing under imaginaire_b200/ imports it.
"""
import math

import torch
import torch.nn.functional as F

from . import spade_oracle as so


# ------------------------------------------------------------------ discriminator

def conv_block(sd, prefix, x, stride, padding, training, act):
    """Conv2dBlock(order='CNA', activation_norm_type='none'): convolution (spectral norm when the keys say so), then
    LeakyReLU(0.2) if `act`.  ref: imaginaire/layers/conv.py:77-104, nonlinearity.py:22"""
    y = so.conv(sd, prefix + 'layers.conv.', x, stride, padding, training)
    return so.leaky_relu(y) if act else y


def patch_discriminator(sd, prefix, x, kernel_size, num_layers, training):
    """NLayerPatchDiscriminator.forward.  ref: imaginaire/discriminators/multires_patch.py:250-313
    layer0 and layers 1 .. num_layers-1 have stride 2, layer num_layers stride 1, all k x k with padding
    floor((k-1)/2) and LeakyReLU; the last layer is a 3x3 stride-1 convolution to one channel with the SAME padding
    and no activation (:281-284).  features = every layer's output but the last (:311-312)."""
    pad = int(math.floor((kernel_size - 1.0) / 2))
    res = [x]
    for n in range(num_layers + 1):
        stride = 2 if n < num_layers else 1
        res.append(conv_block(sd, prefix + 'layer%d.0.' % n, res[-1], stride, pad, training, True))
    out = conv_block(sd, prefix + 'layer%d.0.' % (num_layers + 1), res[-1], 1, pad, training, False)
    return out, res[1:]


def fpse_discriminator(sd, prefix, images, segmaps, kernel_size, training):
    """FPSEDiscriminator.forward.  ref: imaginaire/discriminators/fpse.py:87-132 (layers :58-85)."""
    pad = int(math.ceil((kernel_size - 1.0) / 2))

    def down(name, t):
        return conv_block(sd, prefix + name + '.', t, 2, pad, training, True)

    def lat(name, t):            # 1x1, LeakyReLU (latent_conv2d_block carries the nonlinearity, :48-56)
        return conv_block(sd, prefix + name + '.', t, 1, 0, training, True)

    def final(name, t):
        return conv_block(sd, prefix + name + '.', t, 1, pad, training, True)

    def up(t):                   # nn.Upsample(scale_factor=2, mode='bilinear', align_corners=False), :73-74
        return F.interpolate(t, scale_factor=2, mode='bilinear', align_corners=False)

    # bottom-up: five stride-2 stages; top-down: 1x1 laterals added to the bilinearly upsampled coarser level
    c = [images]
    for i in range(1, 6):
        c.append(down('enc%d' % i, c[-1]))
    top = lat('lat5', c[5])
    pyramid = {}
    for level in (4, 3, 2):
        top = up(top) + lat('lat%d' % level, c[level])
        pyramid[level] = top
    heads = [final('final%d' % level, pyramid[level]) for level in (2, 3, 4)]
    # true/false prediction and the semantic-alignment projection: plain 1x1 convolutions, no weight norm (:82-85)
    logits = [conv_block(sd, prefix + 'output.', h, 1, 0, training, False) for h in heads]
    proj = [conv_block(sd, prefix + 'seg.', h, 1, 0, training, False) for h in heads]
    # label embedding, average-pooled to each head's resolution (1/4, 1/8, 1/16 of the input)
    emb = F.avg_pool2d(conv_block(sd, prefix + 'embedding.', segmaps, 1, 0, training, False), kernel_size=2, stride=2)
    out = []
    for logit, pr in zip(logits, proj):
        emb = F.avg_pool2d(emb, kernel_size=2, stride=2)
        out.append(logit + (emb * pr).sum(dim=1, keepdim=True))
    return out


def discriminator_spec_from_cfg(dis_cfg):
    """The few numbers the functional oracle needs.  ref: imaginaire/discriminators/spade.py:34-41,63-65"""
    return dict(kernel_size=getattr(dis_cfg, 'kernel_size', 3), num_layers=getattr(dis_cfg, 'num_layers', 5),
                num_discriminators=getattr(dis_cfg, 'num_discriminators', 2),
                fpse_kernel_size=getattr(dis_cfg, 'fpse_kernel_size', 3))


def _single_forward(sd, label, image, spec, training):
    """ref: imaginaire/discriminators/spade.py:73-90"""
    x = torch.cat((label, image), 1)
    outputs = fpse_discriminator(sd, 'fpse_discriminator.', image, label, spec['fpse_kernel_size'], training)
    features = []
    for i in range(spec['num_discriminators']):
        o, f = patch_discriminator(sd, 'discriminators.%d.' % i, x, spec['kernel_size'], spec['num_layers'], training)
        outputs.append(o)
        features.append(f)
        x = F.interpolate(x, scale_factor=0.5, mode='bilinear', align_corners=True)
    return outputs, features


def discriminator(sd, data, fake_images, spec, training):
    """Discriminator.forward: the real pass, then the fake pass (so every spectrally normalised layer runs two power
    iterations per call in training mode).  ``sd`` is updated in place (u, v).
    ref: imaginaire/discriminators/spade.py:92-117"""
    out = {}
    out['real_outputs'], out['real_features'] = _single_forward(sd, data['label'], data['images'], spec, training)
    out['fake_outputs'], out['fake_features'] = _single_forward(sd, data['label'], fake_images, spec, training)
    return out


# ------------------------------------------------------------------ losses

def gan_hinge(outputs, t_real, dis_update):
    """GANLoss('hinge') over a list of discriminator outputs: the mean of the per-output losses.
    ref: imaginaire/losses/gan.py:11-25 (the two fused hinge expressions), :52-60 (list), :89-96 (hinge)"""
    total = 0
    for x in outputs:
        if not dis_update:
            loss = -torch.mean(x)
        elif t_real:
            loss = -torch.mean(torch.min(x - 1, x * 0))
        else:
            loss = -torch.mean(torch.min(-x - 1, x * 0))
        total = total + loss
    return total / len(outputs)


def feature_matching(fake_features, real_features):
    """ref: imaginaire/losses/feature_matching.py:19-38 (L1, real side detached, 1 / num_discriminators each)"""
    num_d = len(fake_features)
    loss = fake_features[0][0].new_tensor(0)
    for i in range(num_d):
        for f, r in zip(fake_features[i], real_features[i]):
            loss = loss + (1.0 / num_d) * F.l1_loss(f, r.detach())
    return loss


def gaussian_kl(mu, logvar):
    """ref: imaginaire/losses/kl.py (GaussianKLLoss.forward)"""
    return -0.5 * torch.sum(1 + logvar - mu.pow(2) - logvar.exp())


# torchvision vgg19().features: index of every convolution, and the ReLU outputs the reference names
# (imaginaire/losses/perceptual.py:176-191)
_VGG19_CFG = [64, 64, 'M', 128, 128, 'M', 256, 256, 256, 256, 'M', 512, 512, 512, 512, 'M', 512, 512, 512, 512, 'M']
_VGG19_NAMES = {1: 'relu_1_1', 3: 'relu_1_2', 6: 'relu_2_1', 8: 'relu_2_2', 11: 'relu_3_1', 13: 'relu_3_2', 15: 'relu_3_3',
                17: 'relu_3_4', 20: 'relu_4_1', 22: 'relu_4_2', 24: 'relu_4_3', 26: 'relu_4_4', 29: 'relu_5_1'}


def vgg19_features(vsd, x, layers):
    """torchvision's VGG-19 feature stack up to the last requested layer; vsd has torchvision's keys
    ('<index>.weight' / '<index>.bias' of ``vgg19().features``).  ref: perceptual.py:143-191"""
    out, idx, last = {}, 0, max(k for k, v in _VGG19_NAMES.items() if v in layers)
    for c in _VGG19_CFG:
        if c == 'M':
            x = F.max_pool2d(x, kernel_size=2, stride=2)
            idx += 1
        else:
            x = F.relu(F.conv2d(x, vsd['%d.weight' % idx], vsd['%d.bias' % idx], padding=1))
            idx += 2
            if _VGG19_NAMES.get(idx - 1) in layers:
                out[_VGG19_NAMES[idx - 1]] = x
        if idx > last:
            break
    return out


def imagenet_normalization(x):
    """[-1, 1] images -> ImageNet-normalised.  ref: imaginaire/utils/misc.py:221-236 (three image channels)"""
    mean = x.new_tensor([0.485, 0.456, 0.406]).view(1, 3, 1, 1)
    std = x.new_tensor([0.229, 0.224, 0.225]).view(1, 3, 1, 1)
    return ((x[:, :3] + 1) / 2 - mean) / std


def perceptual(vsd, fake, real, layers, weights):
    """PerceptualLoss.forward with criterion 'l1', one scale, no resize.  ref: perceptual.py:89-140"""
    ff = vgg19_features(vsd, imagenet_normalization(fake), layers)
    with torch.no_grad():
        rf = vgg19_features(vsd, imagenet_normalization(real), layers)
    loss = 0
    for name, w in zip(layers, weights):
        loss = loss + w * F.l1_loss(ff[name], rf[name].detach())
    return loss


# ------------------------------------------------------------------ the two updates

def trainer_spec_from_cfg(cfg, crop_hw):
    t = cfg.trainer
    pl = t.perceptual_loss
    lw = t.loss_weight
    return dict(gen=so.generator_spec_from_cfg(cfg.gen, crop_hw), dis=discriminator_spec_from_cfg(cfg.dis),
                layers=list(pl.layers), layer_weights=[float(w) for w in pl.weights],
                weights={'GAN': float(lw.gan), 'FeatureMatching': float(lw.feature_matching),
                         'GaussianKL': float(lw.kl), 'Perceptual': float(lw.perceptual)},
                gen_opt=dict(lr=float(cfg.gen_opt.lr), betas=(float(cfg.gen_opt.adam_beta1), float(cfg.gen_opt.adam_beta2))),
                dis_opt=dict(lr=float(cfg.dis_opt.lr), betas=(float(cfg.dis_opt.adam_beta1), float(cfg.dis_opt.adam_beta2))))


def is_parameter(key):
    """state_dict entries the optimizer owns (everything but spectral-norm vectors and running statistics)."""
    last = key.rsplit('.', 1)[-1]
    return last in ('weight', 'weight_orig', 'bias')


def _leaves(sd, train):
    """Copy of a state dict whose parameters are autograd leaves (train) or constants."""
    out = {}
    for k, v in sd.items():
        v = v.detach().clone()
        if train and is_parameter(k) and v.is_floating_point():
            v.requires_grad_(True)
        out[k] = v
    return out


def adam_step(params, grads, state, lr, betas, eps=1e-8):
    """torch.optim.Adam (no weight decay, no amsgrad), in place on `params` (a dict of tensors).
    ref: the reference builds torch.optim.Adam in imaginaire/utils/trainer.py:238-252"""
    b1, b2 = betas
    for k, g in grads.items():
        st = state.setdefault(k, {'step': 0, 'm': torch.zeros_like(g), 'v': torch.zeros_like(g)})
        st['step'] += 1
        st['m'].mul_(b1).add_(g, alpha=1 - b1)
        st['v'].mul_(b2).addcmul_(g, g, value=1 - b2)
        bc1, bc2 = 1 - b1 ** st['step'], 1 - b2 ** st['step']
        denom = (st['v'].sqrt() / math.sqrt(bc2)).add_(eps)
        params[k] = params[k].detach() - (lr / bc1) * st['m'] / denom


def dis_update(sd_g, sd_d, data, spec, eps_noise, opt_state):
    """BaseTrainer.dis_update + Trainer.dis_forward: generator forward without autograd (training mode: its running
    statistics and spectral-norm vectors move), discriminator on real and fake, hinge losses, Adam step on D.
    ref: imaginaire/trainers/base.py:635-666, imaginaire/trainers/spade.py:168-187.
    Returns (losses, new sd_g, new sd_d, gradients)."""
    with torch.no_grad():
        g_out = so.generator(sd_g, data, spec['gen'], True, eps_noise)
    d = _leaves(sd_d, True)
    out = discriminator(d, data, g_out['fake_images'].detach(), spec['dis'], True)
    fake = gan_hinge(out['fake_outputs'], False, True)
    true = gan_hinge(out['real_outputs'], True, True)
    losses = {'GAN/fake': fake, 'GAN/true': true, 'GAN': fake + true}
    losses['total'] = losses['GAN'] * spec['weights']['GAN']
    names = [k for k, v in d.items() if v.requires_grad]
    grads = dict(zip(names, torch.autograd.grad(losses['total'], [d[k] for k in names], allow_unused=True)))
    grads = {k: (g if g is not None else torch.zeros_like(d[k])) for k, g in grads.items()}
    new_d = {k: v.detach() for k, v in d.items()}
    adam_step(new_d, grads, opt_state, **spec['dis_opt'])
    return {k: float(v.detach()) for k, v in losses.items()}, {k: v.detach() for k, v in g_out['state_dict'].items()}, new_d, grads


def gen_update(sd_g, sd_d, vsd, data, spec, eps_noise, opt_state):
    """BaseTrainer.gen_update + Trainer.gen_forward: generator with autograd, discriminator (its parameters frozen, its
    spectral-norm vectors still move), GAN + feature matching + KL + perceptual terms, Adam step on G.
    ref: imaginaire/trainers/base.py:594-633, imaginaire/trainers/spade.py:128-166.
    Returns (losses, new sd_g, new sd_d, gradients)."""
    g = _leaves(sd_g, True)
    g_out = so.generator(g, data, spec['gen'], True, eps_noise)
    d = _leaves(sd_d, False)
    out = discriminator(d, data, g_out['fake_images'], spec['dis'], True)
    losses = {'GAN': gan_hinge(out['fake_outputs'], True, False),
              'FeatureMatching': feature_matching(out['fake_features'], out['real_features']),
              'GaussianKL': gaussian_kl(g_out['mu'], g_out['logvar']),
              'Perceptual': perceptual(vsd, g_out['fake_images'], data['images'], spec['layers'], spec['layer_weights'])}
    losses['total'] = sum(losses[k] * spec['weights'][k] for k in ('GAN', 'FeatureMatching', 'GaussianKL', 'Perceptual'))
    names = [k for k, v in g.items() if v.requires_grad]
    grads = dict(zip(names, torch.autograd.grad(losses['total'], [g[k] for k in names], allow_unused=True)))
    grads = {k: (gr if gr is not None else torch.zeros_like(g[k])) for k, gr in grads.items()}
    new_g = {k: v.detach() for k, v in g_out['state_dict'].items()}     # buffers as the forward left them
    adam_step(new_g, grads, opt_state, **spec['gen_opt'])
    return {k: float(v.detach()) for k, v in losses.items()}, new_g, {k: v.detach() for k, v in d.items()}, grads


def model_average_update(ema, sd_g, beta):
    """ModelAverage.update_average: ema = beta * ema + (1 - beta) * value, the value of a spectrally normalised
    filter being W_orig / (u^T W_orig v) with the current vectors (no power iteration); integer buffers are copied.
    ref: imaginaire/utils/model_average.py:87-115,183-197"""
    out = {}
    for k, t in ema.items():
        if k.endswith('weight') and k + '_orig' in sd_g:
            w, u, v = sd_g[k + '_orig'], sd_g[k + '_u'], sd_g[k + '_v']
            val = w / torch.dot(u, torch.mv(w.reshape(w.shape[0], -1), v))
        else:
            val = sd_g[k]
        out[k] = beta * t + (1 - beta) * val if t.is_floating_point() else val.clone()
    return out


def train_iteration(sd_g, sd_d, vsd, data, spec, eps_dis, eps_gen, state=None):
    """dis_update, gen_update, moving average -- one iteration of train.py:76-79.  `state` carries the two Adam states
    and the averaged weights between calls.  Returns (dis losses, gen losses, sd_g, sd_d, state)."""
    state = state if state is not None else {'opt_d': {}, 'opt_g': {}, 'ema': None, 'ema_updates': 0}
    dl, sd_g, sd_d, _ = dis_update(sd_g, sd_d, data, spec, eps_dis, state['opt_d'])
    gl, sd_g, sd_d, _ = gen_update(sd_g, sd_d, vsd, data, spec, eps_gen, state['opt_g'])
    if state['ema'] is not None:
        state['ema_updates'] += 1
        beta = 0.0 if state['ema_updates'] <= state.get('ema_start', 1000) else state.get('ema_beta', 0.9999)
        state['ema'] = model_average_update(state['ema'], sd_g, beta)
    return dl, gl, sd_g, sd_d, state
