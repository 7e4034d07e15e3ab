"""Generates tests/golden/*.npz by running the REFERENCE's own Python modules on CPU.

TEST INFRASTRUCTURE ONLY.  Runs only in the build container, where the reference is
mounted at /root/reference (it imports it through oracle/ref_shims.py, whose stubs do not
alter arithmetic).  The fixtures are committed; nothing on the GPU box reads
/root/reference.

    python -m oracle.make_golden

Fixtures:
  spade_norm_*.npz     SpatiallyAdaptiveNorm + LeakyReLU(0.2)         (reference modules)
  res2d_block_*.npz    Res2dBlock('NACNAC', SPADE, spectral) train-mode, incl. updated buffers
  generator_unit_test.npz  Generator of configs/unit_test/spade.yaml, batch 1 (BASELINE configs[0])
  warp_*.npz           resample() of imaginaire/model_utils/fs_vid2vid.py incl. edge-case flows
"""
import hashlib
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLDEN = os.path.join(ROOT, 'tests', 'golden')
sys.path.insert(0, ROOT)

from oracle import ref_shims  # noqa: E402


def state_dict_digest(sd):
    h = hashlib.sha256()
    for k in sd:
        h.update(k.encode())
        h.update(sd[k].detach().cpu().contiguous().numpy().tobytes())
    return h.hexdigest()


def onehot_label(seed, n, classes, h, w, extra_uniform=0, blocky=1):
    """Synthetic label map: one-hot of randint classes (optionally 'blocky' like real maps)
    plus `extra_uniform` channels of U(0,1) (the edge-map channel of COCO-Stuff)."""
    g = torch.Generator().manual_seed(seed)
    idx = torch.randint(0, classes, (n, 1, h // blocky, w // blocky), generator=g)
    if blocky > 1:
        idx = idx.repeat_interleave(blocky, 2).repeat_interleave(blocky, 3)
    lab = torch.zeros(n, classes, h, w).scatter_(1, idx, 1.0)
    if extra_uniform:
        lab = torch.cat([lab, torch.rand(n, extra_uniform, h, w, generator=g)], 1)
    return lab


def synthetic_flow(seed, b, h, w, scale=8.0):
    g = torch.Generator().manual_seed(seed)
    coarse = torch.randn(b, 2, max(h // 16, 2), max(w // 16, 2), generator=g) * scale
    flow = torch.nn.functional.interpolate(coarse, size=(h, w), mode='bilinear', align_corners=False)
    # edge cases: integer, +-0.5, far out of range
    flow[:, :, 0, :] = 3.0
    flow[:, :, 1, :] = -0.5
    flow[:, :, 2, :] = 2.0 * w
    flow[:, :, 3, :] = -2.0 * h
    flow[:, :, 4, :] = 0.0
    return flow.contiguous()


def save(name, **arrays):
    out = {}
    for k, v in arrays.items():
        if isinstance(v, torch.Tensor):
            v = v.detach().cpu().numpy()
        out[k] = v
    path = os.path.join(GOLDEN, name)
    np.savez_compressed(path, **out)
    print('wrote %s (%.1f KB)' % (path, os.path.getsize(path) / 1024))


def sd_arrays(sd, prefix):
    return {prefix + k: v for k, v in sd.items()}


def main():
    ref_shims.install()
    from imaginaire.layers import Res2dBlock                                   # reference
    from imaginaire.layers.activation_norm import SpatiallyAdaptiveNorm       # reference
    from imaginaire.model_utils.fs_vid2vid import resample                    # reference
    from imaginaire.config import Config as RefConfig                         # reference
    from imaginaire.generators.spade import Generator as RefGenerator         # reference
    from imaginaire.utils.init_weight import weights_init as ref_weights_init  # reference
    os.makedirs(GOLDEN, exist_ok=True)

    # ---- SPADE norm + LeakyReLU --------------------------------------------------
    cases = [
        ('a', dict(num_features=64, cond_dims=5, num_filters=32, kernel_size=3, separate_projection=True,
                   activation_norm_type='instance'), False, (2, 12, 20)),
        ('b', dict(num_features=32, cond_dims=7, num_filters=16, kernel_size=5, separate_projection=False,
                   activation_norm_type='sync_batch',
                   activation_norm_params=types.SimpleNamespace(affine=True)), True, (3, 8, 8)),
        ('c', dict(num_features=64, cond_dims=4, num_filters=0, kernel_size=1, separate_projection=False,
                   activation_norm_type='instance'), False, (1, 16, 16)),
    ]
    for tag, kw, training, (n, h, w) in cases:
        torch.manual_seed(10)
        m = SpatiallyAdaptiveNorm(**kw)
        if m.norm is not None and getattr(m.norm, 'weight', None) is not None:
            with torch.no_grad():
                m.norm.weight.uniform_(0.5, 1.5)
                m.norm.bias.uniform_(-0.5, 0.5)
        m.train(training)
        if not training and hasattr(m.norm, 'running_mean') and m.norm.running_mean is not None:
            m.norm.running_mean.uniform_(-0.3, 0.3)
            m.norm.running_var.uniform_(0.5, 1.5)
        sd_before = {k: v.clone() for k, v in m.state_dict().items()}
        x = torch.randn(n, kw['num_features'], h, w) * 1.5 + 0.3
        label = onehot_label(3, n, kw['cond_dims'], 4 * h, 4 * w)
        with torch.no_grad():
            y = torch.nn.functional.leaky_relu(m(x, label), 0.2)
        sd_after = m.state_dict()
        save('spade_norm_%s.npz' % tag, x=x, label=label, y=y, training=np.array(training),
             kernel_size=np.array(kw['kernel_size']), separate_projection=np.array(kw['separate_projection']),
             num_filters=np.array(kw['num_filters']), norm_kind=np.array(kw['activation_norm_type']),
             **sd_arrays(sd_before, 'sd.'), **sd_arrays(sd_after, 'sd_after.'))

    # ---- Res2dBlock -----------------------------------------------------------------
    for tag, cin, cout, kind, training in [('a', 64, 32, 'sync_batch', True), ('b', 32, 32, 'instance', False)]:
        torch.manual_seed(20)
        an = types.SimpleNamespace(num_filters=16, kernel_size=5, separate_projection=True, cond_dims=6,
                                   weight_norm_type='spectral', activation_norm_type=kind,
                                   activation_norm_params=types.SimpleNamespace(affine=True))
        blk = Res2dBlock(cin, cout, kernel_size=3, padding=1, bias=[True, True, False],
                         weight_norm_type='spectral', activation_norm_type='spatially_adaptive',
                         activation_norm_params=an, skip_activation_norm=True, nonlinearity='leakyrelu',
                         order='NACNAC')
        blk.apply(ref_weights_init('xavier', 0.02))
        blk.train(training)
        sd_before = {k: v.clone() for k, v in blk.state_dict().items()}
        x = torch.randn(2, cin, 16, 24)
        label = onehot_label(4, 2, 6, 64, 96, blocky=4)
        with torch.no_grad():
            y = blk(x, label)
        save('res2d_block_%s.npz' % tag, x=x, label=label, y=y, training=np.array(training),
             norm_kind=np.array(kind), **sd_arrays(sd_before, 'sd.'), **sd_arrays(blk.state_dict(), 'sd_after.'))

    # ---- unit-test generator (BASELINE configs[0]) --------------------------------------
    cfg = RefConfig('/root/reference/configs/unit_test/spade.yaml')
    torch.manual_seed(0)
    net = RefGenerator(cfg.gen, cfg.data)
    net.apply(ref_weights_init(cfg.trainer.init.type, cfg.trainer.init.gain))
    # Training mode, as in the reference's own smoke test (scripts/test_training.sh): spectral
    # norm runs its power iteration (in eval mode right after initialisation sigma = u^T W v
    # of the random u, v is meaningless and the output is NaN); the norms are instance norms.
    net.train()
    label = onehot_label(1, 1, 184, 256, 256, extra_uniform=1)
    g = torch.Generator().manual_seed(2)
    images = torch.randn(1, 3, 256, 256, generator=g).clamp_(-1, 1)
    digest_before = state_dict_digest(net.state_dict())
    torch.manual_seed(3)   # the style encoder draws eps = randn_like(std)
    with torch.no_grad():
        out = net({'label': label, 'images': images})
    torch.manual_seed(3)
    eps = torch.randn(1, 256)
    save('generator_unit_test.npz', fake_images=out['fake_images'].half(), mu=out['mu'], logvar=out['logvar'],
         eps=eps, state_dict_sha256=np.array(digest_before),
         state_dict_after_sha256=np.array(state_dict_digest(net.state_dict())),
         keys=np.array(list(net.state_dict().keys())))

    # ---- resample() ---------------------------------------------------------------------
    for tag, (b, c, h, w) in [('a', (2, 3, 64, 128)), ('b', (1, 4, 33, 47)), ('c', (1, 3, 256, 512))]:
        g = torch.Generator().manual_seed(30)
        img = torch.randn(b, c, h, w, generator=g)
        flow = synthetic_flow(31, b, h, w)
        out = resample(img, flow)
        if tag == 'c':   # keep the fixture small: store a strided subsample of the big case
            save('warp_%s.npz' % tag, shape=np.array([b, c, h, w]), out_sub=out[:, :, ::4, ::4])
        else:
            save('warp_%s.npz' % tag, img=img, flow=flow, out=out)


if __name__ == '__main__':
    main()
