"""Shared test helpers: golden loading and the synthetic-input recipes of
oracle/make_golden.py (re-stated here so the GPU box needs no reference)."""
import hashlib
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def load_golden(name):
    d = np.load(os.path.join(GOLDEN, name), allow_pickle=False)
    return {k: d[k] for k in d.files}


def split_sd(d, prefix):
    return {k[len(prefix):]: torch.from_numpy(np.array(v)) for k, v in d.items() if k.startswith(prefix)}


def state_dict_digest(sd):
    h = hashlib.sha256()
    for k in sd:
        h.update(k.encode())
        h.update(sd[k].detach().cpu().contiguous().numpy().tobytes())
    return h.hexdigest()


def onehot_label(seed, n, classes, h, w, extra_uniform=0, blocky=1):
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
    flow[:, :, 0, :] = 3.0
    flow[:, :, 1, :] = -0.5
    flow[:, :, 2, :] = 2.0 * w
    flow[:, :, 3, :] = -2.0 * h
    flow[:, :, 4, :] = 0.0
    return flow.contiguous()


def rel_err(a, b):
    a = torch.as_tensor(a).double()
    b = torch.as_tensor(b).double()
    return float((a - b).abs().max() / (b.abs().max() + 1e-30))


def unit_test_generator_inputs():
    """Inputs of tests/golden/generator_unit_test.npz (oracle/make_golden.py)."""
    label = onehot_label(1, 1, 184, 256, 256, extra_uniform=1)
    g = torch.Generator().manual_seed(2)
    images = torch.randn(1, 3, 256, 256, generator=g).clamp_(-1, 1)
    return label, images


def build_generator(config_name, seed=0):
    """Seeded build + the trainer's weights_init, exactly as oracle/make_golden.py builds the
    reference generator (same torch RNG stream -> same weights)."""
    from imaginaire_b200.config import Config
    from imaginaire_b200.generators.spade import Generator
    from imaginaire_b200.utils.init_weight import weights_init
    cfg = Config(os.path.join(ROOT, 'configs', config_name))
    torch.manual_seed(seed)
    net = Generator(cfg.gen, cfg.data)
    net.apply(weights_init(cfg.trainer.init.type, cfg.trainer.init.gain))
    return cfg, net


def subsample(t, limit=20000):
    """The rule oracle/make_golden_train.py stores large tensors by: a strided sample of the flattened tensor."""
    flat = t.reshape(-1)
    return flat[::-(-flat.numel() // limit)] if flat.numel() > limit else t


def seeded_vgg19_features(seed=1234):
    """torchvision's VGG-19 with its default initialisation under a fixed seed (oracle/make_golden_train.py builds
    the reference's perceptual network the same way: there is no network to download ImageNet weights from)."""
    from torchvision.models.vgg import vgg19
    with torch.random.fork_rng():
        torch.manual_seed(seed)
        return vgg19(weights=None).features


def c3mini_batch(n=2, h=256, w=256, classes=35):
    """Inputs of tests/golden/{train_step,discriminator,generator}_c3mini.npz."""
    label = onehot_label(5, n, classes, h, w, blocky=8)
    g = torch.Generator().manual_seed(6)
    images = torch.rand(n, 3, h, w, generator=g) * 2 - 1
    return {'label': label, 'images': images}


def style_noise(seed, n, dims):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(n, dims, generator=g)
