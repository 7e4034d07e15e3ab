"""Import shims that let the reference's own Python modules (under
/root/reference) run on a CPU-only box with torch 2.x.

TEST INFRASTRUCTURE ONLY.  This file is used in the build container (where
/root/reference exists) by ``oracle/make_golden.py`` to generate the fixtures
under ``tests/golden/`` and to validate ``oracle/spade_oracle.py``.  Nothing in
the product path imports it, and nothing run on the GPU box imports it
(/root/reference does not exist there).

No shim alters arithmetic:
  * albumentations / lmdb / apex / win32process / torch._six : import stubs
  * ``.cuda()`` / ``.to('cuda')`` : no-ops so constructors work without a GPU
"""
import sys
import types

import torch

REFERENCE_ROOT = '/root/reference'


def _stub(name, **attrs):
    mod = types.ModuleType(name)
    mod.__dict__.update(attrs)
    sys.modules[name] = mod
    return mod


def install():
    """Install the stubs and put the reference on sys.path. Idempotent."""
    if getattr(install, '_done', False):
        return
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)

    class _Anything:
        def __init__(self, *a, **k):
            pass

        def __getattr__(self, item):
            return _Anything()

        def __call__(self, *a, **k):
            return _Anything()

    def _lenient(item):
        # module-level __getattr__ (PEP 562) of a stub: any public name resolves; dunders do not (inspect and
        # importlib probe __file__ / __path__ / __spec__ on everything in sys.modules)
        if item.startswith('__'):
            raise AttributeError(item)
        return _Anything

    import torchvision  # noqa: F401  (before any stub exists: its import walks sys.modules through inspect)

    for name in ('albumentations', 'lmdb', 'win32process', 'imageio',
                 'skimage', 'cv2', 'tensorboard', 'torch.utils.tensorboard',
                 'torch.utils.tensorboard.summary'):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                m = _stub(name)
                m.__getattr__ = _lenient
    if 'torch._six' not in sys.modules:
        six = _stub('torch._six', string_classes=(str, bytes))
        torch._six = six
    if 'apex' not in sys.modules:
        apex = _stub('apex')
        apex.amp = _stub('apex.amp')

        # apex.amp in its pass-through form (opt_level O0: fp32, loss scale 1): the three calls the trainer
        # makes (utils/trainer.py:150-154 initialize, trainers/base.py:614,658 scale_loss, :619 master_params)
        class _ScaleLoss:
            def __init__(self, loss, *a, **k):
                self.loss = loss

            def __enter__(self):
                return self.loss

            def __exit__(self, *exc):
                return False

        def _master_params(opt):
            for group in opt.param_groups:
                for p in group['params']:
                    yield p
        apex.amp.scale_loss = _ScaleLoss
        apex.amp.master_params = _master_params
        apex.amp.initialize = lambda models, optimizers, **k: (models, optimizers)
        apex.optimizers = _stub('apex.optimizers')
        apex.parallel = _stub('apex.parallel')
        apex.__getattr__ = _lenient
        apex.amp.__getattr__ = _lenient
        apex.optimizers.__getattr__ = _lenient
        apex.parallel.__getattr__ = _lenient

    # .cuda() / .to('cuda') are no-ops on the CPU box.
    torch.Tensor.cuda = lambda self, *a, **k: self
    torch.nn.Module.cuda = lambda self, *a, **k: self
    _orig_to = torch.Tensor.to

    def _is_cuda(a):
        return (isinstance(a, str) and a.startswith('cuda')) or (isinstance(a, torch.device) and a.type == 'cuda')

    def _to(self, *args, **kwargs):
        args = tuple('cpu' if _is_cuda(a) else a for a in args)
        if _is_cuda(kwargs.get('device')):
            kwargs['device'] = 'cpu'
        return _orig_to(self, *args, **kwargs)
    torch.Tensor.to = _to
    _orig_module_to = torch.nn.Module.to

    def _module_to(self, *args, **kwargs):
        args = tuple('cpu' if _is_cuda(a) else a for a in args)
        if _is_cuda(kwargs.get('device')):
            kwargs['device'] = 'cpu'
        return _orig_module_to(self, *args, **kwargs)
    torch.nn.Module.to = _module_to
    install._done = True
