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

    for name in ('albumentations', 'lmdb', 'win32process', 'imageio',
                 'skimage', 'cv2'):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                m = _stub(name)
                m.__getattr__ = lambda item: _Anything  # PEP 562
    if 'torch._six' not in sys.modules:
        six = _stub('torch._six', string_classes=(str, bytes))
        torch._six = six
    if 'apex' not in sys.modules:
        apex = _stub('apex')
        apex.amp = _stub('apex.amp')
        apex.optimizers = _stub('apex.optimizers')
        apex.parallel = _stub('apex.parallel')
        apex.__getattr__ = lambda item: _Anything
        apex.amp.__getattr__ = lambda item: _Anything
        apex.optimizers.__getattr__ = lambda item: _Anything
        apex.parallel.__getattr__ = lambda item: _Anything

    # .cuda() / .to('cuda') are no-ops on the CPU box.
    torch.Tensor.cuda = lambda self, *a, **k: self
    torch.nn.Module.cuda = lambda self, *a, **k: self
    _orig_to = torch.Tensor.to

    def _to(self, *args, **kwargs):
        args = tuple('cpu' if (isinstance(a, str) and a.startswith('cuda'))
                     else a for a in args)
        if isinstance(kwargs.get('device'), str) and \
                kwargs['device'].startswith('cuda'):
            kwargs['device'] = 'cpu'
        return _orig_to(self, *args, **kwargs)
    torch.Tensor.to = _to
    install._done = True
