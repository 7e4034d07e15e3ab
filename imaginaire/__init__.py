"""``imaginaire`` -> ``imaginaire_b200``: the reference's import paths.

The reference's configuration files and entry points name modules by path (``cfg.gen.type:
imaginaire.generators.spade``, ``cfg.trainer.type: imaginaire.trainers.spade``; train.py:33-60 resolves them with
importlib).  With this package on ``sys.path`` in place of the reference's, every ``imaginaire.x.y`` import returns
the very module object ``imaginaire_b200.x.y`` -- one set of classes, one set of module-level caches -- so those
files and user code written against the reference run on the native kernels unchanged.
"""
import importlib
import importlib.abc
import importlib.util
import sys

_TARGET = 'imaginaire_b200'


class _AliasLoader(importlib.abc.Loader):
    def __init__(self, target):
        self.target = target

    def create_module(self, spec):
        return importlib.import_module(self.target)

    def exec_module(self, module):
        pass        # already executed under its own name


class _AliasFinder(importlib.abc.MetaPathFinder):
    def find_spec(self, fullname, path=None, target=None):
        if not fullname.startswith(__name__ + '.'):
            return None
        real = _TARGET + fullname[len(__name__):]
        try:
            if importlib.util.find_spec(real) is None:
                return None
        except ModuleNotFoundError:
            return None
        return importlib.util.spec_from_loader(fullname, _AliasLoader(real))


if not any(isinstance(f, _AliasFinder) for f in sys.meta_path):
    sys.meta_path.insert(0, _AliasFinder())

_real = importlib.import_module(_TARGET)
__path__ = []          # submodules come from the finder above, never from this directory
__all__ = getattr(_real, '__all__', [])
