"""ctypes binding of libimx_kernels.so (the C ABI declared in include/imx.h).

The signatures are read from the header itself, so the binding cannot drift from
the declarations.  There is no fallback of any kind: if the library is missing
or a symbol is absent, importing / calling raises.
"""
import ctypes
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'lib', 'libimx_kernels.so')
HEADER_PATH = os.path.join(os.path.dirname(_HERE), 'include', 'imx.h')

_SCALARS = {
    'int': ctypes.c_int,
    'float': ctypes.c_float,
    'double': ctypes.c_double,
    'int64_t': ctypes.c_int64,
    'imx_stream_t': ctypes.c_void_p,
}
_RET = {'int': ctypes.c_int, 'int64_t': ctypes.c_int64, 'const char*': ctypes.c_char_p}


def parse_header(path=HEADER_PATH):
    """Returns {name: (restype, [argtypes], [argnames])} for every IMX_API function."""
    text = open(path).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    decls = {}
    for m in re.finditer(r'IMX_API\s+(const char\*|int64_t|int)\s+(imx_\w+)\s*\(([^)]*)\)\s*;', text):
        ret, name, args = m.group(1), m.group(2), m.group(3)
        argtypes, argnames = [], []
        args = args.strip()
        if args and args != 'void':
            for a in args.split(','):
                a = ' '.join(a.split())
                if '*' in a:
                    argtypes.append(ctypes.c_void_p)
                    argnames.append(a.split('*')[-1].strip())
                else:
                    ty, nm = a.rsplit(' ', 1)
                    argtypes.append(_SCALARS[ty.replace('const ', '')])
                    argnames.append(nm)
        decls[name] = (_RET[ret], argtypes, argnames)
    return decls


class ImxError(RuntimeError):
    pass


class _Lib:
    def __init__(self):
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                'libimx_kernels.so is not built (%s). Run `python -m imaginaire_b200.build` '
                '(or __graft_entry__.build()); there is no fallback path.' % LIB_PATH)
        self._dll = ctypes.CDLL(LIB_PATH)
        self.launches = 0   # kernel launches issued through call(); bench.py reads it
        self.decls = parse_header()
        for name, (ret, argtypes, _) in self.decls.items():
            fn = getattr(self._dll, name)   # AttributeError if the symbol is missing
            fn.restype = ret
            fn.argtypes = argtypes

    def raw(self, name):
        return getattr(self._dll, name)

    def last_error(self):
        return self._dll.imx_last_error().decode('utf-8', 'replace')

    # kernels launched per successful call (everything else launches exactly one)
    _LAUNCHES = {'imx_conv2d_igemm_wgrad_bf16': 2, 'imx_conv2d_igemm_wgrad_dot_bf16': 2, 'imx_conv2d_igemm_wgrad_bf16_noalloc': 2,
                 'imx_conv2d_igemm_fprop_splitk_bf16': 2, 'imx_spectral_norm_sigma_f32': 4, 'imx_spectral_norm_sigma_batched_f32': 4, 'imx_spectral_norm_weight_grad_f32': 2}

    def call(self, name, *args):
        """Calls an int-returning entry point and raises ImxError on a nonzero code."""
        rc = getattr(self._dll, name)(*args)
        if rc != 0:
            raise ImxError('%s failed with code %d: %s' % (name, rc, self.last_error()))
        self.launches += self._LAUNCHES.get(name, 1)


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = _Lib()
    return _lib
