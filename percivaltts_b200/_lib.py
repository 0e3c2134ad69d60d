"""
ctypes binding of libpctts.so.  The prototypes are read from include/pctts.h (the single source of truth for the
C ABI), so a symbol declared there and missing from the library -- or the other way round -- fails at load time.

There is no fallback: if the library is absent it is built with nvcc (percivaltts_b200/build.py); if that fails, or
the device is not a compute-capability-10.x GPU, loading raises.
"""
import ctypes
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
HEADER = os.path.join(os.path.dirname(HERE), 'include', 'pctts.h')
LIBPATH = os.path.join(HERE, 'libpctts.so')

_CTYPES = {
    'int': ctypes.c_int, 'long': ctypes.c_long, 'float': ctypes.c_float, 'double': ctypes.c_double,
    'unsigned long long': ctypes.c_ulonglong, 'void': None,
}

ERRORS = {-1: 'PCTTS_EINVAL (bad argument)', -2: 'PCTTS_EUNSUPPORTED (shape or device not supported)',
          -3: 'PCTTS_ELIMIT (size beyond the fixed capacity of a *_static entry point)'}


class PcttsError(RuntimeError):
    pass


def parse_header(path=HEADER):
    """-> {name: (restype, [argtypes], [argnames])} for every `pctts_*` prototype in the header."""
    with open(path) as f:
        text = f.read()
    text = re.sub(r'/\*.*?\*/', ' ', text, flags=re.S)
    text = re.sub(r'//[^\n]*', ' ', text)
    text = re.sub(r'^\s*#.*$', ' ', text, flags=re.M)
    protos = {}
    for m in re.finditer(r'([A-Za-z_][\w\s\*]*?)\b(pctts_\w+)\s*\(([^)]*)\)\s*;', text):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        argtypes, argnames = [], []
        if args and args != 'void':
            for a in args.split(','):
                a = ' '.join(a.split())
                mm = re.match(r'^(.*?)(\w+)$', a)
                ty, nm = mm.group(1).strip(), mm.group(2)
                if '*' in ty:
                    argtypes.append(ctypes.c_void_p)
                else:
                    argtypes.append(_CTYPES[ty.replace('const ', '').strip()])
                argnames.append(nm)
        if '*' in ret:
            restype = ctypes.c_void_p
        else:
            restype = _CTYPES[ret.replace('const ', '').strip()]
        protos[name] = (restype, argtypes, argnames)
    return protos


_lib = None


def load(build_if_missing=True):
    """Load (building first when needed) libpctts.so and attach the prototypes.  Raises when impossible."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIBPATH):
        if not build_if_missing:
            raise PcttsError('libpctts.so is missing (run `python -m percivaltts_b200.build`); there is no CPU fallback')
        from . import build as _build
        _build.build()
    lib = ctypes.CDLL(LIBPATH)
    for name, (restype, argtypes, _) in parse_header().items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            raise PcttsError('libpctts.so does not export %s declared in include/pctts.h' % name)
        fn.restype = restype
        fn.argtypes = argtypes
    if lib.pctts_abi_version() != 1:
        raise PcttsError('libpctts.so ABI version mismatch')
    _lib = lib
    return lib


def check(rc, what=''):
    if rc == 0:
        return
    if rc < 0:
        raise PcttsError('%s failed: %s' % (what, ERRORS.get(rc, 'error %d' % rc)))
    raise PcttsError('%s failed: cudaError_t %d' % (what, rc))
