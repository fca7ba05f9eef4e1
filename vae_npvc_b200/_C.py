"""ctypes binding of libvae_npvc_b200.so (the C ABI in include/vae_npvc_b200.h).

Prototypes are PARSED from the header, so the Python side cannot drift from the ABI.
``call(name, *args)`` converts torch tensors to raw device pointers (checking device,
contiguity and dtype against the C parameter type), appends the current CUDA stream for the
trailing ``void* stream`` parameter, and raises on a non-zero status.

There is no CPU fallback: if the library is missing or no CUDA device is present the product
path raises.
"""
import ctypes
import os
import re

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
HEADER = os.path.join(os.path.dirname(_HERE), 'include', 'vae_npvc_b200.h')
LIB_PATH = os.path.join(_HERE, 'lib', 'libvae_npvc_b200.so')

BACKEND_AUTO, BACKEND_SIMT, BACKEND_TC = 0, 1, 2
ACT_NONE, ACT_LRELU, ACT_RELU = 0, 1, 2

_PTR_DTYPES = {
    'float*': (torch.float32,), 'const float*': (torch.float32,),
    'int64_t*': (torch.int64,), 'const int64_t*': (torch.int64,),
    'int*': (torch.int32,), 'const int*': (torch.int32,),
    'void*': None, 'const void*': None,
    'double*': (torch.float64,), 'const double*': (torch.float64,),
}
_SCALARS = {'int': ctypes.c_int, 'int64_t': ctypes.c_int64, 'float': ctypes.c_float, 'double': ctypes.c_double,
            'size_t': ctypes.c_size_t}
_RET = {'int': ctypes.c_int, 'int64_t': ctypes.c_int64, 'void': None, 'const char*': ctypes.c_char_p}


def parse_header(path=HEADER):
    """-> {name: (ret_type, [(ctype_str, param_name), ...])} for every vnb_* declaration."""
    src = open(path).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    src = re.sub(r'//[^\n]*', '', src)
    protos = {}
    for m in re.finditer(r'(const char\*|int64_t|int|void)\s+(vnb_\w+)\s*\(([^)]*)\)\s*;', src):
        ret, name, args = m.group(1), m.group(2), m.group(3).strip()
        params = []
        if args and args != 'void':
            for a in args.split(','):
                a = ' '.join(a.split())
                pm = re.match(r'(.*?)(\w+)$', a)
                ctype = pm.group(1).strip().replace(' *', '*')
                params.append((ctype, pm.group(2)))
        protos[name] = (ret, params)
    return protos


PROTOS = parse_header()
_lib = None


def load():
    """Load the shared library (building it in-tree first if nvcc is available and it is
    missing).  Raises if it cannot be loaded - there is no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        from . import build as _build
        _build.build()
    lib = ctypes.CDLL(LIB_PATH)
    for name, (ret, params) in PROTOS.items():
        fn = getattr(lib, name)          # AttributeError here == header/library mismatch
        fn.restype = _RET[ret]
        fn.argtypes = [ctypes.c_void_p if t.endswith('*') else _SCALARS[t] for t, _ in params]
    _lib = lib
    return lib


def last_error():
    return load().vnb_last_error().decode()


def _ptr(name, pname, ctype, v):
    if v is None:
        return None
    if isinstance(v, int):
        return v
    if not isinstance(v, torch.Tensor):
        raise TypeError(f'{name}: argument {pname} must be a tensor or None, got {type(v)}')
    if not v.is_cuda:
        raise RuntimeError(f'{name}: argument {pname} is not a CUDA tensor - vae_npvc_b200 has no CPU path')
    if not v.is_contiguous():
        raise RuntimeError(f'{name}: argument {pname} is not contiguous')
    want = _PTR_DTYPES[ctype]
    if want is not None and v.dtype not in want:
        raise TypeError(f'{name}: argument {pname} has dtype {v.dtype}, expected {want[0]}')
    return v.data_ptr()


def call(name, *args):
    """Invoke ``vnb_<name>``; the trailing stream argument is filled in automatically."""
    lib = load()
    full = 'vnb_' + name
    ret, params = PROTOS[full]
    has_stream = bool(params) and params[-1][1] == 'stream'
    n_user = len(params) - (1 if has_stream else 0)
    if len(args) != n_user:
        raise TypeError(f'{full} takes {n_user} arguments ({[p for _, p in params[:n_user]]}), got {len(args)}')
    cargs = []
    for (ctype, pname), v in zip(params, args):
        if ctype.endswith('*'):
            cargs.append(_ptr(full, pname, ctype, v))
        elif ctype in ('float', 'double'):
            cargs.append(float(v))
        else:
            cargs.append(int(v))
    if has_stream:
        cargs.append(torch.cuda.current_stream().cuda_stream)
    r = getattr(lib, full)(*cargs)
    if name == 'vq_fused' and r == 1:
        return 1                          # shape not taken by the fused kernel: the caller falls back
    if ret == 'int' and name not in ('version', 'get_backend', 'get_option', 'debug_tc_timeouts', 'gn_stats_ws_floats') and r != 0:
        raise RuntimeError(f'{full} failed ({r}): {lib.vnb_last_error().decode()}')
    return r


def set_backend(b):
    load().vnb_set_backend(int(b))


def get_backend():
    return load().vnb_get_backend()


OPT_EW_V2, OPT_FAST_GATE, OPT_WN_V2, OPT_GEMM_V3, OPT_TC3_BN, OPT_TC3_PAIR, OPT_TC3_STAGES, OPT_TC_DBG, OPT_TC3_SMEM_KB, OPT_FAST_EPI, OPT_VQ_FUSED, OPT_WG_FUSE = range(12)


def set_option(which, value):
    load().vnb_set_option(int(which), int(value))


def get_option(which):
    return load().vnb_get_option(int(which))


def launch_count():
    return load().vnb_launch_count()


def launch_count_reset():
    load().vnb_launch_count_reset()


def require_cuda():
    if not torch.cuda.is_available():
        raise RuntimeError('vae_npvc_b200 needs a CUDA device (sm_100a); there is no CPU fallback')
    load()
