"""The drop-in boundary without a GPU: the C-ABI shared library builds in-tree for sm_100a, loads, and exports every
symbol include/vae_npvc_b200.h declares; argument checks fail with VNB_E_ARG and a message before anything is
launched; the product has no CPU fallback and never imports the oracle."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope='module')
def lib():
    from vae_npvc_b200 import _C
    return _C.load()


def test_library_exports_every_declared_symbol(lib):
    from vae_npvc_b200 import _C
    protos = _C.parse_header()
    assert len(protos) >= 50
    header = open(_C.HEADER).read()
    declared = set(re.findall(r'\b(vnb_\w+)\s*\(', header))
    assert declared == set(protos), declared ^ set(protos)          # the parser sees every declaration
    for name in protos:
        assert hasattr(lib, name), name
    # every entry point that launches work takes the stream last, every status is an int
    for name, (ret, params) in protos.items():
        if params and params[-1][1] == 'stream':
            assert ret == 'int' and params[-1][0] == 'void*', name
        for ctype, _ in params:
            assert 'torch' not in ctype and 'Tensor' not in ctype


def test_version_error_string_counter_and_options(lib):
    from vae_npvc_b200 import _C
    assert lib.vnb_version() >= 100
    assert isinstance(_C.last_error(), str)
    _C.launch_count_reset()
    assert _C.launch_count() == 0
    for opt in (_C.OPT_EW_V2, _C.OPT_FAST_GATE, _C.OPT_WN_V2):
        old = _C.get_option(opt)
        _C.set_option(opt, 1 - old)
        assert _C.get_option(opt) == 1 - old
        _C.set_option(opt, old)
    assert _C.get_option(99) == 0                                   # unknown options read as off
    be = _C.get_backend()
    _C.set_backend(_C.BACKEND_SIMT)
    assert _C.get_backend() == _C.BACKEND_SIMT
    _C.set_backend(be)


def test_bad_arguments_fail_before_any_launch(lib):
    """VNB_E_ARG (-1) + thread-local message; nothing is launched (no CUDA context is needed for these calls)."""
    from vae_npvc_b200 import _C
    _C.launch_count_reset()
    null = ctypes.c_void_p(None)
    assert lib.vnb_csid(null, 1, 1, null, null, null, null) == -1
    assert 'csid' in _C.last_error()
    assert lib.vnb_mask_rows(null, null, null, null, 1, 8, 4, 1, null, null) == -1
    assert 'mask_rows' in _C.last_error()
    assert lib.vnb_conv_fwd(null, 1, 1, 8, 8, null, 3, 1, 8, null, 0, null, null, 0, null, 0, null) == -1
    assert 'conv_fwd' in _C.last_error()
    assert lib.vnb_vq_lookup(null, null, null, null, null, null, 1, 8, 4, 16, null, null) == -1
    assert lib.vnb_gate_bwd_rows_p(null, null, null, null, null, null, 1, 8, 4, 2, null, null) == -1
    assert _C.launch_count() == 0


def test_python_binding_rejects_cpu_tensors_and_wrong_dtypes():
    from vae_npvc_b200 import _C
    with pytest.raises(RuntimeError, match='no CPU path'):
        _C.call('csid', torch.zeros(2, 4, dtype=torch.int64), 2, 4, None, torch.zeros(2, 4, dtype=torch.int64),
                torch.zeros(2, dtype=torch.int32))
    with pytest.raises(TypeError):
        _C.call('csid', torch.zeros(2, 4, dtype=torch.int64))      # wrong arity


@pytest.mark.skipif(torch.cuda.is_available(), reason='needs a box without a GPU')
def test_no_cpu_fallback():
    from oracle import synth
    from vae_npvc_b200 import _C
    from vae_npvc_b200.model.vqvae import Model
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        _C.require_cuda()
    arch = synth.small_config()
    m = Model(arch).eval()
    x, y = synth.synthetic_batch(arch, 2, 8, 1)
    with pytest.raises(RuntimeError):
        m.infer((x, y))                                             # CPU tensors into the CUDA-only path
    from vae_npvc_b200.decoder.basic import Decoder
    with pytest.raises(RuntimeError):
        Decoder(dict(arch, use_gpu=False))
    with pytest.raises(RuntimeError):
        Decoder(dict(arch))


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, 'vae_npvc_b200')
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith('.py'):
                src = open(os.path.join(d, f)).read()
                assert not re.search(r'^\s*(from|import)\s+oracle\b', src, flags=re.M), os.path.join(d, f)
                assert '/root/reference' not in src, os.path.join(d, f)
