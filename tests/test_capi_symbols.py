"""CPU checks of the C-ABI boundary: header <-> binding <-> built library."""
import os
import re

import pytest

from atooms_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header='atooms_b200.h'):
    text = open(os.path.join(ROOT, 'include', header)).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return set(re.findall(r'\b(b2_\w+)\s*\(', text))


def test_header_and_binding_agree():
    assert _declared() == set(capi.SYMBOLS)


def test_product_library_builds_and_exports_every_symbol():
    import __graft_entry__
    __graft_entry__.build()
    lib = capi.load()
    assert lib.name == 'cuda-sm_100a'
    for name in _declared():
        assert hasattr(lib.dll, name), name


def test_product_library_exports_the_multi_gpu_entry_points():
    """include/atooms_b200_nccl.h: CUDA library only (NCCL and peer-memory transports, device-side
    slab refresh); every declared symbol must be exported and bound."""
    import __graft_entry__
    __graft_entry__.build()
    lib = capi.load()
    names = _declared('atooms_b200_nccl.h')
    assert {'b2_nccl_run', 'b2_peer_run', 'b2_peer_refresh', 'b2_peer_set_ids', 'b2_peer_state'} <= names
    assert lib.has_nccl
    for name in names:
        fn = getattr(lib.dll, name)
        assert fn.argtypes is not None, name + ' is exported but not bound in capi.Library'


def test_oracle_library_exports_every_symbol(oracle_lib):
    assert oracle_lib.name == 'oracle-c'
    for name in _declared():
        assert hasattr(oracle_lib.dll, name), name


def test_no_cpu_fallback():
    """Without a GPU the product must fail loudly, not compute on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip('a GPU is present')
    import __graft_entry__
    __graft_entry__.build()
    with pytest.raises(capi.B2Error) as err:
        capi.Engine(capi.FP64)
    assert err.value.code == -6


def test_product_package_never_imports_oracle():
    pkg = os.path.join(ROOT, 'atooms_b200')
    for base, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                text = open(os.path.join(base, f)).read()
                assert 'import oracle' not in text and 'from oracle' not in text, f
                assert 'liboracle' not in text, f
