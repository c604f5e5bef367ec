"""The C-ABI library loads on a machine without a GPU and exports every symbol include/plxbo.h
declares; the ctypes binding declares a prototype for each of them; compute calls fail loudly
without a device (there is no CPU path)."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, 'include', 'plxbo.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(plx_[a-z0-9_]+)\s*\(', text)) - {'plx_fg_callback'})


def test_library_exports_every_declared_symbol():
    from polyaxon_b200 import _lib
    lib = _lib.load()
    names = declared_symbols()
    assert len(names) >= 15
    for name in names:
        assert hasattr(lib, name), name
        assert name in _lib.SIGNATURES, 'no ctypes prototype for ' + name
    assert sorted(_lib.SIGNATURES) == names
    assert b'sm_100a' in lib.plx_version()


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip('a GPU is present')
    from polyaxon_b200 import _lib
    with pytest.raises(_lib.PlxError, match='no CPU path'):
        _lib.Handle(0)


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, 'polyaxon_b200')
    for base, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                src = open(os.path.join(base, f)).read()
                assert not re.search(r'^\s*(from|import)\s+oracle\b', src, flags=re.M), os.path.join(base, f)
                assert 'sklearn' not in src or f in ('acquisition_function.py', 'hptuning.py') or f.endswith(('.cu', '.cuh', '.h')), f
