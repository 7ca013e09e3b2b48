"""CPU-side checks of the drop-in boundary: the shared library loads, exports every symbol
include/kvarq_b200.h declares, and refuses to work without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, 'include', 'kvarq_b200.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(kvq_[a-z_0-9]+)\s*\(', src)))


def test_library_exports_every_header_symbol():
    from kvarq_b200 import _lib
    lib = _lib.load()
    syms = header_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(lib, s), 'libkvarq_b200.so does not export ' + s
    assert sorted(_lib.SIGNATURES) == syms, 'ctypes table and header disagree'
    assert b'sm_100a' in lib.kvq_version()


def test_struct_layouts_match_header():
    from kvarq_b200 import _lib
    assert ctypes.sizeof(_lib.HitRec) == 24
    assert ctypes.sizeof(_lib.Config) == 20
    assert ctypes.sizeof(_lib.Stats) == 1024 * 8 + 5 * 8 + 8
    assert ctypes.sizeof(_lib.Timing) == 9 * 8


def test_no_cpu_fallback():
    from kvarq_b200 import _lib, engine
    lib = _lib.load()
    if lib.kvq_device_count() > 0:
        pytest.skip('a GPU is present')
    with pytest.raises(RuntimeError):
        engine.Engine(0)
    h = ctypes.c_void_p()
    assert lib.kvq_create(ctypes.byref(h), 0) == _lib.KVQ_ERR_RUNTIME and not h.value


def test_product_never_touches_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, 'kvarq_b200')):
        if 'build' in dirpath:
            continue
        for f in files:
            if f.endswith(('.py', '.cu', '.cpp', '.h', '.cuh')) or f == 'Makefile':
                txt = open(os.path.join(dirpath, f), errors='replace').read()
                assert 'oracle' not in txt.replace('no CPU matcher', ''), f + ' references oracle/'


def test_module_surface():
    from kvarq_b200 import engine
    for name in ('config', 'get_config', 'findseqs', 'stats', 'stop', 'test', 'Hit'):
        assert hasattr(engine, name)
    assert engine.Hit._fields == ('seq_nr', 'file_pos', 'seq_pos', 'length', 'readlength')
