"""ctypes declarations for libkvarq_b200.so (include/kvarq_b200.h).

The library is the product; this file only declares its C ABI.  Importing it fails loudly
when the shared library has not been built -- there is no Python or CPU fallback.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('KVQ_B200_LIB') or os.path.join(HERE, 'libkvarq_b200.so')   # env: kernel-variant sweeps only

KVQ_MAX_READLENGTH = 1024

KVQ_OK, KVQ_ERR_IO, KVQ_ERR_FASTQ_FORMAT, KVQ_ERR_RUNTIME, KVQ_ERR_MEMORY, KVQ_ERR_TYPE = range(6)


class Config(C.Structure):
    _fields_ = [('maxerrors', C.c_int32), ('minoverlap', C.c_int32), ('minreadlength', C.c_int32),
                ('nthreads', C.c_int32), ('Amin', C.c_char), ('Azero', C.c_char)]


class HitRec(C.Structure):
    _fields_ = [('seq_nr', C.c_int32), ('seq_pos', C.c_int32), ('file_pos', C.c_int64),
                ('length', C.c_int32), ('readlength', C.c_int32)]


class Stats(C.Structure):
    _fields_ = [('readlengths', C.c_int64 * KVQ_MAX_READLENGTH), ('rls_longest', C.c_int64),
                ('records_parsed', C.c_int64), ('parsed', C.c_int64), ('total', C.c_int64),
                ('sigints', C.c_int64), ('progress', C.c_float), ('nseqs', C.c_int32)]


class Timing(C.Structure):
    _fields_ = [('kernel_ms', C.c_double), ('wall_ms', C.c_double), ('launches', C.c_int64),
                ('h2d_bytes', C.c_int64), ('d2h_bytes', C.c_int64), ('bases', C.c_int64),
                ('stream_bytes', C.c_int64), ('exhaustive_reads', C.c_int64),
                ('candidates', C.c_int64)]


# every symbol include/kvarq_b200.h declares: name -> (restype, argtypes)
P = C.POINTER
VP = C.c_void_p
SIGNATURES = {
    'kvq_create': (C.c_int, [P(VP), C.c_int]),
    'kvq_destroy': (None, [VP]),
    'kvq_config': (C.c_int, [VP, P(Config)]),
    'kvq_get_config': (C.c_int, [VP, P(Config)]),
    'kvq_findseqs': (C.c_int, [VP, P(C.c_char_p), C.c_int, P(C.c_char_p), P(C.c_int32), C.c_int,
                               C.c_int, P(VP)]),
    'kvq_findseqs_mem': (C.c_int, [VP, P(C.c_void_p), P(C.c_int64), C.c_int, P(C.c_char_p),
                                   P(C.c_int32), C.c_int, C.c_int, P(VP)]),
    'kvq_stage_mem': (C.c_int, [VP, P(C.c_void_p), P(C.c_int64), C.c_int, P(VP)]),
    'kvq_scan_staged': (C.c_int, [VP, VP, P(C.c_char_p), P(C.c_int32), C.c_int, C.c_int, C.c_int,
                                  P(VP)]),
    'kvq_scan_staged_reduce': (C.c_int, [VP, VP, P(C.c_char_p), P(C.c_int32), C.c_int, C.c_int, C.c_int,
                                         P(VP), P(C.c_double)]),
    'kvq_staged_free': (None, [VP]),
    'kvq_host_alloc': (VP, [C.c_int64]),
    'kvq_host_free': (None, [VP]),
    'kvq_result_nhits': (C.c_int64, [VP]),
    'kvq_result_hits': (P(HitRec), [VP]),
    'kvq_result_hitseq_blob': (VP, [VP]),
    'kvq_result_hitseq_offsets': (P(C.c_int64), [VP]),
    'kvq_result_stats': (P(Stats), [VP]),
    'kvq_result_nseqhits': (P(C.c_int64), [VP]),
    'kvq_result_nseqbasehits': (P(C.c_int64), [VP]),
    'kvq_result_coverage': (P(C.c_int64), [VP, P(C.c_int64)]),
    'kvq_result_mutcount': (P(C.c_int64), [VP, P(C.c_int64)]),
    'kvq_result_timing': (P(Timing), [VP]),
    'kvq_result_free': (None, [VP]),
    'kvq_stats_snapshot': (C.c_int, [VP, P(Stats), P(C.c_int64), P(C.c_int64), C.c_int]),
    'kvq_stop': (C.c_int, [VP]),
    'kvq_last_error': (C.c_char_p, [VP, P(C.c_int)]),
    'kvq_comm_unique_id': (C.c_int, [C.c_char_p]),
    'kvq_comm_init': (C.c_int, [VP, C.c_char_p, C.c_int, C.c_int]),
    'kvq_comm_destroy': (C.c_int, [VP]),
    'kvq_comm_info': (C.c_int, [VP, P(C.c_int), P(C.c_int), P(C.c_int)]),
    'kvq_allreduce': (C.c_int, [VP, VP, P(C.c_double)]),
    'kvq_job_begin': (C.c_int, [VP, C.c_int, C.c_int64]),
    'kvq_job_row_words': (C.c_int64, [VP]),
    'kvq_job_store': (C.c_int, [VP, C.c_int, VP]),
    'kvq_job_allreduce': (C.c_int, [VP, P(C.c_int64), C.c_int64, P(C.c_double)]),
    'kvq_order_hits_static': (C.c_int, [C.c_int, VP, C.c_int64, VP, C.c_int32, VP, VP, VP]),
    'kvq_static_last_error': (C.c_char_p, []),
    'kvq_version': (C.c_char_p, []),
    'kvq_device_count': (C.c_int, []),
}

_lib = None


def _pick_nccl():
    """The library binds NCCL at run time (csrc/nccl_shim.h).  A process can only hold one libnccl.so.2
    (the loader matches by SONAME), so when PyTorch's bundled copy is installed that is the one to load
    -- whichever of the two libraries comes first -- unless $KVQ_NCCL_LIB says otherwise."""
    if os.environ.get('KVQ_NCCL_LIB'):
        return
    try:
        import importlib.util
        spec = importlib.util.find_spec('nvidia.nccl')
        for d in (spec.submodule_search_locations if spec else []):
            path = os.path.join(d, 'lib', 'libnccl.so.2')
            if os.path.exists(path):
                os.environ['KVQ_NCCL_LIB'] = path
                return
    except Exception:
        pass


def load():
    """Load libkvarq_b200.so and bind every entry point; raises if it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError('%s is missing: build it with `make -C kvarq_b200/csrc` '
                              '(or __graft_entry__.build()); there is no fallback' % LIB_PATH)
        _pick_nccl()
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)     # AttributeError = library does not match the header
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib
