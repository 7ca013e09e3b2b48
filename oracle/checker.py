"""TEST INFRASTRUCTURE ONLY -- ctypes front-end for the two CPU checkers.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module; the product (kvarq_b200/) never does.

  impl="ref"     oracle/_ref/libkvarq_ref.so   reference csrc/workhorse.c compiled unmodified
  impl="oracle"  oracle/libkvarq_oracle.so     plain-C restatement (kvarq_oracle.c)

Both return the same dict shape as the reference's engine.findseqs
(csrc/workhorse.c:1434-1437): {'hits': tuple[Hit], 'stats': {...}, 'hitseqs': [bytes]}.
"""
import collections
import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))

Hit = collections.namedtuple('Hit', 'seq_nr file_pos seq_pos length readlength')

ERR_NAMES = {1: 'IOError', 2: 'FastqFileFormatException', 3: 'RuntimeError', 4: 'MemoryError'}


class OracleError(Exception):
    def __init__(self, kind, msg):
        Exception.__init__(self, '%s: %s' % (kind, msg))
        self.kind = kind
        self.msg = msg


class _Hit(C.Structure):
    _fields_ = [('seq_nr', C.c_int32), ('seq_pos', C.c_int32), ('file_pos', C.c_int64),
                ('length', C.c_int32), ('readlength', C.c_int32)]


class _Result(C.Structure):
    _fields_ = [('nhits', C.c_int64), ('hits', C.POINTER(_Hit)),
                ('hitseq_blob', C.POINTER(C.c_char)), ('hitseq_off', C.POINTER(C.c_int64)),
                ('readlengths', C.c_int64 * 1024), ('rls_longest', C.c_int64),
                ('records_parsed', C.c_int64), ('parsed', C.c_int64), ('total', C.c_int64),
                ('nseqs', C.c_int32), ('nseqhits', C.POINTER(C.c_int64)),
                ('nseqbasehits', C.POINTER(C.c_int64)), ('error', C.c_int32),
                ('errstr', C.c_char * 1024)]


_libs = {}


def lib_path(impl):
    if impl == 'ref':
        return os.path.join(HERE, '_ref', 'libkvarq_ref.so')
    return os.path.join(HERE, 'libkvarq_oracle.so')


def build(quiet=True):
    """(Re)build both checkers; the ref target is a no-op without /root/reference."""
    out = subprocess.run(['make', '-C', HERE, 'all'], capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError('oracle build failed:\n' + out.stdout + out.stderr)
    if not quiet:
        print(out.stdout)


def available(impl):
    return os.path.exists(lib_path(impl))


def _load(impl):
    if impl not in _libs:
        lib = C.CDLL(lib_path(impl))
        prefix = 'ref' if impl == 'ref' else 'orc'
        fn = getattr(lib, prefix + '_findseqs')
        fn.restype = C.c_int
        fn.argtypes = [C.POINTER(C.c_char_p), C.POINTER(C.c_char_p), C.POINTER(C.c_int),
                       C.c_int, C.c_int, C.c_int, C.c_int, C.c_char, C.POINTER(_Result)]
        fr = getattr(lib, prefix + '_free')
        fr.restype = None
        fr.argtypes = [C.POINTER(_Result)]
        _libs[impl] = (fn, fr)
    return _libs[impl]


def findseqs(fnames, seqs, maxerrors=0, minoverlap=20, minreadlength=10, nthreads=1,
             Amin=b'!', impl='ref', want_hits=True):
    """Run one CPU scan.  Defaults are the reference's module defaults (workhorse.c:71-75)."""
    if isinstance(fnames, (str, bytes)):
        fnames = [fnames]
    fn, fr = _load(impl)
    bf = [f.encode() if isinstance(f, str) else f for f in fnames]
    bs = [s.encode('latin-1') if isinstance(s, str) else s for s in seqs]
    c_fn = (C.c_char_p * (len(bf) + 1))(*bf, None)
    c_sq = (C.c_char_p * (len(bs) + 1))(*bs, None)
    c_ln = (C.c_int * max(1, len(bs)))(*[len(s) for s in bs])
    if isinstance(Amin, str):
        Amin = Amin.encode('latin-1')
    res = _Result()
    rc = fn(c_fn, c_sq, c_ln, maxerrors, minoverlap, minreadlength, nthreads, Amin, C.byref(res))
    try:
        if rc != 0:
            raise OracleError(ERR_NAMES.get(rc, 'RuntimeError'), res.errstr.decode('latin-1'))
        n = res.nhits
        hits, hitseqs = (), []
        if want_hits:
            blob = C.string_at(res.hitseq_blob, res.hitseq_off[n]) if n else b''
            hits = tuple(Hit(res.hits[i].seq_nr, res.hits[i].file_pos, res.hits[i].seq_pos,
                             res.hits[i].length, res.hits[i].readlength) for i in range(n))
            hitseqs = [blob[res.hitseq_off[i]:res.hitseq_off[i + 1]] for i in range(n)]
        total = res.total
        progress = 0.0
        if total > 0:
            # workhorse.c:1230-1232 -- single precision
            progress = float(C.c_float(float(C.c_float(min(res.parsed, total)).value) / total).value)
        stats = {
            'readlengths': tuple(res.readlengths[i] for i in range(min(res.rls_longest + 1, 1024))),  # the reference reads past rls_buf beyond 1024 (workhorse.c:1214-1216)
            'progress': progress,
            'nseqbasehits': tuple(res.nseqbasehits[i] for i in range(res.nseqs)),
            'nseqhits': tuple(res.nseqhits[i] for i in range(res.nseqs)),
            'parsed': res.parsed, 'total': total, 'sigints': 0,
            'records_parsed': res.records_parsed,
        }
        return {'hits': hits, 'stats': stats, 'hitseqs': hitseqs, 'nhits': n}
    finally:
        fr(C.byref(res))
