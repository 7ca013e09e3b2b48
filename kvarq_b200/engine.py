"""Drop-in for the reference's ``kvarq.engine`` extension module (csrc/workhorse.c:1524-1596).

Same module surface -- ``config``, ``get_config``, ``findseqs``, ``stats``, ``stop``, ``test``
and the ``Hit`` namedtuple -- with the scan done by the sm_100a kernels of
libkvarq_b200.so.  Module-level state (the six config knobs, the running flag) mirrors the
reference's module globals (workhorse.c:69-75); :class:`Engine` is the explicit-handle form the
benchmark and multi-GPU driver use.

No GPU, no library => every scanning call raises; nothing here computes on the CPU.
"""
import collections
import ctypes as C
import os

from . import _lib as L
from .fastq import FastqFileFormatException

#: kvarq.engine.Hit (workhorse.c:1579-1586)
Hit = collections.namedtuple('Hit', 'seq_nr file_pos seq_pos length readlength')
Hit.__doc__ = (
    "seq_nr : refers to the list of sequences in call to engine.findseqs\n"
    "file_pos : beginning of read (within decompressed data)\n"
    "seq_pos : places the beginning of the read relative to the beginning of the sequence\n"
    "length : gives the number of overlapping basepairs\n"
    "readlength : length of the (quality trimmed) read containing the hit\n")

_EXC = {L.KVQ_ERR_IO: IOError, L.KVQ_ERR_FASTQ_FORMAT: FastqFileFormatException,
        L.KVQ_ERR_RUNTIME: RuntimeError, L.KVQ_ERR_MEMORY: MemoryError, L.KVQ_ERR_TYPE: TypeError}


def _as_bytes(s, what):
    if isinstance(s, bytes):
        return s
    if isinstance(s, str):
        return s.encode('latin-1')
    raise TypeError(what)


def _as_char(v):
    if isinstance(v, str):
        v = v.encode('latin-1')
    if not isinstance(v, bytes) or len(v) != 1:
        raise TypeError('must be char')         # PyArg "c" format, workhorse.c:1503
    return v


class Result(dict):
    """The dict findseqs returns, plus the fused outputs as attributes: ``coverage`` (int64
    numpy array over the concatenated plus templates) and ``mutcount`` ([..][6] alternate-base
    counts A,C,G,T,N,other), ``timing``, ``nhits``."""
    coverage = None
    mutcount = None
    timing = None


class Engine(object):
    """One engine handle on one GPU (kvq_create .. kvq_destroy)."""

    def __init__(self, device=None):
        self._lib = L.load()
        if device is None:
            device = int(os.environ.get('KVARQ_B200_DEVICE', os.environ.get('LOCAL_RANK', '0')))
            device %= max(1, self._lib.kvq_device_count())
        h = C.c_void_p()
        rc = self._lib.kvq_create(C.byref(h), device)
        if rc != 0:
            raise RuntimeError('kvarq_b200: no usable CUDA device %d (sm_100 required); '
                               'there is no CPU fallback' % device)
        self._h = h
        self.device = device

    def close(self):
        if getattr(self, '_h', None):
            self._free_kept()
            self._lib.kvq_destroy(self._h)
            self._h = None

    __del__ = close

    # -- error plumbing -------------------------------------------------------------------
    def _raise(self, rc):
        cls = C.c_int()
        msg = self._lib.kvq_last_error(self._h, C.byref(cls))
        raise _EXC.get(rc, RuntimeError)((msg or b'').decode('latin-1'))

    # -- engine.config / get_config (workhorse.c:1498-1507 / 1484-1493) --------------------
    def config(self, **kw):
        cfg = L.Config()
        self._lib.kvq_get_config(self._h, C.byref(cfg))
        for k, v in kw.items():
            if k in ('maxerrors', 'minoverlap', 'minreadlength', 'nthreads'):
                if isinstance(v, bool) or not isinstance(v, int):
                    raise TypeError('an integer is required')
                setattr(cfg, k, v)
            elif k in ('Amin', 'Azero'):
                setattr(cfg, k, _as_char(v))
            else:
                raise TypeError("'%s' is an invalid keyword argument for this function" % k)
        self._lib.kvq_config(self._h, C.byref(cfg))

    def get_config(self):
        cfg = L.Config()
        self._lib.kvq_get_config(self._h, C.byref(cfg))
        return {'maxerrors': cfg.maxerrors, 'minoverlap': cfg.minoverlap,
                'minreadlength': cfg.minreadlength, 'nthreads': cfg.nthreads,
                'Amin': cfg.Amin.decode('latin-1'), 'Azero': cfg.Azero.decode('latin-1')}

    # -- helpers ---------------------------------------------------------------------------
    def _seq_args(self, seqlist):
        key = tuple(seqlist) if isinstance(seqlist, (list, tuple)) else None
        cached = getattr(self, '_seq_cache', None)
        if key is not None and cached is not None and cached[0] == key:
            return cached[1]
        out = self._seq_args_build(seqlist)
        if key is not None:
            self._seq_cache = (key, out)
        return out

    @staticmethod
    def _seq_args_build(seqlist):
        try:
            iter(seqlist)
            n = len(seqlist)
        except TypeError:
            raise TypeError('seqlist must be sequence of strings')
        if isinstance(seqlist, (str, bytes)):
            seqlist = list(seqlist) if isinstance(seqlist, str) else [bytes([b]) for b in seqlist]
        bs = [_as_bytes(s, 'seqlist must be list of strings') for s in seqlist]
        arr = (C.c_char_p * max(1, n))(*bs)
        lens = (C.c_int32 * max(1, n))(*[len(b) for b in bs])
        return bs, arr, lens, n

    def _stats_dict(self, st, seqhits, seqbase):
        n = st.rls_longest + 1
        return {
            'readlengths': tuple(st.readlengths[i] for i in range(min(n, L.KVQ_MAX_READLENGTH))),
            'progress': float(st.progress),
            'nseqbasehits': tuple(seqbase),
            'nseqhits': tuple(seqhits),
            'parsed': int(st.parsed), 'total': int(st.total), 'sigints': int(st.sigints),
            'records_parsed': int(st.records_parsed),
        }

    def _free_kept(self):
        if getattr(self, '_kept', None):
            self._lib.kvq_result_free(self._kept)
            self._kept = None

    def _take_result(self, res, want_hits=True):
        lib = self._lib
        self._free_kept()
        self._kept = C.c_void_p(res.value)       # the C result of the last scan stays alive for allreduce() / job_store()
        try:
            n = lib.kvq_result_nhits(res)
            if want_hits == 'timing':           # bench loop: no Python objects beyond the counters
                out = Result(hits=(), stats={'records_parsed': lib.kvq_result_stats(res).contents.records_parsed},
                             hitseqs=[])
                t = lib.kvq_result_timing(res).contents
                out.timing = {k: getattr(t, k) for k, _ in L.Timing._fields_}
                out.nhits = n
                return out
            hits, hitseqs = (), []
            if want_hits and n:
                recs = lib.kvq_result_hits(res)
                offs = lib.kvq_result_hitseq_offsets(res)
                blob = C.string_at(lib.kvq_result_hitseq_blob(res), offs[n])
                hits = tuple(Hit(recs[i].seq_nr, recs[i].file_pos, recs[i].seq_pos,
                                 recs[i].length, recs[i].readlength) for i in range(n))
                hitseqs = [blob[offs[i]:offs[i + 1]].decode('latin-1') for i in range(n)]
            st = lib.kvq_result_stats(res).contents
            ns = st.nseqs
            sh = lib.kvq_result_nseqhits(res)
            sb = lib.kvq_result_nseqbasehits(res)
            out = Result(hits=hits, stats=self._stats_dict(st, [sh[i] for i in range(ns)],
                                                           [sb[i] for i in range(ns)]),
                         hitseqs=hitseqs)
            cn = C.c_int64()
            cov = lib.kvq_result_coverage(res, C.byref(cn))
            if cov and cn.value:
                import numpy as np
                out.coverage = np.ctypeslib.as_array(cov, shape=(cn.value,)).copy()
                mut = lib.kvq_result_mutcount(res, C.byref(cn))
                out.mutcount = np.ctypeslib.as_array(mut, shape=(cn.value // 6, 6)).copy()
            t = lib.kvq_result_timing(res).contents
            out.timing = {k: getattr(t, k) for k, _ in L.Timing._fields_}
            out.nhits = n
            return out
        finally:
            pass

    # -- engine.findseqs (workhorse.c:1249-1464) --------------------------------------------
    def findseqs(self, fname, seqlist, nplus=0, convert=True):
        """convert=False: the hit list is produced and copied back but not turned into Python tuples
        (result.nhits counts them); the scan itself is identical."""
        if isinstance(fname, (str, bytes)):
            fnames = [fname]
        else:
            try:
                fnames = list(fname)
            except TypeError:
                raise TypeError('fname must be [sequence of] string[s]')
        bf = [os.fsencode(f) if isinstance(f, (str, bytes)) else _as_bytes(f, 'fname must be [sequence of] string[s]')
              for f in fnames]
        _, arr, lens, n = self._seq_args(seqlist)
        c_fn = (C.c_char_p * max(1, len(bf)))(*bf)
        res = C.c_void_p()
        rc = self._lib.kvq_findseqs(self._h, c_fn, len(bf), arr, lens, n, nplus, C.byref(res))
        if rc != 0:
            self._raise(rc)
        return self._take_result(res, convert)

    def findseqs_mem(self, buffers, seqlist, nplus=0, want_hits=True):
        """Scan in-memory FastQ 'files' (bytes / numpy uint8 / anything with a buffer).
        want_hits=False only skips the Python-side conversion of the hit list (result.nhits
        still counts them); the scan itself is identical."""
        views, ptrs, lens64 = self._buffers(buffers)
        _, arr, lens, n = self._seq_args(seqlist)
        res = C.c_void_p()
        rc = self._lib.kvq_findseqs_mem(self._h, ptrs, lens64, len(views), arr, lens, n, nplus, C.byref(res))
        if rc != 0:
            self._raise(rc)
        return self._take_result(res, want_hits)

    @staticmethod
    def _buffers(buffers):
        import numpy as np
        arrs = [np.frombuffer(b, dtype=np.uint8) if not isinstance(b, np.ndarray)
                else np.ascontiguousarray(b).view(np.uint8).reshape(-1) for b in buffers]
        ptrs = (C.c_void_p * max(1, len(arrs)))(*[a.ctypes.data if a.size else None for a in arrs])
        lens64 = (C.c_int64 * max(1, len(arrs)))(*[a.size for a in arrs])
        return arrs, ptrs, lens64

    def pinned_copy(self, data):
        """Copy `data` (bytes / uint8 array) into page-locked host memory (kvq_host_alloc) and
        return it as a numpy uint8 array; free with release_pinned()."""
        import numpy as np
        src = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data.reshape(-1).view(np.uint8)
        ptr = self._lib.kvq_host_alloc(max(1, src.size))
        if not ptr:
            raise MemoryError('kvq_host_alloc failed')
        arr = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(max(1, src.size),))[:src.size]
        arr[:] = src
        self._pinned = getattr(self, '_pinned', {})
        self._pinned[arr.ctypes.data] = ptr
        return arr

    def pinned_empty(self, nbytes):
        """Page-locked host memory as a numpy uint8 array (kvq_host_alloc); free with release_pinned()."""
        import numpy as np
        ptr = self._lib.kvq_host_alloc(max(1, nbytes))
        if not ptr:
            raise MemoryError('kvq_host_alloc failed')
        arr = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_uint8)), shape=(max(1, nbytes),))
        self._pinned = getattr(self, '_pinned', {})
        self._pinned[arr.ctypes.data] = ptr
        return arr

    def release_pinned(self, arr):
        ptr = getattr(self, '_pinned', {}).pop(arr.ctypes.data, None)
        if ptr:
            self._lib.kvq_host_free(ptr)

    def stage(self, buffers):
        """Copy FastQ buffers into HBM once; returns a handle for scan_staged()."""
        views, ptrs, lens64 = self._buffers(buffers)
        st = C.c_void_p()
        rc = self._lib.kvq_stage_mem(self._h, ptrs, lens64, len(views), C.byref(st))
        if rc != 0:
            self._raise(rc)
        return Staged(self, st)

    def scan_staged(self, staged, seqlist, nplus=0, want_hits=True, convert=None, reduce=False):
        """want_hits=False: the kernels only count hits (no hit records / hitseqs written);
        convert=False: hits are produced and copied back but not turned into Python tuples;
        convert='timing': only nhits, records_parsed and the timing block are read;
        reduce=True (every rank of the communicator must make the call): the counters of the result
        are the sums over all ranks -- the all-reduce runs behind the scan kernels, beside the
        ordering of this rank's hits (kvq_scan_staged_reduce); timing['reduce_ms'] = its device time."""
        _, arr, lens, n = self._seq_args(seqlist)
        res = C.c_void_p()
        if reduce:
            ms = C.c_double()
            rc = self._lib.kvq_scan_staged_reduce(self._h, staged._h, arr, lens, n, nplus, 1 if want_hits else 0,
                                                  C.byref(res), C.byref(ms))
        else:
            rc = self._lib.kvq_scan_staged(self._h, staged._h, arr, lens, n, nplus, 1 if want_hits else 0,
                                           C.byref(res))
        if rc != 0:
            self._raise(rc)
        out = self._take_result(res, want_hits if convert is None else convert)
        if reduce:
            out.timing['reduce_ms'] = ms.value
        return out

    # -- multi-GPU: one engine per process and GPU, NCCL inside the library -------
    @staticmethod
    def unique_id():
        """128-byte NCCL id made by one rank; every rank passes it to comm_init()."""
        buf = C.create_string_buffer(128)
        if L.load().kvq_comm_unique_id(buf) != 0:
            raise RuntimeError('kvarq_b200: NCCL is not available (libnccl.so.2)')
        return buf.raw

    def comm_init(self, unique_id, rank, world):
        rc = self._lib.kvq_comm_init(self._h, unique_id, rank, world)
        if rc != 0:
            self._raise(rc)

    def comm_info(self):
        r, w, v = C.c_int(), C.c_int(), C.c_int()
        self._lib.kvq_comm_info(self._h, C.byref(r), C.byref(w), C.byref(v))
        return {'rank': r.value, 'world': w.value, 'nccl_version': v.value}

    def allreduce(self, result=None, convert=True):
        """Sum the counters of the last scan over all ranks (one ncclAllReduce of the device
        accumulators, int64) and return them as a Result (no hits: those stay per rank).
        result.timing['reduce_ms'] = device time of the collective."""
        if not getattr(self, '_kept', None):
            raise RuntimeError('allreduce() follows a scan')
        ms = C.c_double()
        rc = self._lib.kvq_allreduce(self._h, self._kept, C.byref(ms))
        if rc != 0:
            self._raise(rc)
        kept, self._kept = self._kept, None
        out = self._take_result(kept, 'timing' if convert == 'timing' else False)
        out.timing['reduce_ms'] = ms.value
        return out

    def job_begin(self, nslots):
        """Batch of isolates: one row of counters per isolate in HBM (call after the first scan of
        the template set, which fixes the row length)."""
        self._job = (nslots, self._lib.kvq_job_row_words(self._h))
        rc = self._lib.kvq_job_begin(self._h, nslots, self._job[1])
        if rc != 0:
            self._raise(rc)

    def job_store(self, slot):
        rc = self._lib.kvq_job_store(self._h, slot, self._kept)
        if rc != 0:
            self._raise(rc)

    def job_allreduce(self):
        """One ncclAllReduce over the whole [nslots][row] matrix; returns (matrix, reduce_ms)."""
        import numpy as np
        n, w = self._job
        out = np.zeros((n, w), dtype=np.int64)
        ms = C.c_double()
        rc = self._lib.kvq_job_allreduce(self._h, out.ctypes.data_as(C.POINTER(C.c_int64)), out.size, C.byref(ms))
        if rc != 0:
            self._raise(rc)
        return out, ms.value

    # -- engine.stats / engine.stop (workhorse.c:1205-1244 / 1469-1479) ----------------------
    def stats(self):
        st = L.Stats()
        cap = 1 << 16
        a = (C.c_int64 * cap)()
        b = (C.c_int64 * cap)()
        self._lib.kvq_stats_snapshot(self._h, C.byref(st), a, b, cap)
        n = min(st.nseqs, cap)
        return self._stats_dict(st, a[:n], b[:n])

    def stop(self):
        self._lib.kvq_stop(self._h)


class Staged(object):
    def __init__(self, engine, handle):
        self._e = engine
        self._h = handle

    def free(self):
        if self._h:
            self._e._lib.kvq_staged_free(self._h)
            self._h = None

    __del__ = free


# ------------------------------------------------------------------ module-level drop-in API --
_default = None


def _engine():
    global _default
    if _default is None:
        _default = Engine()
    return _default


def config(**kw):
    """config(**kwargs) -- configure the engine (maxerrors, minoverlap, minreadlength,
    nthreads, Amin, Azero)."""
    _engine().config(**kw)


def get_config():
    """get_config() -- get the current config as dictionary."""
    return _engine().get_config()


def findseqs(fname, sequences):
    """findseqs(fname, sequences) -- finds occurences of base sequences in fastq files.

    returns {'hits': tuple of Hit, 'stats': same dict as stats(), 'hitseqs': list of str}"""
    return _engine().findseqs(fname, sequences)


def stats():
    """stats() -- get statistics during scanning process."""
    return _engine().stats()


def stop():
    """stop() -- stops the scanning process."""
    _engine().stop()


def test():
    """test() -- profile dump; a no-op in the shipped reference build (workhorse.c:1512-1517)."""
    return None
