"""Chunk-parallel inflate (kvarq_b200/csrc/gz_parallel.h, N4) on the CPU: the bytes and the read
positions (GzEvent: they feed the reference's size estimate, csrc/workhorse.c:883-884) of a parallel
decode must equal those of the sequential reader -- the round-1 code path, which the GPU tests pin
to the reference -- and the bytes must equal zlib's, for single- and multi-member files, every
compression level, stored blocks, incompressible stretches, truncated and tiny files."""
import gzip
import os
import subprocess
import zlib

import numpy as np
import pytest

from kvarq_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope='module')
def harness(tmp_path_factory):
    d = tmp_path_factory.mktemp('gzh')
    exe = str(d / 'gz_harness')
    subprocess.run(['g++', '-O2', '-std=c++17', '-o', exe, os.path.join(ROOT, 'tests', 'gz_harness.cpp'), '-lz', '-lpthread'],
                   check=True)
    return exe


def run(exe, threads, fn, out):
    p = subprocess.run([exe, str(threads), fn, out], capture_output=True, text=True, check=True)
    par = int(p.stderr.strip().split()[-1])
    return open(out, 'rb').read(), p.stdout, par


def inflate_members(blob):
    """All members of a gzip file, trailing junk ignored (what the reference reads)."""
    out = []
    while blob:
        d = zlib.decompressobj(31)
        try:
            out.append(d.decompress(blob))
        except zlib.error:
            break
        if not d.eof:
            break
        blob = d.unused_data
        k = blob.find(b'\x1f\x8b')
        if k < 0 or k > 10:
            break
        blob = blob[k:]
    return b''.join(out)


def test_parallel_equals_sequential(harness, tmp_path):
    g = synth.genome(500000, seed=3)
    fq = synth.fastq_reads(g, 90000, readlen=250, seed=5, qmean0=37.0, qdrop=20.0).tobytes()    # ~47 MB
    rnd = np.random.default_rng(1).integers(0, 256, 9000000, dtype=np.uint8).tobytes()
    cut = fq.index(b'\n@SYN', len(fq) // 3) + 1
    cases = {
        'level6': gzip.compress(fq, 6),
        'level1': gzip.compress(fq, 1),
        'level9': gzip.compress(fq[:len(fq) // 2], 9),
        'members': gzip.compress(fq[:cut], 6) + gzip.compress(b'', 6) + gzip.compress(fq[cut:], 6) + b'junk',
        'stored': gzip.compress(fq[:20000000], 0),
        'mixed': gzip.compress(rnd + fq[:30000000] + rnd[:3000000], 6),
        'truncated': gzip.compress(fq, 6)[:-2000000],
        'small': gzip.compress(fq[:2000000], 6),
        'empty': gzip.compress(b'', 6),
    }
    used_parallel = 0
    for name, blob in cases.items():
        fn = str(tmp_path / (name + '.fastq.gz'))
        open(fn, 'wb').write(blob)
        seq, ev_seq, _ = run(harness, 1, fn, str(tmp_path / 'seq.bin'))
        par, ev_par, nb = run(harness, 4, fn, str(tmp_path / 'par.bin'))
        assert par == seq, name + ': bytes differ'
        assert ev_par == ev_seq, name + ': read events differ'
        assert 'ERROR' not in ev_seq, name       # (a truncated stream just ends, as in the reference: workhorse.c:885)
        assert seq == inflate_members(blob), name + ': sequential reader differs from zlib'
        used_parallel += nb > 0
        if name in ('level6', 'level1', 'level9', 'members'):
            assert nb > 0.4 * len(seq), name + ': the parallel path was not taken'
    assert used_parallel >= 5


def bgzf(data, block=65280, junk=b''):
    """BGZF: one gzip member per <= 64 KiB of input, each with the 'BC' extra field (BSIZE) of the SAM spec;
    `junk` goes between the members (the reference tolerates <= 10 bytes incl. the 8-byte trailer)."""
    import struct
    out = bytearray()
    for i in range(0, len(data), block):
        c = zlib.compressobj(6, zlib.DEFLATED, -15)
        body = c.compress(data[i:i + block]) + c.flush()
        bsize = 12 + 6 + len(body) + 8 - 1
        out += b'\x1f\x8b\x08\x04' + b'\0' * 4 + b'\x00\xff' + struct.pack('<H', 6) + b'BC' + struct.pack('<HH', 2, bsize & 0xFFFF)
        out += body + struct.pack('<II', zlib.crc32(data[i:i + block]), len(data[i:i + block]) & 0xFFFFFFFF) + junk
    return bytes(out)


def test_files_of_many_short_members(harness, tmp_path):
    """BGZF-like files (the .fastq.gz of Illumina's converters): the chunk rounds give up after two members, the
    member rounds decode the members side by side; bytes and read positions as the sequential reader's"""
    g = synth.genome(300000, seed=4)
    fq = synth.fastq_reads(g, 100000, readlen=150, seed=6).tobytes()                   # ~33 MB, ~500 members
    plain = b''.join(gzip.compress(fq[i:i + 65280], 6) for i in range(0, len(fq), 65280))
    cases = {
        'bgzf': bgzf(fq) + bgzf(b''),                                                    # (with the empty EOF member)
        'members64k': plain,
        'junk2': bgzf(fq, junk=b'\x00\x01'),
        'members3m': b''.join(gzip.compress(fq[i:i + 3000000], 6) for i in range(0, len(fq), 3000000)),     # (> 1 MiB compressed each)
        'uneven': b''.join(gzip.compress(fq[a:b], 6) for a, b in zip([0, 30000, 9000000, 9100000, 20000000], [30000, 9000000, 9100000, 20000000, len(fq)])),
        'magic': b''.join(gzip.compress(fq[i:i + 65280].replace(b'@SYN', b'\x1f\x8b\x08\x00'), 0) for i in range(0, 8000000, 65280)) + plain,
    }
    for name, blob in cases.items():
        fn = str(tmp_path / (name + '.fastq.gz'))
        open(fn, 'wb').write(blob)
        seq, ev_seq, _ = run(harness, 1, fn, str(tmp_path / 'seq.bin'))
        for threads in (2, 5):
            par, ev_par, nb = run(harness, threads, fn, str(tmp_path / 'par.bin'))
            assert par == seq, name + ': bytes differ'
            assert ev_par == ev_seq, name + ': read events differ'
            if name in ('bgzf', 'members64k', 'junk2', 'members3m'):
                assert nb > 0.5 * len(seq), name + ': the member rounds were not taken'
        assert seq == inflate_members(blob), name + ': sequential reader differs from zlib'


def test_corrupt_inputs_end_the_same_way(harness, tmp_path):
    """Flipped bits, overwritten stretches, truncation, a stray member header: the parallel decoder never trusts
    what it cannot chain, so it ends like the sequential reader -- same bytes and read positions where the stream
    stays decodable, the same error where it does not (the bytes delivered in front of an error may differ: the
    engine raises IOError for the whole scan either way, csrc/workhorse.c:819-826)."""
    import random
    g = synth.genome(200000, seed=8)
    fq = synth.fastq_reads(g, 50000, readlen=150, seed=9).tobytes()                    # ~16 MB
    base = {'single': gzip.compress(fq, 6),
            'blocks': b''.join(gzip.compress(fq[i:i + 65280], 6) for i in range(0, len(fq), 65280)),
            'members': b''.join(gzip.compress(fq[i:i + 3000000], 6) for i in range(0, len(fq), 3000000))}
    rng = random.Random(5)
    for name, blob in base.items():
        for kind in range(4):
            b = bytearray(blob)
            if kind == 0:
                for _ in range(rng.randint(1, 5)):
                    b[rng.randrange(len(b))] ^= 1 << rng.randrange(8)
            elif kind == 1:
                a = rng.randrange(len(b) - 5000)
                b[a:a + rng.randint(1, 4000)] = bytes(rng.randrange(256) for _ in range(rng.randint(1, 4000)))
            elif kind == 2:
                b = b[:rng.randrange(len(b) // 2, len(b))]
            else:
                a = rng.randrange(len(b))
                b[a:a] = b'\x1f\x8b\x08\x00' + bytes(rng.randrange(256) for _ in range(20))
            fn = str(tmp_path / ('%s_%d.fastq.gz' % (name, kind)))
            open(fn, 'wb').write(bytes(b))
            seq, ev_seq, _ = run(harness, 1, fn, str(tmp_path / 'seq.bin'))
            par, ev_par, _ = run(harness, 4, fn, str(tmp_path / 'par.bin'))
            err_seq = [l for l in ev_seq.splitlines() if l.startswith('ERROR')]
            err_par = [l for l in ev_par.splitlines() if l.startswith('ERROR')]
            assert err_seq == err_par, (name, kind)
            if not err_seq:
                assert par == seq and ev_par == ev_seq, (name, kind)
