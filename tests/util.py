"""Shared helpers of the parity tests."""


def canon(result):
    """(hit tuple, hitseq) pairs sorted -- the multiset the north star compares."""
    hs = result['hitseqs']
    hs = [h.decode('latin-1') if isinstance(h, bytes) else h for h in hs]
    return sorted((tuple(h), s) for h, s in zip(result['hits'], hs))


def assert_same_result(got, want, order=True, what=''):
    """Bit-exact comparison of an engine result with a checker result (or golden case)."""
    g_hits = [tuple(h) for h in got['hits']]
    w_hits = [tuple(h) for h in want['hits']]
    g_seq = [h.decode('latin-1') if isinstance(h, bytes) else h for h in got['hitseqs']]
    w_seq = [h.decode('latin-1') if isinstance(h, bytes) else h for h in want['hitseqs']]
    assert len(g_hits) == len(w_hits), '%s: %d hits, expected %d' % (what, len(g_hits), len(w_hits))
    assert canon(got) == canon(want), what + ': hit multiset differs'
    if order:
        assert g_hits == w_hits, what + ': hit order differs from the reference nthreads=1 order'
        assert g_seq == w_seq, what + ': hitseq order differs'
    gs, ws = got['stats'], want['stats']
    for k in ('readlengths', 'nseqhits', 'nseqbasehits'):
        assert list(gs[k]) == list(ws[k]), '%s: stats[%s] differs' % (what, k)
    for k in ('records_parsed', 'parsed', 'total'):
        assert gs[k] == ws[k], '%s: stats[%s] %r != %r' % (what, k, gs[k], ws[k])
    assert abs(gs['progress'] - ws['progress']) < 1e-6, what + ': progress differs'


def coverage_from_hits(hits, hitseqs, seqs, nplus):
    """Py3 restatement of Coverage.apply_hit over all hits (kvarq/analyse.py:57-78, with
    Sequence.plus_idx/plus_base, kvarq/genes.py:264-276): per plus-strand base the depth and the
    alternate-base counts in columns A,C,G,T,N,other."""
    pairs = {'A': 'T', 'T': 'A', 'G': 'C', 'C': 'G', 'N': 'N'}
    col = {'A': 0, 'C': 1, 'G': 2, 'T': 3, 'N': 4}
    off = [0]
    for i in range(nplus):
        off.append(off[-1] + len(seqs[i]))
    cov = [0] * off[-1]
    mut = [[0] * 6 for _ in range(off[-1])]
    for hit, hs in zip(hits, hitseqs):
        seq_nr, _, seq_pos, length, _ = hit
        if isinstance(hs, bytes):
            hs = hs.decode('latin-1')
        seq = seqs[seq_nr]
        if isinstance(seq, bytes):
            seq = seq.decode('latin-1')
        plus = seq_nr < nplus
        tpl = seq_nr if plus else seq_nr - nplus
        start = max(0, seq_pos)
        for i, j in enumerate(range(start, start + length)):
            c_j = j if plus else len(seq) - j - 1
            cov[off[tpl] + c_j] += 1
            if hs[i] != seq[j]:
                b = hs[i] if plus else pairs.get(hs[i], hs[i])
                mut[off[tpl] + c_j][col.get(b, 5)] += 1
    return cov, mut
