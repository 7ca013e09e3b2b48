"""`python -m kvarq_b200.scan` -- the wiring of the reference's `kvarq scan` (kvarq/cli.py:57-196)
around the B200 engine: Fastq pre-flight -> engine.config(Amin=Q2A(Q), Azero, ...)
-> Analyser.scan in a worker thread while this thread polls engine.stats() (progress; a second
Ctrl-C within 2 s calls engine.stop()) -> testsuite calls -> JSON result file.

Testsuites are not discovered on disk and exec'd like the reference's (`kvarq/testsuites.py`): they
are built from tables (kvarq_b200/testsuites.py).  `-l` takes
    MTBC/phylo  MTBC/resistance  MTBC/spoligo  MTBC        with --genome <ancestor .bases / FASTA | synthetic>
    ebola  coll14  CanSNPer/<organism>                     from the bundled tables (testsuite_tables.json) or
                                                           a --tables JSON of the same layout
"""
import argparse
import json
import os
import signal
import sys
import threading
import time
from collections import OrderedDict

ERROR_COMMAND_LINE_SWITCH = -1      # kvarq/cli.py:25-27
ERROR_FASTQ_FORMAT_ERROR = -2
ERROR_FILE_EXISTS = -3

DEFAULTS = dict(quality=13, errors=2, readlength=25, overlap=25, threads=8, spacing=25)     # kvarq/config.py:2-10


def load_suites(names, genome=None, tables=None):
    """{name: Testsuite} for the -l selections"""
    from . import genes, synth, testsuites
    out = OrderedDict()
    G = None
    if genome == 'synthetic':
        G = genes.Genome.from_bases(synth.mtbc_genome(), 'MTB ancestor (synthetic)')
    elif genome:
        G = genes.Genome(genome, 'MTB ancestor')
    if not tables:                          # the tables generated from the reference's data files, shipped with the package
        bundled = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'testsuite_tables.json')
        tables = bundled if os.path.exists(bundled) else None
    data = json.load(open(tables)) if tables else None
    for name in names:
        if name == 'MTBC' or name.startswith('MTBC/'):
            if G is None and name != 'MTBC/spoligo':
                raise ValueError('%s needs --genome (the MTB ancestor genome is not shipped)' % name)
            if name == 'MTBC/spoligo':
                out[name] = testsuites.spoligo()
                continue
            suites = testsuites.mtbc(G)
            for k, v in suites.items():
                if name in ('MTBC', k):
                    out[k] = v
        elif data is None:
            raise ValueError('%s needs --tables' % name)
        elif name == 'ebola':
            g = genes.Genome.from_bases(data['ebola']['genome'], data['ebola']['identifier'])
            out[name] = testsuites.count_genotype(g, data['ebola']['snps'])
        elif name == 'coll14':
            if G is None:
                raise ValueError('coll14 needs --genome')
            out[name] = testsuites.barcode(G, data['coll14'], force=(genome == 'synthetic'))
        elif name.startswith('CanSNPer/'):
            d = data['cansnper'][name.split('/', 1)[1]]
            if not d['windows']:
                raise ValueError('%s: genome not available' % name)
            gs = dict((s, genes.SparseGenome(w['windows'], s, size=w['size'])) for s, w in d['windows'].items())
            out[name] = testsuites.snp_tree(gs, d['snps'], d['tree'])
        else:
            raise ValueError('unknown testsuite "%s"' % name)
    if len(out) == 0:
        raise ValueError('you must specify at least one testsuite (-l)')
    return out


def scan(args, engine=None, out=sys.stderr):
    """kvarq/cli.py:57-196.  Returns the Analyser."""
    from . import analyse
    from .fastq import Fastq, FastqFileFormatException
    if engine is None:
        from . import engine as engine_module
        engine = engine_module._engine()
    try:
        suites = load_suites(args.select or [], args.genome, args.tables)
    except ValueError as e:
        out.write('\n*** %s ***\n\n' % e)
        sys.exit(ERROR_COMMAND_LINE_SWITCH)
    try:
        fastq = Fastq(args.fastq, paired=not args.no_paired, variant=args.variant, quiet=True)
    except FastqFileFormatException as e:
        out.write('cannot open file %s : %s\n' % (args.fastq, e))
        sys.exit(ERROR_FASTQ_FORMAT_ERROR)
    engine.config(nthreads=args.threads, maxerrors=args.errors, Amin=fastq.Q2A(args.quality), Azero=fastq.Azero,
                  minreadlength=args.readlength, minoverlap=args.overlap)                  # cli.py:78-85
    if not args.force and os.path.exists(args.json):
        out.write('will not overwrite file %s\n' % args.json)
        sys.exit(ERROR_FILE_EXISTS)
    analyser = analyse.Analyser(spacing=args.spacing)
    out.write('scanning %s (%s)...\n' % (', '.join(fastq.filenames()),
                                         ', '.join('%.2f MB' % (s / 1024. ** 2) for s in fastq.filesizes())))
    t0 = time.time()
    state = {'finished': False, 'exception': None}

    def work():                                                                             # AnalyseThread, cli.py:106-126
        try:
            analyser.scan(fastq, suites, do_reverse=not args.no_reverse, engine=engine)
            state['finished'] = True
        except Exception as e:
            state['exception'] = e

    sig = {'n': 0, 't': 0.0}
    old_handler = None
    if threading.current_thread() is threading.main_thread():
        def on_sigint(signum, frame):                                                       # the reference counts SIGINT in C (:133-136)
            sig['n'] += 1
        old_handler = signal.signal(signal.SIGINT, on_sigint)
    th = threading.Thread(target=work, name='analyse-thread')
    th.start()
    seen, stats = 0, None
    try:
        while not state['finished'] and state['exception'] is None:                         # cli.py:134-175
            th.join(args.poll)
            stats = engine.stats()
            if not stats['records_parsed']:
                continue
            if args.progress:
                out.write('\r%5.1f%% %d records' % (100 * stats['progress'], stats['records_parsed']))
            if sig['n'] > seen:
                if time.time() - sig['t'] < 2.:
                    out.write('\n\n*** caught multiple <CTRL-C> within 2s : abort scanning ***\n')
                    engine.stop()
                    th.join()
                    break
                seen, sig['t'] = sig['n'], time.time()
                out.write('\nreadlengths: %s\n' % (stats['readlengths'],))
    finally:
        th.join()
        if old_handler is not None:
            signal.signal(signal.SIGINT, old_handler)
    if state['exception'] is not None:
        out.write('could not scan %s : %s\n' % (args.fastq, state['exception']))
        sys.exit(ERROR_FASTQ_FORMAT_ERROR)
    stats = analyser.stats
    out.write('\nperformed scanning of %.2f%% (%dmb/%dmb, %d records) in %.3f seconds\n' % (
        1e2 * stats['progress'], stats['parsed'] // 1024 ** 2, stats['total'] // 1024 ** 2, stats['records_parsed'], time.time() - t0))
    analyser.update_testsuites()
    with open(args.json, 'w', encoding='utf-8') as f:                                       # cli.py:189-193
        analyse.json_dump(analyser.encode(hits=args.hits), f)
    return analyser


def parser():
    p = argparse.ArgumentParser(prog='python -m kvarq_b200.scan', description='scan a .fastq[.gz] file and write the result .json')
    p.add_argument('-p', '--progress', action='store_true')
    p.add_argument('-l', '--select', action='append', help='testsuite to load (see module docstring); may be repeated')
    p.add_argument('--genome', help='MTB ancestor genome (.bases / FASTA) for the MTBC suites, or "synthetic"')
    p.add_argument('--tables', help='JSON with testsuite tables (ebola, coll14, CanSNPer); default: the bundled ones')
    p.add_argument('-t', '--threads', type=int, default=DEFAULTS['threads'])
    p.add_argument('-Q', '--quality', type=int, default=DEFAULTS['quality'])
    p.add_argument('-e', '--errors', type=int, default=DEFAULTS['errors'])
    p.add_argument('-r', '--readlength', type=int, default=DEFAULTS['readlength'])
    p.add_argument('-o', '--overlap', type=int, default=DEFAULTS['overlap'])
    p.add_argument('-s', '--spacing', type=int, default=DEFAULTS['spacing'])
    p.add_argument('-1', '--no-reverse', action='store_true')
    p.add_argument('-P', '--no-paired', action='store_true')
    p.add_argument('--variant', help='.fastq variant, in case the heuristic fails')
    p.add_argument('-f', '--force', action='store_true', help='overwrite an existing .json file')
    p.add_argument('-H', '--hits', action='store_true', help='save all hits in the .json file')
    p.add_argument('--poll', type=float, default=1.0, help='seconds between engine.stats() polls')
    p.add_argument('fastq')
    p.add_argument('json')
    return p


def main(argv=None):
    scan(parser().parse_args(argv))


if __name__ == '__main__':
    main()
