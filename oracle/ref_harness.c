/*
 * TEST INFRASTRUCTURE ONLY -- not part of the product.
 *
 * Drives the reference's csrc/workhorse.c, compiled UNMODIFIED from where it lies
 * under /root/reference (see oracle/Makefile: -I$(REF)/csrc), exactly the way
 * engine_findseqs does (workhorse.c:1343-1408): fastq_open -> init_stats ->
 * nthreads x scan_filepart -> walk root->next.  Output goes to oracle/_ref/.
 */
#include "workhorse.c"      /* the reference translation unit, found via -I */
#include "oracle_abi.h"

PyObject stub_exc_io, stub_exc_runtime, stub_exc_memory, stub_exc_type;
static PyObject stub_exc_fastq, stub_hitseq_list;

/* captured hitseq bytes (add_hit holds ll_mutex while calling these) */
static char *cap_blob; static int64_t cap_blob_n, cap_blob_cap;
static int64_t *cap_off; static int64_t cap_n, cap_cap;

PyObject *PyString_FromStringAndSize(const char *s, long n)
{
    PyObject *o = (PyObject *) malloc(sizeof(PyObject));
    o->kind = 1; o->size = n; o->data = (char *) malloc(n > 0 ? n : 1);
    if (n > 0) memcpy(o->data, s, n);
    return o;
}

int PyList_Append(PyObject *list, PyObject *item)
{
    (void) list;
    if (cap_blob_n + item->size > cap_blob_cap) {
        cap_blob_cap = (cap_blob_n + item->size) * 2 + 4096;
        cap_blob = (char *) realloc(cap_blob, cap_blob_cap);
    }
    if (cap_n + 2 > cap_cap) {
        cap_cap = cap_cap * 2 + 1024;
        cap_off = (int64_t *) realloc(cap_off, cap_cap * sizeof(int64_t));
    }
    if (cap_n == 0) cap_off[0] = 0;
    memcpy(cap_blob + cap_blob_n, item->data, item->size);
    cap_blob_n += item->size;
    cap_off[++cap_n] = cap_blob_n;
    return 0;
}

void stub_decref(PyObject *o)
{
    if (o && o->kind == 1) { free(o->data); free(o); }
}

static int map_exception(void)
{
    if (exception == NULL) return ORC_OK;
    if (exception == &stub_exc_io) return ORC_ERR_IO;
    if (exception == &stub_exc_fastq) return ORC_ERR_FASTQ_FORMAT;
    if (exception == &stub_exc_memory) return ORC_ERR_MEMORY;
    return ORC_ERR_RUNTIME;
}

int ref_findseqs(const char **fnames, char **seqs, int *seqlens,
        int cfg_maxerrors, int cfg_minoverlap, int cfg_minreadlength, int cfg_nthreads,
        char cfg_Amin, orc_result *out)
{
    struct scanargs args;
    pthread_t *threads;
    ll_item *item;
    int t; int64_t i;

    memset(out, 0, sizeof(*out));
    maxerrors = cfg_maxerrors; minoverlap = cfg_minoverlap; minreadlength = cfg_minreadlength;
    nthreads = cfg_nthreads; Amin = cfg_Amin;
    fastq_exception = &stub_exc_fastq;
    exception = NULL; errstr[0] = 0; stop = 0; lo_log_i = 0;
    cap_blob_n = 0; cap_n = 0;

    args.seqlist = seqs; args.seqlengths = seqlens;
    args.fastq = fastq_open(fnames);
    if (args.fastq == NULL) {
        out->error = map_exception(); exception = NULL;
        strncpy(out->errstr, errstr, ORC_ERRSTR - 1);
        return out->error;
    }
    args.root = (ll_item *) malloc(sizeof(ll_item));
    args.root->fpos = -1L; args.root->next = NULL;
    args.pyhitseqs = &stub_hitseq_list;
    init_stats(args.seqlist);

    threads = (pthread_t *) malloc(sizeof(pthread_t) * nthreads);
    for (t = 0; t < nthreads; t++)
        pthread_create(threads + t, NULL, (void *(*)(void *)) scan_filepart, (void *) &args);
    for (t = 0; t < nthreads; t++)
        pthread_join(threads[t], NULL);
    free(threads);

    out->error = map_exception();
    strncpy(out->errstr, errstr, ORC_ERRSTR - 1);
    if (exception == NULL) {
        for (i = 0, item = args.root->next; item != NULL; item = item->next) i++;
        out->nhits = i;
        out->hits = (orc_hit *) malloc(sizeof(orc_hit) * (i > 0 ? i : 1));
        for (i = 0, item = args.root->next; item != NULL; item = item->next, i++) {
            out->hits[i].seq_nr = item->seqi; out->hits[i].file_pos = item->fpos;
            out->hits[i].seq_pos = item->spos; out->hits[i].length = item->length;
            out->hits[i].readlength = item->readlength;
        }
        out->hitseq_blob = (char *) malloc(cap_blob_n > 0 ? cap_blob_n : 1);
        memcpy(out->hitseq_blob, cap_blob, cap_blob_n);
        out->hitseq_off = (int64_t *) malloc(sizeof(int64_t) * (out->nhits + 1));
        out->hitseq_off[0] = 0;
        for (i = 1; i <= out->nhits; i++) out->hitseq_off[i] = cap_off[i];
        for (i = 0; i < MAX_READLENGTH; i++) out->readlengths[i] = rls_buf[i];
        out->rls_longest = rls_longest;
        out->records_parsed = records_parsed;
        out->parsed = (int64_t) fastq_parsed;
        out->total = (int64_t) fastq_size_estimated;
        out->nseqs = nseqs;
        out->nseqhits = (int64_t *) malloc(sizeof(int64_t) * (nseqs > 0 ? nseqs : 1));
        out->nseqbasehits = (int64_t *) malloc(sizeof(int64_t) * (nseqs > 0 ? nseqs : 1));
        for (i = 0; i < nseqs; i++) { out->nseqhits[i] = seqhits[i]; out->nseqbasehits[i] = seqbasehits[i]; }
    }
    exception = NULL;
    for (i = 0; i < lo_log_i; i++) free(lo_log_bufs[i]);
    lo_log_i = 0;
    free_ll(args.root);
    fastq_close(args.fastq);
    free(seqhits); free(seqbasehits); seqhits = seqbasehits = NULL;
    return out->error;
}

void ref_free(orc_result *r)
{
    free(r->hits); free(r->hitseq_blob); free(r->hitseq_off);
    free(r->nseqhits); free(r->nseqbasehits);
    memset(r, 0, sizeof(*r));
}
