/*
 * TEST INFRASTRUCTURE ONLY -- not part of the product, never linked into or called by it.
 *
 * Plain-C restatement of the KvarQ read-scanning path (reference: csrc/workhorse.c).
 * It exists to check the CUDA engine (tests/, __graft_entry__.smoke()) and as the "port"
 * CPU baseline of bench.py.  Parity of this restatement is PINNED: tests/test_oracle.py
 * checks it against every known answer of the reference's tests/test_engine.py and,
 * where oracle/_ref/libkvarq_ref.so (the reference C compiled unmodified) is present,
 * differentially against the reference itself.
 *
 * Differences from the reference by design:
 *   - system zlib (raw inflate) instead of the vendored miniz 1.15 (csrc/gz/miniz.c);
 *   - with nthreads > 1 hits are gathered per chunk and concatenated in chunk order, so the
 *     result is deterministic and equals the nthreads=1 result (the reference races,
 *     workhorse.c:953-955 vs :756-770);
 *   - rl == 0 records are never matched (the reference reads an uninitialised
 *     `startlongest` there, workhorse.c:1064-1070; only reachable with minreadlength <= 0).
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <zlib.h>
#include "oracle_abi.h"

#define CHUNK (1024 * 1024)     /* SCANBUFSIZE, workhorse.c:15 */

/* ------------------------------------------------------------------ config ---- */
typedef struct {
    int maxerrors, minoverlap, minreadlength, nthreads;
    signed char amin;
} cfg_t;

/* ------------------------------------------------------------------ reader ---- */
typedef struct {
    const char **names; int next_name;
    FILE *fd; int gz; int at_eof;
    int64_t size_all;       /* sum of on-disk sizes        (fastq_file.size,   :43) */
    int64_t done_disk;      /* on-disk bytes of closed files (ftell0,          :44) */
    int64_t fpos;           /* inflated stream offset, never reset (fpos,      :45) */
    char *tail; int64_t tail_n;   /* deferred partial record (buf/buf_size,    :48-49) */
    unsigned char *zin; z_stream zs; int zs_live; int64_t remaining;
    /* stats mirrored from the reference globals (:100) */
    int64_t parsed, total;
    int err; char errstr[ORC_ERRSTR];
    pthread_mutex_t mu;
    int64_t chunk_seq;
} reader_t;

/* gzip member header, restating skip_gz_header (workhorse.c:482-541) */
static const char *gz_skip_header(FILE *fd, int dist)
{
    int st = 0, junk = 0, c = fgetc(fd);
    while (st != 2 && junk <= dist && c != EOF) {
        if (st == 0 && c == 0x1F) st = 1;
        else if (st == 1 && c == 0x8B) st = 2;
        else { st = 0; junk++; }
        c = fgetc(fd);
    }
    if (st != 2) return "magic bytes not found";
    if (c != 8) return "expected method==DEFLATED";
    int flags = fgetc(fd);
    if (flags & (0x02 | 0x20 | 0xC0))
        return "unsupported flags (CONTINUATION or ENCRYPTED or RESERVED)";
    for (int i = 0; i < 6; i++) (void) fgetc(fd);
    if (flags & 0x04) { int n = fgetc(fd); n |= fgetc(fd) << 8; while (n--) (void) fgetc(fd); }
    if (flags & 0x08) { do c = fgetc(fd); while (c > 0); }
    if (flags & 0x10) { do c = fgetc(fd); while (c > 0); }
    return NULL;
}

static int has_gz_suffix(const char *s)
{
    size_t n = strlen(s);
    return n >= 3 && strcmp(s + n - 3, ".gz") == 0;   /* workhorse.c:582 */
}

/* fastq_open_next (workhorse.c:559-629) */
static int reader_next_file(reader_t *r)
{
    if (r->next_name > 0) { r->done_disk += ftell(r->fd); fclose(r->fd); r->fd = NULL; }
    const char *name = r->names[r->next_name++];
    r->fd = fopen(name, "rb");
    if (!r->fd) { r->err = ORC_ERR_IO; snprintf(r->errstr, ORC_ERRSTR, "cannot open file"); return -1; }
    if (has_gz_suffix(name)) {
        r->gz = 1;      /* sticky, as fastq->compressed is never cleared (:585) */
        if (r->zs_live) { inflateEnd(&r->zs); r->zs_live = 0; }
        memset(&r->zs, 0, sizeof(r->zs));
        if (inflateInit2(&r->zs, -15) != Z_OK) {
            r->err = ORC_ERR_RUNTIME; snprintf(r->errstr, ORC_ERRSTR, "cannot mz_inflateInit()"); return -1;
        }
        r->zs_live = 1;
        if (!r->zin) r->zin = (unsigned char *) malloc(CHUNK);
        fseek(r->fd, 0, SEEK_END); r->remaining = ftell(r->fd); fseek(r->fd, 0, SEEK_SET);
        const char *msg = gz_skip_header(r->fd, 0);
        if (msg) {
            r->err = ORC_ERR_IO;
            snprintf(r->errstr, ORC_ERRSTR, "no valid gzip header found at beginning of file : %s", msg);
            return -1;
        }
        r->remaining -= ftell(r->fd);
        r->total *= 3;                                  /* "random guess" (:625) */
    }
    return 0;
}

/* fastq_open (workhorse.c:641-686) */
static int reader_open(reader_t *r, const char **names)
{
    memset(r, 0, sizeof(*r));
    pthread_mutex_init(&r->mu, NULL);
    r->names = names;
    for (int i = 0; names[i]; i++) {
        FILE *fd = fopen(names[i], "rb");
        if (!fd) {
            r->err = ORC_ERR_IO;
            snprintf(r->errstr, ORC_ERRSTR, "cannot open file '%s' for getting filesize", names[i]);
            return -1;
        }
        fseek(fd, 0, SEEK_END); r->size_all += ftell(fd); fclose(fd);
    }
    r->parsed = 0; r->total = r->size_all;
    return reader_next_file(r);
}

static void reader_close(reader_t *r)
{
    if (r->fd) fclose(r->fd);
    if (r->zs_live) inflateEnd(&r->zs);
    free(r->zin); free(r->tail);
    pthread_mutex_destroy(&r->mu);
}

/* fastq_rewind (workhorse.c:696-718): scanning backwards, a line starting with '+' arms the
 * search and the next line start (further back) beginning with '@' is the cut.  Returns the
 * number of trailing bytes to defer, or -1. */
static int64_t tail_length(const char *buf, int64_t n)
{
    int armed = 0;
    for (int64_t back = 1; back + 1 < n; back++) {
        char c = buf[n - back], prev = buf[n - back - 1];
        int line_start = (prev == '\n' || prev == '\r');
        if (c == '+' && line_start) armed = 1;
        else if (armed && c == '@' && line_start) return back;
    }
    return -1;
}

/* inflate up to `cap` bytes of the current .gz file (workhorse.c:781-885) */
static int64_t fill_gz(reader_t *r, char *dst, int64_t cap)
{
    int64_t got = 0;
    r->zs.next_out = (unsigned char *) dst;
    r->zs.avail_out = (unsigned int) cap;
    do {
        if (r->zs.avail_in == 0) {
            int64_t m = r->remaining < CHUNK ? r->remaining : CHUNK;
            if ((int64_t) fread(r->zin, 1, m, r->fd) != m) {
                r->err = ORC_ERR_IO;
                snprintf(r->errstr, ORC_ERRSTR, "could not read enough bytes from .fastq.gz");
                return -1;
            }
            r->zs.next_in = r->zin; r->zs.avail_in = (unsigned int) m; r->remaining -= m;
        }
        unsigned int before = r->zs.avail_out;
        int st = inflate(&r->zs, Z_SYNC_FLUSH);
        if (st == Z_BUF_ERROR && r->zs.avail_out == 0) st = Z_OK;     /* output full is not an error */
        if (st != Z_OK && st != Z_STREAM_END) {
            r->err = ORC_ERR_IO;
            snprintf(r->errstr, ORC_ERRSTR, "error while inflating compressed data : status=%d fpos=%ld",
                    st, (long) r->fpos);
            return -1;
        }
        got += before - r->zs.avail_out;
        /* multi-member files: another header may follow after <= 10 junk bytes (:842-866) */
        if (st == Z_STREAM_END && r->remaining + r->zs.avail_in > 10) {
            fseek(r->fd, -((long) r->zs.avail_in), SEEK_CUR);
            r->remaining += r->zs.avail_in;
            long pos = ftell(r->fd);
            if (gz_skip_header(r->fd, 10) != NULL) {
                /* reference logs "cannot read next deflated stream" and gives up on the file */
                r->remaining = 0; r->zs.avail_in = 0;
                break;
            }
            r->remaining -= ftell(r->fd) - pos;
            inflateReset2(&r->zs, -15);
            r->zs.avail_in = 0;
        } else if (st == Z_STREAM_END) {
            break;      /* only the trailer is left: nothing more to inflate */
        }
    } while (r->remaining + r->zs.avail_in > 10 && r->zs.avail_out > 0);
    if (r->zs.avail_out > 0) r->at_eof = 1;
    /* single-precision size estimate (:883-884) */
    float est = (float) r->size_all * (float) (uint64_t) (r->fpos + got)
              / (float) (uint64_t) (r->done_disk + ftell(r->fd));
    r->total = (int64_t) (uint64_t) est;
    return got;
}

/* fastq_read (workhorse.c:737-956).  Returns chunk length (0 = end of scan, -1 = error) and
 * the stream offset of buf[0]; *seq numbers the chunks in stream order. */
static int64_t reader_read(reader_t *r, char *buf, int64_t *first_off, int64_t *seq)
{
    int64_t ret;
    pthread_mutex_lock(&r->mu);
    if (r->err) { pthread_mutex_unlock(&r->mu); return -1; }
    if (r->at_eof && r->names[r->next_name] != NULL && reader_next_file(r) != 0) {
        pthread_mutex_unlock(&r->mu); return -1;
    }
    int64_t left = r->tail_n;
    if (left > 0) { memcpy(buf, r->tail, left); free(r->tail); r->tail = NULL; r->tail_n = 0; }
    *first_off = r->fpos - left;
    *seq = r->chunk_seq++;
    r->at_eof = 0;
    int64_t n;
    if (r->gz) {
        n = fill_gz(r, buf + left, CHUNK - left);
        if (n < 0) { pthread_mutex_unlock(&r->mu); return -1; }
    } else {
        n = (int64_t) fread(buf + left, 1, CHUNK - left, r->fd);
        if (ferror(r->fd)) {
            r->err = ORC_ERR_IO; snprintf(r->errstr, ORC_ERRSTR, "error while reading from file in fastq_read");
            pthread_mutex_unlock(&r->mu); return -1;
        }
        if (feof(r->fd)) r->at_eof = 1;
    }
    if (n == 0) { pthread_mutex_unlock(&r->mu); return left; }
    r->fpos += n; r->parsed += n;
    ret = left + n;
    if (!r->at_eof) {
        int64_t t = tail_length(buf, left + n);
        if (t < 0) {
            r->err = ORC_ERR_RUNTIME;
            snprintf(r->errstr, ORC_ERRSTR, "could find beginning of record; read %ld bytes up to %ld",
                    (long) n, ftell(r->fd));
            pthread_mutex_unlock(&r->mu); return -1;
        }
        r->tail = (char *) malloc(t); r->tail_n = t;
        memcpy(r->tail, buf + left + n - t, t);
        ret -= t;
    }
    pthread_mutex_unlock(&r->mu);
    return ret;
}

/* ------------------------------------------------------------------- hits ---- */
typedef struct {
    orc_hit *hits; int64_t n, cap;
    char *blob; int64_t blob_n, blob_cap;
    int64_t *off;                   /* [cap+1] */
} hitlist_t;

static void hl_push(hitlist_t *h, int seq_nr, int64_t file_pos, int seq_pos, int length, int rl,
        const char *bytes)
{
    if (h->n == h->cap) {
        h->cap = h->cap * 2 + 256;
        h->hits = (orc_hit *) realloc(h->hits, h->cap * sizeof(orc_hit));
        h->off = (int64_t *) realloc(h->off, (h->cap + 1) * sizeof(int64_t));
        h->off[0] = 0;
    }
    if (h->blob_n + length > h->blob_cap) {
        h->blob_cap = (h->blob_n + length) * 2 + 4096;
        h->blob = (char *) realloc(h->blob, h->blob_cap);
    }
    orc_hit *x = &h->hits[h->n];
    x->seq_nr = seq_nr; x->file_pos = file_pos; x->seq_pos = seq_pos; x->length = length; x->readlength = rl;
    if (length > 0) memcpy(h->blob + h->blob_n, bytes, length);
    h->blob_n += length;
    h->off[++h->n] = h->blob_n;
}

/* mismatches between a[0..n) and b[0..n), raw byte inequality (workhorse.c:1118-1120) */
static inline int within(const char *a, const char *b, int n, int maxerr)
{
    int e = 0;
    for (int j = 0; j < n; j++) {
        e += (a[j] != b[j]);
        if (e > maxerr) return 0;
    }
    return e <= maxerr;
}

/* The matcher for one trimmed read against one template (workhorse.c:1112-1174).
 * Emission order: (a) read tail on template head, longest shift first; (b) read head on
 * template tail; then (c) template inside read or (d) read inside template. */
static void match_one(const cfg_t *c, const char *read, int rl, int64_t read_off,
        const char *seq, int sl, int seq_nr, hitlist_t *out)
{
    const int O = c->minoverlap, E = c->maxerrors;
    if (rl > O && sl > O) {
        for (int i = rl - O; i > 0 && rl - i <= sl - 1; i--)            /* (a) :1116 */
            if (within(read + i, seq, rl - i, E))
                hl_push(out, seq_nr, read_off, -i, rl - i, rl, read + i);
        for (int i = sl - O; i > 0 && sl - i <= rl; i--)                /* (b) :1130 */
            if (within(seq + i, read, sl - i, E))
                hl_push(out, seq_nr, read_off, i, sl - i, rl, read);
    }
    if (rl > sl) {
        for (int i = 0; i <= rl - sl; i++)                              /* (c) :1147 */
            if (within(read + i, seq, sl, E))
                hl_push(out, seq_nr, read_off, -i, sl, rl, read + i);
    } else {
        for (int i = 0; i <= sl - rl; i++)                              /* (d) :1163 */
            if (within(seq + i, read, rl, E))
                hl_push(out, seq_nr, read_off, i, rl, rl, read);
    }
}

/* ------------------------------------------------------------------ chunks ---- */
typedef struct chunk_out_ {
    int64_t seq; hitlist_t hl; struct chunk_out_ *next;
} chunk_out;

typedef struct {
    const cfg_t *cfg; reader_t *rd;
    char **seqs; const int *seqlens; int nseqs;
    pthread_mutex_t mu;             /* guards everything below */
    chunk_out *done;
    int64_t rls[ORC_MAX_READLENGTH]; int64_t longest; int64_t records;
    int err; char errstr[ORC_ERRSTR]; int64_t err_chunk;
} job_t;

/* Record split + sanity checks + Phred trim + matcher for one chunk
 * (workhorse.c:1010-1187).  Returns 0, or -1 after setting *errstr. */
static int scan_chunk(job_t *job, const char *buf, int64_t n, int64_t first_off, hitlist_t *out,
        int64_t *rls, int64_t *longest, int64_t *records, char *errstr)
{
    const cfg_t *c = job->cfg;
    int64_t p = 0;
    while (p < n) {
        int64_t nl[4]; int k = 0; int64_t q = p;
        while (q < n && k < 4) { if (buf[q] == '\n') nl[k++] = q; q++; }
        if (k < 4) break;                                   /* partial record: dropped (:1033) */
        if (buf[p] != '@') {
            snprintf(errstr, ORC_ERRSTR, "record must start with '@' (and not '%c') fpos=%ld",
                    buf[p], (long) (first_off + p));
            return -1;
        }
        if (buf[nl[1] + 1] != '+') {
            snprintf(errstr, ORC_ERRSTR, "3rd line of record must start with '+' fpos=%ld",
                    (long) (first_off + nl[1] + 1));
            return -1;
        }
        (*records)++;
        /* Phred trim (:1055-1068): first longest run of quality bytes >= Amin; the byte that
         * closes a run is any byte < Amin up to and including the line's '\n'. */
        int64_t qs = nl[2] + 1, run = qs, best = 0; int rl = 0; int open = 1;
        for (int64_t t = qs; t <= nl[3]; t++) {
            if ((signed char) buf[t] >= c->amin) {
                if (!open) { open = 1; run = t; }
            } else if (open) {
                if (t - run > rl) { rl = (int) (t - run); best = run; }
                open = 0;
            }
        }
        if (rl < ORC_MAX_READLENGTH) rls[rl]++;             /* add_rl (:394-402) */
        if (rl > *longest) *longest = rl;
        p = q;
        if (rl < c->minreadlength || rl == 0) continue;     /* (:1100) */
        int64_t rd0 = nl[0] + 1 + (best - qs);              /* startread += startlongest-startscore (:1070) */
        for (int s = 0; s < job->nseqs; s++)
            match_one(c, buf + rd0, rl, first_off + rd0, job->seqs[s], job->seqlens[s], s, out);
    }
    return 0;
}

static void *worker(void *arg)
{
    job_t *job = (job_t *) arg;
    char *buf = (char *) malloc(CHUNK);
    int64_t n, off, seq;
    while ((n = reader_read(job->rd, buf, &off, &seq)) > 0) {
        chunk_out *co = (chunk_out *) calloc(1, sizeof(chunk_out));
        int64_t rls[ORC_MAX_READLENGTH]; int64_t longest = -1, records = 0;
        char errstr[ORC_ERRSTR];
        memset(rls, 0, sizeof(rls));
        co->seq = seq;
        int rc = scan_chunk(job, buf, n, off, &co->hl, rls, &longest, &records, errstr);
        pthread_mutex_lock(&job->mu);
        if (rc != 0 && (job->err == 0 || seq < job->err_chunk)) {
            job->err = ORC_ERR_FASTQ_FORMAT; job->err_chunk = seq;
            memcpy(job->errstr, errstr, ORC_ERRSTR);
        }
        for (int i = 0; i < ORC_MAX_READLENGTH; i++) job->rls[i] += rls[i];
        if (longest > job->longest) job->longest = longest;
        job->records += records;
        co->next = job->done; job->done = co;
        int stop = job->err != 0;
        pthread_mutex_unlock(&job->mu);
        if (stop) break;
    }
    free(buf);
    return NULL;
}

/* ------------------------------------------------------------------ driver ---- */
static int by_seq(const void *a, const void *b)
{
    int64_t x = (*(chunk_out * const *) a)->seq, y = (*(chunk_out * const *) b)->seq;
    return x < y ? -1 : x > y;
}

int orc_findseqs(const char **fnames, char **seqs, int *seqlens,
        int maxerrors, int minoverlap, int minreadlength, int nthreads, char Amin, orc_result *out)
{
    cfg_t cfg = { maxerrors, minoverlap, minreadlength, nthreads < 1 ? 1 : nthreads, (signed char) Amin };
    reader_t rd; job_t job;
    memset(out, 0, sizeof(*out));
    out->rls_longest = -1;
    if (reader_open(&rd, fnames) != 0) {
        out->error = rd.err; memcpy(out->errstr, rd.errstr, ORC_ERRSTR);
        if (rd.fd) { fclose(rd.fd); rd.fd = NULL; }
        reader_close(&rd);
        return out->error;
    }
    memset(&job, 0, sizeof(job));
    job.cfg = &cfg; job.rd = &rd; job.seqs = seqs; job.seqlens = seqlens; job.longest = -1;
    for (job.nseqs = 0; seqs[job.nseqs]; job.nseqs++);
    pthread_mutex_init(&job.mu, NULL);

    pthread_t *th = (pthread_t *) malloc(sizeof(pthread_t) * cfg.nthreads);
    for (int t = 0; t < cfg.nthreads; t++) pthread_create(&th[t], NULL, worker, &job);
    for (int t = 0; t < cfg.nthreads; t++) pthread_join(th[t], NULL);
    free(th);

    /* order the per-chunk hit lists by stream position */
    int64_t nchunks = 0;
    for (chunk_out *c = job.done; c; c = c->next) nchunks++;
    chunk_out **order = (chunk_out **) malloc(sizeof(chunk_out *) * (nchunks + 1));
    nchunks = 0;
    for (chunk_out *c = job.done; c; c = c->next) order[nchunks++] = c;
    qsort(order, nchunks, sizeof(chunk_out *), by_seq);

    if (rd.err) { out->error = rd.err; memcpy(out->errstr, rd.errstr, ORC_ERRSTR); }
    else if (job.err) { out->error = job.err; memcpy(out->errstr, job.errstr, ORC_ERRSTR); }

    if (!out->error) {
        int64_t nh = 0, nb = 0;
        for (int64_t i = 0; i < nchunks; i++) { nh += order[i]->hl.n; nb += order[i]->hl.blob_n; }
        out->nhits = nh;
        out->hits = (orc_hit *) malloc(sizeof(orc_hit) * (nh + 1));
        out->hitseq_blob = (char *) malloc(nb + 1);
        out->hitseq_off = (int64_t *) malloc(sizeof(int64_t) * (nh + 1));
        out->nseqs = job.nseqs;
        out->nseqhits = (int64_t *) calloc(job.nseqs + 1, sizeof(int64_t));
        out->nseqbasehits = (int64_t *) calloc(job.nseqs + 1, sizeof(int64_t));
        nh = 0; nb = 0; out->hitseq_off[0] = 0;
        for (int64_t i = 0; i < nchunks; i++) {
            hitlist_t *h = &order[i]->hl;
            for (int64_t k = 0; k < h->n; k++) {
                out->hits[nh] = h->hits[k];
                int64_t len = h->off[k + 1] - h->off[k];
                memcpy(out->hitseq_blob + nb, h->blob + h->off[k], len);
                nb += len; out->hitseq_off[++nh] = nb;
                out->nseqhits[h->hits[k].seq_nr]++;                  /* add_hit (:434-435) */
                out->nseqbasehits[h->hits[k].seq_nr] += h->hits[k].length;
            }
        }
        memcpy(out->readlengths, job.rls, sizeof(job.rls));
        out->rls_longest = job.longest;
        out->records_parsed = job.records;
        out->parsed = rd.parsed; out->total = rd.total;
    }
    for (int64_t i = 0; i < nchunks; i++) {
        free(order[i]->hl.hits); free(order[i]->hl.blob); free(order[i]->hl.off); free(order[i]);
    }
    free(order);
    pthread_mutex_destroy(&job.mu);
    reader_close(&rd);
    return out->error;
}

void orc_free(orc_result *r)
{
    free(r->hits); free(r->hitseq_blob); free(r->hitseq_off);
    free(r->nseqhits); free(r->nseqbasehits);
    memset(r, 0, sizeof(*r));
}
