/*
 * kvarq_b200.h -- C ABI of libkvarq_b200.so, the B200-native (sm_100a) replacement for the
 * read-scanning engine of KvarQ (reference: csrc/workhorse.c, Python module `kvarq.engine`).
 *
 * Plain C, plain pointers and sizes.  No Python, no PyTorch.  Every entry point names the
 * reference interface it replaces (file:line relative to the reference tree).  A Python-3
 * ctypes binding with the reference's exact module surface lives in kvarq_b200/engine.py;
 * INTEGRATION.md shows the binding a reference maintainer would add.
 *
 * There is NO CPU fallback: every scanning call fails with KVQ_ERR_RUNTIME when no CUDA
 * device (sm_100) is usable.
 */
#ifndef KVARQ_B200_H
#define KVARQ_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KVQ_MAX_READLENGTH 1024      /* MAX_READLENGTH, csrc/workhorse.c:105 */
#define KVQ_ERRSTR_LENGTH  1024      /* ERRSTR_LENGTH,  csrc/workhorse.c:96  */

/* error classes = the Python exception the reference raises (csrc/workhorse.c:577,1038,1454-1458) */
enum {
    KVQ_OK = 0,
    KVQ_ERR_IO = 1,            /* IOError                                  */
    KVQ_ERR_FASTQ_FORMAT = 2,  /* kvarq.fastq.FastqFileFormatException     */
    KVQ_ERR_RUNTIME = 3,       /* RuntimeError (also: CUDA failures)       */
    KVQ_ERR_MEMORY = 4,        /* MemoryError                              */
    KVQ_ERR_TYPE = 5           /* TypeError (bad arguments)                */
};

/* engine.config()/get_config() knobs (csrc/workhorse.c:69-75,1484-1507).  Defaults are the
 * reference's module defaults: 0, 20, 10, 1, '!', '!'.  `nthreads` here sizes the host
 * reader/inflate side; the scan itself runs on the GPU. */
typedef struct {
    int32_t maxerrors;
    int32_t minoverlap;
    int32_t minreadlength;
    int32_t nthreads;
    char Amin;
    char Azero;
} kvq_config_t;

/* kvarq.engine.Hit (csrc/workhorse.c:29-36,1579-1586) */
typedef struct {
    int32_t seq_nr;
    int32_t seq_pos;
    int64_t file_pos;
    int32_t length;
    int32_t readlength;
} kvq_hit_t;

/* engine.stats() (csrc/workhorse.c:1205-1244) */
typedef struct {
    int64_t readlengths[KVQ_MAX_READLENGTH];
    int64_t rls_longest;        /* len(stats['readlengths']) - 1; -1 before the first record */
    int64_t records_parsed;
    int64_t parsed;             /* inflated bytes consumed (fastq_parsed, :914)             */
    int64_t total;              /* size estimate (fastq_size_estimated, :677,:883)          */
    int64_t sigints;
    float progress;             /* min(parsed,total)/total in single precision (:1230-1232) */
    int32_t nseqs;
} kvq_stats_t;

/* timing / launch accounting of the last scan, for bench.py */
typedef struct {
    double kernel_ms;           /* sum of CUDA-event durations of the scan kernels          */
    double wall_ms;             /* host wall clock of the whole call                        */
    int64_t launches;           /* kernels launched                                          */
    int64_t h2d_bytes, d2h_bytes;
    int64_t bases;              /* untrimmed bases in parsed records (sequence-line bytes)  */
    int64_t stream_bytes;       /* inflated FastQ bytes scanned                              */
    int64_t exhaustive_reads;   /* reads that took the exhaustive (non-seeded) matcher      */
    int64_t candidates;         /* seed candidates verified                                  */
} kvq_timing_t;

typedef struct kvq_engine kvq_engine;      /* replaces the module globals of workhorse.c:69-110 */
typedef struct kvq_result kvq_result;      /* replaces the dict built at workhorse.c:1413-1441  */
typedef struct kvq_staged kvq_staged;      /* a FastQ stream resident in HBM (bench `value`)    */

/* module init / teardown (initengine, csrc/workhorse.c:1567).  device = CUDA ordinal. */
int kvq_create(kvq_engine **out, int device);
void kvq_destroy(kvq_engine *e);

/* engine.config / engine.get_config (csrc/workhorse.c:1498-1507 / 1484-1493) */
int kvq_config(kvq_engine *e, const kvq_config_t *cfg);
int kvq_get_config(kvq_engine *e, kvq_config_t *cfg);

/* engine.findseqs(fnames, seqlist) (csrc/workhorse.c:1249-1464).
 * fnames[nfiles]: paths, ".gz" suffix selects inflate (:582); scanned in order with a continuing
 * file_pos (:45).  seqs[nseqs]/seqlens[nseqs]: raw template bytes (any bytes; compared bytewise).
 * nplus > 0 declares seqs[nplus+i] == reverse-complement(seqs[i]) for i < nplus == nseqs/2
 * (kvarq/analyse.py:352-354) and enables the fused coverage output; 0 disables it.
 * Returns KVQ_OK and a result, or an error class (message via kvq_last_error). */
int kvq_findseqs(kvq_engine *e, const char *const *fnames, int nfiles,
                 const char *const *seqs, const int32_t *seqlens, int nseqs, int nplus,
                 kvq_result **out);

/* Same scan over in-memory FastQ "files" (host buffers; plain, already inflated).  The buffers
 * are scanned in place: batches are DMA'd from the caller's memory (fast when it is page-locked,
 * see kvq_host_alloc) and must stay valid until the call returns. */
int kvq_findseqs_mem(kvq_engine *e, const char *const *bufs, const int64_t *buflens, int nbufs,
                     const char *const *seqs, const int32_t *seqlens, int nseqs, int nplus,
                     kvq_result **out);

/* Page-locked host memory for kvq_findseqs_mem inputs: buffers allocated here are copied to the
 * GPU by DMA straight from where they lie (no staging copy on the host). */
void *kvq_host_alloc(int64_t nbytes);
void kvq_host_free(void *p);

/* Device-resident variant: stage once (host->HBM, not timed), scan many times. */
int kvq_stage_mem(kvq_engine *e, const char *const *bufs, const int64_t *buflens, int nbufs,
                  kvq_staged **out);
int kvq_scan_staged(kvq_engine *e, kvq_staged *st,
                    const char *const *seqs, const int32_t *seqlens, int nseqs, int nplus,
                    int want_hits, kvq_result **out);
void kvq_staged_free(kvq_staged *st);

/* result accessors (dict built at csrc/workhorse.c:1413-1441).  Hits are in the reference's
 * nthreads=1 emission order. */
int64_t kvq_result_nhits(const kvq_result *r);
const kvq_hit_t *kvq_result_hits(const kvq_result *r);
const char *kvq_result_hitseq_blob(const kvq_result *r);
const int64_t *kvq_result_hitseq_offsets(const kvq_result *r);     /* [nhits+1] */
const kvq_stats_t *kvq_result_stats(const kvq_result *r);
const int64_t *kvq_result_nseqhits(const kvq_result *r);           /* [nseqs]   */
const int64_t *kvq_result_nseqbasehits(const kvq_result *r);       /* [nseqs]   */
/* fused Coverage.apply_hit (kvarq/analyse.py:57-78): coverage[sum(len of plus seqs)],
 * mutcount[...][6] = alternate-base counts A,C,G,T,N,other on the plus strand; NULL if nplus==0 */
const int64_t *kvq_result_coverage(const kvq_result *r, int64_t *n);
const int64_t *kvq_result_mutcount(const kvq_result *r, int64_t *n);
const kvq_timing_t *kvq_result_timing(const kvq_result *r);
void kvq_result_free(kvq_result *r);

/* engine.stats() / engine.stop() -- callable from any thread while findseqs runs
 * (csrc/workhorse.c:1205-1244 / 1469-1479).  nseqhits/nseqbasehits may be NULL. */
int kvq_stats_snapshot(kvq_engine *e, kvq_stats_t *out, int64_t *nseqhits, int64_t *nseqbasehits,
                       int cap);
int kvq_stop(kvq_engine *e);

/* class + message of the last failed call on this engine (PyErr_SetString(exception, errstr),
 * csrc/workhorse.c:1454-1458) */
const char *kvq_last_error(kvq_engine *e, int *cls);

/* ---- multi-GPU: one process and one engine per GPU.  The reference has no counterpart --
 * its parallelism is nthreads pthreads sharing one file handle (csrc/workhorse.c:1375-1406); here the
 * read stream is dealt to the ranks (records are independent, templates replicated) and the
 * per-template counters of Coverage.apply_hit / add_hit (kvarq/analyse.py:57-78, workhorse.c:408-443)
 * are summed with one ncclAllReduce(int64) over NVLink.  NCCL is bound at run time (the copy already
 * in the process, else $KVQ_NCCL_LIB, else libnccl.so.2). */
#define KVQ_COMM_ID_BYTES 128
int kvq_comm_unique_id(char *id128);                       /* rank 0 makes it, the launcher hands it to every rank */
int kvq_comm_init(kvq_engine *e, const char *id128, int rank, int world);
int kvq_comm_destroy(kvq_engine *e);
int kvq_comm_info(kvq_engine *e, int *rank, int *world, int *nccl_version);
/* Sums the counters of the engine's last scan (`r`) over all ranks, in place on the device, and
 * writes the sums into r: coverage, mutcount, nseqhits, nseqbasehits, readlengths, records_parsed,
 * parsed, bases.  Hits stay per rank.  reduce_ms (may be NULL): device time of the collective. */
int kvq_allreduce(kvq_engine *e, kvq_result *r, double *reduce_ms);
/* kvq_scan_staged + kvq_allreduce in one call: the collective is enqueued right behind the scan kernels on
 * its own stream (the tail of the row -- readlengths, records, bases -- is written by a kernel from the
 * batch's device counters) and runs beside the kernels that order the hits; one host synchronisation per
 * scan instead of two.  *out holds this rank's hits and the counters summed over all ranks.  Every rank of
 * the communicator must make the call.  Should a rank have to run its batch again (survivor queue, line
 * phases), every rank sees that in the sum and the collective is repeated.  Without a communicator: the
 * same as kvq_scan_staged. */
int kvq_scan_staged_reduce(kvq_engine *e, kvq_staged *st,
                           const char *const *seqs, const int32_t *seqlens, int nseqs, int nplus,
                           int want_hits, kvq_result **out, double *reduce_ms);
/* Batch of isolates (docs/tutorial.rst:378 "hundreds of genomes"): [nslots][row] int64 matrix in HBM,
 * row = counters of one scan (kvq_job_row_words: nseqhits, nseqbasehits, coverage, mutcount[6],
 * readlengths[1024], records, bases, parsed, longest+1, 0); each rank stores the isolates it scanned,
 * one all-reduce at the end gives every rank the whole matrix. */
int kvq_job_begin(kvq_engine *e, int nslots, int64_t row_words);
int64_t kvq_job_row_words(kvq_engine *e);
int kvq_job_store(kvq_engine *e, int slot, const kvq_result *r);
int kvq_job_allreduce(kvq_engine *e, int64_t *out, int64_t cap_words, double *reduce_ms);

/* ---- ordering hits without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving
 * loops).  Puts `nhits` hits (device memory, any order) into the reference's nthreads=1 emission order
 * (file order, template, rules (a) (b) shifts descending, (c)/(d) ascending): d_sorted[r] = the hit of
 * rank r, d_order[r] (may be NULL) = its index in d_hits -- use it to gather the hitseqs.  d_seqlens[nseqs]
 * = template lengths (device memory).  One kernel on `stream` (a cudaStream_t, NULL = default stream),
 * not synchronised; a single-block variant of the kernels behind kvq_scan_staged whose scratch is a
 * file-scope device array of 2 MiB that every GPU loading the library holds until exit.  At most
 * KVQ_ORDER_STATIC_MAX hits (more: KVQ_ERR_TYPE, message via kvq_static_last_error).  Calls no
 * cudaMalloc / cudaFree.  All calls share the scratch: calls that overlap on the GPU race. */
#define KVQ_ORDER_STATIC_MAX 65536
int kvq_order_hits_static(int device, const kvq_hit_t *d_hits, int64_t nhits, const int32_t *d_seqlens,
                          int32_t nseqs, kvq_hit_t *d_sorted, uint32_t *d_order, void *stream);
const char *kvq_static_last_error(void);

/* build/device info, "sm_100a" etc. */
const char *kvq_version(void);
int kvq_device_count(void);

#ifdef __cplusplus
}
#endif
#endif
