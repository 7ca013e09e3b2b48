/*
 * TEST INFRASTRUCTURE ONLY -- not part of the product.
 *
 * Result layout shared by the two CPU checkers:
 *   oracle/_ref/libkvarq_ref.so   = the reference's csrc/workhorse.c, unmodified (ref_harness.c)
 *   oracle/libkvarq_oracle.so     = plain-C restatement of the same algorithm (kvarq_oracle.c)
 * Both export   int <prefix>_findseqs(...)  /  void <prefix>_free(...)  with this struct.
 */
#ifndef KVQ_ORACLE_ABI_H
#define KVQ_ORACLE_ABI_H

#include <stdint.h>

#define ORC_MAX_READLENGTH 1024   /* workhorse.c:105 */
#define ORC_ERRSTR 1024           /* workhorse.c:96  */

enum { ORC_OK = 0, ORC_ERR_IO = 1, ORC_ERR_FASTQ_FORMAT = 2, ORC_ERR_RUNTIME = 3, ORC_ERR_MEMORY = 4 };

typedef struct {
    int32_t seq_nr;      /* ll_item.seqi       workhorse.c:30 */
    int32_t seq_pos;     /* ll_item.spos       workhorse.c:32 */
    int64_t file_pos;    /* ll_item.fpos       workhorse.c:31 */
    int32_t length;      /* ll_item.length     workhorse.c:33 */
    int32_t readlength;  /* ll_item.readlength workhorse.c:34 */
} orc_hit;

typedef struct {
    int64_t nhits;
    orc_hit *hits;             /* [nhits] in emission order                       */
    char *hitseq_blob;         /* concatenated hitseq bytes                       */
    int64_t *hitseq_off;       /* [nhits+1]                                       */
    int64_t readlengths[ORC_MAX_READLENGTH];
    int64_t rls_longest;       /* -1 when no record parsed                        */
    int64_t records_parsed;
    int64_t parsed;            /* fastq_parsed            workhorse.c:914         */
    int64_t total;             /* fastq_size_estimated    workhorse.c:677,883     */
    int32_t nseqs;
    int64_t *nseqhits;         /* [nseqs]                                         */
    int64_t *nseqbasehits;     /* [nseqs]                                         */
    int32_t error;             /* ORC_*                                           */
    char errstr[ORC_ERRSTR];
} orc_result;

#endif
