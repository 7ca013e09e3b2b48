// Structures shared between the host engine (engine.cpp) and the sm_100a kernels (scan_kernel.cu).
// All pointers are device pointers unless a field name ends in _h.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

#define KVQ_SEED_K 8                    // seed length in bases (16-bit keys)
#define KVQ_NKEYS (1 << (2 * KVQ_SEED_K))
#define KVQ_BITMAP_WORDS (KVQ_NKEYS / 32)
#define KVQ_T1_LONG 16                  // small sets: pieces at least this long fire with their first 16 bases (two 8-mers 8 apart)
#define KVQ_T1X_K 12                   // T1 second level: a piece's first 12 bases (24-bit key), bitmap in L2
#define KVQ_T1X_MIN_PIECES 2048         // template sets with more T1 pieces run the kernel variant with the second level
#define KVQ_T1X_WORDS ((1 << (2 * KVQ_T1X_K)) / 32)
#define KVQ_HEAD_MORE 127u              // bucket-head size field [30:24]; 127 = read it from the overflow array
#define KVQ_HEAD_SHORT 0x80000000u      // bucket holds a piece shorter than KVQ_T1X_K: no second level for this key
#define KVQ_MAX_PIECES 8                // seeded path needs maxerrors+1 <= 8
#define KVQ_MAX_SEEDED_SEQLEN 32000     // diagonal must fit the candidate encoding
#define KVQ_HIST 1024                   // MAX_READLENGTH, workhorse.c:105
#define KVQ_MAXRL_FAST 512              // longer trimmed reads take the long-record kernel
#define KVQ_P2W (5 * KVQ_MAX_PIECES)     // fixed read offsets probed against T2, per read length (3 classes; 5 for large sets)
#define KVQ_X_HASH_BITS 20              // large sets: heads of the 12-mer table, addressed by a hash of the 24-bit key
#define KVQ_X_HASH(k24) ((uint32_t) ((k24) * 0x9E3779B1u) >> (32 - KVQ_X_HASH_BITS))
#define KVQ_P2_POS 0x3FFF               // p2_tab entry: read offset ...
#define KVQ_P2_LONG 0x4000              // ... | flag: only templates longer than the short class can match here
#define KVQ_ROW 1024                    // bytes one warp loads per step (32 lanes x 32 bytes)
#ifndef KVQ_OWN_ROWS
#define KVQ_OWN_ROWS 10
#endif
#ifndef KVQ_T2L
#define KVQ_T2L 1                       // separate T2 bitmap of the long templates for whole-read probes
#endif
#ifndef KVQ_TPK_SMEM_WORDS
#define KVQ_TPK_SMEM_WORDS 2048          // packed template codes up to this size are staged in shared memory
#endif
#define KVQ_TILE (KVQ_OWN_ROWS * KVQ_ROW) // bytes of a segment owned by one tile (host builds the tile table)

// 2-bit code of a raw byte: A->0 C->1 T->2 G->3; any other byte maps to *some* code.  Equal
// bytes always give equal codes, so exact-match seeds found in code space are complete; every
// accepted alignment is confirmed on the raw bytes afterwards.
#define KVQ_CODE(c) ((((unsigned) (c)) >> 1) & 3u)

struct KvqParams {
    int32_t maxerrors, minoverlap, minreadlength;
    int32_t amin;                       // signed char value of Amin
    int32_t t1x;                        // 0, or (many templates) KVQ_T1X_K * npieces: T1 seeds of templates at least this long pass the 12-mer second level
    int32_t seed_ok;                    // floor(minoverlap/(E+1)) >= K and E+1 <= KVQ_MAX_PIECES
    int32_t npieces;                    // E+1
    uint32_t pieces_inv;                // 2^32 / npieces + 1: floor(x / npieces) == __umulhi(x, pieces_inv) for x < 2^19 (x = j * seqlen)
    int32_t max_seqlen;                 // longest template (decides the whole-read T2 probes)
    int32_t want_hits;                  // 0: count only (no hit records / blob written)
    int32_t min_seed_rl;                // KVQ_SEED_K * npieces: shortest read the seed scheme covers
    int32_t o_piece[KVQ_MAX_PIECES];    // floor(j * minoverlap / npieces)
    int32_t nl_byte;                    // '\n' (a parameter, not an immediate: see nl_bits in scan_kernel.cu)
    int32_t n2max;                      // T2 probe slots per read: 3 * npieces (5 * npieces for large sets)
    int32_t accumulate;                 // 0: re-run after a buffer overflow, counters already taken
    int32_t exact_phase;                // 1: every tile waits for its line phase from the look-back chain (re-run after a
                                        //    tile's locally inferred phase was contradicted: malformed input)
};

struct KvqTemplates {
    const uint8_t *bytes;               // concatenated raw template bytes
    const uint32_t *byte_off;           // [nseqs+1]
    const uint32_t *packed;             // 2-bit codes, 16 bases per word, 2 zero pad words per seq
    const uint32_t *packed_off;         // [nseqs+1] word offsets
    int32_t packed_words;               // words in packed[]
    const int32_t *len;                 // [nseqs]
    const int32_t *exh_list;            // indices of non-seedable seqs
    int32_t nseqs, n_exh;
    // seed tables: presence per 16-bit key (T1: one byte per key, T2: one bit per key), CSR
    // offsets per key and entries {seq, toff | len << 16, packed_off, byte_off}
    // *_first[key] = first bucket entry {seq | more << 24, toff | len << 16, packed_off, overflow index}:
    // `more` further entries of the bucket start at *_ent[overflow index] (more == 255: the count is
    // in *_ent[overflow index - 1].x)
    const uint4 *t1_first, *t2_first;
    const uint8_t *t1_present; const uint4 *t1_ent;   // template-piece seeds
    const uint32_t *t2l_bitmap;         // T2 presence restricted to templates longer than short_max
    const uint32_t *t1x_bitmap;         // 2^24 bits: 12-mers that start a piece of a template with pieces >= 12 bases
    // large sets (KvqParams.t1x != 0): t2_first / t2_ent hold every T2 class back to back --
    // [W: all offsets, 65536 heads][H_j: head slots, 65536 each][T_j: tail slots, 65536 each]
    // [X: 12-mers at all offsets, 2^KVQ_X_HASH_BITS heads]; overflow indices are global in t2_ent.
    const uint32_t *t2e_bitmap;         // union of the H_j / T_j keys (staged in shared memory)
    const uint32_t *t2x_bitmap;         // 2^24 bits: 12-mers at any template offset
    const uint32_t *t2_bitmap; const uint4 *t2_ent;                  // every template offset
    const int16_t *p2_tab;              // [KVQ_MAXRL_FAST+1][KVQ_P2W] T2 probe offsets per read length, -1 = unused
    const uint32_t *cov_off;            // [nplus+1] offset of each plus template in coverage[]
    int32_t nplus;                      // 0 = no fused coverage
};

struct KvqBatch {
    const uint8_t *data;                // batch bytes, 16-byte aligned, followed by (KVQ_OWN_ROWS + 3) KiB of zeroed slack
    int64_t nbytes;
    const int64_t *seg_off;             // [nseg+1] record-aligned cuts (reference chunk emulation)
    int32_t nseg;
    int64_t base_fpos;                  // stream offset of data[0]
    const longlong2 *tile_info;         // [ntiles] {start of the tile (KVQ_TILE bytes of one segment),
                                        //           end of its segment | first tile of the segment << 62}
    uint32_t *tile_status;              // [ntiles] look-back words, zeroed before the launch
    int32_t ntiles;
};

struct KvqHitRec {                      // 40 bytes
    int32_t seq_nr, seq_pos;
    int64_t file_pos;
    int32_t length, readlength;
    int64_t blob_off;
    int32_t rule;                       // 0 = (a) read tail on template head, 1 = (b) read head on
    int32_t pad;                        //     template tail, 2 = (c)/(d) containment (emission order)
};

struct KvqLongRec {                     // work for the long-record kernel
    int64_t start;                      // batch offset: record start (kind 0) or trimmed read start (kind 1)
    int64_t seg_end;                    // batch offset of the end of the owning segment
    int32_t kind;                       // 0 = parse + account + match, 1 = match only
    int32_t rl;                         // kind 1: trimmed read length
};

struct KvqSurvRec {                     // work for the emit kernel, 32 bytes
    int64_t read_off;                   // batch offset of the trimmed read
    int32_t rl;
    int32_t seq;                        // >= 0: candidate alignment that passed the code-space check;
                                        // -1: match the read exhaustively against every seq; -2: against exh_list
    int32_t d;                          // diagonal (seq index aligned with read[0])
    int32_t kindpos;                    // seed that found it: kind << 16 | pos (T1 read offset / T2 slot)
    int32_t pad[2];
};

struct KvqCounters {                    // zeroed before every batch launch (except the accumulators)
    unsigned long long hit_count;       // hits wanted (may exceed capacity)
    unsigned long long blob_used;
    unsigned long long long_count;
    unsigned long long err_key;         // min over format errors, ~0 = none (see scan_kernel.cu)
    unsigned long long records, bases, exhaustive_reads, candidates;
    unsigned long long surv_count;      // survivor records wanted (may exceed capacity: then the emit kernel does nothing)
    unsigned long long rls[KVQ_HIST];
    long long longest;
    unsigned int tile_next;             // dynamic tile queue (tiles are claimed in stream order)
    unsigned int long_next;
    unsigned int overflow_rounds;
    unsigned int phase_retry;           // tiles whose inferred line phase was contradicted by the newline counts: the emit and
                                        // long-record kernels do nothing, the host re-runs the batch with exact_phase
};

struct KvqOut {
    KvqCounters *ctr;
    KvqHitRec *hits; unsigned long long hit_cap;
    uint8_t *blob; unsigned long long blob_cap;
    KvqLongRec *longs; unsigned long long long_cap;
    KvqSurvRec *surv; unsigned long long surv_cap;
    unsigned long long *nseqhits, *nseqbasehits;        // [nseqs]
    unsigned long long *coverage;                       // [cov_total]
    unsigned long long *mutcount;                       // [cov_total][6]
};

// launch wrappers (scan_kernel.cu); return cudaError_t as int
int kvq_launch_scan(const KvqBatch &b, const KvqTemplates &t, const KvqParams &p, const KvqOut &o,
                    int sm_count, void *stream);
int kvq_launch_long(const KvqBatch &b, const KvqTemplates &t, const KvqParams &p, const KvqOut &o,
                    int sm_count, void *stream);
int kvq_launch_emit(const KvqBatch &b, const KvqTemplates &t, const KvqParams &p, const KvqOut &o,
                    int sm_count, void *stream);
int kvq_scan_smem_bytes(void);

// ---- result marshalling on the device (HBM-resident scans): the hits of a batch in the reference's
// nthreads=1 order (workhorse.c:1107-1175: file order, template, rules (a) (b) shifts descending,
// (c)/(d) ascending), hitseqs back to back in that order.  Bucket sort over the stream position
// (as many buckets as hits), rank inside the (tiny) buckets by comparing full keys.
#define KVQ_ORDER_MAX 65536             // more hits than this: the host sorts (engine.cpp, finish_result)

struct KvqHitOut {                      // == kvq_hit_t (include/kvarq_b200.h)
    int32_t seq_nr, seq_pos;
    int64_t file_pos;
    int32_t length, readlength;
};

struct KvqOrder {
    ulonglong2 *keys;                   // [cap] {file_pos, seq_nr << 34 | rule << 32 | shift key}
    uint32_t *bkt;                      // [cap] bucket of hit i
    uint32_t *start;                    // [cap + 1] bucket starts (exclusive scan of the counts; zeroed before the launch)
    uint32_t *cursor;                   // [cap + 1]
    uint32_t *members;                  // [cap] hit indices grouped by bucket
    uint32_t *lens;                     // [cap + 1] hitseq offsets by rank (scan of the lengths), 32-bit: cap * 512 bytes at most
    unsigned long long *src;            // [cap] blob offset of the hit with rank r
    KvqHitOut *hits;                    // [cap] by rank
    long long *offsets;                 // [cap + 1] by rank
    uint8_t *blob;                      // [blob_cap] hitseqs by rank
    // page-locked host memory (mapped under unified addressing): the three arrays above are exported there by one
    // kernel with wide coalesced stores -- exactly the bytes of this scan, no copy sized on the host
    KvqHitOut *out_hits; long long *out_offsets; uint8_t *out_blob;
    uint32_t *status;                   // [0] number of ordered hits, 0xFFFFFFFF = not ordered (overflow, re-run pending, too many)
    uint32_t cap; uint32_t pad;
    unsigned long long blob_cap;
    long long span;                     // stream bytes of the batch (bucket scale)
};

int kvq_launch_order(const KvqBatch &b, const KvqOut &o, const KvqOrder &ord, void *stream);
// tail of the accumulator row from the batch's device counters (see scan_kernel.cu); tail[KVQ_HIST + 4] = 1 when the
// batch has to be run again (survivor queue overflow, contradicted line phase)
int kvq_launch_tail(const KvqCounters *ctr, unsigned long long *tail, long long parsed, unsigned long long surv_cap, void *stream);
// zeroed counters (err_key = ~0, longest = -1) and tile status words: what a batch starts from, one launch
int kvq_launch_reset(KvqCounters *ctr, uint32_t *tile_status, int ntiles, void *stream);
