// sm_100a kernels of the KvarQ read-scanning path.
//
// One persistent CTA takes record-aligned segments of the FastQ stream from a queue and walks
// each of them tile by tile (32 KiB + spill margin, staged in shared memory).  Per tile:
//   P1  newline bitmap (SWAR compare on 128-bit loads) + block prefix scan   [workhorse.c:1018-1030]
//   P2  record table: record k of a segment = its lines 4k..4k+3             [workhorse.c:1010-1052]
//   P3  '@' / '+' sanity checks, Phred trim (first longest run >= Amin),
//       read-length histogram, length filter                                [workhorse.c:1037-1103]
//   P4  2-bit packing of the trimmed reads
//   P5  seed probes: every read offset against the template-piece table (T1), a few fixed read
//       offsets against the all-offsets table (T2); bitmaps live in shared memory
//   P6  bucket expansion -> candidate diagonals
//   P7  XOR/popc Hamming verify on packed words, byte-exact confirm, hit emission through
//       atomics, fused coverage                                             [workhorse.c:1107-1175,408-443]
// Reads or templates the seed scheme cannot cover take the exhaustive matcher (same acceptance
// rule, every diagonal).  Acceptance rule: SURVEY.md Appendix B / workhorse.c:1112-1174.
#include <cuda_runtime.h>
#include <stdint.h>
#include "kvq_device.h"

namespace {

constexpr int THREADS = 512;
constexpr int NWARPS = THREADS / 32;
constexpr int TILE = 32768;                 // bytes of a segment owned by one tile
constexpr int MARGIN = 8192;                // spill: records starting in the tile may end here
constexpr int LOADMAX = TILE + MARGIN + 16; // + alignment slack of the first 16-byte chunk
constexpr int NLW = LOADMAX / 32 + 2;       // newline bitmap words
constexpr int CAP = 256;                    // records per round
constexpr int MAXRL_FAST = 512;             // longer trimmed reads go to the long-record kernel
constexpr int SLOTW = MAXRL_FAST / 16 + 2;  // packed words per read slot
constexpr int SEEDQ = 4096;
constexpr int CANDQ = 4096;
constexpr int MAXP2 = 3 * KVQ_MAX_PIECES;

struct Rec { uint16_t start; uint16_t nl[4]; };
struct Rd { uint16_t pos; uint16_t rl; };

struct Smem {
    uint4 tile[LOADMAX / 16 + 4];      // + slack for word-aligned over-reads
    uint32_t nlmask[NLW];
    uint32_t nlpref[NLW + 1];
    Rec rec[CAP];
    Rd rd[CAP];
    uint32_t packed[CAP * SLOTW];
    uint32_t bm1[KVQ_BITMAP_WORDS];
    uint32_t bm2[KVQ_BITMAP_WORDS];
    uint32_t seedq[SEEDQ];
    unsigned long long candq[CANDQ];
    uint32_t hist[KVQ_HIST];
    uint16_t exh_reads[CAP];
    uint16_t seed_reads[CAP];
    uint32_t warp_tot[NWARPS];
    // per-round / per-tile scalars
    int nrd, nseed, ncand, n_exh_reads, n_seed_reads, overflow;
    int seg_cur;
    // per-CTA accumulators
    unsigned long long records, bases, exh_count, cand_count;
    int longest;
};

__device__ __forceinline__ uint32_t funnel_r(uint32_t lo, uint32_t hi, uint32_t s)
{
    return __funnelshift_r(lo, hi, s);      // (hi:lo >> (s & 31)) low word
}

// 4-bit mask of the bytes of w equal to '\n'
__device__ __forceinline__ uint32_t nl_nibble(uint32_t w)
{
    uint32_t t = w ^ 0x0A0A0A0Au;
    uint32_t z = ((t & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | t;     // bit7 set <=> byte != 0
    uint32_t m = (~z & 0x80808080u) >> 7;
    return (m * 0x10204080u) >> 28;
}

// 4-bit mask of bytes b of w with (signed char) b >= amin.  cl = (amin & 0x7f) * 0x01010101,
// chi = amin has bit 7 set (negative).
__device__ __forceinline__ uint32_t ge_nibble(uint32_t w, uint32_t cl, bool chi)
{
    // compare (b ^ 0x80) >= (amin ^ 0x80) as unsigned bytes
    uint32_t t = (w | 0x80808080u) - cl;                    // bit7: low 7 bits of b >= low 7 bits of amin
    uint32_t uhi = ~w;                                      // bit7 of (b ^ 0x80)
    // amin^0x80 has bit7 = !chi
    uint32_t ge = chi ? (uhi | t) : (uhi & t);
    uint32_t m = (ge & 0x80808080u) >> 7;
    return (m * 0x10204080u) >> 28;
}

// 4 bases (bytes of w) -> 8 bits of 2-bit codes, base 0 in the low bits
__device__ __forceinline__ uint32_t pack4(uint32_t w)
{
    return (((w >> 1) & 0x03030303u) * 0x01041040u) >> 24;
}

// 16 bases starting at base offset `b` of a packed array (needs word b/16+1 readable)
__device__ __forceinline__ uint32_t extract16(const uint32_t *pk, int b)
{
    int w = b >> 4;
    return funnel_r(pk[w], pk[w + 1], (uint32_t) ((b & 15) << 1));
}

__device__ __forceinline__ uint32_t ldg32(const uint32_t *p) { return __ldg(p); }

__device__ __forceinline__ uint32_t extract16_g(const uint32_t *pk, int b)
{
    int w = b >> 4;
    return funnel_r(__ldg(pk + w), __ldg(pk + w + 1), (uint32_t) ((b & 15) << 1));
}

// Is diagonal d (= seq index aligned with read[0]) one the reference evaluates?  Returns the
// number of hits an accepted alignment yields there (0, 1 or 2 -- the (b)+(d) duplicate,
// workhorse.c:1130-1141 + 1163-1173) and the overlap geometry.
__device__ __forceinline__ int diag_rule(int d, int rl, int sl, int O, int &ro, int &to, int &L, int &rule)
{
    rule = 2;
    ro = d < 0 ? -d : 0;
    to = d > 0 ? d : 0;
    int lr = rl - ro, ls = sl - to;
    L = lr < ls ? lr : ls;
    if (L <= 0) return 0;
    if (d <= 0 && d + rl >= sl) {
        // template inside read: rule (c) needs rl > sl, else d == 0 && rl == sl is rule (d)
        return 1;
    }
    if (d >= 0 && d + rl <= sl) {
        // read inside template, rule (d); flush-right also satisfies rule (b)
        if (d > 0 && d + rl == sl && rl > O && sl > O) return 2;
        return 1;
    }
    // partial overlaps, rules (a) / (b)
    rule = d < 0 ? 0 : 1;
    if (rl > O && sl > O && L >= O) return 1;
    return 0;
}

} // namespace

namespace {

struct TileCtx {
    int64_t a0;          // batch offset of smem byte 0 (16-byte aligned)
    int shift;           // smem coordinate of the first byte of the tile
    int tile_end;        // smem coordinate one past the bytes owned by the tile
    int reg_end;         // smem coordinate one past the loaded region
    int first;           // first tile of its segment
    int at_seg_end;      // the region reaches the end of the segment
    int phase_in;        // newlines of the segment before the tile, mod 4
    int64_t seg_end;     // batch offset of the segment end
};

// block-wide exclusive scan of one value per thread; returns the exclusive prefix, *total = sum
__device__ __forceinline__ uint32_t block_exscan(uint32_t v, uint32_t *warp_tot, uint32_t *total)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    if (lane == 31) warp_tot[wid] = inc;
    __syncthreads();
    uint32_t base = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < NWARPS; w++) {
        uint32_t t = warp_tot[w];
        if (w < wid) base += t;
        tot += t;
    }
    __syncthreads();
    *total = tot;
    return base + inc - v;
}

// P1: stage the region in shared memory and build the newline bitmap + prefix counts.
// Returns (to every thread) the number of newlines in the whole region.
__device__ uint32_t load_and_index(Smem &s, const KvqBatch &b, const TileCtx &t)
{
    const int nchunks = (t.reg_end + 15) >> 4;
    const uint4 *src = reinterpret_cast<const uint4 *>(b.data + t.a0);
    uint16_t *nl16 = reinterpret_cast<uint16_t *>(s.nlmask);
    for (int c = threadIdx.x; c < NLW * 2; c += THREADS) {
        uint32_t m = 0;
        if (c < nchunks) {
            uint4 v = __ldg(src + c);
            s.tile[c] = v;
            m = nl_nibble(v.x) | (nl_nibble(v.y) << 4) | (nl_nibble(v.z) << 8) | (nl_nibble(v.w) << 12);
            int lo = t.shift - (c << 4), hi = t.reg_end - (c << 4);
            if (lo > 0) m &= lo >= 16 ? 0u : (0xFFFFu << lo);
            if (hi < 16) m &= hi <= 0 ? 0u : ((1u << hi) - 1u);
        }
        nl16[c] = (uint16_t) m;
    }
    __syncthreads();
    // three bitmap words per thread
    const int w0 = threadIdx.x * 3;
    uint32_t c0 = 0, c1 = 0, c2 = 0;
    if (w0 < NLW) c0 = __popc(s.nlmask[w0]);
    if (w0 + 1 < NLW) c1 = __popc(s.nlmask[w0 + 1]);
    if (w0 + 2 < NLW) c2 = __popc(s.nlmask[w0 + 2]);
    uint32_t total;
    uint32_t ex = block_exscan(c0 + c1 + c2, s.warp_tot, &total);
    if (w0 < NLW) s.nlpref[w0] = ex;
    if (w0 + 1 < NLW) s.nlpref[w0 + 1] = ex + c0;
    if (w0 + 2 < NLW) s.nlpref[w0 + 2] = ex + c0 + c1;
    if (threadIdx.x == 0) s.nlpref[NLW] = total;
    __syncthreads();
    return total;
}

// newlines at smem coordinates < pos
__device__ __forceinline__ uint32_t nl_before(const Smem &s, int pos)
{
    int w = pos >> 5;
    uint32_t m = s.nlmask[w] & ((1u << (pos & 31)) - 1u);
    return s.nlpref[w] + __popc(m);
}

// P2: fill the record table for records [k0, k0+CAP) of the tile.  `off` shifts newline ranks
// to record-relative numbering: newline v = rank - off is line (v & 3) of record (v >> 2); the
// byte after newline v = 4k-1 is the start of record k.
__device__ void fill_records(Smem &s, const TileCtx &t, int off, int k0)
{
    if (threadIdx.x == 0 && t.first && k0 == 0) s.rec[0].start = (uint16_t) t.shift;
    const int vlo = 4 * k0 - 1, vhi = 4 * (k0 + CAP) - 1;       // inclusive range of interest
    const int w0 = threadIdx.x * 3;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        int w = w0 + i;
        if (w >= NLW) break;
        uint32_t m = s.nlmask[w];
        if (!m) continue;
        int v = (int) s.nlpref[w] - off;
        if (v > vhi || v + 32 < vlo) continue;
        while (m) {
            int bit = __ffs(m) - 1;
            m &= m - 1;
            int pos = (w << 5) + bit;
            if (v >= vlo && v <= vhi) {
                if (v >= 0 && (v >> 2) < k0 + CAP && (v >> 2) >= k0)
                    s.rec[(v >> 2) - k0].nl[v & 3] = (uint16_t) pos;
                if (((v + 1) & 3) == 0) {
                    int k = (v + 1) >> 2;
                    if (k >= k0 && k < k0 + CAP) s.rec[k - k0].start = (uint16_t) (pos + 1);
                }
            }
            v++;
        }
    }
}

} // namespace

namespace {

// format-error key: smaller = earlier in the stream.  [63:22] stream offset of the record start,
// [21] 0 = bad '@' / 1 = bad '+', [20:0] payload (the offending byte, or plus_pos - record_start)
__device__ __forceinline__ void report_format_error(const KvqOut &o, int64_t rec_fpos, int type, uint32_t payload)
{
    unsigned long long key = ((unsigned long long) rec_fpos << 22) | ((unsigned long long) type << 21)
                           | (payload & 0x1FFFFFu);
    atomicMin(&o.ctr->err_key, key);
}

__device__ __forceinline__ void push_long(const KvqOut &o, int64_t start, int64_t seg_end, int kind, int rl)
{
    unsigned long long i = atomicAdd(&o.ctr->long_count, 1ULL);
    if (i < o.long_cap) {
        KvqLongRec r; r.start = start; r.seg_end = seg_end; r.kind = kind; r.rl = rl;
        o.longs[i] = r;
    }
}

// Phred trim of the quality line [qs, qe] (qe = its '\n'), workhorse.c:1055-1068: the first
// longest run of bytes >= Amin that is closed by a byte < Amin at or before qe.
__device__ __forceinline__ void phred_trim(const Smem &s, int qs, int qe, uint32_t cl, bool chi,
                                           int &rl, int &best)
{
    int runstart = qs;
    rl = 0; best = qs;
    for (int gb = qs & ~31; gb <= qe; gb += 32) {
        uint4 a = s.tile[gb >> 4], b = s.tile[(gb >> 4) + 1];
        uint32_t good = ge_nibble(a.x, cl, chi) | (ge_nibble(a.y, cl, chi) << 4)
                      | (ge_nibble(a.z, cl, chi) << 8) | (ge_nibble(a.w, cl, chi) << 12)
                      | (ge_nibble(b.x, cl, chi) << 16) | (ge_nibble(b.y, cl, chi) << 20)
                      | (ge_nibble(b.z, cl, chi) << 24) | (ge_nibble(b.w, cl, chi) << 28);
        uint32_t bad = ~good;
        if (gb < qs) bad &= 0xFFFFFFFFu << (qs - gb);
        if (qe - gb < 31) bad &= (2u << (qe - gb)) - 1u;
        while (bad) {
            int z = gb + __ffs(bad) - 1;
            bad &= bad - 1;
            int len = z - runstart;
            if (len > rl) { rl = len; best = runstart; }
            runstart = z + 1;
        }
    }
}

// P3: one thread per record of the round.
__device__ void process_records(Smem &s, const KvqBatch &b, const TileCtx &t, const KvqParams &p,
                                const KvqOut &o, int k0, int nrec, int kcomplete)
{
    const uint8_t *tb = reinterpret_cast<const uint8_t *>(s.tile);
    const uint32_t cl = (uint32_t) (p.amin & 0x7F) * 0x01010101u;
    const bool chi = (p.amin & 0x80) != 0;
    for (int i = threadIdx.x; i < nrec; i += THREADS) {
        const int k = k0 + i;
        const Rec r = s.rec[i];
        if (k > kcomplete) {
            // the record runs past the staged region: partial at the end of the segment is
            // dropped (workhorse.c:1033), anything else is a long record
            if (!t.at_seg_end) push_long(o, t.a0 + r.start, t.seg_end, 0, 0);
            continue;
        }
        const int S = r.start, n0 = r.nl[0], n1 = r.nl[1], n2 = r.nl[2], n3 = r.nl[3];
        if (tb[S] != '@')
            report_format_error(o, b.base_fpos + t.a0 + S, 0, tb[S]);
        else if (tb[n1 + 1] != '+')
            report_format_error(o, b.base_fpos + t.a0 + S, 1, (uint32_t) (n1 + 1 - S));
        atomicAdd(&s.records, 1ULL);
        atomicAdd(&s.bases, (unsigned long long) (n1 - n0 - 1));
        int rl, best;
        phred_trim(s, n2 + 1, n3, cl, chi, rl, best);
        if (rl < KVQ_HIST) atomicAdd(&s.hist[rl], 1u);                  // add_rl, workhorse.c:394-402
        atomicMax(&s.longest, rl);
        if (rl < p.minreadlength || rl == 0) continue;                  // workhorse.c:1100
        const int rd0 = n0 + 1 + (best - (n2 + 1));                     // workhorse.c:1070
        if (rl > MAXRL_FAST) { push_long(o, t.a0 + rd0, t.seg_end, 1, rl); continue; }
        int idx = atomicAdd(&s.nrd, 1);
        Rd q; q.pos = (uint16_t) rd0; q.rl = (uint16_t) rl;
        s.rd[idx] = q;
    }
}

// P4: 2-bit pack the trimmed reads of the round, 16 bases per thread-step
__device__ void pack_reads(Smem &s)
{
    const uint32_t *tw = reinterpret_cast<const uint32_t *>(s.tile);
    const int nrd = s.nrd;
    for (int item = threadIdx.x; item < nrd * SLOTW; item += THREADS) {
        const int r = item / SLOTW, j = item - r * SLOTW;
        const int rl = s.rd[r].rl;
        const int nvalid = rl - 16 * j;
        if (nvalid <= -16) continue;                 // words beyond ceil(rl/16) are never read
        uint32_t codes = 0;
        if (nvalid > 0) {
            const int pos = s.rd[r].pos + 16 * j;
            const int a = pos >> 2; const uint32_t sh = (uint32_t) (pos & 3) << 3;
            uint32_t w0 = tw[a], w1 = tw[a + 1], w2 = tw[a + 2], w3 = tw[a + 3], w4 = tw[a + 4];
            codes = pack4(funnel_r(w0, w1, sh)) | (pack4(funnel_r(w1, w2, sh)) << 8)
                  | (pack4(funnel_r(w2, w3, sh)) << 16) | (pack4(funnel_r(w3, w4, sh)) << 24);
            if (nvalid < 16) codes &= (1u << (2 * nvalid)) - 1u;
        }
        s.packed[r * SLOTW + j] = codes;
    }
}

} // namespace

namespace {

// ---------------------------------------------------------------- emission / coverage --------
__device__ __forceinline__ int base_column(uint8_t c)
{
    return c == 'A' ? 0 : c == 'C' ? 1 : c == 'G' ? 2 : c == 'T' ? 3 : c == 'N' ? 4 : 5;
}
__device__ __forceinline__ uint8_t complement(uint8_t c)     // Sequence.pairs, kvarq/genes.py:204
{
    return c == 'A' ? 'T' : c == 'T' ? 'A' : c == 'C' ? 'G' : c == 'G' ? 'C' : c;
}

// add_hit (workhorse.c:408-443) + Coverage.apply_hit (kvarq/analyse.py:57-78).
// `rd` points at the trimmed read's bytes (shared or global), `tp` at the template's bytes.
__device__ void emit_hit(const KvqTemplates &t, const KvqParams &p, const KvqOut &o, int seq, int d,
                         int ro, int to, int L, int rl, int64_t file_pos, const uint8_t *rd,
                         const uint8_t *tp, int copies, int rule)
{
    if (p.accumulate) {
        atomicAdd(&o.nseqhits[seq], (unsigned long long) copies);
        atomicAdd(&o.nseqbasehits[seq], (unsigned long long) copies * (unsigned long long) L);
    }
    unsigned long long slot = atomicAdd(&o.ctr->hit_count, (unsigned long long) copies);
    if (p.want_hits) {
        unsigned long long boff = atomicAdd(&o.ctr->blob_used, (unsigned long long) copies * (unsigned long long) L);
        for (int c = 0; c < copies; c++) {
            if (slot + c < o.hit_cap) {
                KvqHitRec h;
                h.seq_nr = seq; h.seq_pos = d; h.file_pos = file_pos; h.length = L; h.readlength = rl;
                h.blob_off = (long long) (boff + (unsigned long long) c * L);
                // the flush-right duplicate: first copy is rule (b), second rule (d)
                h.rule = (copies == 2) ? (c == 0 ? 1 : 2) : rule; h.pad = 0;
                o.hits[slot + c] = h;
            }
            if (boff + (unsigned long long) (c + 1) * L <= o.blob_cap)
                for (int j = 0; j < L; j++) o.blob[boff + (unsigned long long) c * L + j] = rd[ro + j];
        }
    }
    if (t.nplus > 0 && p.accumulate) {
        const bool plus = seq < t.nplus;
        const int tpl = plus ? seq : seq - t.nplus;
        const int sl = t.len[seq];
        const unsigned long long cb = t.cov_off[tpl];
        for (int j = 0; j < L; j++) {
            const int sidx = to + j;
            const unsigned long long pidx = cb + (unsigned long long) (plus ? sidx : sl - 1 - sidx);
            atomicAdd(&o.coverage[pidx], (unsigned long long) copies);
            const uint8_t hb = rd[ro + j];
            if (hb != tp[sidx])
                atomicAdd(&o.mutcount[pidx * 6 + base_column(plus ? hb : complement(hb))],
                          (unsigned long long) copies);
        }
    }
}

__device__ __forceinline__ int byte_mismatches(const uint8_t *a, const uint8_t *b, int n, int maxerr)
{
    int e = 0;
    for (int j = 0; j < n; j++) {
        e += (a[j] != b[j]);
        if (e > maxerr) break;
    }
    return e;
}

// ---------------------------------------------------------------- seed positions -------------
// Fixed read offsets probed against the all-offsets table T2: E+1 pieces of the whole read
// (rule d), of its first minoverlap bases (rule b) and of its last minoverlap bases (rule a).
__device__ __forceinline__ int build_p2(int rl, int O, int np, int max_sl, int *out)
{
    int n = 0;
#define KVQ_ADD_P2(v) do { int _v = (v); bool dup = false; \
        for (int _i = 0; _i < n; _i++) dup |= (out[_i] == _v); if (!dup) out[n++] = _v; } while (0)
    if (max_sl >= rl)
        for (int j = 0; j < np; j++) KVQ_ADD_P2(j * rl / np);
    if (rl > O) {
        for (int j = 0; j < np; j++) KVQ_ADD_P2(j * O / np);
        for (int j = 0; j < np; j++) KVQ_ADD_P2(rl - O + j * O / np);
    }
#undef KVQ_ADD_P2
    return n;
}

__device__ __forceinline__ bool read_seedable(const KvqParams &p, int rl)
{
    return p.seed_ok && (rl / p.npieces) >= KVQ_SEED_K;
}

__device__ __forceinline__ void push_seeds(Smem &s, int r, int pbase, uint32_t mask, int kind)
{
    int cnt = __popc(mask);
    int at = atomicAdd(&s.nseed, cnt);
    while (mask) {
        int j = __ffs(mask) - 1;
        mask &= mask - 1;
        if (at < SEEDQ) s.seedq[at] = ((uint32_t) r << 16) | ((uint32_t) (pbase + j) << 1) | (uint32_t) kind;
        at++;
    }
}

// P5: probes.  T1: every read offset, 16 offsets per thread-step.  T2: a few fixed offsets.
__device__ void probe_reads(Smem &s, const KvqParams &p, const KvqTemplates &t)
{
    constexpr int MAXBLK = (MAXRL_FAST - KVQ_SEED_K + 16) / 16;
    const int nrd = s.nrd;
    for (int item = threadIdx.x; item < nrd * MAXBLK; item += THREADS) {
        const int r = item / MAXBLK, q = item - r * MAXBLK;
        const int rl = s.rd[r].rl;
        const int npos = rl - KVQ_SEED_K + 1 - 16 * q;      // probe offsets left from 16q on
        if (npos <= 0 || !read_seedable(p, rl)) continue;
        const uint32_t *pk = s.packed + r * SLOTW + q;
        const uint32_t lo = pk[0], hi = pk[1];
        uint32_t mask = 0;
#pragma unroll
        for (int j = 0; j < 16; j++) {
            uint32_t key = funnel_r(lo, hi, 2 * j) & 0xFFFFu;
            mask |= ((s.bm1[key >> 5] >> (key & 31)) & 1u) << j;
        }
        if (npos < 16) mask &= (1u << npos) - 1u;
        if (mask) push_seeds(s, r, 16 * q, mask, 0);
    }
    for (int r = threadIdx.x; r < nrd; r += THREADS) {
        const int rl = s.rd[r].rl;
        if (!read_seedable(p, rl)) continue;
        int p2[MAXP2];
        const int n2 = build_p2(rl, p.minoverlap, p.npieces, p.max_seqlen, p2);
        const uint32_t *pk = s.packed + r * SLOTW;
        for (int i = 0; i < n2; i++) {
            uint32_t key = extract16(pk, p2[i]) & 0xFFFFu;
            if ((s.bm2[key >> 5] >> (key & 31)) & 1u) push_seeds(s, r, p2[i], 1u, 1);
        }
    }
}

// P6: walk the bucket of every seed and queue the diagonals the reference would evaluate
__device__ void expand_seeds(Smem &s, const KvqParams &p, const KvqTemplates &t)
{
    const int nseed = s.nseed;
    for (int i = threadIdx.x; i < nseed; i += THREADS) {
        const uint32_t sd = s.seedq[i];
        const int r = sd >> 16, pos = (sd >> 1) & 0x7FFF, kind = sd & 1;
        const int rl = s.rd[r].rl;
        const uint32_t key = extract16(s.packed + r * SLOTW, pos) & 0xFFFFu;
        const uint32_t *off = kind ? t.t2_off : t.t1_off;
        const uint2 *ent = kind ? t.t2_ent : t.t1_ent;
        const uint32_t e0 = __ldg(off + key), e1 = __ldg(off + key + 1);
        for (uint32_t e = e0; e < e1; e++) {
            const uint2 en = __ldg(ent + e);
            const int sq = (int) en.x, toff = (int) en.y;
            const int d = toff - pos;
            int ro, to, L, rule;
            if (!diag_rule(d, rl, __ldg(t.len + sq), p.minoverlap, ro, to, L, rule)) continue;
            int at = atomicAdd(&s.ncand, 1);
            if (at < CANDQ)
                s.candq[at] = (unsigned long long) r | ((unsigned long long) sq << 8)
                            | ((unsigned long long) (d + 32768) << 32)
                            | ((unsigned long long) pos << 48) | ((unsigned long long) kind << 58);
        }
    }
}

// Is (kind, pos) the first seed, in the fixed order T1 pieces then T2 offsets, that fires on
// diagonal d of (read, seq)?  Every firing seed queues the same diagonal; only this one emits.
__device__ bool canonical_seed(const KvqParams &p, const uint32_t *rpk, const uint32_t *tpk,
                               int rl, int sl, int d, int kind, int pos)
{
    for (int j = 0; j < p.npieces; j++) {
        const int toff = j * sl / p.npieces;
        const int pj = toff - d;
        if (pj < 0 || pj > rl - KVQ_SEED_K) continue;
        if (((extract16(rpk, pj) ^ extract16_g(tpk, toff)) & 0xFFFFu) == 0)
            return kind == 0 && pos == pj;
    }
    int p2[MAXP2];
    const int n2 = build_p2(rl, p.minoverlap, p.npieces, p.max_seqlen, p2);
    for (int i = 0; i < n2; i++) {
        const int toff = p2[i] + d;
        if (toff < 0 || toff > sl - KVQ_SEED_K) continue;
        if (((extract16(rpk, p2[i]) ^ extract16_g(tpk, toff)) & 0xFFFFu) == 0)
            return kind == 1 && pos == p2[i];
    }
    return false;
}

// P7: one thread per candidate diagonal
__device__ void verify_candidates(Smem &s, const KvqBatch &b, const TileCtx &tc, const KvqParams &p,
                                  const KvqTemplates &t, const KvqOut &o)
{
    const uint8_t *tb = reinterpret_cast<const uint8_t *>(s.tile);
    const int ncand = s.ncand;
    for (int i = threadIdx.x; i < ncand; i += THREADS) {
        const unsigned long long c = s.candq[i];
        const int r = (int) (c & 0xFF), sq = (int) ((c >> 8) & 0xFFFFFF);
        const int d = (int) ((c >> 32) & 0xFFFF) - 32768;
        const int pos = (int) ((c >> 48) & 0x3FF), kind = (int) ((c >> 58) & 1);
        const int rl = s.rd[r].rl, sl = __ldg(t.len + sq);
        int ro, to, L, rule;
        const int copies = diag_rule(d, rl, sl, p.minoverlap, ro, to, L, rule);
        const uint32_t *rpk = s.packed + r * SLOTW;
        const uint32_t *tpk = t.packed + __ldg(t.packed_off + sq);
        // Hamming distance in code space (a lower bound of the byte distance)
        int e = 0;
        for (int j = 0; j < L && e <= p.maxerrors; j += 16) {
            uint32_t x = extract16(rpk, ro + j) ^ extract16_g(tpk, to + j);
            uint32_t m = (x | (x >> 1)) & 0x55555555u;
            if (L - j < 16) m &= (1u << (2 * (L - j))) - 1u;
            e += __popc(m);
        }
        if (e > p.maxerrors) continue;
        const uint8_t *rd = tb + s.rd[r].pos;
        const uint8_t *tp = t.bytes + __ldg(t.byte_off + sq);
        if (byte_mismatches(rd + ro, tp + to, L, p.maxerrors) > p.maxerrors) continue;
        if (!canonical_seed(p, rpk, tpk, rl, sl, d, kind, pos)) continue;
        emit_hit(t, p, o, sq, d, ro, to, L, rl, b.base_fpos + tc.a0 + s.rd[r].pos, rd, tp, copies, rule);
    }
}

// Exhaustive matcher: one warp per (read, seq) pair, lanes over all diagonals.
__device__ void exhaustive_pair(const KvqParams &p, const KvqTemplates &t, const KvqOut &o,
                                const uint8_t *rd, int rl, int64_t file_pos, int sq)
{
    const int lane = threadIdx.x & 31;
    const int sl = __ldg(t.len + sq);
    const uint8_t *tp = t.bytes + __ldg(t.byte_off + sq);
    for (int d = -(rl - 1) + lane; d <= sl - 1; d += 32) {
        int ro, to, L, rule;
        const int copies = diag_rule(d, rl, sl, p.minoverlap, ro, to, L, rule);
        if (!copies) continue;
        if (byte_mismatches(rd + ro, tp + to, L, p.maxerrors) > p.maxerrors) continue;
        emit_hit(t, p, o, sq, d, ro, to, L, rl, file_pos, rd, tp, copies, rule);
    }
}

} // namespace

namespace {

// One round = up to CAP owned records of the current tile.
__device__ void run_round(Smem &s, const KvqBatch &b, const TileCtx &tc, const KvqParams &p,
                          const KvqTemplates &t, const KvqOut &o, int off, int k0, int nrec, int kcomplete)
{
    if (threadIdx.x == 0) {
        s.nrd = 0; s.nseed = 0; s.ncand = 0; s.n_exh_reads = 0; s.n_seed_reads = 0; s.overflow = 0;
    }
    fill_records(s, tc, off, k0);
    __syncthreads();
    process_records(s, b, tc, p, o, k0, nrec, kcomplete);
    __syncthreads();
    const int nrd = s.nrd;
    __syncthreads();                            // everyone has read nrd before the next round resets it
    if (nrd == 0) return;                       // uniform
    const uint8_t *tb = reinterpret_cast<const uint8_t *>(s.tile);
    bool seeded = p.seed_ok && t.n_exh < t.nseqs;
    if (seeded) {
        pack_reads(s);
        __syncthreads();
        probe_reads(s, p, t);
        __syncthreads();
        if (s.nseed > SEEDQ) seeded = false;    // uniform: queue overflow -> exhaustive round
        if (seeded) {
            expand_seeds(s, p, t);
            __syncthreads();
            if (s.ncand > CANDQ) seeded = false;
        }
        if (seeded) {
            if (threadIdx.x == 0) s.cand_count += (unsigned long long) s.ncand;
            verify_candidates(s, b, tc, p, t, o);
        } else if (threadIdx.x == 0) {
            atomicAdd(&o.ctr->overflow_rounds, 1u);
        }
    }
    // split the reads of the round: seeded ones still owe the non-seedable templates
    for (int r = threadIdx.x; r < nrd; r += THREADS) {
        if (seeded && read_seedable(p, s.rd[r].rl)) s.seed_reads[atomicAdd(&s.n_seed_reads, 1)] = (uint16_t) r;
        else s.exh_reads[atomicAdd(&s.n_exh_reads, 1)] = (uint16_t) r;
    }
    __syncthreads();
    const int wid = threadIdx.x >> 5;
    const int ne = s.n_exh_reads, ns = s.n_seed_reads;
    if (threadIdx.x == 0) s.exh_count += (unsigned long long) ne;
    for (int pair = wid; pair < ne * t.nseqs; pair += NWARPS) {
        const int r = s.exh_reads[pair / t.nseqs], sq = pair % t.nseqs;
        exhaustive_pair(p, t, o, tb + s.rd[r].pos, s.rd[r].rl, b.base_fpos + tc.a0 + s.rd[r].pos, sq);
    }
    for (int pair = wid; pair < ns * t.n_exh; pair += NWARPS) {
        const int r = s.seed_reads[pair / t.n_exh], sq = __ldg(t.exh_list + pair % t.n_exh);
        exhaustive_pair(p, t, o, tb + s.rd[r].pos, s.rd[r].rl, b.base_fpos + tc.a0 + s.rd[r].pos, sq);
    }
    __syncthreads();
}

__global__ void __launch_bounds__(THREADS, 1)
kvq_scan_kernel(KvqBatch b, KvqTemplates t, KvqParams p, KvqOut o)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Smem &s = *reinterpret_cast<Smem *>(smem_raw);

    for (int i = threadIdx.x; i < KVQ_BITMAP_WORDS; i += THREADS) {
        s.bm1[i] = __ldg(t.t1_bitmap + i);
        s.bm2[i] = __ldg(t.t2_bitmap + i);
    }
    for (int i = threadIdx.x; i < KVQ_HIST; i += THREADS) s.hist[i] = 0;
    if (threadIdx.x == 0) { s.records = 0; s.bases = 0; s.exh_count = 0; s.cand_count = 0; s.longest = -1; }
    __syncthreads();

    for (;;) {
        if (threadIdx.x == 0) s.seg_cur = (int) atomicAdd(&o.ctr->seg_next, 1u);
        __syncthreads();
        const int seg = s.seg_cur;
        __syncthreads();
        if (seg >= b.nseg) break;
        const int64_t seg_beg = b.seg_off[seg], seg_end = b.seg_off[seg + 1];
        int phase = 0;
        for (int64_t t0 = seg_beg; t0 < seg_end; t0 += TILE) {
            TileCtx tc;
            tc.a0 = t0 & ~(int64_t) 15;
            tc.shift = (int) (t0 - tc.a0);
            const int64_t t1 = t0 + TILE < seg_end ? t0 + TILE : seg_end;
            const int64_t g1 = t1 + MARGIN < seg_end ? t1 + MARGIN : seg_end;
            tc.tile_end = (int) (t1 - tc.a0);
            tc.reg_end = (int) (g1 - tc.a0);
            tc.first = (t0 == seg_beg);
            tc.at_seg_end = (g1 == seg_end);
            tc.phase_in = phase;
            tc.seg_end = seg_end;
            const int n_all = (int) load_and_index(s, b, tc);
            const int n_tile = (int) nl_before(s, tc.tile_end);
            // newline rank -> record-relative numbering (see fill_records)
            const int off = tc.first ? 0 : ((3 - phase) & 3) + 1;
            const int vmax_tile = n_tile - 1 - off, vmax_all = n_all - 1 - off;
            int nown = 0;
            if (tc.first) nown = (vmax_tile + 1) / 4 + 1;
            else if (vmax_tile >= -1) nown = (vmax_tile + 1) / 4 + 1;
            // record k is complete iff its 4th newline v = 4k+3 is inside the region
            const int kcomplete = vmax_all >= 3 ? (vmax_all - 3) / 4 : -1;
            for (int k0 = 0; k0 < nown; k0 += CAP) {
                const int nrec = nown - k0 < CAP ? nown - k0 : CAP;
                run_round(s, b, tc, p, t, o, off, k0, nrec, kcomplete);
            }
            phase = (phase + n_tile) & 3;
            __syncthreads();
        }
    }

    // flush the per-CTA statistics
    for (int i = threadIdx.x; i < KVQ_HIST; i += THREADS)
        if (s.hist[i]) atomicAdd(&o.ctr->rls[i], (unsigned long long) s.hist[i]);
    if (threadIdx.x == 0) {
        atomicAdd(&o.ctr->records, s.records);
        atomicAdd(&o.ctr->bases, s.bases);
        atomicAdd(&o.ctr->exhaustive_reads, s.exh_count);
        atomicAdd(&o.ctr->candidates, s.cand_count);
        atomicMax(&o.ctr->longest, (long long) s.longest);
    }
}

} // namespace

namespace {

// Records that do not fit a tile's spill margin, or whose trimmed read exceeds the packed slot:
// one warp per record straight from global memory, exhaustive matcher.  Rare by construction.
__global__ void __launch_bounds__(256)
kvq_long_kernel(KvqBatch b, KvqTemplates t, KvqParams p, KvqOut o)
{
    const int lane = threadIdx.x & 31;
    const unsigned long long n = o.ctr->long_count < o.long_cap ? o.ctr->long_count : o.long_cap;
    for (;;) {
        unsigned int i = 0;
        if (lane == 0) i = atomicAdd(&o.ctr->long_next, 1u);
        i = __shfl_sync(0xffffffffu, i, 0);
        if (i >= n) break;
        const KvqLongRec lr = o.longs[i];
        long long rd0 = lr.start; int rl = lr.rl;
        if (lr.kind == 0) {
            int ok = 0;
            if (lane == 0) {
                long long nl[4]; int k = 0;
                for (long long q = lr.start; q < lr.seg_end && k < 4; q++)
                    if (b.data[q] == '\n') nl[k++] = q;
                if (k == 4) {
                    if (b.data[lr.start] != '@')
                        report_format_error(o, b.base_fpos + lr.start, 0, b.data[lr.start]);
                    else if (b.data[nl[1] + 1] != '+')
                        report_format_error(o, b.base_fpos + lr.start, 1, (uint32_t) (nl[1] + 1 - lr.start));
                    atomicAdd(&o.ctr->records, 1ULL);
                    atomicAdd(&o.ctr->bases, (unsigned long long) (nl[1] - nl[0] - 1));
                    long long qs = nl[2] + 1, run = qs, best = qs; int open = 1; rl = 0;
                    for (long long q = qs; q <= nl[3]; q++) {
                        if ((int) (signed char) b.data[q] >= p.amin) { if (!open) { open = 1; run = q; } }
                        else if (open) { if (q - run > rl) { rl = (int) (q - run); best = run; } open = 0; }
                    }
                    if (rl < KVQ_HIST) atomicAdd(&o.ctr->rls[rl], 1ULL);
                    atomicMax(&o.ctr->longest, (long long) rl);
                    rd0 = nl[0] + 1 + (best - qs);
                    ok = (rl >= p.minreadlength && rl > 0);
                }
            }
            ok = __shfl_sync(0xffffffffu, ok, 0);
            rl = __shfl_sync(0xffffffffu, rl, 0);
            rd0 = __shfl_sync(0xffffffffu, rd0, 0);
            if (!ok) continue;
        }
        if (lane == 0) atomicAdd(&o.ctr->exhaustive_reads, 1ULL);
        for (int sq = 0; sq < t.nseqs; sq++)
            exhaustive_pair(p, t, o, b.data + rd0, rl, b.base_fpos + rd0, sq);
    }
}

} // namespace

int kvq_scan_smem_bytes(void) { return (int) sizeof(Smem); }

int kvq_launch_scan(const KvqBatch &b, const KvqTemplates &t, const KvqParams &p, const KvqOut &o,
                    int sm_count, void *stream)
{
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(kvq_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int) sizeof(Smem));
        if (e != cudaSuccess) return (int) e;
        attr_set = true;
    }
    int grid = b.nseg < sm_count ? b.nseg : sm_count;
    if (grid < 1) grid = 1;
    kvq_scan_kernel<<<grid, THREADS, sizeof(Smem), (cudaStream_t) stream>>>(b, t, p, o);
    return (int) cudaGetLastError();
}

int kvq_launch_long(const KvqBatch &b, const KvqTemplates &t, const KvqParams &p, const KvqOut &o,
                    int sm_count, void *stream)
{
    kvq_long_kernel<<<sm_count, 256, 0, (cudaStream_t) stream>>>(b, t, p, o);
    return (int) cudaGetLastError();
}
