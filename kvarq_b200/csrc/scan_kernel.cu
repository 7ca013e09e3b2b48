// sm_100a kernels of the KvarQ read-scanning path (reference: csrc/workhorse.c:976-1197).
//
// Persistent kernel, one CTA per SM.  A CTA holds G independent warp groups; every group claims
// 32 KiB tiles of the FastQ stream in stream order from a global queue and runs the whole path
// on its tile in shared memory, synchronising only inside the group (named barriers):
//   P1  stage tile (+ spill margin) with 128-bit loads, SWAR newline bitmap, group prefix scan;
//       the tile's newline count is published and the line phase of the tile (newlines of its
//       segment before it, mod 4) resolved by decoupled look-back      [workhorse.c:1018-1030]
//   P2  one thread per record: select its 4 newlines from the bitmap, '@' / '+' sanity checks,
//       Phred trim (first longest run >= Amin), histogram, length filter [workhorse.c:1037-1103]
//   P3  2-bit packing of the trimmed reads (dense work lists, 16 bases per step)
//   P4  seed probes against bitmaps in shared memory: every read offset vs the template-piece
//       table T1, a few fixed read offsets vs the all-offsets table T2
//   P5  bucket expansion (tables in L2) -> candidate diagonals
//   P6  XOR/popc Hamming verify on packed words, byte-exact confirm, canonical-seed dedupe,
//       hit emission through atomics, fused coverage            [workhorse.c:1107-1175,408-443]
// Reads or templates the seed scheme cannot cover take the exhaustive matcher (same acceptance
// rule, every diagonal).  Acceptance rule: SURVEY.md Appendix B / workhorse.c:1112-1174.
#include <cuda_runtime.h>
#include <stdint.h>
#include "kvq_device.h"

namespace {

constexpr int G = 2;                        // warp groups per CTA
constexpr int GT = 256;                     // threads per group
constexpr int THREADS = G * GT;
constexpr int GW = GT / 32;                 // warps per group
constexpr int TILE = KVQ_TILE;
constexpr int MARGIN = 3072;                // spill: records starting in the tile may end here
constexpr int LOADMAX = TILE + MARGIN + 32; // + alignment slack of the first 32-byte word
constexpr int NLW = LOADMAX / 32 + 1;       // newline bitmap words (32 bytes each)
constexpr int WPT = (NLW + GT - 1) / GT;    // bitmap words per thread in the prefix scan
constexpr int CAP = GT;                     // records per round (one thread per record)
constexpr int MAXRL_FAST = KVQ_MAXRL_FAST;   // longer trimmed reads go to the long-record kernel
constexpr int PACKW = 4096;                 // packed words per round (16 bases each)
constexpr int QCAP = 4096;                  // T1 probe items per round (16 read offsets each)
constexpr int SEEDQ = 2048;
constexpr int CANDQ = 1536;
constexpr int ACCQ = 256;                   // code-space survivors per round (warp-cooperative confirm)

struct Rd { uint16_t pos, rl, poff, flags; };   // flags bit0: seedable

struct GroupSmem {
    uint4 tile[LOADMAX / 16 + 4];           // + slack for word-aligned over-reads
    uint32_t nlmask[NLW + 1];
    uint32_t nlpref[NLW + 1];
    Rd rd[CAP];
    uint32_t packed[PACKW + 2];
    uint8_t pitem[PACKW];                   // packed word -> read
    uint16_t qitem[QCAP];                   // probe item -> read << 6 | block
    uint32_t seedq[SEEDQ];
    unsigned long long candq[CANDQ];
    uint16_t exh_reads[CAP];
    uint16_t seed_reads[CAP];
    uint16_t accq[ACCQ];
    uint32_t warp_tot[GW];
    int nrd, nseed, ncand, nacc, n_exh_reads, n_seed_reads;
    int tile_cur, phase_in;
};

struct Smem {
    uint32_t bm1[KVQ_BITMAP_WORDS];
    uint32_t bm2[KVQ_BITMAP_WORDS];
    uint32_t hist[KVQ_HIST];
    int longest;
    GroupSmem g[G];
};
static_assert(sizeof(Smem) <= 232448, "shared memory budget (227 KB) exceeded");

struct ThreadStats { unsigned long long records, bases, exh, cands; };

__device__ __forceinline__ void group_sync(int gid)
{
    asm volatile("bar.sync %0, %1;" ::"r"(gid + 1), "r"(GT) : "memory");
}

__device__ __forceinline__ uint32_t funnel_r(uint32_t lo, uint32_t hi, uint32_t s)
{
    return __funnelshift_r(lo, hi, s);
}

// bit7 of every byte of w that equals '\n' (other bits 0)
__device__ __forceinline__ uint32_t nl_bits(uint32_t w)
{
    uint32_t t = w ^ 0x0A0A0A0Au;
    uint32_t z = ((t & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | t;     // bit7 set <=> byte != 0
    return ~z & 0x80808080u;
}
// bits 7,15,23,31 -> 4-bit value in the top nibble (no carries: all partial products distinct)
__device__ __forceinline__ uint32_t top_nibble(uint32_t e) { return e * 0x00204081u; }

// 32-bit newline mask of 32 bytes
__device__ __forceinline__ uint32_t nl_mask32(uint4 a, uint4 b)
{
    uint32_t m = top_nibble(nl_bits(a.x)) >> 28;
    m |= (top_nibble(nl_bits(a.y)) >> 24) & 0x000000F0u;
    m |= (top_nibble(nl_bits(a.z)) >> 20) & 0x00000F00u;
    m |= (top_nibble(nl_bits(a.w)) >> 16) & 0x0000F000u;
    m |= (top_nibble(nl_bits(b.x)) >> 12) & 0x000F0000u;
    m |= (top_nibble(nl_bits(b.y)) >> 8) & 0x00F00000u;
    m |= (top_nibble(nl_bits(b.z)) >> 4) & 0x0F000000u;
    m |= top_nibble(nl_bits(b.w)) & 0xF0000000u;
    return m;
}

// bit7 of every byte b of w with (signed char) b >= amin.  cl = (amin & 0x7f) * 0x01010101,
// chi = amin is negative.  Compares (b ^ 0x80) >= (amin ^ 0x80) as unsigned bytes.
__device__ __forceinline__ uint32_t ge_bits(uint32_t w, uint32_t cl, bool chi)
{
    uint32_t t = (w | 0x80808080u) - cl;            // bit7: low 7 bits of b >= low 7 bits of amin
    uint32_t uhi = ~w;                              // bit7 of (b ^ 0x80)
    return (chi ? (uhi | t) : (uhi & t)) & 0x80808080u;
}
__device__ __forceinline__ uint32_t ge_mask32(uint4 a, uint4 b, uint32_t cl, bool chi)
{
    uint32_t m = top_nibble(ge_bits(a.x, cl, chi)) >> 28;
    m |= (top_nibble(ge_bits(a.y, cl, chi)) >> 24) & 0x000000F0u;
    m |= (top_nibble(ge_bits(a.z, cl, chi)) >> 20) & 0x00000F00u;
    m |= (top_nibble(ge_bits(a.w, cl, chi)) >> 16) & 0x0000F000u;
    m |= (top_nibble(ge_bits(b.x, cl, chi)) >> 12) & 0x000F0000u;
    m |= (top_nibble(ge_bits(b.y, cl, chi)) >> 8) & 0x00F00000u;
    m |= (top_nibble(ge_bits(b.z, cl, chi)) >> 4) & 0x0F000000u;
    m |= top_nibble(ge_bits(b.w, cl, chi)) & 0xF0000000u;
    return m;
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
__device__ __forceinline__ uint32_t extract16_g(const uint32_t *pk, int b)
{
    int w = b >> 4;
    return funnel_r(__ldg(pk + w), __ldg(pk + w + 1), (uint32_t) ((b & 15) << 1));
}

// Is diagonal d (= seq index aligned with read[0]) one the reference evaluates?  Returns the
// number of hits an accepted alignment yields there (0, 1 or 2 -- the (b)+(d) duplicate,
// workhorse.c:1130-1141 + 1163-1173), the overlap geometry and the rule (emission phase).
__device__ __forceinline__ int diag_rule(int d, int rl, int sl, int O, int &ro, int &to, int &L, int &rule)
{
    rule = 2;
    ro = d < 0 ? -d : 0;
    to = d > 0 ? d : 0;
    int lr = rl - ro, ls = sl - to;
    L = lr < ls ? lr : ls;
    if (L <= 0) return 0;
    if (d <= 0 && d + rl >= sl) return 1;           // template inside read: rule (c) / (d) at d == 0
    if (d >= 0 && d + rl <= sl) {                   // read inside template: rule (d)
        if (d > 0 && d + rl == sl && rl > O && sl > O) return 2;    // flush right: rule (b) fires too
        return 1;
    }
    rule = d < 0 ? 0 : 1;                           // partial overlaps: rules (a) / (b)
    if (rl > O && sl > O && L >= O) return 1;
    return 0;
}

} // namespace

namespace {

struct TileCtx {
    int64_t a0;          // batch offset of smem byte 0 (32-byte aligned)
    int shift;           // smem coordinate of the first byte of the tile
    int tile_end;        // smem coordinate one past the bytes owned by the tile
    int reg_end;         // smem coordinate one past the staged region
    int first;           // first tile of its segment
    int at_seg_end;      // the region reaches the end of the segment
    int64_t seg_end;     // batch offset of the segment end
};

// group-wide exclusive scan of one value per thread; returns the exclusive prefix, *total = sum
__device__ __forceinline__ uint32_t group_exscan(uint32_t v, uint32_t *warp_tot, uint32_t *total, int gid, int gt)
{
    const int lane = gt & 31, wid = gt >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    if (lane == 31) warp_tot[wid] = inc;
    group_sync(gid);
    uint32_t base = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < GW; w++) {
        uint32_t t = warp_tot[w];
        if (w < wid) base += t;
        tot += t;
    }
    group_sync(gid);
    *total = tot;
    return base + inc - v;
}

// P1: stage the region in shared memory, build the newline bitmap + prefix counts.
// Returns (to every thread of the group) the number of newlines in the whole region.
__device__ uint32_t load_and_index(GroupSmem &s, const KvqBatch &b, const TileCtx &t, int gid, int gt)
{
    const int nwords = (t.reg_end + 31) >> 5;
    const uint4 *src = reinterpret_cast<const uint4 *>(b.data + t.a0);
    constexpr int NI = (NLW + 1 + GT - 1) / GT;
    uint4 va[NI], vb[NI];
#pragma unroll
    for (int i = 0; i < NI; i++) {                  // every load of the tile in flight at once
        const int w = gt + i * GT;
        if (w < nwords) { va[i] = __ldg(src + 2 * w); vb[i] = __ldg(src + 2 * w + 1); }
    }
#pragma unroll
    for (int i = 0; i < NI; i++) {
        const int w = gt + i * GT;
        if (w >= NLW + 1) break;
        uint32_t m = 0;
        if (w < nwords) {
            s.tile[2 * w] = va[i]; s.tile[2 * w + 1] = vb[i];
            m = nl_mask32(va[i], vb[i]);
            const int lo = t.shift - (w << 5), hi = t.reg_end - (w << 5);
            if (lo > 0) m &= lo >= 32 ? 0u : (0xFFFFFFFFu << lo);
            if (hi < 32) m &= hi <= 0 ? 0u : ((1u << hi) - 1u);
        }
        s.nlmask[w] = m;
    }
    group_sync(gid);
    const int w0 = gt * WPT;
    uint32_t c[WPT]; uint32_t sum = 0;
#pragma unroll
    for (int i = 0; i < WPT; i++) {
        c[i] = (w0 + i < NLW) ? __popc(s.nlmask[w0 + i]) : 0;
        sum += c[i];
    }
    uint32_t total;
    uint32_t ex = group_exscan(sum, s.warp_tot, &total, gid, gt);
#pragma unroll
    for (int i = 0; i < WPT; i++) {
        if (w0 + i < NLW) s.nlpref[w0 + i] = ex;
        ex += c[i];
    }
    if (gt == 0) s.nlpref[NLW] = total;
    group_sync(gid);
    return total;
}

// newlines at smem coordinates < pos
__device__ __forceinline__ uint32_t nl_before(const GroupSmem &s, int pos)
{
    int w = pos >> 5;
    uint32_t m = s.nlmask[w] & ((1u << (pos & 31)) - 1u);
    return s.nlpref[w] + __popc(m);
}

// Decoupled look-back over the tiles of one segment: every tile publishes its newline count
// (mod 4); a tile's phase is the sum over its predecessors.  Status word: bit3 = inclusive
// prefix, bit2 = tile aggregate, bits 1:0 = value.  Tiles are claimed in stream order, so a
// predecessor is always resident or finished.
__device__ int resolve_phase(const KvqBatch &b, int tile, int first, int n_tile)
{
    volatile uint32_t *st = b.tile_status;
    if (first) {
        st[tile] = 8u | (uint32_t) (n_tile & 3);
        return 0;
    }
    st[tile] = 4u | (uint32_t) (n_tile & 3);
    int acc = 0;
    for (int j = tile - 1;; j--) {
        uint32_t v;
        while ((v = st[j]) == 0) __nanosleep(20);
        acc += (int) (v & 3);
        if (v & 8u) break;
    }
    acc &= 3;
    st[tile] = 8u | (uint32_t) ((acc + n_tile) & 3);
    return acc;
}

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
__device__ __forceinline__ void phred_trim(const GroupSmem &s, int qs, int qe, uint32_t cl, bool chi,
                                           int &rl, int &best)
{
    int runstart = qs;
    rl = 0; best = qs;
    for (int gb = qs & ~31; gb <= qe; gb += 32) {
        uint32_t bad = ~ge_mask32(s.tile[gb >> 4], s.tile[(gb >> 4) + 1], cl, chi);
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

// P2: one thread per record of the round.  Newline ranks are record-relative: newline
// v = rank - off is line (v & 3) of record (v >> 2); the byte after newline 4k-1 starts record k.
__device__ void process_records(GroupSmem &s, Smem &sh, const KvqBatch &b, const TileCtx &t,
                                const KvqParams &p, const KvqOut &o, int off, int k0, int nrec,
                                int kcomplete, int gt, ThreadStats &ts)
{
    if (gt >= nrec) return;
    const uint8_t *tb = reinterpret_cast<const uint8_t *>(s.tile);
    const int k = k0 + gt;
    // locate the newline that precedes the record (or the record's first newline for k == 0
    // of a segment's first tile) by binary search over the per-word prefix counts
    const bool has_prev = !(t.first && k == 0);
    const int r0 = 4 * k + off - (has_prev ? 1 : 0);        // rank of the first newline we need
    int lo = 0, hi = NLW;                                   // largest w with nlpref[w] <= r0
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if ((int) s.nlpref[mid] <= r0) lo = mid; else hi = mid;
    }
    int w = lo;
    uint32_t m = s.nlmask[w];
    for (int skip = r0 - (int) s.nlpref[w]; skip > 0; skip--) m &= m - 1;
    // walk to the next newline of the bitmap
    auto next_nl = [&]() -> int {
        while (m == 0) m = s.nlmask[++w];
        const int pos = (w << 5) + __ffs(m) - 1;
        m &= m - 1;
        return pos;
    };
    int S = t.shift;
    if (has_prev) S = next_nl() + 1;
    if (k > kcomplete) {
        // the record runs past the staged region: a partial record at the end of the segment
        // is dropped (workhorse.c:1033), anything else is a long record
        if (!t.at_seg_end) push_long(o, t.a0 + S, t.seg_end, 0, 0);
        return;
    }
    const int n0 = next_nl(), n1 = next_nl(), n2 = next_nl(), n3 = next_nl();
    if (tb[S] != '@')
        report_format_error(o, b.base_fpos + t.a0 + S, 0, tb[S]);
    else if (tb[n1 + 1] != '+')
        report_format_error(o, b.base_fpos + t.a0 + S, 1, (uint32_t) (n1 + 1 - S));
    ts.records++;
    ts.bases += (unsigned long long) (n1 - n0 - 1);
    const uint32_t cl = (uint32_t) (p.amin & 0x7F) * 0x01010101u;
    const bool chi = (p.amin & 0x80) != 0;
    int rl, best;
    phred_trim(s, n2 + 1, n3, cl, chi, rl, best);
    if (rl < KVQ_HIST) atomicAdd(&sh.hist[rl], 1u);                 // add_rl, workhorse.c:394-402
    atomicMax(&sh.longest, rl);
    if (rl < p.minreadlength || rl == 0) return;                    // workhorse.c:1100
    const int rd0 = n0 + 1 + (best - (n2 + 1));                     // workhorse.c:1070
    if (rl > MAXRL_FAST) { push_long(o, t.a0 + rd0, t.seg_end, 1, rl); return; }
    const int idx = atomicAdd(&s.nrd, 1);
    Rd q; q.pos = (uint16_t) rd0; q.rl = (uint16_t) rl; q.poff = 0;
    q.flags = (p.seed_ok && rl >= p.min_seed_rl) ? 1 : 0;
    s.rd[idx] = q;
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

// P3a: per-read offsets into the packed area and the probe work list (one exclusive scan)
__device__ void layout_reads(GroupSmem &s, const KvqOut &o, const TileCtx &t, int gid, int gt, int nrd,
                             uint32_t &totw, uint32_t &totq)
{
    uint32_t nw = 0, nq = 0; int rl = 0;
    if (gt < nrd) {
        rl = s.rd[gt].rl;
        nw = (uint32_t) ((rl + 15) >> 4) + 1;
        if (s.rd[gt].flags & 1) nq = (uint32_t) ((rl - KVQ_SEED_K + 16) >> 4);
    }
    uint32_t total;
    const uint32_t ex = group_exscan(nw | (nq << 16), s.warp_tot, &total, gid, gt);
    uint32_t pw = ex & 0xFFFFu, pq = ex >> 16;
    if (gt < nrd) {
        if (pw + nw > PACKW || pq + nq > QCAP) {
            // does not fit this round's work lists: hand the read to the long-record kernel
            push_long(o, t.a0 + s.rd[gt].pos, t.seg_end, 1, rl);
            s.rd[gt].flags = 2;                                     // dead for the rest of the round
        }
        s.rd[gt].poff = (uint16_t) min(pw, (uint32_t) PACKW);
        for (uint32_t j = 0; j < nw && pw + j < PACKW; j++) s.pitem[pw + j] = (uint8_t) gt;
        for (uint32_t q = 0; q < nq && pq + q < QCAP; q++) s.qitem[pq + q] = (uint16_t) ((gt << 6) | q);
    }
    totw = min(total & 0xFFFFu, (uint32_t) PACKW);
    totq = min(total >> 16, (uint32_t) QCAP);
    group_sync(gid);
}

// P3b: 2-bit pack, one packed word (16 bases) per step
__device__ void pack_reads(GroupSmem &s, int gt, int totw)
{
    const uint32_t *tw = reinterpret_cast<const uint32_t *>(s.tile);
    for (int x = gt; x < totw; x += GT) {
        const int r = s.pitem[x];
        const Rd rd = s.rd[r];
        if (rd.flags & 2) continue;                                 // dead read (prefix holes)
        const int j = x - rd.poff;
        const int nvalid = rd.rl - 16 * j;
        uint32_t codes = 0;
        if (nvalid > 0) {
            const int pos = rd.pos + 16 * j;
            const int a = pos >> 2; const uint32_t sh = (uint32_t) (pos & 3) << 3;
            uint32_t w0 = tw[a], w1 = tw[a + 1], w2 = tw[a + 2], w3 = tw[a + 3], w4 = tw[a + 4];
            codes = pack4(funnel_r(w0, w1, sh)) | (pack4(funnel_r(w1, w2, sh)) << 8)
                  | (pack4(funnel_r(w2, w3, sh)) << 16) | (pack4(funnel_r(w3, w4, sh)) << 24);
            if (nvalid < 16) codes &= (1u << (2 * nvalid)) - 1u;
        }
        s.packed[x] = codes;
    }
}

__device__ __forceinline__ void push_seeds(GroupSmem &s, int r, int pbase, uint32_t mask, int kind)
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

// P4: probes.  T1: every read offset, 16 offsets per work item.  T2: a few fixed offsets per read.
__device__ void probe_reads(GroupSmem &s, const Smem &sh, const KvqParams &p, const KvqTemplates &t,
                            int gt, int nrd, int totq)
{
    for (int i = gt; i < totq; i += GT) {
        const int it = s.qitem[i];
        const int r = it >> 6, q = it & 63;
        const Rd rd = s.rd[r];
        if (rd.flags & 2) continue;
        const int npos = rd.rl - KVQ_SEED_K + 1 - 16 * q;           // probe offsets left from 16q on
        const uint32_t *pk = s.packed + rd.poff + q;
        const uint32_t lo = pk[0], hi = pk[1];
        uint32_t mask = 0;
#pragma unroll
        for (int j = 0; j < 16; j++) {
            uint32_t key = funnel_r(lo, hi, 2 * j) & 0xFFFFu;
            mask |= ((sh.bm1[key >> 5] >> (key & 31)) & 1u) << j;
        }
        if (npos < 16) mask &= (1u << npos) - 1u;
        if (mask) push_seeds(s, r, 16 * q, mask, 0);
    }
    // T2: a few fixed offsets per read (whole-read / head / tail pieces), table per read length
    const int n2 = p.n2max;
    for (int x = gt; x < nrd * n2; x += GT) {
        const int r = (int) (((uint32_t) x * p.div_magic_n2) >> 16), i = x - r * n2;
        const Rd rd = s.rd[r];
        if ((rd.flags & 3) != 1) continue;
        const int pos = __ldg(t.p2_tab + rd.rl * KVQ_P2W + i);
        if (pos < 0) continue;
        const uint32_t key = extract16(s.packed + rd.poff, pos) & 0xFFFFu;
        if ((sh.bm2[key >> 5] >> (key & 31)) & 1u) push_seeds(s, r, pos, 1u, 1);
    }
}

// P5: walk the bucket of every seed and queue the diagonals the reference would evaluate
__device__ void expand_seeds(GroupSmem &s, const KvqParams &p, const KvqTemplates &t, int gt)
{
    const int nseed = s.nseed;
    for (int i = gt; i < nseed; i += GT) {
        const uint32_t sd = s.seedq[i];
        const int r = sd >> 16, pos = (sd >> 1) & 0x7FFF, kind = sd & 1;
        const int rl = s.rd[r].rl;
        const uint32_t key = extract16(s.packed + s.rd[r].poff, pos) & 0xFFFFu;
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
__device__ bool canonical_seed(const KvqParams &p, const KvqTemplates &t, const uint32_t *rpk, const uint32_t *tpk,
                               int rl, int sl, int d, int kind, int pos)
{
    for (int j = 0; j < p.npieces; j++) {
        const int toff = j * sl / p.npieces;
        const int pj = toff - d;
        if (pj < 0 || pj > rl - KVQ_SEED_K) continue;
        if (((extract16(rpk, pj) ^ extract16_g(tpk, toff)) & 0xFFFFu) == 0)
            return kind == 0 && pos == pj;
    }
    const int16_t *p2 = t.p2_tab + rl * KVQ_P2W;
    for (int i = 0; i < p.n2max; i++) {
        const int pp = __ldg(p2 + i);
        if (pp < 0) continue;
        const int toff = pp + d;
        if (toff < 0 || toff > sl - KVQ_SEED_K) continue;
        if (((extract16(rpk, pp) ^ extract16_g(tpk, toff)) & 0xFFFFu) == 0)
            return kind == 1 && pos == pp;
    }
    return false;
}

// Warp-cooperative add_hit + Coverage.apply_hit: lane 0 reserves the output slots, all lanes copy
// the hitseq bytes and update the coverage arrays.
__device__ void emit_hit_warp(const KvqTemplates &t, const KvqParams &p, const KvqOut &o, int seq, int d,
                              int ro, int to, int L, int rl, int64_t file_pos, const uint8_t *rd,
                              const uint8_t *tp, int copies, int rule)
{
    const int lane = threadIdx.x & 31;
    unsigned long long slot = 0, boff = 0;
    if (lane == 0) {
        if (p.accumulate) {
            atomicAdd(&o.nseqhits[seq], (unsigned long long) copies);
            atomicAdd(&o.nseqbasehits[seq], (unsigned long long) copies * (unsigned long long) L);
        }
        slot = atomicAdd(&o.ctr->hit_count, (unsigned long long) copies);
        if (p.want_hits) boff = atomicAdd(&o.ctr->blob_used, (unsigned long long) copies * (unsigned long long) L);
    }
    slot = __shfl_sync(0xffffffffu, slot, 0);
    boff = __shfl_sync(0xffffffffu, boff, 0);
    if (p.want_hits) {
        for (int c = 0; c < copies; c++) {
            if (lane == 0 && slot + c < o.hit_cap) {
                KvqHitRec h;
                h.seq_nr = seq; h.seq_pos = d; h.file_pos = file_pos; h.length = L; h.readlength = rl;
                h.blob_off = (long long) (boff + (unsigned long long) c * L);
                h.rule = (copies == 2) ? (c == 0 ? 1 : 2) : rule; h.pad = 0;
                o.hits[slot + c] = h;
            }
            if (boff + (unsigned long long) (c + 1) * L <= o.blob_cap)
                for (int j = lane; j < L; j += 32) o.blob[boff + (unsigned long long) c * L + j] = rd[ro + j];
        }
    }
    if (t.nplus > 0 && p.accumulate) {
        const bool plus = seq < t.nplus;
        const int tpl = plus ? seq : seq - t.nplus;
        const int sl = __ldg(t.len + seq);
        const unsigned long long cb = __ldg(t.cov_off + tpl);
        for (int j = lane; j < L; j += 32) {
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

// P6a: one thread per candidate diagonal: Hamming distance in code space (XOR/popc on packed
// words), a lower bound of the byte distance.  Survivors are queued for P6b.
__device__ void filter_candidates(GroupSmem &s, const KvqParams &p, const KvqTemplates &t, int gt)
{
    const int ncand = s.ncand;
    for (int i = gt; i < ncand; i += GT) {
        const unsigned long long c = s.candq[i];
        const int r = (int) (c & 0xFF), sq = (int) ((c >> 8) & 0xFFFFFF);
        const int d = (int) ((c >> 32) & 0xFFFF) - 32768;
        const Rd rd = s.rd[r];
        const int rl = rd.rl, sl = __ldg(t.len + sq);
        int ro, to, L, rule;
        diag_rule(d, rl, sl, p.minoverlap, ro, to, L, rule);
        const uint32_t *rpk = s.packed + rd.poff;
        const uint32_t *tpk = t.packed + __ldg(t.packed_off + sq);
        int e = 0;
        for (int j = 0; j < L && e <= p.maxerrors; j += 16) {
            uint32_t x = extract16(rpk, ro + j) ^ extract16_g(tpk, to + j);
            uint32_t m = (x | (x >> 1)) & 0x55555555u;
            if (L - j < 16) m &= (1u << (2 * (L - j))) - 1u;
            e += __popc(m);
        }
        if (e > p.maxerrors) continue;
        const int at = atomicAdd(&s.nacc, 1);
        if (at < ACCQ) s.accq[at] = (uint16_t) i;
    }
}

// P6b: one warp per survivor: byte-exact confirm, canonical-seed dedupe, emission
__device__ void confirm_candidates(GroupSmem &s, const KvqBatch &b, const TileCtx &tc, const KvqParams &p,
                                   const KvqTemplates &t, const KvqOut &o, int gt)
{
    const uint8_t *tb = reinterpret_cast<const uint8_t *>(s.tile);
    const int lane = gt & 31;
    const int nacc = s.nacc < ACCQ ? s.nacc : ACCQ;
    for (int a = gt >> 5; a < nacc; a += GW) {
        const unsigned long long c = s.candq[s.accq[a]];
        const int r = (int) (c & 0xFF), sq = (int) ((c >> 8) & 0xFFFFFF);
        const int d = (int) ((c >> 32) & 0xFFFF) - 32768;
        const int pos = (int) ((c >> 48) & 0x3FF), kind = (int) ((c >> 58) & 1);
        const Rd rd = s.rd[r];
        const int rl = rd.rl, sl = __ldg(t.len + sq);
        int ro, to, L, rule;
        const int copies = diag_rule(d, rl, sl, p.minoverlap, ro, to, L, rule);
        const uint8_t *rdp = tb + rd.pos;
        const uint8_t *tp = t.bytes + __ldg(t.byte_off + sq);
        int e = 0;
        for (int j = lane; j < L; j += 32) e += (rdp[ro + j] != tp[to + j]);
        e = __reduce_add_sync(0xffffffffu, e);
        if (e > p.maxerrors) continue;
        const uint32_t *rpk = s.packed + rd.poff;
        const uint32_t *tpk = t.packed + __ldg(t.packed_off + sq);
        if (!canonical_seed(p, t, rpk, tpk, rl, sl, d, kind, pos)) continue;      // warp-uniform
        emit_hit_warp(t, p, o, sq, d, ro, to, L, rl, b.base_fpos + tc.a0 + rd.pos, rdp, tp, copies, rule);
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
__device__ void run_round(GroupSmem &s, Smem &sh, const KvqBatch &b, const TileCtx &tc, const KvqParams &p,
                          const KvqTemplates &t, const KvqOut &o, int off, int k0, int nrec, int kcomplete,
                          int gid, int gt, ThreadStats &ts)
{
    if (gt == 0) { s.nrd = 0; s.nseed = 0; s.ncand = 0; s.nacc = 0; s.n_exh_reads = 0; s.n_seed_reads = 0; }
    group_sync(gid);
    process_records(s, sh, b, tc, p, o, off, k0, nrec, kcomplete, gt, ts);
    group_sync(gid);
    const int nrd = s.nrd;
    group_sync(gid);                                // everyone has read nrd before it is reset
    if (nrd == 0) return;                           // uniform
    const uint8_t *tb = reinterpret_cast<const uint8_t *>(s.tile);
    bool seeded = p.seed_ok && t.n_exh < t.nseqs;
    if (seeded) {
        uint32_t totw, totq;
        layout_reads(s, o, tc, gid, gt, nrd, totw, totq);
        pack_reads(s, gt, (int) totw);
        group_sync(gid);
        probe_reads(s, sh, p, t, gt, nrd, (int) totq);
        group_sync(gid);
        if (s.nseed > SEEDQ) seeded = false;        // uniform: queue overflow -> exhaustive round
        if (seeded) {
            expand_seeds(s, p, t, gt);
            group_sync(gid);
            if (s.ncand > CANDQ) seeded = false;
        }
        if (seeded) {
            filter_candidates(s, p, t, gt);
            group_sync(gid);
            if (s.nacc > ACCQ) seeded = false;      // uniform: too many survivors -> exhaustive round
        }
        if (seeded) {
            if (gt == 0) ts.cands += (unsigned long long) s.ncand;
            confirm_candidates(s, b, tc, p, t, o, gt);
        } else if (gt == 0) {
            atomicAdd(&o.ctr->overflow_rounds, 1u);
        }
    }
    // split the reads of the round: seeded ones still owe the non-seedable templates
    if (gt < nrd) {
        const int fl = s.rd[gt].flags;
        if (!(fl & 2)) {
            if (seeded && (fl & 1)) s.seed_reads[atomicAdd(&s.n_seed_reads, 1)] = (uint16_t) gt;
            else s.exh_reads[atomicAdd(&s.n_exh_reads, 1)] = (uint16_t) gt;
        }
    }
    group_sync(gid);
    const int wid = gt >> 5;
    const int ne = s.n_exh_reads, ns = s.n_seed_reads;
    if (gt == 0) ts.exh += (unsigned long long) ne;
    for (int pair = wid; pair < ne * t.nseqs; pair += GW) {
        const int r = s.exh_reads[pair / t.nseqs], sq = pair % t.nseqs;
        exhaustive_pair(p, t, o, tb + s.rd[r].pos, s.rd[r].rl, b.base_fpos + tc.a0 + s.rd[r].pos, sq);
    }
    if (t.n_exh > 0)
        for (int pair = wid; pair < ns * t.n_exh; pair += GW) {
            const int r = s.seed_reads[pair / t.n_exh], sq = __ldg(t.exh_list + pair % t.n_exh);
            exhaustive_pair(p, t, o, tb + s.rd[r].pos, s.rd[r].rl, b.base_fpos + tc.a0 + s.rd[r].pos, sq);
        }
}

__global__ void __launch_bounds__(THREADS, 1)
kvq_scan_kernel(KvqBatch b, KvqTemplates t, KvqParams p, KvqOut o)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Smem &sh = *reinterpret_cast<Smem *>(smem_raw);
    const int gid = threadIdx.x / GT, gt = threadIdx.x % GT;
    GroupSmem &s = sh.g[gid];

    for (int i = threadIdx.x; i < KVQ_BITMAP_WORDS; i += THREADS) {
        sh.bm1[i] = __ldg(t.t1_bitmap + i);
        sh.bm2[i] = __ldg(t.t2_bitmap + i);
    }
    for (int i = threadIdx.x; i < KVQ_HIST; i += THREADS) sh.hist[i] = 0;
    if (threadIdx.x == 0) sh.longest = -1;
    __syncthreads();
    ThreadStats ts{0, 0, 0, 0};

    for (;;) {
        if (gt == 0) s.tile_cur = (int) atomicAdd(&o.ctr->tile_next, 1u);
        group_sync(gid);
        const int tile = s.tile_cur;
        if (tile >= b.ntiles) break;
        const int seg = __ldg(b.tile_seg + tile);
        const int64_t t0 = __ldg(b.tile_off + tile);
        const int64_t seg_beg = __ldg(b.seg_off + seg), seg_end = __ldg(b.seg_off + seg + 1);
        TileCtx tc;
        tc.a0 = t0 & ~(int64_t) 31;
        tc.shift = (int) (t0 - tc.a0);
        const int64_t t1 = t0 + TILE < seg_end ? t0 + TILE : seg_end;
        const int64_t g1 = t1 + MARGIN < seg_end ? t1 + MARGIN : seg_end;
        tc.tile_end = (int) (t1 - tc.a0);
        tc.reg_end = (int) (g1 - tc.a0);
        tc.first = (t0 == seg_beg);
        tc.at_seg_end = (g1 == seg_end);
        tc.seg_end = seg_end;
        const int n_all = (int) load_and_index(s, b, tc, gid, gt);
        const int n_tile = (int) nl_before(s, tc.tile_end);
        if (gt == 0) s.phase_in = resolve_phase(b, tile, tc.first, n_tile);
        group_sync(gid);
        const int phase = s.phase_in;
        // newline rank -> record-relative numbering (see process_records)
        const int off = tc.first ? 0 : ((3 - phase) & 3) + 1;
        const int vmax_tile = n_tile - 1 - off, vmax_all = n_all - 1 - off;
        int nown = 0;
        if (tc.first || vmax_tile >= -1) nown = (vmax_tile + 1) / 4 + 1;
        // record k is complete iff its 4th newline v = 4k+3 is inside the staged region
        const int kcomplete = vmax_all >= 3 ? (vmax_all - 3) / 4 : -1;
        for (int k0 = 0; k0 < nown; k0 += CAP) {
            const int nrec = nown - k0 < CAP ? nown - k0 : CAP;
            run_round(s, sh, b, tc, p, t, o, off, k0, nrec, kcomplete, gid, gt, ts);
        }
        group_sync(gid);                            // tile buffers are reused by the next tile
    }

    // flush the statistics: warp-reduce the per-thread counters, one atomic per warp
#pragma unroll
    for (int o2 = 16; o2 > 0; o2 >>= 1) {
        ts.records += __shfl_down_sync(0xffffffffu, ts.records, o2);
        ts.bases += __shfl_down_sync(0xffffffffu, ts.bases, o2);
        ts.exh += __shfl_down_sync(0xffffffffu, ts.exh, o2);
        ts.cands += __shfl_down_sync(0xffffffffu, ts.cands, o2);
    }
    if ((threadIdx.x & 31) == 0) {
        if (ts.records) atomicAdd(&o.ctr->records, ts.records);
        if (ts.bases) atomicAdd(&o.ctr->bases, ts.bases);
        if (ts.exh) atomicAdd(&o.ctr->exhaustive_reads, ts.exh);
        if (ts.cands) atomicAdd(&o.ctr->candidates, ts.cands);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < KVQ_HIST; i += THREADS)
        if (sh.hist[i]) atomicAdd(&o.ctr->rls[i], (unsigned long long) sh.hist[i]);
    if (threadIdx.x == 0) atomicMax(&o.ctr->longest, (long long) sh.longest);
}

// Records that do not fit a tile's spill margin, or whose trimmed read exceeds the packed work
// lists: one warp per record straight from global memory, exhaustive matcher.  Rare by construction.
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
int kvq_scan_groups(void) { return G; }

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
    int grid = (b.ntiles + G - 1) / G;
    if (grid > sm_count) grid = sm_count;
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
