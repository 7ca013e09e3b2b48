// sm_100a kernels of the KvarQ read-scanning path (reference: csrc/workhorse.c:976-1197).
//
// kvq_scan_kernel: persistent, one CTA per SM, every WARP an independent worker (no block-wide barrier
// in the steady state).  A warp claims 10 KiB tiles of the FastQ stream in stream order from a global
// queue and runs the path on its tile:
//   P1  coalesced 128-bit loads, 32 bytes per lane, rows through three register sets; straight from
//       registers every 32-byte word is reduced to bit planes in shared memory: quality-pass mask
//       (byte >= Amin) and 2-bit base codes; the newline mask stays in registers.  One packed warp scan
//       ranks the newlines of three rows and their positions are scattered without divergent branches.
//       The line phase of the tile (newlines of its segment before it, mod 4) is INFERRED from the
//       tile's own line starts ('@' / '+' pattern), published at once and checked against the
//       predecessor's prefix afterwards; an undecided tile walks back to the nearest prefix
//       (decoupled look-back); a contradiction makes the host re-run the batch with exact phases
//                                                                              [workhorse.c:1018-1030]
//   P2  one lane per record: its 4 newlines, '@' / '+' sanity checks, Phred trim (first longest
//       run >= Amin) on the quality-pass plane, histogram, length filter  [workhorse.c:1037-1103]
//   P3  seed probes on the code plane: every read offset against the byte-per-key role table T1
//       (template pieces) in shared memory, a few fixed read offsets against the bitmap T2 (all
//       template offsets); seeds go to a small per-warp stack that is drained 32 at a time
//   P4  one candidate per seed and pass (table entries in L2), XOR/popc Hamming verify on packed
//       words; the few survivors are queued for the emit kernel
// kvq_emit_kernel: one warp per survivor -- byte-exact confirm, canonical-seed dedupe, hit emission
// through atomics and the fused coverage update                    [workhorse.c:1107-1175,408-443]
// kvq_long_kernel: records longer than a tile's spill rows / reads longer than 512 bases, exhaustively.
// kvq_order_*: the hits of an HBM-resident scan in the reference's nthreads=1 order, exported into the
// result's page-locked buffer.  kvq_reset_kernel / kvq_tail_kernel: a batch's start state; the tail of
// the counter row for the all-reduce.
// Reads or templates the seed scheme cannot cover take the exhaustive matcher (same acceptance
// rule, every diagonal).  Acceptance rule: SURVEY.md Appendix B / workhorse.c:1112-1174.
#include <cuda_runtime.h>
#include <stdint.h>
#include "kvq_device.h"

namespace {

#ifndef KVQ_NWARPS
#define KVQ_NWARPS 23                   // warps per CTA (one CTA per SM): 80 registers, 216 KB shared memory
#endif
#ifndef KVQ_PREFETCH
#define KVQ_PREFETCH 1
#endif
#ifndef KVQ_NS
#define KVQ_NS 1
#endif
#ifndef KVQ_FMASK
#define KVQ_FMASK 1
#endif
#ifndef KVQ_OLD_P1
#define KVQ_OLD_P1 0                    // 1: per-row newline ranking (round 1), for A/B timing
#endif
#ifndef KVQ_ROW_SETS
#define KVQ_ROW_SETS 3                  // register sets the rows of a tile rotate through (rows in flight + 1)
#endif
constexpr int NWARPS = KVQ_NWARPS;          // warps per CTA (shared memory bound)
constexpr int THREADS = NWARPS * 32;
constexpr int OWN_ROWS = KVQ_OWN_ROWS;
constexpr int EXTRA_ROWS = 2;               // spill: records starting in the tile may end here
constexpr int MAX_ROWS = OWN_ROWS + EXTRA_ROWS;
constexpr int NWORDS = MAX_ROWS * 32;       // 32-byte words of the staged region
constexpr int NLCAP = 144;                  // newline positions kept per round (32 records + slack)
constexpr int MAXRL_FAST = KVQ_MAXRL_FAST;  // longer trimmed reads go to the long-record kernel
constexpr int NS = KVQ_NS;                  // seeds per lane and drain
constexpr int DRAIN = 32 * NS;
constexpr int SEEDQ = 2 * DRAIN;

struct WarpSmem {
    uint32_t ge[NWORDS + 1];                // quality plane, bit j of word w <=> byte 32 w + j is >= Amin
    alignas(16) uint32_t pk[2 * NWORDS + 4]; // code plane: 2 bits per byte, contiguous stream (read as uint2)
    uint16_t nlpos[NLCAP + 2];              // positions of the newlines ranked [base, base + NLCAP); [NLCAP]: spare slot (store_two)
    uint32_t rd[32];                        // reads of the round: start | length << 16 (0 = none)
    uint32_t seedq[SEEDQ];
    uint32_t qn;                            // seeds reserved on the stack (may exceed SEEDQ until drained)
    uint32_t pad;
};

struct Smem {
    uint8_t t1[KVQ_NKEYS];                  // T1 presence, one byte per key
    uint32_t bm2[KVQ_BITMAP_WORDS];         // T2 presence, one bit per key
#if KVQ_T2L
    uint32_t bm2l[KVQ_BITMAP_WORDS];        // T2 restricted to the long templates (whole-read probes)
#endif
    uint32_t hist[KVQ_HIST];
    uint32_t tpk[KVQ_TPK_SMEM_WORDS];       // packed template codes when they fit (else read through L1/L2)
    int longest;
    int pad[3];
    WarpSmem w[NWARPS];
};
static_assert(sizeof(Smem) <= 232448, "shared memory budget (227 KB) exceeded");

struct ThreadStats { unsigned long long records, bases, exh, cands; };

__device__ __forceinline__ uint32_t funnel_r(uint32_t lo, uint32_t hi, uint32_t s)
{
    return __funnelshift_r(lo, hi, s);
}

// ---- byte planes: 4 bytes -> flags in bit 7 of every byte, gathered with dot products -------
// bit7 of every byte of w that equals '\n'
// `c` = 0x0A0A0A0A held in a register (not an immediate: a LOP3 takes one immediate, with two the
// xor and the and are two instructions)
__device__ __forceinline__ uint32_t nl_bits(uint32_t w, uint32_t c)
{
    uint32_t s = ((w ^ c) & 0x7F7F7F7Fu) + 0x7F7F7F7Fu;             // bit7: low 7 bits differ from '\n'
    return ~(s | w) & 0x80808080u;
}
// bit7 of every byte b of w with (signed char) b >= amin, amin >= 0.  cl = amin * 0x01010101
__device__ __forceinline__ uint32_t ge_bits_pos(uint32_t w, uint32_t cl)
{
    uint32_t t = (w | 0x80808080u) - cl;            // bit7: low 7 bits of b >= amin
    return ~w & t & 0x80808080u;
}
// amin < 0: cl = (amin & 0x7f) * 0x01010101; bytes 0..127 always pass
__device__ __forceinline__ uint32_t ge_bits_neg(uint32_t w, uint32_t cl)
{
    uint32_t t = (w | 0x80808080u) - cl;
    return (~w | t) & 0x80808080u;
}
// 8 flag bytes (0x80 / 0x00) of two words -> 8 mask bits at [7, 15)
__device__ __forceinline__ uint32_t gather2(uint32_t e0, uint32_t e1)
{
    return __dp4a(e0, 0x08040201u, __dp4a(e1, 0x80402010u, 0u));
}
__device__ __forceinline__ uint32_t gather32(uint32_t p0, uint32_t p1, uint32_t p2, uint32_t p3)
{
    return (p0 >> 7) | (p1 << 1) | (p2 << 9) | (p3 << 17);
}
// 8 bases (two words) -> 16 bits of 2-bit codes, base 0 in the low bits.  Code = (byte >> 1) & 3.
__device__ __forceinline__ uint32_t pack8(uint32_t w0, uint32_t w1)
{
    uint32_t d = __dp4a(w0 & 0x06060606u, 0x40100401u, 0u) + (__dp4a(w1 & 0x06060606u, 0x40100401u, 0u) << 8);
    return d >> 1;
}

// 16 bases starting at base offset `b` of a packed stream (needs word b/16+1 readable)
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

struct TileCtx {
    int64_t a0;          // batch offset of plane coordinate 0 (32-byte aligned)
    int shift;           // coordinate of the first byte of the tile
    int tile_end;        // coordinate one past the bytes owned by the tile
    int lim;             // coordinate one past the last byte of the segment (clamped to the planes)
    int first;           // first tile of its segment
    int64_t seg_end;     // batch offset of the segment end
};

// format-error key: smaller = earlier in the stream.  [63:22] stream offset of the record start,
// [21] 0 = bad '@' / 1 = bad '+', [20:0] payload (the offending byte, or plus_pos - record_start)
__device__ __forceinline__ void report_format_error(const KvqOut &o, int64_t rec_fpos, int type, uint32_t payload)
{
    unsigned long long key = ((unsigned long long) rec_fpos << 22) | ((unsigned long long) type << 21)
                           | (payload & 0x1FFFFFu);
    atomicMin(&o.ctr->err_key, key);
}

// Work for the emit kernel: the lanes with `want` append one record each (one atomic per warp).
// Nothing cold runs inside the scan kernel -- no calls, no spills on its hot paths.
__device__ __forceinline__ void push_survivors(const KvqOut &o, bool want, int64_t read_off, int rl, int seq, int d,
                                               int kindpos, int lane)
{
    const uint32_t m = __ballot_sync(0xffffffffu, want);
    if (!m) return;
    unsigned long long base = 0;
    if (lane == __ffs(m) - 1) base = atomicAdd(&o.ctr->surv_count, (unsigned long long) __popc(m));
    base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
    const unsigned long long at = base + (unsigned long long) __popc(m & ((1u << lane) - 1u));
    if (want && at < o.surv_cap) {
        uint4 *q = reinterpret_cast<uint4 *>(o.surv + at);
        q[0] = make_uint4((uint32_t) read_off, (uint32_t) ((unsigned long long) read_off >> 32), (uint32_t) rl, (uint32_t) seq);
        q[1] = make_uint4((uint32_t) d, (uint32_t) kindpos, 0u, 0u);
    }
}

__device__ __forceinline__ void push_long(const KvqOut &o, int64_t start, int64_t seg_end, int kind, int rl)
{
    unsigned long long i = atomicAdd(&o.ctr->long_count, 1ULL);
    if (i < o.long_cap) {
        KvqLongRec r; r.start = start; r.seg_end = seg_end; r.kind = kind; r.rl = rl;
        o.longs[i] = r;
    }
}

} // namespace

namespace {

// ---- P1 ---------------------------------------------------------------------------------------
struct Row { uint4 a, b; };

__device__ __forceinline__ Row load_row(const KvqBatch &b, const TileCtx &t, int row, int lane)
{
    // unconditional: the batch is followed by zeroed slack for a whole tile (engine.cpp, DEV_PAD); bytes
    // beyond the segment only reach the planes, never a result (newlines there are masked, reads end before)
    Row r;
    const uint4 *src = reinterpret_cast<const uint4 *>(b.data + t.a0) + 2 * (row * 32 + lane);
    r.a = __ldg(src); r.b = __ldg(src + 1);
    return r;
}

// Reduce one 32-byte word per lane to the planes: quality-pass mask and 2-bit codes go to their
// planes in shared memory, the newline mask (restricted to the tile's segment) is returned;
// rank_rows() turns the masks of three rows into ranked newline positions.
template <bool NEG>
__device__ __forceinline__ uint32_t reduce_row(WarpSmem &s, const TileCtx &t, const Row &r, int row, int lane,
                                               uint32_t cl, uint32_t cnl)
{
    const int w = row * 32 + lane;
    uint32_t nl = gather32(gather2(nl_bits(r.a.x, cnl), nl_bits(r.a.y, cnl)), gather2(nl_bits(r.a.z, cnl), nl_bits(r.a.w, cnl)),
                           gather2(nl_bits(r.b.x, cnl), nl_bits(r.b.y, cnl)), gather2(nl_bits(r.b.z, cnl), nl_bits(r.b.w, cnl)));
    uint32_t ge;
    if (NEG)
        ge = gather32(gather2(ge_bits_neg(r.a.x, cl), ge_bits_neg(r.a.y, cl)), gather2(ge_bits_neg(r.a.z, cl), ge_bits_neg(r.a.w, cl)),
                      gather2(ge_bits_neg(r.b.x, cl), ge_bits_neg(r.b.y, cl)), gather2(ge_bits_neg(r.b.z, cl), ge_bits_neg(r.b.w, cl)));
    else
        ge = gather32(gather2(ge_bits_pos(r.a.x, cl), ge_bits_pos(r.a.y, cl)), gather2(ge_bits_pos(r.a.z, cl), ge_bits_pos(r.a.w, cl)),
                      gather2(ge_bits_pos(r.b.x, cl), ge_bits_pos(r.b.y, cl)), gather2(ge_bits_pos(r.b.z, cl), ge_bits_pos(r.b.w, cl)));
    uint2 pk;
    pk.x = pack8(r.a.x, r.a.y) | (pack8(r.a.z, r.a.w) << 16);
    pk.y = pack8(r.b.x, r.b.y) | (pack8(r.b.z, r.b.w) << 16);
    if (row == 0 || (row + 1) * KVQ_ROW > t.lim) {  // warp-uniform: newlines outside [shift, lim) belong to other tiles' segments
        const int lo = t.shift - (w << 5), hi = t.lim - (w << 5);
        if (lo > 0) nl &= lo >= 32 ? 0u : (0xFFFFFFFFu << lo);
        if (hi < 32) nl &= hi <= 0 ? 0u : ((1u << hi) - 1u);
    }
    s.ge[w] = ge;
    reinterpret_cast<uint2 *>(s.pk)[w] = pk;
    return nl;
}

// the first two newlines of a word get their positions stored straight away (most words hold
// none or one, "\n+\n" gives two); what is left stays in `nl`
__device__ __forceinline__ void store_two(WarpSmem &s, uint32_t &nl, uint32_t rank, int wb)
{
    // unconditional stores (no divergent branches): what has no newline, or no room, lands in the
    // spare slot behind the list
    const uint32_t l0 = nl & (0u - nl);
    s.nlpos[(nl != 0u && rank < NLCAP) ? rank : NLCAP] = (uint16_t) (wb + 31 - __clz(l0));
    nl ^= l0;
    const uint32_t l1 = nl & (0u - nl);
    s.nlpos[(nl != 0u && rank + 1 < NLCAP) ? rank + 1 : NLCAP] = (uint16_t) (wb + 31 - __clz(l1));
    nl ^= l1;
}

// Newline masks of three rows (one word per lane and row) -> ranks and positions.  One shuffle
// scan serves the three rows: their per-lane counts (<= 32) travel in 11-bit fields of one word.
// Updates the running newline count of the tile and the lane's share of the newlines the tile
// owns (coordinates below tile_end; only the last rows of a tile can lie beyond it).
__device__ __forceinline__ void rank_rows(WarpSmem &s, const TileCtx &t, uint32_t nl0, uint32_t nl1, uint32_t nl2, int row0,
                                          bool last, int lane, uint32_t &nl_total, uint32_t &nl_owned)
{
    const uint32_t c0 = __popc(nl0), c1 = __popc(nl1), c2 = __popc(nl2);
    const uint32_t packed = c0 | (c1 << 11) | (c2 << 22);
    uint32_t inc = packed;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    const uint32_t tot = __shfl_sync(0xffffffffu, inc, 31), excl = inc - packed;
    const uint32_t t0 = tot & 0x7FFu, t1 = (tot >> 11) & 0x7FFu, t2 = tot >> 22;
    const uint32_t r0 = nl_total + (excl & 0x7FFu), r1 = nl_total + t0 + ((excl >> 11) & 0x7FFu),
                   r2 = nl_total + t0 + t1 + (excl >> 22);
    nl_total += t0 + t1 + t2;
    const int wb = (row0 * 32 + lane) << 5;         // tile coordinate of the lane's word in row0
    if (!last) nl_owned += c0 + c1 + c2;            // rows below OWN_ROWS - 1: beyond tile_end only what lim has masked already
    else {
        const int e0 = t.tile_end - wb;
        nl_owned += (uint32_t) (__popc(nl0 & __funnelshift_rc(0xFFFFFFFFu, 0u, (uint32_t) max(32 - max(e0, 0), 0)))
                              + __popc(nl1 & __funnelshift_rc(0xFFFFFFFFu, 0u, (uint32_t) max(32 - max(e0 - KVQ_ROW, 0), 0)))
                              + __popc(nl2 & __funnelshift_rc(0xFFFFFFFFu, 0u, (uint32_t) max(32 - max(e0 - 2 * KVQ_ROW, 0), 0))));
    }
    // positions of the first NLCAP newlines (round 0)
    store_two(s, nl0, r0, wb);
    store_two(s, nl1, r1, wb + KVQ_ROW);
    store_two(s, nl2, r2, wb + 2 * KVQ_ROW);
    if (__any_sync(0xffffffffu, (nl0 | nl1 | nl2) != 0)) {      // short lines: more than two newlines in 32 bytes
        uint32_t q = r0 + 2;
        while (nl0 && q < NLCAP) { s.nlpos[q++] = (uint16_t) (wb + 31 - __clz(nl0 & (0u - nl0))); nl0 &= nl0 - 1; }
        q = r1 + 2;
        while (nl1 && q < NLCAP) { s.nlpos[q++] = (uint16_t) (wb + KVQ_ROW + 31 - __clz(nl1 & (0u - nl1))); nl1 &= nl1 - 1; }
        q = r2 + 2;
        while (nl2 && q < NLCAP) { s.nlpos[q++] = (uint16_t) (wb + 2 * KVQ_ROW + 31 - __clz(nl2 & (0u - nl2))); nl2 &= nl2 - 1; }
    }
}

#if KVQ_OLD_P1
// Reduce one 32-byte word per lane to the three planes and rank its newlines (exclusive prefix
// kept per word, running total in `nl_total`, those owned by the tile in `nl_owned`).
template <bool NEG>
__device__ __forceinline__ void reduce_row_old(WarpSmem &s, const TileCtx &t, const Row &r, int row, int lane,
                                           uint32_t cl, uint32_t cnl, uint32_t &nl_total, uint32_t &nl_owned)
{
    const int w = row * 32 + lane;
    uint32_t nl = gather32(gather2(nl_bits(r.a.x, cnl), nl_bits(r.a.y, cnl)), gather2(nl_bits(r.a.z, cnl), nl_bits(r.a.w, cnl)),
                           gather2(nl_bits(r.b.x, cnl), nl_bits(r.b.y, cnl)), gather2(nl_bits(r.b.z, cnl), nl_bits(r.b.w, cnl)));
    uint32_t ge;
    if (NEG)
        ge = gather32(gather2(ge_bits_neg(r.a.x, cl), ge_bits_neg(r.a.y, cl)), gather2(ge_bits_neg(r.a.z, cl), ge_bits_neg(r.a.w, cl)),
                      gather2(ge_bits_neg(r.b.x, cl), ge_bits_neg(r.b.y, cl)), gather2(ge_bits_neg(r.b.z, cl), ge_bits_neg(r.b.w, cl)));
    else
        ge = gather32(gather2(ge_bits_pos(r.a.x, cl), ge_bits_pos(r.a.y, cl)), gather2(ge_bits_pos(r.a.z, cl), ge_bits_pos(r.a.w, cl)),
                      gather2(ge_bits_pos(r.b.x, cl), ge_bits_pos(r.b.y, cl)), gather2(ge_bits_pos(r.b.z, cl), ge_bits_pos(r.b.w, cl)));
    uint2 pk;
    pk.x = pack8(r.a.x, r.a.y) | (pack8(r.a.z, r.a.w) << 16);
    pk.y = pack8(r.b.x, r.b.y) | (pack8(r.b.z, r.b.w) << 16);
    uint32_t cnt;
    const int end = t.lim < t.tile_end ? t.lim : t.tile_end;
    if (row > 0 && (row + 1) * KVQ_ROW <= end) {    // warp-uniform: every byte of the row is owned
        cnt = __popc(nl);
        nl_owned += cnt;
    } else {
        // newlines outside [shift, lim) belong to other tiles' segments
        const int lo = t.shift - (w << 5), hi = t.lim - (w << 5);
        if (lo > 0) nl &= lo >= 32 ? 0u : (0xFFFFFFFFu << lo);
        if (hi < 32) nl &= hi <= 0 ? 0u : ((1u << hi) - 1u);
        const int own = t.tile_end - (w << 5);
        nl_owned += __popc(own >= 32 ? nl : own <= 0 ? 0u : (nl & ((1u << own) - 1u)));
        cnt = __popc(nl);
    }
    s.ge[w] = ge;                                   // (the newline mask is not kept: ranks + the first NLCAP positions are)
    reinterpret_cast<uint2 *>(s.pk)[w] = pk;
    uint32_t inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    uint32_t rank = nl_total + inc - cnt;
    nl_total += __shfl_sync(0xffffffffu, inc, 31);
    while (nl && rank < NLCAP) {                    // positions of the first NLCAP newlines (round 0)
        s.nlpos[rank++] = (uint16_t) ((w << 5) + __ffs(nl) - 1);
        nl &= nl - 1;
    }
}

#endif

// positions of the newlines ranked [base, base + NLCAP) -- only tiles with more than NLCAP
// newlines (very short records) come here: the masks are computed again from the bytes (L1/L2)
__device__ __forceinline__ void scatter_newlines(WarpSmem &s, const KvqBatch &b, const TileCtx &t, int nrows,
                                                 int base, int lane, uint32_t cnl)
{
    int total = 0;
#pragma unroll 1
    for (int row = 0; row < nrows; row++) {
        const int w = row * 32 + lane;
        const Row r = load_row(b, t, row, lane);
        uint32_t m = gather32(gather2(nl_bits(r.a.x, cnl), nl_bits(r.a.y, cnl)), gather2(nl_bits(r.a.z, cnl), nl_bits(r.a.w, cnl)),
                              gather2(nl_bits(r.b.x, cnl), nl_bits(r.b.y, cnl)), gather2(nl_bits(r.b.z, cnl), nl_bits(r.b.w, cnl)));
        const int lo = t.shift - (w << 5), hi = t.lim - (w << 5);
        if (lo > 0) m &= lo >= 32 ? 0u : (0xFFFFFFFFu << lo);
        if (hi < 32) m &= hi <= 0 ? 0u : ((1u << hi) - 1u);
        const int cnt = __popc(m);
        int inc = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int n = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += n;
        }
        int rank = total + inc - cnt - base;
        total += __shfl_sync(0xffffffffu, inc, 31);
        while (m && rank < NLCAP) {
            if (rank >= 0) s.nlpos[rank] = (uint16_t) ((w << 5) + __ffs(m) - 1);
            rank++;
            m &= m - 1;
        }
    }
}

// Decoupled look-back over the tiles of one segment, 32 predecessors per step.  Status word:
// bit3 = inclusive prefix, bit2 = tile aggregate, bits 1:0 = newline count mod 4.  Tiles are
// claimed in stream order, so a predecessor is always resident or finished.
__device__ __forceinline__ int resolve_phase(const KvqBatch &b, int tile, int first, int n_tile, int lane)
{
    volatile uint32_t *st = b.tile_status;
    if (first) return 0;                            // (the first tile of a segment: nothing in front of it; its owner publishes the inclusive prefix)
    int acc = 0;
    for (int j = tile - 1;;) {
        const int i = j - lane;
        const uint32_t v = i >= 0 ? st[i] : 8u;
        const uint32_t incl = __ballot_sync(0xffffffffu, (v & 8u) != 0);
        const uint32_t zero = __ballot_sync(0xffffffffu, v == 0);
        const uint32_t upto = incl ? ((2u << (__ffs(incl) - 1)) - 1u) : 0xFFFFFFFFu;    // lanes up to the nearest prefix
        if (zero & upto) { __nanosleep(40); continue; }
        acc += __reduce_add_sync(0xffffffffu, ((upto >> lane) & 1u) ? (v & 3u) : 0u);
        if (incl) break;
        j -= 32;
    }
    acc &= 3;
    if (lane == 0) st[tile] = 8u | (uint32_t) ((acc + n_tile) & 3);
    return acc;
}

// Line phase of a tile from its own line starts.  In a stream the reference accepts, every line
// 0 of a record starts with '@' and every line 2 with '+' (workhorse.c:1037-1048 checks nothing
// else), so the true phase passes this test on the first lines of the tile; when it is the ONLY
// phase that passes, it is the phase -- no waiting for the predecessors' newline counts.
// Returns -1 when the test is not decisive (fewer than a record's worth of lines, or quality
// lines that happen to start with '@').  The caller still checks the phase against the counts
// after the tile is done; a contradiction (malformed input: the reference raises) re-runs the
// batch with exact phases, so a wrong guess never reaches a result.
__device__ __forceinline__ int infer_phase(const WarpSmem &s, const KvqBatch &b, const TileCtx &t, int n_all, int lane)
{
    __syncwarp();                                   // nlpos[] of this tile
    const int m = n_all < 32 ? n_all : 32;
    uint8_t ch = 0; bool v = false;
    if (lane < m) {
        const int64_t at = t.a0 + (int64_t) s.nlpos[lane] + 1;      // first byte of the line after newline `lane`
        if (at < t.seg_end) { ch = __ldg(b.data + at); v = true; }
    }
    const uint32_t valid = __ballot_sync(0xffffffffu, v);
    const uint32_t not_at = valid & ~__ballot_sync(0xffffffffu, ch == '@');
    const uint32_t not_plus = valid & ~__ballot_sync(0xffffffffu, ch == '+');
    int found = -1, nfound = 0;
#pragma unroll
    for (int ph = 0; ph < 4; ph++) {
        // the line after newline i is line (ph + i + 1) & 3 of its record
        const uint32_t l0 = 0x11111111u << ((3 - ph) & 3), l2 = 0x11111111u << ((1 - ph) & 3);
        if (!((l0 & not_at) | (l2 & not_plus))) { found = ph; nfound++; }
    }
    return nfound == 1 ? found : -1;
}

// ---- P2 ---------------------------------------------------------------------------------------
// Phred trim of the quality line [qs, qe] (qe = its '\n'), workhorse.c:1055-1068: the first
// longest run of bytes >= Amin that is closed by a byte < Amin at or before qe.
__device__ __forceinline__ void phred_trim(const WarpSmem &s, int qs, int qe, int &rl, int &best)
{
    int runstart = qs;
    rl = 0; best = qs;
    for (int w = qs >> 5; w <= (qe >> 5); w++) {
        uint32_t bad = ~s.ge[w];
        const int gb = w << 5;
        if (gb < qs) bad &= 0xFFFFFFFFu << (qs - gb);
        if (qe - gb < 31) bad &= (2u << (qe - gb)) - 1u;
        while (bad) {
            const int z = gb + __ffs(bad) - 1;
            bad &= bad - 1;
            const int len = z - runstart;
            if (len > rl) { rl = len; best = runstart; }
            runstart = z + 1;
        }
    }
}

} // namespace

namespace {

// ---- emission / coverage ----------------------------------------------------------------------
__device__ __forceinline__ int base_column(uint8_t c)
{
    return c == 'A' ? 0 : c == 'C' ? 1 : c == 'G' ? 2 : c == 'T' ? 3 : c == 'N' ? 4 : 5;
}
__device__ __forceinline__ uint8_t complement(uint8_t c)     // Sequence.pairs, kvarq/genes.py:204
{
    return c == 'A' ? 'T' : c == 'T' ? 'A' : c == 'C' ? 'G' : c == 'G' ? 'C' : c;
}

// add_hit (workhorse.c:408-443) + Coverage.apply_hit (kvarq/analyse.py:57-78), one thread.
// `rd` points at the trimmed read's bytes, `tp` at the template's bytes (both global).
__device__ __noinline__ void emit_hit(const KvqTemplates &t, const KvqParams &p, const KvqOut &o, int seq, int d,
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

// Cooperative variant for a group of G lanes (gmask = the group's lanes, sub = lane within the group, first =
// the group's first lane): the first lane reserves the output slots, all lanes copy the hitseq bytes and update
// the coverage arrays.
__device__ __noinline__ void emit_hit_warp(const KvqTemplates &t, const KvqParams &p, const KvqOut &o, int seq, int d,
                              int ro, int to, int L, int rl, int sl, int64_t file_pos, const uint8_t *rd,
                              const uint8_t *tp, int copies, int rule, int lane, int G, uint32_t gmask, int first)
{
    unsigned long long slot = 0, boff = 0;
    if (lane == 0) {
        if (p.accumulate) {
            atomicAdd(&o.nseqhits[seq], (unsigned long long) copies);
            atomicAdd(&o.nseqbasehits[seq], (unsigned long long) copies * (unsigned long long) L);
        }
        slot = atomicAdd(&o.ctr->hit_count, (unsigned long long) copies);
        if (p.want_hits) boff = atomicAdd(&o.ctr->blob_used, (unsigned long long) copies * (unsigned long long) L);
    }
    slot = __shfl_sync(gmask, slot, first);
    boff = __shfl_sync(gmask, boff, first);
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
                for (int j = lane; j < L; j += G) o.blob[boff + (unsigned long long) c * L + j] = rd[ro + j];
        }
    }
    if (t.nplus > 0 && p.accumulate) {
        const bool plus = seq < t.nplus;
        const int tpl = plus ? seq : seq - t.nplus;
        const unsigned long long cb = __ldg(t.cov_off + tpl);
        for (int j = lane; j < L; j += G) {
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

// Exhaustive matcher: one warp per (read, seq) pair, lanes over all diagonals, raw bytes.
__device__ __noinline__ void exhaustive_pair(const KvqParams &p, const KvqTemplates &t, const KvqOut &o,
                                const uint8_t *rd, int rl, int64_t file_pos, int sq, int lane, int G = 32)
{
    const int sl = __ldg(t.len + sq);
    const uint8_t *tp = t.bytes + __ldg(t.byte_off + sq);
    for (int d = -(rl - 1) + lane; d <= sl - 1; d += G) {
        int ro, to, L, rule;
        const int copies = diag_rule(d, rl, sl, p.minoverlap, ro, to, L, rule);
        if (!copies) continue;
        if (byte_mismatches(rd + ro, tp + to, L, p.maxerrors) > p.maxerrors) continue;
        emit_hit(t, p, o, sq, d, ro, to, L, rl, file_pos, rd, tp, copies, rule);
    }
}

// Is (kind, pos) the first seed, in the fixed order T1 pieces then T2 offsets, that fires on
// diagonal d of (read, seq)?  Every firing seed reaches the same diagonal; only this one emits.
// rpk = code plane, rb = base offset of the read in it.
template <bool BIG>
__device__ __noinline__ bool canonical_seed(const KvqParams &p, const KvqTemplates &t, const uint32_t *rpk, int rb,
                               const uint32_t *tpk, int rl, int sl, int d, int kind, int pos)
{
    for (int j = 0; j < p.npieces; j++) {
        const int toff = (int) __umulhi((unsigned) (j * sl), p.pieces_inv);       // floor(j * sl / npieces) without the division (KvqParams)
        const int pj = toff - d;
        if (pj < 0 || pj > rl - KVQ_SEED_K) continue;
        const uint32_t x16 = extract16(rpk, rb + pj) ^ extract16_g(tpk, toff);
        if (x16 & 0xFFFFu) continue;
        // second level (drain_seeds): where 12 bases fit both the piece and the read they must match too
        if (BIG && sl >= p.t1x && (pj + KVQ_T1X_K > rl || (x16 & 0xFFFFFFu))) continue;
        // small sets (probe_roles): a piece of 16 bases or more fires with its first 16 bases in the read
        if (!BIG && sl >= KVQ_T1_LONG * p.npieces && (pj + KVQ_T1_LONG > rl || x16)) continue;
        return kind == 0 && pos == pj;
    }
    const int16_t *p2 = t.p2_tab + rl * KVQ_P2W;
    if (BIG) {
        // slot = class * npieces + j, classes W, H, T (8-mers), HX, TX (12-mers); `pos` is the slot.
        // A slot fires where the keys agree and (template, offset) is in the slot's table
        // (engine.cpp, TemplateSet::build).
        const int P = p.npieces, O = p.minoverlap, OX = KVQ_T1X_K * P;
        for (int i = 0; i < p.n2max; i++) {
            const int pv = __ldg(p2 + i);
            if (pv < 0) continue;
            const int pp = pv & KVQ_P2_POS;
            const int toff = pp + d;
            const int cls = i < P ? 0 : i < 2 * P ? 1 : i < 3 * P ? 2 : 3;
            const int kc = cls == 3 ? KVQ_T1X_K : KVQ_SEED_K;
            if (toff < 0 || toff > sl - kc) continue;
            const uint32_t x16 = extract16(rpk, rb + pp) ^ extract16_g(tpk, toff);
            if (x16 & (cls == 3 ? 0xFFFFFFu : 0xFFFFu)) continue;
            if (cls == 0 && (pv & KVQ_P2_LONG)) {
                const uint32_t key = extract16_g(tpk, toff) & 0xFFFFu;
                if (!((__ldg(t.t2l_bitmap + (key >> 5)) >> (key & 31)) & 1u)) continue;
            }
            if (cls == 1 && !(sl > O && toff >= sl - OX + 1 + pp && toff <= sl - O + pp)) continue;
            if (cls == 2) {
                const int oj = p.o_piece[i - 2 * P];
                if (!(sl > O && toff >= oj && toff <= OX - 1 - O + oj && toff <= sl - 1 - O + oj)) continue;
            }
            return kind == 1 && pos == i;
        }
        return false;
    }
    for (int i = 0; i < p.n2max; i++) {
        const int pv = __ldg(p2 + i);
        if (pv < 0) continue;
        const int pp = pv & KVQ_P2_POS;
        const int toff = pp + d;
        if (toff < 0 || toff > sl - KVQ_SEED_K) continue;
        const uint32_t key = extract16(rpk, rb + pp) & 0xFFFFu;
        if (key != (extract16_g(tpk, toff) & 0xFFFFu)) continue;
        if ((pv & KVQ_P2_LONG) && !((__ldg(t.t2l_bitmap + (key >> 5)) >> (key & 31)) & 1u)) continue;
        return kind == 1 && pos == pp;
    }
    return false;
}

} // namespace

namespace {

struct Ctx {
    const KvqBatch &b; const KvqTemplates &t; const KvqParams &p; const KvqOut &o; const TileCtx &tc;
    const uint32_t *tpk;                    // packed template codes: shared copy when they fit, else global
};

__device__ __forceinline__ uint32_t extract16_t(const uint32_t *pk, int b)      // generic address space
{
    int w = b >> 4;
    return funnel_r(pk[w], pk[w + 1], (uint32_t) ((b & 15) << 1));
}

// ---- P4: seeds -> bucket entries -> code-space verify -> confirm + emit ------------------------
// Seeds s.seedq[base, base + n), n <= 32.  Seed = read | pos << 5 | kind << 15.
// Stage A, one lane per seed: the key's bucket head (a table addressed by the key itself holds the
// first entry and the size; further entries follow in the overflow array).  A T1 seed whose 12
// bases fit the read must also be in the 12-mer bitmap of piece starts (L2), unless its bucket
// holds a piece shorter than 12 bases: at thousands of templates nearly every 8-mer is some
// piece's start, the 12-mer is not.  Stage B, one lane per CANDIDATE: the (seed, entry) pairs of
// the surviving buckets are numbered through a warp prefix sum and dealt to the lanes 32 at a
// time, so a long bucket is verified by many lanes instead of keeping 31 lanes waiting for one.
template <bool BIG>
__device__ __forceinline__ int drain_seeds(WarpSmem &s, const Ctx &c, int base, int n, int lane)
{
    static_assert(NS == 1, "one seed per lane");
    const KvqTemplates &t = c.t; const KvqParams &p = c.p;
    uint4 head = make_uint4(0, 0, 0, 0);
    uint32_t sa = 0, sb = 0;                // rd0 | rl << 16;  pos | kind << 10
    int cnt = 0;
    if (lane < n) {
        const uint32_t sd = s.seedq[base + lane];
        const uint32_t rdv = s.rd[sd & 31];
        int pos = (int) ((sd >> 5) & 1023);
        const int kind = (int) (sd >> 15);
        const int rd0 = (int) (rdv & 0xFFFFu), rl = (int) (rdv >> 16);
        const int slot = pos;
        if (kind) pos = __ldg(t.p2_tab + rl * KVQ_P2W + pos) & KVQ_P2_POS;      // T2 seeds carry their probe slot
        const uint32_t r16 = extract16(s.pk, rd0 + pos);
        uint32_t hidx = r16 & 0xFFFFu;
        if (BIG && kind && slot >= p.npieces)   // table of the slot's class, all back to back in t2_first
            hidx = slot < 3 * p.npieces ? hidx + ((uint32_t) (slot - p.npieces + 1) << 16)
                                        : ((uint32_t) (2 * p.npieces + 1) << 16) + KVQ_X_HASH(r16 & 0xFFFFFFu);
        head = __ldg((kind ? t.t2_first : t.t1_first) + hidx);
        bool pass = true;
        if (BIG && !kind) {                     // pieces of >= 12 bases fire only with their 12 bases inside the read
            const uint32_t k24 = r16 & 0xFFFFFFu;
            pass = pos + KVQ_T1X_K <= rl && ((__ldg(t.t1x_bitmap + (k24 >> 5)) >> (k24 & 31)) & 1u);
        }
        uint32_t more = BIG ? (head.x >> 24) & KVQ_HEAD_MORE : head.x >> 24;   // (the flag is only set for BIG)
        if (more == KVQ_HEAD_MORE) more = __ldg(&(kind ? t.t2_ent : t.t1_ent)[head.w - 1].x);
        cnt = 1 + (int) more;
        if (BIG && ((!pass && !(head.x & KVQ_HEAD_SHORT)) || head.y == 0)) cnt = 0;
        sa = rdv; sb = (uint32_t) pos | ((uint32_t) kind << 10) | ((uint32_t) slot << 11);
    }
    // BIG: all entries are dealt out.  Otherwise (a few hundred templates: most buckets hold one
    // entry) round -1 verifies every bucket's head on its own lane and only the overflow entries
    // are dealt out.
    const int mine = BIG ? cnt : (cnt > 0 ? cnt - 1 : 0);
    int inc = mine;                         // inclusive prefix sum
    if (BIG || __any_sync(0xffffffffu, mine != 0)) {
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += v;
        }
    }
    const int total = __shfl_sync(0xffffffffu, inc, 31);
    const int exc = inc - mine;
    int ncand = 0;
    for (int x0 = BIG ? 0 : -32; x0 < total; x0 += 32) {
        const int x = x0 + lane;
        uint32_t a = sa, b = sb;
        uint4 en = head;
        bool active = lane < n;
        if (BIG || x0 >= 0) {
            int src = 0;                    // the bucket entry x belongs to: first lane with inc > x
#pragma unroll
            for (int step = 16; step; step >>= 1) {
                const int v = __shfl_sync(0xffffffffu, inc, src + step - 1);
                if (v <= x) src += step;
            }
            src &= 31;
            int k = x - __shfl_sync(0xffffffffu, exc, src);
            a = __shfl_sync(0xffffffffu, sa, src); b = __shfl_sync(0xffffffffu, sb, src);
            if (BIG) {
                en.x = __shfl_sync(0xffffffffu, head.x, src); en.y = __shfl_sync(0xffffffffu, head.y, src);
                en.z = __shfl_sync(0xffffffffu, head.z, src);
                k--;                        // entry 0 is the head itself
            }
            const uint32_t ovf = __shfl_sync(0xffffffffu, head.w, src);
            active = x < total;
            if (active && (!BIG || k >= 0)) en = __ldg(((b >> 10) & 1u ? t.t2_ent : t.t1_ent) + ovf + k);
        }
        const int pos = (int) (b & 1023u), kind = (int) ((b >> 10) & 1u);
        const int rd0 = (int) (a & 0xFFFFu), rl = (int) (a >> 16);
        int d = 0; bool surv = false;
        if (active) {
            ncand++;
            const int sl = (int) (en.y >> 16);
            d = (int) (en.y & 0xFFFFu) - pos;
            // overlap of the diagonal; Hamming distance in code space first (a lower bound of the
            // byte distance, most candidates die on the first word), the acceptance rule after
            const int ro = d < 0 ? -d : 0, to = d > 0 ? d : 0;
            const int L = min(rl - ro, sl - to);
            const uint32_t *tpk = c.tpk + en.z;
            int e = 0;
            for (int j = 0; j < L && e <= p.maxerrors; j += 16) {
                uint32_t xw = extract16(s.pk, rd0 + ro + j) ^ extract16_t(tpk, to + j);
                uint32_t m = (xw | (xw >> 1)) & 0x55555555u;
                if (L - j < 16) m &= (1u << (2 * (L - j))) - 1u;
                e += __popc(m);
            }
            if (e <= p.maxerrors) {
                int ro2, to2, L2, rule;
                surv = diag_rule(d, rl, sl, p.minoverlap, ro2, to2, L2, rule) != 0;
                // template inside read (rule c): some T1 piece fires for every accepted alignment and
                // T1 comes first in the canonical order, so a T2 seed never emits such a diagonal
                if (kind && d <= 0 && d + rl >= sl) surv = false;
            }
        }
        // survivors go to the emit kernel (byte-exact confirm, canonical-seed dedupe, emission)
        push_survivors(c.o, surv, c.tc.a0 + rd0, rl, (int) (en.x & 0xFFFFFFu), d,
                       (kind << 16) | (BIG && kind ? (int) (b >> 11) : pos), lane);
    }
    return ncand;
}

// Small sets (a few hundred templates, most buckets hold one entry): one CANDIDATE per seed and
// pass.  A seed word carries the index k of the bucket entry it stands for (bits 31:16, 0 = the
// head); a seed whose bucket has further entries goes back on the stack as (seed, k + 1) -- into
// the slots this pass has just emptied -- and is verified together with whatever comes next, so
// the (few) overflow entries never cost a pass of their own with most lanes idle.
__device__ __forceinline__ int drain_lean(WarpSmem &s, const Ctx &c, int base, int n, int lane, ThreadStats &ts)
{
    const KvqTemplates &t = c.t; const KvqParams &p = c.p;
    const bool act = lane < n;
    uint32_t sd = 0;
    int d = 0, rd0 = 0, rl = 0, pos = 0, kind = 0, seq = 0;
    bool surv = false, again = false;
    if (act) {
        sd = s.seedq[base + lane];
        const uint32_t rdv = s.rd[sd & 31];
        pos = (int) ((sd >> 5) & 1023);
        kind = (int) ((sd >> 15) & 1u);
        const uint32_t k = sd >> 16;
        rd0 = (int) (rdv & 0xFFFFu); rl = (int) (rdv >> 16);
        if (kind) pos = __ldg(t.p2_tab + rl * KVQ_P2W + pos) & KVQ_P2_POS;      // T2 seeds carry their probe slot
        const uint32_t hidx = extract16(s.pk, rd0 + pos) & 0xFFFFu;
        const uint4 *ent = kind ? t.t2_ent : t.t1_ent;
        uint4 en = __ldg((kind ? t.t2_first : t.t1_first) + hidx);
        uint32_t more = en.x >> 24;
        if (more == KVQ_HEAD_MORE) more = __ldg(&ent[en.w - 1].x);
        again = k < more;
        if (k) en = __ldg(ent + en.w + k - 1);
        ts.cands++;
        const int sl = (int) (en.y >> 16);
        d = (int) (en.y & 0xFFFFu) - pos;
        seq = (int) (en.x & 0xFFFFFFu);
        // overlap of the diagonal; Hamming distance in code space first (a lower bound of the
        // byte distance, most candidates die on the first word), the acceptance rule after
        const int ro = d < 0 ? -d : 0, to = d > 0 ? d : 0;
        const int L = min(rl - ro, sl - to);
        const uint32_t *tpk = c.tpk + en.z;
        int e = 0;
        for (int j = 0; j < L && e <= p.maxerrors; j += 16) {
            uint32_t xw = extract16(s.pk, rd0 + ro + j) ^ extract16_t(tpk, to + j);
            uint32_t m = (xw | (xw >> 1)) & 0x55555555u;
            if (L - j < 16) m &= (1u << (2 * (L - j))) - 1u;
            e += __popc(m);
        }
        if (e <= p.maxerrors) {
            int ro2, to2, L2, rule;
            surv = diag_rule(d, rl, sl, p.minoverlap, ro2, to2, L2, rule) != 0;
            // template inside read (rule c): some T1 piece fires for every accepted alignment and
            // T1 comes first in the canonical order, so a T2 seed never emits such a diagonal
            if (kind && d <= 0 && d + rl >= sl) surv = false;
        }
    }
    // survivors go to the emit kernel (byte-exact confirm, canonical-seed dedupe, emission)
    push_survivors(c.o, surv, c.tc.a0 + rd0, rl, seq, d, (kind << 16) | pos, lane);
    const uint32_t m = __ballot_sync(0xffffffffu, again);
    __syncwarp();                                   // every lane has read its seed
    if (again) s.seedq[base + __popc(m & ((1u << lane) - 1u))] = sd + 0x10000u;
    __syncwarp();
    return __popc(m);
}

// Push the set bits of every lane's `hits` (bit j = read offset posbase + j) on the seed stack:
// a lane reserves room with one shared-memory atomic.  Whenever 32 seeds are waiting the top 32
// are drained with all lanes busy; what does not fit the stack waits for the drain.
template <bool BIG>
__device__ __forceinline__ void push_seeds(WarpSmem &s, const Ctx &c, uint32_t hits, uint32_t tag,
                                           int posbase, bool flush, int lane, ThreadStats &ts)
{
    for (;;) {
        if (hits) {
            uint32_t at = atomicAdd(&s.qn, (uint32_t) __popc(hits));
            while (hits && at < SEEDQ) {
                const int j = __ffs(hits) - 1;
                hits &= hits - 1;
                s.seedq[at++] = tag | ((uint32_t) (posbase + j) << 5);
            }
        }
        __syncwarp();
        const uint32_t q = s.qn;
        const uint32_t left = __ballot_sync(0xffffffffu, hits != 0);       // only when the stack is full
        if (q < DRAIN && !left && !(flush && q)) break;
        int stored = q < SEEDQ ? (int) q : SEEDQ;
        const bool all = flush && !left;
        while (stored >= DRAIN || (all && stored > 0)) {
            const int n = stored < DRAIN ? stored : DRAIN, base = stored - n;
            if (BIG) { ts.cands += (unsigned long long) drain_seeds<true>(s, c, base, n, lane); stored = base; }
            else stored = base + drain_lean(s, c, base, n, lane, ts);
        }
        __syncwarp();
        if (lane == 0) s.qn = (uint32_t) stored;
        __syncwarp();
        if (!left) break;
    }
}

// ---- P3: T1 probes of one 32-byte word: bit j <=> the 8-mer at base offset j is a template piece
// BIG: the byte of a key is a mask over the 3 code bits that follow the 8-mer in the piece (all
// ones for pieces shorter than 12 bases), so most chance 8-mer matches never become seeds.
template <bool BIG>
__device__ __forceinline__ uint32_t probe_word(const uint8_t *t1, uint32_t x0, uint32_t x1, uint32_t x2)
{
    uint32_t hits = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const uint32_t sa = j ? funnel_r(x0, x1, 2 * j) : x0;       // bases j .. j+15
        const uint32_t sb = j ? funnel_r(x1, x2, 2 * j) : x1;       // bases 16+j .. 31+j
        if (BIG) {
            const uint32_t sc = x2 >> (2 * j);                      // bases 32+j .. (low bits only)
            hits += (((uint32_t) t1[sa & 0xFFFFu] >> ((sa >> 16) & 7u)) & 1u) << j;
            hits += (((uint32_t) t1[sa >> 16] >> (sb & 7u)) & 1u) << (j + 8);
            hits += (((uint32_t) t1[sb & 0xFFFFu] >> ((sb >> 16) & 7u)) & 1u) << (j + 16);
            hits += (((uint32_t) t1[sb >> 16] >> (sc & 7u)) & 1u) << (j + 24);
            continue;
        }
        hits += (uint32_t) t1[sa & 0xFFFFu] << j;
        hits += (uint32_t) t1[sa >> 16] << (j + 8);
        hits += (uint32_t) t1[sb & 0xFFFFu] << (j + 16);
        hits += (uint32_t) t1[sb >> 16] << (j + 24);
    }
    return hits;
}

// Small sets: the byte of a key holds the ROLES of the 8-mer -- bit 0: first 8-mer of a piece
// shorter than 16 bases (fires alone), bit 1: first 8-mer of a longer piece, bit 2: second
// 8-mer (bases 8..15) of a longer piece.  A longer piece fires only where both its 8-mers are
// found 8 offsets apart (its first 16 bases lie in the read and match: see canonical_seed), which
// removes nearly every chance 8-mer match before it becomes a seed.  Roles of the 32 offsets of
// one 32-byte word, one 4-bit field per offset, 8 offsets per register.
struct Roles { uint32_t g0, g1, g2, g3; };
__device__ __forceinline__ Roles probe_roles(const uint8_t *t1, uint32_t x0, uint32_t x1, uint32_t x2)
{
    Roles r{0u, 0u, 0u, 0u};
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const uint32_t sa = j ? funnel_r(x0, x1, 2 * j) : x0;       // bases j .. j+15
        const uint32_t sb = j ? funnel_r(x1, x2, 2 * j) : x1;       // bases 16+j .. 31+j
        r.g0 += (uint32_t) t1[sa & 0xFFFFu] << (4 * j);
        r.g1 += (uint32_t) t1[sa >> 16] << (4 * j);
        r.g2 += (uint32_t) t1[sb & 0xFFFFu] << (4 * j);
        r.g3 += (uint32_t) t1[sb >> 16] << (4 * j);
    }
    return r;
}
// 8 offsets (fields of a) -> 8 bits: fires alone, or longer-piece key with the second 8-mer 8 offsets on (same field of nxt)
__device__ __forceinline__ uint32_t fire8(uint32_t a, uint32_t nxt)
{
    const uint32_t f = (a | ((a >> 1) & (nxt >> 2))) & 0x11111111u;
    return __dp4a(((f | (f >> 3)) & 0x03030303u), 0x40100401u, 0u);
}

} // namespace

namespace {

// ---- P2: one lane per record of the round.  Newline ranks are record-relative: newline
// v = rank - off is line (v & 3) of record (v >> 2); the byte after newline 4k-1 starts record k.
// Returns start | length << 16 of the lane's trimmed read if it is to be matched here, else 0.
__device__ __forceinline__ uint32_t parse_record(WarpSmem &s, Smem &sh, const Ctx &c, int k, int idx, bool act,
                                                 int n_all, int off, bool at_seg_end, int lane, ThreadStats &ts)
{
    const TileCtx &tc = c.tc; const KvqParams &p = c.p;
    int S = tc.shift, rl = -1;
    uint32_t rdv = 0;
    if (act) {
        if (!(tc.first && k == 0)) S = (int) s.nlpos[idx] + 1;
        if (4 * k + 3 + off >= n_all) {
            // the record runs past the staged region: a partial record at the end of the segment
            // is dropped (workhorse.c:1033), anything else is a long record
            if (!at_seg_end) push_long(c.o, tc.a0 + S, tc.seg_end, 0, 0);
            act = false;
        }
    }
    if (act) {
        const int n0 = s.nlpos[idx + 1], n1 = s.nlpos[idx + 2], n2 = s.nlpos[idx + 3], n3 = s.nlpos[idx + 4];
        const uint8_t *rec = c.b.data + tc.a0;
        const uint8_t c0 = __ldg(rec + S), c2 = __ldg(rec + n1 + 1);
        if (c0 != '@')
            report_format_error(c.o, c.b.base_fpos + tc.a0 + S, 0, c0);
        else if (c2 != '+')
            report_format_error(c.o, c.b.base_fpos + tc.a0 + S, 1, (uint32_t) (n1 + 1 - S));
        ts.records++;
        ts.bases += (unsigned long long) (n1 - n0 - 1);
        int best;
        phred_trim(s, n2 + 1, n3, rl, best);
        if (rl >= p.minreadlength && rl > 0) {                          // workhorse.c:1100
            const int rd0 = n0 + 1 + (best - (n2 + 1));                 // workhorse.c:1070
            if (rl > MAXRL_FAST) push_long(c.o, tc.a0 + rd0, tc.seg_end, 1, rl);
            else rdv = (uint32_t) rd0 | ((uint32_t) rl << 16);
        }
    }
    // add_rl (workhorse.c:394-402); reads of one round mostly share their length
    const uint32_t m = __ballot_sync(0xffffffffu, rl >= 0);
    if (m) {
        const int leader = __ffs(m) - 1;
        const int rl0 = __shfl_sync(0xffffffffu, rl, leader);
        if (__ballot_sync(0xffffffffu, rl == rl0) == m) {
            if (lane == leader) {
                if (rl0 < KVQ_HIST) atomicAdd(&sh.hist[rl0], (uint32_t) __popc(m));
                atomicMax(&sh.longest, rl0);
            }
        } else if (rl >= 0) {
            if (rl < KVQ_HIST) atomicAdd(&sh.hist[rl], 1u);
            atomicMax(&sh.longest, rl);
        }
    }
    return rdv;
}

// One round = up to 32 owned records of the tile, one per lane.
template <bool BIG>
__device__ void run_round(WarpSmem &s, Smem &sh, const Ctx &c, int k0, int nown, int base, int n_all, int off,
                          bool at_seg_end, int lane, ThreadStats &ts)
{
    const KvqTemplates &t = c.t; const KvqParams &p = c.p; const TileCtx &tc = c.tc;
    const int k = k0 + lane;
    const uint32_t rdv = parse_record(s, sh, c, k, 4 * k + off - 1 - base, k < nown, n_all, off, at_seg_end, lane, ts);
    s.rd[lane] = rdv;
#ifdef KVQ_ONLY_P2
    return;
#endif
    __syncwarp();
    if (!__any_sync(0xffffffffu, rdv != 0)) return;
    const int rd0 = (int) (rdv & 0xFFFFu), rl = (int) (rdv >> 16);
    const bool seeded = p.seed_ok && t.n_exh < t.nseqs;
    const bool seedable = seeded && rdv != 0 && rl >= p.min_seed_rl;
    if (seeded) {
        // T2: a few fixed offsets per read (whole-read / head / tail pieces), table per read length
        uint32_t hit2 = 0;
        if (seedable) {
            const int16_t *p2 = t.p2_tab + rl * KVQ_P2W;
            for (int i = 0; i < p.n2max; i++) {
                const int pv = __ldg(p2 + i);
                if (pv < 0) continue;
                const uint32_t r16 = extract16(s.pk, rd0 + (pv & KVQ_P2_POS));
                const uint32_t key = r16 & 0xFFFFu;
#if KVQ_T2L
                const uint32_t *bm = (pv & KVQ_P2_LONG) ? sh.bm2l : sh.bm2;
#else
                const uint32_t *bm = sh.bm2;
#endif
                if (BIG && i >= p.npieces) {    // slot classes: W | H, T (8-mers, own tables) | HX, TX (12-mers)
                    if (i >= 3 * p.npieces) {
                        const uint32_t k24 = r16 & 0xFFFFFFu;
                        hit2 |= ((__ldg(t.t2x_bitmap + (k24 >> 5)) >> (k24 & 31)) & 1u) << i;
                        continue;
                    }
                    bm = sh.tpk;
                }
                hit2 |= ((bm[key >> 5] >> (key & 31)) & 1u) << i;
            }
        }
        // T1: every read offset rd0 .. rd0 + rl - 8, one 32-byte word of the code plane per step;
        // the last step pushes the T2 seeds and flushes the stack (one push / drain site).
        // Small sets decide the last 8 offsets of a word one step later (they need the roles of
        // the next word's first 8 offsets): the window of a step starts 8 offsets before its word.
        const int w0 = rd0 >> 5, plast = rd0 + rl - KVQ_SEED_K;
        const int nw = seedable ? (plast >> 5) - w0 + 1 : 0;
        const int maxnw = __reduce_max_sync(0xffffffffu, nw);
        const int last = BIG ? maxnw : maxnw + 1;
        uint32_t carry = 0;
        for (int q = 0; q <= last; q++) {
            uint32_t hits = 0, tag = (uint32_t) lane;
            const int w = w0 + q;
            const int wb = BIG ? (w << 5) : (w << 5) - 8;           // coordinate of bit 0 of `hits`
            int posbase = wb - rd0;
            if (BIG ? q < nw : (nw > 0 && q <= nw)) {
                if (BIG) {
                    const uint2 x = reinterpret_cast<const uint2 *>(s.pk)[w];
                    hits = probe_word<BIG>(sh.t1, x.x, x.y, s.pk[2 * w + 2]);
                } else {
                    Roles r{0u, 0u, 0u, 0u};
                    if (q < nw) {
                        const uint2 x = reinterpret_cast<const uint2 *>(s.pk)[w];
                        r = probe_roles(sh.t1, x.x, x.y, s.pk[2 * w + 2]);
                    }
                    hits = fire8(carry, r.g0) | (fire8(r.g0, r.g1) << 8) | (fire8(r.g1, r.g2) << 16) | (fire8(r.g2, r.g3) << 24);
                    carry = r.g3;
                }
                const int lo = rd0 - wb, hi = plast - wb;
#if KVQ_FMASK
                // bits lo .. hi of the window (clamping funnel shifts: a shift of 32 or more gives 0)
                hits &= __funnelshift_lc(0u, 0xFFFFFFFFu, (uint32_t) max(lo, 0))
                      & __funnelshift_rc(0xFFFFFFFFu, 0u, (uint32_t) max(31 - hi, 0));
#else
                if (lo > 0) hits &= lo >= 32 ? 0u : 0xFFFFFFFFu << lo;
                if (hi < 31) hits &= hi < 0 ? 0u : (2u << hi) - 1u;
#endif
            }
            if (q == last) { hits = hit2; tag |= 1u << 15; posbase = 0; }
            push_seeds<BIG>(s, c, hits, tag, posbase, q == last, lane, ts);
        }
    }
    // exhaustive matcher (emit kernel): reads the seed scheme does not cover against every
    // template, the others against the templates it does not cover
    ts.exh += (unsigned long long) (rdv != 0 && !seedable);
    push_survivors(c.o, rdv != 0 && !seedable, tc.a0 + rd0, rl, -1, 0, 0, lane);
    if (t.n_exh > 0) push_survivors(c.o, seedable, tc.a0 + rd0, rl, -2, 0, 0, lane);
}

} // namespace

namespace {

template <bool NEG, bool BIG>
__global__ void __launch_bounds__(THREADS, 1)
kvq_scan_kernel(const __grid_constant__ KvqBatch b, const __grid_constant__ KvqTemplates t,
                const __grid_constant__ KvqParams p, const __grid_constant__ KvqOut o)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Smem &sh = *reinterpret_cast<Smem *>(smem_raw);
    const int lane = threadIdx.x & 31;
    WarpSmem &s = sh.w[threadIdx.x >> 5];

    {   // tables: T1 presence bytes, T2 bitmap
        const uint4 *src = reinterpret_cast<const uint4 *>(t.t1_present);
        uint4 *dst = reinterpret_cast<uint4 *>(sh.t1);
        for (int i = threadIdx.x; i < KVQ_NKEYS / 16; i += THREADS) dst[i] = __ldg(src + i);
        for (int i = threadIdx.x; i < KVQ_BITMAP_WORDS; i += THREADS) {
            sh.bm2[i] = __ldg(t.t2_bitmap + i);
#if KVQ_T2L
            sh.bm2l[i] = __ldg(t.t2l_bitmap + i);
#endif
        }
        if (lane == 0) s.qn = 0;
        for (int i = threadIdx.x; i < KVQ_HIST; i += THREADS) sh.hist[i] = 0;
        static_assert(KVQ_TPK_SMEM_WORDS == KVQ_BITMAP_WORDS, "BIG keeps the head/tail-slot bitmap where the template codes would be");
        if (BIG)
            for (int i = threadIdx.x; i < KVQ_BITMAP_WORDS; i += THREADS) sh.tpk[i] = __ldg(t.t2e_bitmap + i);
        else if (t.packed_words <= KVQ_TPK_SMEM_WORDS)
            for (int i = threadIdx.x; i < t.packed_words; i += THREADS) sh.tpk[i] = __ldg(t.packed + i);
        if (threadIdx.x == 0) sh.longest = -1;
    }
    __syncthreads();
    ThreadStats ts{0, 0, 0, 0};
    const uint32_t cl = (uint32_t) (p.amin & 0x7F) * 0x01010101u;
    const uint32_t cnl = (uint32_t) p.nl_byte * 0x01010101u;        // '\n' x 4, from a parameter so that it lives in a register
    const uint32_t *tpk = (!BIG && t.packed_words <= KVQ_TPK_SMEM_WORDS) ? sh.tpk : t.packed;

    // Tile geometry: claim the next tile and start its bytes on their way to L2.
    struct Claim { int tile; int first; int64_t t0, seg_end; };
    auto claim = [&]() -> Claim {
        int tile = 0;
        if (lane == 0) tile = (int) atomicAdd(&o.ctr->tile_next, 1u);
        tile = __shfl_sync(0xffffffffu, tile, 0);
        if (tile >= b.ntiles) return Claim{-1, 0, 0, 0};
        const longlong2 ti = __ldg(b.tile_info + tile);            // {start, segment end | first << 62}
        const int64_t t0 = ti.x, seg_end = ti.y & ~(1ll << 62);
        const int first = (int) ((ti.y >> 62) & 1);
#if KVQ_PREFETCH
        {
            const int64_t a0 = t0 & ~(int64_t) 31;
            const int64_t t1 = t0 + KVQ_TILE + KVQ_ROW < seg_end ? t0 + KVQ_TILE + KVQ_ROW : seg_end;
            for (int64_t x = a0 + lane * 128; x < t1; x += 32 * 128)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(b.data + x));
        }
#endif
        return Claim{tile, first, t0, seg_end};
    };

    Claim cur{-2, 0, 0, 0};                         // -2: nothing claimed yet, -1: queue exhausted
    for (;;) {
        const Claim nxt = claim();
        if (cur.tile >= 0) {
        const int tile = cur.tile;
        const int64_t t0 = cur.t0, seg_end = cur.seg_end;
        TileCtx tc;
        tc.a0 = t0 & ~(int64_t) 31;
        tc.shift = (int) (t0 - tc.a0);
        const int64_t t1 = t0 + KVQ_TILE < seg_end ? t0 + KVQ_TILE : seg_end;
        tc.tile_end = (int) (t1 - tc.a0);
        const int64_t lim = seg_end - tc.a0;
        tc.lim = lim < (int64_t) MAX_ROWS * KVQ_ROW ? (int) lim : MAX_ROWS * KVQ_ROW;
        tc.first = cur.first;
        tc.seg_end = seg_end;
        const Ctx c{b, t, p, o, tc, tpk};

        // P1: rows 0 .. OWN_ROWS (the tile plus one row of spill) through three register sets
        // taking turns, one row being reduced while two are in flight.  After row OWN_ROWS
        // the newlines the tile owns are counted: the count is published for the look-back
        // chain and the line phase of the tile settled (see settle_phase); further spill rows
        // follow while the last owned record is incomplete.
        static_assert(OWN_ROWS % 2 == 0 && EXTRA_ROWS == 2, "row pairs");
        uint32_t n_all = 0, n_own = 0;                  // n_own: per-lane partial counts
        int nrows = 0, off = 0, nown = 0, phase = 0, n_tile = 0;
        bool inferred = false;
        // settles the phase after row OWN_ROWS; true if spill row `next` is needed
        auto settle = [&](int next) -> bool {
            // line phase of the tile = newlines of its segment before it, mod 4.  A tile that
            // infers its phase publishes phase + count as its inclusive prefix right away
            // (status bit 3) and compares its phase with the predecessor's prefix when it is
            // done: every neighbouring pair consistent <=> every inferred phase right (the
            // first tile of a segment is exact).  A tile that cannot infer publishes its
            // count (bit 2) and walks back to the nearest prefix.
            n_tile = (int) __reduce_add_sync(0xffffffffu, n_own);
            if (!tc.first && !p.exact_phase) phase = infer_phase(s, b, tc, (int) n_all, lane);
            inferred = !tc.first && !p.exact_phase && phase >= 0;
            if (lane == 0)
                ((volatile uint32_t *) b.tile_status)[tile] =
                    (tc.first || inferred) ? 8u | (uint32_t) ((phase + n_tile) & 3) : 4u | (uint32_t) (n_tile & 3);
            if (!tc.first && !inferred) phase = resolve_phase(b, tile, 0, n_tile, lane);
            // newline rank -> record-relative numbering (see parse_record)
            off = tc.first ? 0 : ((3 - phase) & 3) + 1;
            const int vmax_tile = n_tile - 1 - off;
            if (tc.first || vmax_tile >= -1) nown = (vmax_tile + 1) / 4 + 1;
            const int need = nown > 0 ? 4 * (nown - 1) + 3 + off : -1;  // rank of the last newline the tile needs
            return !(need < (int) n_all || next * KVQ_ROW >= tc.lim);
        };
#if KVQ_OLD_P1
        {
            Row ra = load_row(b, tc, 0, lane), rb = load_row(b, tc, 1, lane), rc = load_row(b, tc, 2, lane);
#pragma unroll 1
            for (int row = 0; row < MAX_ROWS; row += 3) {
                reduce_row_old<NEG>(s, tc, ra, row, lane, cl, cnl, n_all, n_own);
                if (row + 3 <= OWN_ROWS) ra = load_row(b, tc, row + 3, lane);
                reduce_row_old<NEG>(s, tc, rb, row + 1, lane, cl, cnl, n_all, n_own);
                nrows = row + 2;
                if (row + 4 <= OWN_ROWS) rb = load_row(b, tc, row + 4, lane);
                if (row + 1 == OWN_ROWS) {
                    if (!settle(row + 2)) break;
                    rc = load_row(b, tc, row + 2, lane);
                }
                reduce_row_old<NEG>(s, tc, rc, row + 2, lane, cl, cnl, n_all, n_own);
                nrows = row + 3;
                if (row + 5 <= OWN_ROWS) rc = load_row(b, tc, row + 5, lane);
            }
        }
#else
        {
            static_assert((OWN_ROWS + 2) % 3 == 0 && KVQ_ROW_SETS == 3, "row triples");
            Row ra = load_row(b, tc, 0, lane), rb = load_row(b, tc, 1, lane), rc = load_row(b, tc, 2, lane);
#pragma unroll 1
            for (int row = 0; row < MAX_ROWS; row += 3) {
                const uint32_t nl0 = reduce_row<NEG>(s, tc, ra, row, lane, cl, cnl);
                if (row + 3 <= OWN_ROWS) ra = load_row(b, tc, row + 3, lane);
                const uint32_t nl1 = reduce_row<NEG>(s, tc, rb, row + 1, lane, cl, cnl);
                nrows = row + 2;
                if (row + 4 <= OWN_ROWS) rb = load_row(b, tc, row + 4, lane);
                if (row + 1 == OWN_ROWS) {          // the last two mandatory rows: rank them, settle the phase
                    rank_rows(s, tc, nl0, nl1, 0u, row, true, lane, n_all, n_own);
                    if (!settle(row + 2)) break;
                    rc = load_row(b, tc, row + 2, lane);
                    const uint32_t nl2 = reduce_row<NEG>(s, tc, rc, row + 2, lane, cl, cnl);
                    nrows = row + 3;
                    rank_rows(s, tc, 0u, 0u, nl2, row, true, lane, n_all, n_own);
                    break;
                }
                const uint32_t nl2 = reduce_row<NEG>(s, tc, rc, row + 2, lane, cl, cnl);
                nrows = row + 3;
                if (row + 5 <= OWN_ROWS) rc = load_row(b, tc, row + 5, lane);
                rank_rows(s, tc, nl0, nl1, nl2, row, false, lane, n_all, n_own);
            }
        }
#endif
        const bool at_seg_end = tc.a0 + (int64_t) nrows * KVQ_ROW >= seg_end;
#ifdef KVQ_ONLY_P1
        if (n_all == 0x7fffffff)
#endif
        for (int k0 = 0; k0 < nown; k0 += 32) {
            const int base = k0 > 0 ? 4 * k0 + off - 1 : 0;
            __syncwarp();
            if (k0 > 0) scatter_newlines(s, b, tc, nrows, base, lane, cnl);
            __syncwarp();
            run_round<BIG>(s, sh, c, k0, nown, base, (int) n_all, off, at_seg_end, lane, ts);
        }
        // an inferred phase is checked against the predecessor's prefix (published long ago:
        // nobody waits here); a contradiction means malformed input -- the batch is run again
        // with exact phases
        if (inferred) {
            uint32_t v;
            while (!((v = ((volatile uint32_t *) b.tile_status)[tile - 1]) & 8u)) __nanosleep(40);
            if ((int) (v & 3u) != phase && lane == 0) atomicAdd(&o.ctr->phase_retry, 1u);
        }
        __syncwarp();                               // planes are reused by the next tile
        }
        cur = nxt;
        if (cur.tile < 0) break;
    }

    // flush the statistics: warp-reduce the per-thread counters, one atomic per warp
#pragma unroll
    for (int o2 = 16; o2 > 0; o2 >>= 1) {
        ts.records += __shfl_down_sync(0xffffffffu, ts.records, o2);
        ts.bases += __shfl_down_sync(0xffffffffu, ts.bases, o2);
        ts.exh += __shfl_down_sync(0xffffffffu, ts.exh, o2);
        ts.cands += __shfl_down_sync(0xffffffffu, ts.cands, o2);
    }
    if (lane == 0) {
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

} // namespace

namespace {

// Records that do not fit a tile's spill rows, or whose trimmed read is longer than the fast
// path handles: one warp per record straight from global memory, exhaustive matcher.  Rare by
// construction (such a record is longer than 1 KiB).
__global__ void __launch_bounds__(256)
kvq_long_kernel(const __grid_constant__ KvqBatch b, const __grid_constant__ KvqTemplates t,
                const __grid_constant__ KvqParams p, const __grid_constant__ KvqOut o)
{
    const int lane = threadIdx.x & 31;
    const unsigned long long n = o.ctr->long_count < o.long_cap ? o.ctr->long_count : o.long_cap;
    if (o.ctr->surv_count > o.surv_cap || o.ctr->phase_retry) return;   // the batch is run again
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
            exhaustive_pair(p, t, o, b.data + rd0, rl, b.base_fpos + rd0, sq, lane);
    }
}

// ---- emit kernel: one warp per survivor record of the scan kernel ------------------------------
// seq >= 0: the candidate passed the Hamming check on 2-bit codes; confirm it on the raw bytes,
// keep it if its seed is the canonical one of the diagonal, emit (hit record, hitseq, per-seq
// counters, fused coverage).  seq < 0: exhaustive matching of the read.  Nothing happens when the
// queue overflowed: the host re-runs the batch with a larger one.
#ifndef KVQ_EMIT_BLOCKS
#define KVQ_EMIT_BLOCKS 8                 // resident blocks per SM (latency-bound: dependent loads per survivor)
#endif
#ifndef KVQ_EMIT_G
#define KVQ_EMIT_G 32                     // lanes per survivor in the emit kernel (32 / G survivors per warp side by side);
                                          // measured, kernels of one config-2 step: G = 32: 0.7157 ms, 16: 0.7220, 8: 0.7255
#endif
template <bool BIG>
__global__ void __launch_bounds__(256, KVQ_EMIT_BLOCKS)
kvq_emit_kernel(const __grid_constant__ KvqBatch b, const __grid_constant__ KvqTemplates t,
                const __grid_constant__ KvqParams p, const __grid_constant__ KvqOut o)
{
    // G lanes per survivor, 32 / G survivors per warp side by side.  Every warp walks ~6 survivors in a row, each a
    // chain of dependent loads (record -> template -> read bytes -> codes -> tables -> atomics); several survivors per
    // warp shorten the row but run their (divergent) chains through one instruction stream -- slower, see KVQ_EMIT_G
    constexpr int G = KVQ_EMIT_G, NG = 32 / G;
    __shared__ uint32_t rpk_all[8 * NG][KVQ_MAXRL_FAST / 16 + 4];
    const int wlane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int grp = wlane / G, lane = wlane % G, first = grp * G;
    const uint32_t gmask = (G == 32 ? 0xFFFFFFFFu : ((1u << G) - 1u) << first);
    uint32_t *rpk = rpk_all[wib * NG + grp];
    const unsigned long long n = o.ctr->surv_count;
    if (n > o.surv_cap || o.ctr->phase_retry) return;
    const unsigned long long nw = (unsigned long long) gridDim.x * 8 * NG;
    for (unsigned long long i = ((unsigned long long) blockIdx.x * 8 + wib) * NG + grp; i < n; i += nw) {
        const uint4 q0 = __ldg(reinterpret_cast<const uint4 *>(o.surv + i));
        const uint4 q1 = __ldg(reinterpret_cast<const uint4 *>(o.surv + i) + 1);
        const long long rd0 = (long long) (((unsigned long long) q0.y << 32) | q0.x);
        const int rl = (int) q0.z, seq = (int) q0.w, d = (int) q1.x;
        const uint8_t *rdp = b.data + rd0;
        if (seq < 0) {
            if (seq == -1)
                for (int sq = 0; sq < t.nseqs; sq++) exhaustive_pair(p, t, o, rdp, rl, b.base_fpos + rd0, sq, lane, G);
            else
                for (int k = 0; k < t.n_exh; k++)
                    exhaustive_pair(p, t, o, rdp, rl, b.base_fpos + rd0, __ldg(t.exh_list + k), lane, G);
            continue;
        }
        const int sl = __ldg(t.len + seq);
        int ro, to, L, rule;
        const int copies = diag_rule(d, rl, sl, p.minoverlap, ro, to, L, rule);
        const uint8_t *tp = t.bytes + __ldg(t.byte_off + seq);
        int e = 0;
        for (int j = lane; j < L; j += G) e += (rdp[ro + j] != tp[to + j]);
#pragma unroll
        for (int x = G / 2; x > 0; x >>= 1) e += __shfl_xor_sync(gmask, e, x);
        if (e > p.maxerrors || !copies) continue;
        // 2-bit codes of the read for the canonical-seed rule: every lane packs 8 bases (16 bits), neighbours are
        // joined to words -- 8 G bases per pass; bytes behind the read are loaded (the batch is followed by zeroed
        // slack) and masked away
        __syncwarp(gmask);
        const int nwords = (rl + 15) / 16 + 2;
        for (int w0 = 0; w0 < nwords; w0 += G / 2) {
            const int at = 16 * w0 + 8 * lane;
            uint32_t v = 0;
#pragma unroll
            for (int k = 0; k < 8; k++) v |= KVQ_CODE(rdp[at + k]) << (2 * k);
            const int valid = min(max(rl - at, 0), 8);
            v &= (1u << (2 * valid)) - 1u;
            v |= __shfl_down_sync(gmask, v, 1, G) << 16;
            if (!(lane & 1) && w0 + (lane >> 1) < nwords) rpk[w0 + (lane >> 1)] = v;
        }
        __syncwarp(gmask);
        if (!canonical_seed<BIG>(p, t, rpk, 0, t.packed + __ldg(t.packed_off + seq), rl, sl, d,
                                 (int) (q1.y >> 16), (int) (q1.y & 0xFFFFu))) continue;
        emit_hit_warp(t, p, o, seq, d, ro, to, L, rl, sl, b.base_fpos + rd0, rdp, tp, copies, rule, lane, G, gmask, first);
    }
}

} // namespace

int kvq_scan_smem_bytes(void) { return (int) sizeof(Smem); }
int kvq_scan_groups(void) { return NWARPS; }

int kvq_launch_scan(const KvqBatch &b, const KvqTemplates &t, const KvqParams &p, const KvqOut &o,
                    int sm_count, void *stream)
{
    void (*k)(const KvqBatch, const KvqTemplates, const KvqParams, const KvqOut) =
        p.amin < 0 ? (p.t1x ? kvq_scan_kernel<true, true> : kvq_scan_kernel<true, false>)
                   : (p.t1x ? kvq_scan_kernel<false, true> : kvq_scan_kernel<false, false>);
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) sizeof(Smem));
    if (e != cudaSuccess) return (int) e;
    int grid = (b.ntiles + NWARPS - 1) / NWARPS;
    if (grid > sm_count) grid = sm_count;
    if (grid < 1) grid = 1;
    k<<<grid, THREADS, sizeof(Smem), (cudaStream_t) stream>>>(b, t, p, o);
    return (int) cudaGetLastError();
}

int kvq_launch_emit(const KvqBatch &b, const KvqTemplates &t, const KvqParams &p, const KvqOut &o,
                    int sm_count, void *stream)
{
    if (p.t1x) kvq_emit_kernel<true><<<KVQ_EMIT_BLOCKS * sm_count, 256, 0, (cudaStream_t) stream>>>(b, t, p, o);
    else kvq_emit_kernel<false><<<KVQ_EMIT_BLOCKS * sm_count, 256, 0, (cudaStream_t) stream>>>(b, t, p, o);
    return (int) cudaGetLastError();
}

int kvq_launch_long(const KvqBatch &b, const KvqTemplates &t, const KvqParams &p, const KvqOut &o,
                    int sm_count, void *stream)
{
    kvq_long_kernel<<<sm_count, 256, 0, (cudaStream_t) stream>>>(b, t, p, o);
    return (int) cudaGetLastError();
}

// ---- hits in the reference's order, on the device (see KvqOrder) -------------------------------
namespace {

// hits to order, or 0 when the host has to do it (buffers overflowed, batch will be run again, too many)
__device__ __forceinline__ uint32_t order_n(const KvqOut &o, const KvqOrder &ord)
{
    const KvqCounters &c = *o.ctr;
    if (c.surv_count > o.surv_cap || c.phase_retry || c.hit_count > o.hit_cap || c.blob_used > o.blob_cap ||
        c.hit_count > ord.cap || c.blob_used > ord.blob_cap)
        return 0;
    return (uint32_t) c.hit_count;
}

__device__ __forceinline__ bool key_before(const ulonglong2 &a, uint32_t ia, const ulonglong2 &b, uint32_t ib)
{
    return a.x != b.x ? a.x < b.x : a.y != b.y ? a.y < b.y : ia < ib;
}

__global__ void __launch_bounds__(256)
kvq_order_count(const __grid_constant__ KvqBatch b, const __grid_constant__ KvqOut o, const __grid_constant__ KvqOrder ord)
{
    const uint32_t n = order_n(o, ord);
    if (blockIdx.x == 0 && threadIdx.x == 0) ord.status[0] = n ? n : (o.ctr->hit_count == 0 ? 0u : 0xFFFFFFFFu);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const KvqHitRec h = o.hits[i];
        const int sh = h.rule == 0 ? h.seq_pos : h.rule == 1 ? -h.seq_pos : (h.seq_pos < 0 ? -h.seq_pos : h.seq_pos);
        ulonglong2 k;
        k.x = (unsigned long long) h.file_pos;
        k.y = ((unsigned long long) (uint32_t) h.seq_nr << 34) | ((unsigned long long) (uint32_t) h.rule << 32)
            | (unsigned long long) (uint32_t) (sh + (1 << 30));
        ord.keys[i] = k;
        const unsigned long long rel = (unsigned long long) (h.file_pos - b.base_fpos);
        unsigned long long bk = rel * n / (unsigned long long) (ord.span + 1);
        if (bk >= n) bk = n - 1;
        ord.bkt[i] = (uint32_t) bk;
        atomicAdd(&ord.start[bk + 1], 1u);
    }
}

// one block of 1024 threads: x[0 .. n] -> inclusive prefix sums in place (x[0] stays), copy to y
// when given.  Every thread takes a contiguous run of ceil((n + 1) / 1024) elements (sum, one block
// scan over the 1024 run totals, prefix written back): three barriers whatever n is.
__device__ __forceinline__ void block_scan(uint32_t *x, uint32_t *y, uint32_t n)
{
    __shared__ uint32_t wsum[32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const uint32_t per = (n + 1 + 1023) / 1024;
    const uint32_t lo = threadIdx.x * per, hi = min(lo + per, n + 1);
    uint32_t sum = 0;
    for (uint32_t i = lo; i < hi; i++) sum += x[i];
    uint32_t inc = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t u = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += u;
    }
    if (lane == 31) wsum[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        uint32_t w = wsum[lane], winc = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, winc, d);
            if (lane >= d) winc += u;
        }
        wsum[lane] = winc - w;                      // exclusive prefix of the warp totals
    }
    __syncthreads();
    uint32_t run = wsum[wid] + inc - sum;           // everything in front of this thread's run
    for (uint32_t i = lo; i < hi; i++) {
        run += x[i];
        x[i] = run;
        if (y) y[i] = run;
    }
    __syncthreads();
}

__global__ void __launch_bounds__(1024)
kvq_order_scan_starts(const __grid_constant__ KvqOut o, const __grid_constant__ KvqOrder ord)
{
    const uint32_t n = order_n(o, ord);
    if (n) block_scan(ord.start, ord.cursor, n);
}

__global__ void __launch_bounds__(256)
kvq_order_place(const __grid_constant__ KvqOut o, const __grid_constant__ KvqOrder ord)
{
    const uint32_t n = order_n(o, ord);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        ord.members[atomicAdd(&ord.cursor[ord.bkt[i]], 1u)] = i;
}

__global__ void __launch_bounds__(256)
kvq_order_rank(const __grid_constant__ KvqOut o, const __grid_constant__ KvqOrder ord)
{
    const uint32_t n = order_n(o, ord);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t bk = ord.bkt[i];
        const ulonglong2 me = ord.keys[i];
        uint32_t r = ord.start[bk];
        for (uint32_t m = ord.start[bk]; m < ord.start[bk + 1]; m++) {
            const uint32_t j = ord.members[m];
            if (j != i && key_before(ord.keys[j], j, me, i)) r++;
        }
        const KvqHitRec h = o.hits[i];
        KvqHitOut q; q.seq_nr = h.seq_nr; q.seq_pos = h.seq_pos; q.file_pos = h.file_pos; q.length = h.length; q.readlength = h.readlength;
        ord.hits[r] = q;
        ord.src[r] = (unsigned long long) h.blob_off;
        ord.lens[r + 1] = (uint32_t) h.length;
        if (i == 0) ord.lens[0] = 0;
    }
}

__global__ void __launch_bounds__(1024)
kvq_order_scan_lens(const __grid_constant__ KvqOut o, const __grid_constant__ KvqOrder ord)
{
    const uint32_t n = order_n(o, ord);
    if (n) block_scan(ord.lens, nullptr, n);
}

__global__ void __launch_bounds__(256)
kvq_order_copy(const __grid_constant__ KvqOut o, const __grid_constant__ KvqOrder ord)
{
    const uint32_t n = order_n(o, ord);
    const int lane = threadIdx.x & 31;
    const uint32_t nw = gridDim.x * (blockDim.x >> 5);
    for (uint32_t r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); r < n; r += nw) {
        const uint32_t at = ord.lens[r], len = ord.lens[r + 1] - at;
        const uint8_t *from = o.blob + ord.src[r];
        for (uint32_t x = lane; x < len; x += 32) ord.blob[at + x] = from[x];
        if (lane == 0) {
            ord.offsets[r] = (long long) at;
            if (r + 1 == n) ord.offsets[n] = (long long) ord.lens[n];
        }
    }
}

} // namespace

namespace {

__device__ __forceinline__ void export16(void *dst, const void *src, size_t nbytes, size_t tid, size_t nthreads)
{
    const uint4 *s4 = reinterpret_cast<const uint4 *>(src);
    uint4 *d4 = reinterpret_cast<uint4 *>(dst);
    for (size_t i = tid; i < (nbytes + 15) / 16; i += nthreads) d4[i] = s4[i];       // (both buffers are padded to 16 bytes)
}

__global__ void __launch_bounds__(256)
kvq_order_export(const __grid_constant__ KvqOut o, const __grid_constant__ KvqOrder ord)
{
    const uint32_t n = order_n(o, ord);
    if (!n) return;
    const size_t tid = (size_t) blockIdx.x * blockDim.x + threadIdx.x, nt = (size_t) gridDim.x * blockDim.x;
    export16(ord.out_hits, ord.hits, (size_t) n * sizeof(KvqHitOut), tid, nt);
    export16(ord.out_offsets, ord.offsets, ((size_t) n + 1) * 8, tid, nt);
    export16(ord.out_blob, ord.blob, (size_t) ord.lens[n], tid, nt);
}

} // namespace

int kvq_launch_order(const KvqBatch &b, const KvqOut &o, const KvqOrder &ord, void *stream)
{
    cudaStream_t st = (cudaStream_t) stream;
    cudaError_t e = cudaMemsetAsync(ord.start, 0, ((size_t) ord.cap + 1) * 4, st);
    if (e != cudaSuccess) return (int) e;
    const int g = (int) ((ord.cap + 255) / 256);
    kvq_order_count<<<g, 256, 0, st>>>(b, o, ord);
    kvq_order_scan_starts<<<1, 1024, 0, st>>>(o, ord);
    kvq_order_place<<<g, 256, 0, st>>>(o, ord);
    kvq_order_rank<<<g, 256, 0, st>>>(o, ord);
    kvq_order_scan_lens<<<1, 1024, 0, st>>>(o, ord);
    kvq_order_copy<<<296, 256, 0, st>>>(o, ord);
    if (ord.out_hits) kvq_order_export<<<148, 256, 0, st>>>(o, ord);
    return (int) cudaGetLastError();
}

// ---- the state a batch starts from, in one launch (instead of four memsets in a row in front of every scan kernel) ----
namespace {
__global__ void kvq_reset_kernel(KvqCounters *ctr, uint32_t *tile_status, int ntiles)
{
    const int tid = blockIdx.x * blockDim.x + threadIdx.x, nt = gridDim.x * blockDim.x;
    uint32_t *c = reinterpret_cast<uint32_t *>(ctr);
    constexpr int ERR = (int) (offsetof(KvqCounters, err_key) / 4), LON = (int) (offsetof(KvqCounters, longest) / 4);
    // all zero but err_key = ~0 (no format error yet) and longest = -1 (rls_longest, workhorse.c:332)
    for (int i = tid; i < (int) (sizeof(KvqCounters) / 4); i += nt)
        c[i] = (i == ERR || i == ERR + 1 || i == LON || i == LON + 1) ? 0xFFFFFFFFu : 0u;
    for (int i = tid; i < ntiles; i += nt) tile_status[i] = 0;
}
} // namespace

int kvq_launch_reset(KvqCounters *ctr, uint32_t *tile_status, int ntiles, void *stream)
{
    static_assert(sizeof(KvqCounters) % 4 == 0, "KvqCounters is zeroed word by word");
    kvq_reset_kernel<<<64, 256, 0, (cudaStream_t) stream>>>(ctr, tile_status, ntiles < 1 ? 1 : ntiles);
    return (int) cudaGetLastError();
}

// ---- tail of the accumulator row (multi-GPU scans that reduce their counters right behind the kernels) ----
// [readlengths KVQ_HIST][records][bases][parsed][longest + 1][this rank's batch has to be run again]: what the
// host adds to the row before a reduce (engine.cpp:upload_tail), written from the batch's device counters so
// that the all-reduce can be enqueued before the host has seen them.
namespace {
__global__ void kvq_tail_kernel(const KvqCounters *ctr, unsigned long long *tail, long long parsed, unsigned long long surv_cap)
{
    for (int i = threadIdx.x; i < KVQ_HIST; i += blockDim.x) tail[i] = ctr->rls[i];
    if (threadIdx.x == 0) {
        tail[KVQ_HIST] = ctr->records; tail[KVQ_HIST + 1] = ctr->bases; tail[KVQ_HIST + 2] = (unsigned long long) parsed;
        tail[KVQ_HIST + 3] = (unsigned long long) (ctr->longest + 1);
        tail[KVQ_HIST + 4] = (ctr->surv_count > surv_cap || ctr->phase_retry) ? 1ull : 0ull;
    }
}
} // namespace

int kvq_launch_tail(const KvqCounters *ctr, unsigned long long *tail, long long parsed, unsigned long long surv_cap, void *stream)
{
    kvq_tail_kernel<<<1, 256, 0, (cudaStream_t) stream>>>(ctr, tail, parsed, surv_cap);
    return (int) cudaGetLastError();
}

// ---- hit ordering without any allocation (kvq_order_hits_static, include/kvarq_b200.h) ----------
// No counterpart in the reference; added for callers that
// cannot call cudaMalloc (CUDA graph capture, tight serving loops).  A single-block variant of the
// kvq_order_* kernels above: same bucket sort over the stream position, same rank-by-full-key
// inside the (tiny) buckets, but its scratch -- keys, bucket numbers, bucket starts, cursors, member
// lists -- lives in ONE file-scope device array of fixed size instead of cudaMalloc'ed buffers, and
// the rule tag of a hit (which the public kvq_hit_t does not carry) is derived from the tuple and the
// template length.  Every GPU that loads the library holds the array (2 MiB) until exit; all calls
// share it, so calls that overlap on the GPU race (serialise them on one stream).
#include <cstddef>
#include <mutex>
#include <string>
#include "../../include/kvarq_b200.h"

namespace {

struct OrderScratch {
    ulonglong2 keys[KVQ_ORDER_STATIC_MAX];
    uint32_t bkt[KVQ_ORDER_STATIC_MAX];             // bucket | flush-right flag << 31
    uint32_t start[KVQ_ORDER_STATIC_MAX + 1];
    uint32_t cursor[KVQ_ORDER_STATIC_MAX + 1];
    uint32_t members[KVQ_ORDER_STATIC_MAX];
};

static __device__ __align__(256) unsigned char g_order_scratch[sizeof(OrderScratch)];

__device__ __forceinline__ unsigned long long order_key_y(int seq_nr, int rule, int seq_pos)
{
    const int sh = rule == 0 ? seq_pos : rule == 1 ? -seq_pos : (seq_pos < 0 ? -seq_pos : seq_pos);
    return ((unsigned long long) (uint32_t) seq_nr << 34) | ((unsigned long long) (uint32_t) rule << 32)
         | (unsigned long long) (uint32_t) (sh + (1 << 30));
}

// one block of 1024 threads; `sc` is the device address of g_order_scratch
__global__ void __launch_bounds__(1024)
kvq_order_static_kernel(OrderScratch *sc, const KvqHitOut *hits, uint32_t n, const int32_t *seqlens, int32_t nseqs,
                        KvqHitOut *sorted, uint32_t *order)
{
    __shared__ unsigned long long lo_s, hi_s;
    const uint32_t tid = threadIdx.x;
    if (tid == 0) { lo_s = ~0ULL; hi_s = 0ULL; }
    for (uint32_t i = tid; i <= n; i += 1024) sc->start[i] = 0;      // the scratch keeps the last call's contents
    __syncthreads();
    unsigned long long lo = ~0ULL, hi = 0ULL;
    for (uint32_t i = tid; i < n; i += 1024) {
        const unsigned long long f = (unsigned long long) hits[i].file_pos;
        lo = f < lo ? f : lo; hi = f > hi ? f : hi;
    }
    if (lo != ~0ULL) { atomicMin(&lo_s, lo); atomicMax(&hi_s, hi); }
    __syncthreads();
    lo = lo_s; hi = hi_s;
    // keys and buckets.  Rule of a hit (workhorse.c:1112-1174) from its tuple: seq_pos < 0: template
    // inside the read (c) when the whole template matched, else read tail on template head (a);
    // seq_pos > 0: read inside the template (d) when the whole read matched, else read head on
    // template tail (b).  A flush-right (d) hit may be one of the reference's two identical copies
    // (rules b + d): flagged, settled below.
    for (uint32_t i = tid; i < n; i += 1024) {
        const KvqHitOut h = hits[i];
        const int sl = (h.seq_nr >= 0 && h.seq_nr < nseqs) ? seqlens[h.seq_nr] : 0;
        int rule = 2; uint32_t flag = 0;
        if (h.seq_pos < 0) rule = h.length == sl ? 2 : 0;
        else if (h.seq_pos > 0) {
            rule = h.length == h.readlength ? 2 : 1;
            flag = (rule == 2 && h.seq_pos + h.readlength == sl) ? 0x80000000u : 0u;
        }
        ulonglong2 k; k.x = (unsigned long long) h.file_pos; k.y = order_key_y(h.seq_nr, rule, h.seq_pos);
        sc->keys[i] = k;
        unsigned long long bk = (k.x - lo) * n / (hi - lo + 1);
        if (bk >= n) bk = n - 1;
        sc->bkt[i] = (uint32_t) bk | flag;
        atomicAdd(&sc->start[bk + 1], 1u);
    }
    __syncthreads();
    if (n) block_scan(sc->start, sc->cursor, n);
    __syncthreads();
    for (uint32_t i = tid; i < n; i += 1024) sc->members[atomicAdd(&sc->cursor[sc->bkt[i] & 0x7FFFFFFFu], 1u)] = i;
    __syncthreads();
    // the two copies of a flush-right duplicate: the first is the reference's rule (b) hit, the second its rule (d) hit
    for (uint32_t i = tid; i < n; i += 1024) {
        if (!(sc->bkt[i] & 0x80000000u)) continue;
        const uint32_t bk = sc->bkt[i] & 0x7FFFFFFFu;
        const KvqHitOut me = hits[i];
        uint32_t same = 0, before = 0;
        for (uint32_t m = sc->start[bk]; m < sc->start[bk + 1]; m++) {
            const uint32_t j = sc->members[m];
            const KvqHitOut o = hits[j];
            if (o.file_pos == me.file_pos && o.seq_nr == me.seq_nr && o.seq_pos == me.seq_pos && o.length == me.length) {
                same++; before += j < i;
            }
        }
        if (same >= 2 && (before & 1u) == 0) sc->keys[i].y = order_key_y(me.seq_nr, 1, me.seq_pos);
    }
    __syncthreads();
    for (uint32_t i = tid; i < n; i += 1024) {
        const uint32_t bk = sc->bkt[i] & 0x7FFFFFFFu;
        const ulonglong2 me = sc->keys[i];
        uint32_t r = sc->start[bk];
        for (uint32_t m = sc->start[bk]; m < sc->start[bk + 1]; m++) {
            const uint32_t j = sc->members[m];
            if (j != i && key_before(sc->keys[j], j, me, i)) r++;
        }
        sorted[r] = hits[i];
        if (order) order[r] = i;
    }
}

std::mutex g_static_mu;
OrderScratch *g_static_ptr[64];                     // device address of g_order_scratch, per device
thread_local std::string g_static_err;

} // namespace

extern "C" const char *kvq_static_last_error(void) { return g_static_err.c_str(); }

extern "C" int kvq_order_hits_static(int device, const kvq_hit_t *d_hits, int64_t nhits, const int32_t *d_seqlens,
                                     int32_t nseqs, kvq_hit_t *d_sorted, uint32_t *d_order, void *stream)
{
    static_assert(sizeof(KvqHitOut) == sizeof(kvq_hit_t), "device hit layout == kvq_hit_t");
    g_static_err.clear();
    if (nhits < 0 || nhits > KVQ_ORDER_STATIC_MAX) {
        g_static_err = "kvq_order_hits_static: " + std::to_string((long long) nhits) + " hits, the static scratch takes at most "
                     + std::to_string(KVQ_ORDER_STATIC_MAX) + " (KVQ_ORDER_STATIC_MAX)";
        return KVQ_ERR_TYPE;
    }
    if (device < 0 || device >= 64 || (nhits > 0 && (!d_hits || !d_sorted || !d_seqlens)) || nseqs < 0) {
        g_static_err = "kvq_order_hits_static: bad arguments";
        return KVQ_ERR_TYPE;
    }
    if (nhits == 0) return KVQ_OK;
    int prev = -1;
    cudaError_t e = cudaGetDevice(&prev);
    if (e == cudaSuccess && prev != device) e = cudaSetDevice(device);
    OrderScratch *sc = nullptr;
    if (e == cudaSuccess) {
        // the array's name is a host-side stand-in, not a device address: ask for the address once per
        // device, with that device current
        std::lock_guard<std::mutex> g(g_static_mu);
        if (!g_static_ptr[device]) {
            void *p = nullptr;
            e = cudaGetSymbolAddress(&p, g_order_scratch);
            if (e == cudaSuccess) g_static_ptr[device] = static_cast<OrderScratch *>(p);
        }
        sc = g_static_ptr[device];
    }
    if (e == cudaSuccess) {
        kvq_order_static_kernel<<<1, 1024, 0, (cudaStream_t) stream>>>(sc, reinterpret_cast<const KvqHitOut *>(d_hits), (uint32_t) nhits,
                                                                      d_seqlens, nseqs, reinterpret_cast<KvqHitOut *>(d_sorted), d_order);
        e = cudaGetLastError();
    }
    if (prev >= 0 && prev != device) cudaSetDevice(prev);
    if (e != cudaSuccess) {
        g_static_err = std::string("kvq_order_hits_static: CUDA failure: ") + cudaGetErrorString(e);
        return KVQ_ERR_RUNTIME;
    }
    return KVQ_OK;
}
