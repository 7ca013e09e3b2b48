// Chunk-parallel inflate of one gzip file held in memory (N4: the step in front of the scan).
//
// The reference inflates a .gz FastQ with one miniz stream under the reader mutex
// (csrc/workhorse.c:781-885); round 1 of this engine used one zlib stream per file: 0.26 GB/s of
// FastQ, two orders of magnitude below what the GPU scans.  A deflate stream has no index, but it
// can still be decoded by many threads (the idea of pugz / rapidgzip):
//
//   1. the compressed bytes are cut into chunks of 2 MiB, two per thread; every thread looks for the first
//      deflate block that starts in its chunks (bit by bit: BFINAL = 0 / dynamic / HLIT, HDIST in range and a
//      complete code-length code are tested in line -- one position in ~1 500 survives --, then a header
//      whose literal and distance codes are complete, whose block decodes to its end-of-block and is
//      followed by another valid block header);
//   2. it decodes from there WITHOUT the 32 KiB of history in front of it: output goes to 16-bit
//      cells, a back-reference into the unknown history becomes a marker (0x8000 | window offset) and is
//      copied along like any other cell (in a FastQ the markers never die out: every header line copies
//      from the one before it), until the chunk lands exactly on the start another thread found.  The
//      first chunk of a round knows its history and writes bytes into the same cells.  The symbol loop is
//      branch-free (hot1 / hot2 below: a literal is a copy of one cell; literal-or-copy is a coin toss in
//      DNA, and so is every branch on it) and runs over the thread's two chunks together, so that one
//      chunk's chain of dependent table look-ups overlaps the other's;
//   3. the chunks are chained in stream order (a chunk start nobody lands on was a false positive and
//      its output is dropped); cells become bytes through a 64 KiB table made of the 32 KiB in front of
//      the chunk (the tail of the chunk before it, resolved first).
//
// Nothing is trusted: a chain that does not close, a candidate that fails, invalid data -> the
// member is decoded by the sequential reader below, which is the round-1 code on a memory buffer
// and reproduces the reference byte for byte, including the positions of its 1 MiB compressed
// reads (GzEvent: they feed the size estimate of engine.stats(), workhorse.c:883-884).
#pragma once
#include <stdint.h>
#include <string.h>
#include <zlib.h>
#include <stdio.h>
#include <stdlib.h>
#include <math.h>
#include <chrono>
#include <algorithm>
#include <atomic>
#include <functional>
#include <memory>
#include <mutex>
#include <condition_variable>
#include <stdexcept>
#include <string>
#include <thread>
#if defined(__x86_64__)
#include <immintrin.h>
#include <emmintrin.h>
#endif
#include <vector>

namespace kvq_gz {

// File position of the reference's reader as a function of the inflated bytes delivered: `pos`
// holds from output offset `out` on (at_end: the reposition behind a member's end, which happens in
// the call that delivers the member's last byte).
struct Event { int64_t out, pos; bool at_end; };

struct Error : std::runtime_error {
    bool io;        // IOError of the reference (else RuntimeError)
    Error(bool io_, const std::string &m) : std::runtime_error(m), io(io_) {}
};

// Buffers whose bytes are written before they are read: growing them must not zero-fill (a memset of the
// whole output, on one thread, costs as much as decoding it on sixteen).
template <class T> struct NoInit {
    using value_type = T;
    NoInit() = default;
    template <class U> NoInit(const NoInit<U> &) {}
    T *allocate(size_t n) { return static_cast<T *>(::operator new(n * sizeof(T))); }
    void deallocate(T *p, size_t) { ::operator delete(p); }
    template <class U> void construct(U *p) { ::new ((void *) p) U; }
    template <class U, class A0, class... A> void construct(U *p, A0 &&a0, A &&... a) { ::new ((void *) p) U(std::forward<A0>(a0), std::forward<A>(a)...); }
    template <class U> bool operator==(const NoInit<U> &) const { return true; }
    template <class U> bool operator!=(const NoInit<U> &) const { return false; }
};
using Bytes = std::vector<uint8_t, NoInit<uint8_t>>;
using Cells = std::vector<uint16_t, NoInit<uint16_t>>;

constexpr int64_t PIECE = 1024 * 1024;      // compressed bytes per read of the reference (SCANBUFSIZE, workhorse.c:15)
constexpr int WSIZE = 32768;

// ---------------------------------------------------------------------------------- bit access
struct Bits {
    const uint8_t *d; int64_t nbits;        // readable bits; d is readable 8 bytes past the end (callers pad)
    inline uint64_t peek(int64_t pos) const {       // >= 57 valid bits from bit position pos
        uint64_t v;
        memcpy(&v, d + (pos >> 3), 8);
        return v >> (pos & 7);
    }
};

// ---------------------------------------------------------------------------------- Huffman codes
// Canonical decoder: 10-bit direct table for the short codes, count/offset search for longer ones.
struct Huff {
    static constexpr int FAST = 10;
    uint16_t fast[1 << FAST];               // sym << 4 | len, 0 = not a short code
    uint16_t count[16], symbol[288];
    int maxlen = 0;

    // returns 0 complete, 1 incomplete, -1 over-subscribed
    int build(const uint8_t *len, int n)
    {
        memset(count, 0, sizeof(count));
        for (int i = 0; i < n; i++) count[len[i]]++;
        maxlen = 0;
        for (int l = 15; l > 0; l--) if (count[l]) { maxlen = l; break; }
        int left = 1;
        for (int l = 1; l <= 15; l++) {
            left <<= 1; left -= count[l];
            if (left < 0) return -1;
        }
        uint16_t offs[16]; offs[1] = 0;
        for (int l = 1; l < 15; l++) offs[l + 1] = offs[l] + count[l];
        for (int i = 0; i < n; i++) if (len[i]) symbol[offs[len[i]]++] = (uint16_t) i;
        memset(fast, 0, sizeof(fast));
        // canonical codes, bit-reversed into the table
        int code = 0, idx = 0;
        for (int l = 1; l <= FAST && l <= maxlen; l++) {
            for (int k = 0; k < count[l]; k++, idx++, code++) {
                int rev = 0;
                for (int b = 0; b < l; b++) rev |= ((code >> b) & 1) << (l - 1 - b);
                const uint16_t e = (uint16_t) (symbol[idx] << 4 | l);
                for (int f = rev; f < (1 << FAST); f += 1 << l) fast[f] = e;
            }
            code <<= 1;
        }
        count[0] = 0;
        return left > 0 ? 1 : 0;
    }

    // decodes one symbol from the low bits of v; returns the symbol and its length in `len`, -1 = invalid
    inline int decode(uint64_t v, int &len) const
    {
        const uint16_t e = fast[v & ((1 << FAST) - 1)];
        if (e) { len = e & 15; return e >> 4; }
        int code = 0, first = 0, index = 0;
        for (int l = 1; l <= maxlen; l++) {
            code |= (int) (v & 1); v >>= 1;
            const int c = count[l];
            if (code - c < first) { len = l; return symbol[index + (code - first)]; }
            index += c; first += c; first <<= 1; code <<= 1;
        }
        return -1;
    }
};

static const uint16_t LEN_BASE[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
static const uint8_t LEN_EXTRA[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
static const uint16_t DIST_BASE[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
static const uint8_t DIST_EXTRA[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};

// Packed entries for the decoder's hot loop, indexed by the next FAST bits:
//   bits 0-7 code length + extra bits, 8-11 code length, 14 a length (a copy follows), 15 not for the hot loop
//   (end of block, a code longer than FAST bits, an invalid code), 16-31 the literal or the base of the length /
//   distance.  A distance entry of 0 (long or invalid code) yields distance 0, which the hot loop stops at.
struct BlockCodes {
    Huff lit, dist;
    uint32_t lt[1 << Huff::FAST], dt[1 << Huff::FAST];
    void pack();
};

// Header of a dynamic block at bit position pos (behind BFINAL / BTYPE): code lengths -> codes.
// Strict: every code must be complete (a lone distance code excepted, as zlib allows), 256 must
// have a code.  Returns the bit position behind the header, -1 if invalid.
inline int64_t read_dynamic_header(const Bits &b, int64_t pos, BlockCodes &bc)
{
    static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
    if (pos + 14 > b.nbits) return -1;
    uint64_t v = b.peek(pos);
    const int nlen = (int) (v & 31) + 257, ndist = (int) ((v >> 5) & 31) + 1, ncode = (int) ((v >> 10) & 15) + 4;
    pos += 14;
    if (nlen > 286 || ndist > 30) return -1;
    if (pos + 3 * ncode > b.nbits) return -1;
    uint8_t cl[19] = {0};
    v = b.peek(pos);
    for (int i = 0; i < ncode; i++) { cl[order[i]] = (uint8_t) (v & 7); v >>= 3; }      // <= 57 bits
    pos += 3 * ncode;
    {   // complete code-length code?  (cheap: rejects nearly every wrong bit position of the block search)
        int kraft = 0;
        for (int i = 0; i < 19; i++) if (cl[i]) kraft += 128 >> cl[i];
        if (kraft != 128) return -1;
    }
    Huff lencode;
    if (lencode.build(cl, 19) != 0) return -1;
    uint8_t lens[320];
    int idx = 0;
    while (idx < nlen + ndist) {
        if (pos + 16 > b.nbits) return -1;
        v = b.peek(pos);
        int l;
        const int sym = lencode.decode(v, l);
        if (sym < 0) return -1;
        pos += l; v >>= l;
        if (sym < 16) lens[idx++] = (uint8_t) sym;
        else {
            int prev = 0, rep;
            if (sym == 16) {
                if (idx == 0) return -1;
                prev = lens[idx - 1]; rep = 3 + (int) (v & 3); pos += 2;
            } else if (sym == 17) { rep = 3 + (int) (v & 7); pos += 3; }
            else { rep = 11 + (int) (v & 127); pos += 7; }
            if (idx + rep > nlen + ndist) return -1;
            while (rep--) lens[idx++] = (uint8_t) prev;
        }
    }
    if (lens[256] == 0) return -1;
    if (bc.lit.build(lens, nlen) != 0) return -1;
    const int dr = bc.dist.build(lens + nlen, ndist);
    if (dr < 0 || (dr > 0 && bc.dist.count[1] + bc.dist.count[2] + bc.dist.count[3] + bc.dist.count[4] + bc.dist.count[5]
                   + bc.dist.count[6] + bc.dist.count[7] + bc.dist.count[8] + bc.dist.count[9] + bc.dist.count[10]
                   + bc.dist.count[11] + bc.dist.count[12] + bc.dist.count[13] + bc.dist.count[14] + bc.dist.count[15] > 1))
        return -1;
    return pos;
}

inline void BlockCodes::pack()
{
    for (int i = 0; i < (1 << Huff::FAST); i++) {
        uint32_t e = lit.fast[i], x = 0x8000;
        if (e) {
            const uint32_t sym = e >> 4, l = e & 15;
            if (sym < 256) x = sym << 16 | l << 8 | l;
            else if (sym > 256 && sym < 286)
                x = (uint32_t) LEN_BASE[sym - 257] << 16 | 0x4000 | l << 8 | (l + LEN_EXTRA[sym - 257]);
        }
        lt[i] = x;
        e = dist.fast[i]; x = 0;
        if (e && (e >> 4) < 30) x = (uint32_t) DIST_BASE[e >> 4] << 16 | (e & 15) << 8 | ((e & 15) + DIST_EXTRA[e >> 4]);
        dt[i] = x;
    }
}

inline void fixed_codes(BlockCodes &bc)
{
    uint8_t l[288];
    for (int i = 0; i < 144; i++) l[i] = 8;
    for (int i = 144; i < 256; i++) l[i] = 9;
    for (int i = 256; i < 280; i++) l[i] = 7;
    for (int i = 280; i < 288; i++) l[i] = 8;
    bc.lit.build(l, 288);
    uint8_t d[30];
    for (int i = 0; i < 30; i++) d[i] = 5;
    bc.dist.build(d, 30);
}

// ------------------------------------------------------------------------ decoding without history
// Output cells: 0..255 a byte, 0x8000 | k the byte at offset k of the (unknown) 32 KiB in front of the chunk.
struct MarkerOut {
    Cells cells;                            // capacity; `n` cells are valid
    int64_t n = 0;
    inline void room(int64_t more) { if (n + more > (int64_t) cells.size()) cells.resize((size_t) std::max<int64_t>(2 * (int64_t) cells.size(), n + more + (1 << 20))); }
};

// crossings of the reference's 1 MiB input pieces: (piece boundary as file byte offset, output produced by then)
struct Crossing { int64_t boundary, out; };

constexpr int64_t NO_BOUNDARY = INT64_MAX / 16;

struct CrossTracker {
    int64_t next;                           // next piece boundary (file byte offset) not yet crossed; NO_BOUNDARY: none
    int64_t step_base;                      // boundaries are step_base + k * PIECE
    std::vector<Crossing> list;
    void init(int64_t base, int64_t from_bit, int64_t file_size)
    {
        step_base = base;
        // first boundary b with 8 b >= from_bit (a symbol that ends behind 8 b crosses it)
        int64_t k = (std::max<int64_t>(0, (from_bit + 7) / 8 - base) + PIECE - 1) / PIECE;
        if (k == 0) k = 1;                  // the piece that starts at `base` is read before anything is decoded
        next = base + k * PIECE;
        if (next >= file_size) next = NO_BOUNDARY;
        fsize = file_size;
    }
    inline void symbol_end(int64_t end_bit, int64_t out_before)
    {
        while (end_bit > 8 * next) {
            list.push_back(Crossing{next, out_before});
            next += PIECE;
            if (next >= fsize) next = NO_BOUNDARY;
        }
    }
    int64_t fsize = 0;
};

struct LitCells { uint16_t cell[256 + 16]; LitCells() { for (int i = 0; i < 256 + 16; i++) cell[i] = (uint16_t) i; } };
static const LitCells LIT;

// One deflate block of one stream, resumable: begin() reads the header, hot1() / hot2() run the
// branch-free symbol loop (for one stream, or for two independent streams in the same loop: one
// symbol costs a chain of ~28 dependent cycles -- table look-up, shift, table look-up, shift --
// that a second stream's chain overlaps), careful() takes one symbol with every check: the ones the
// hot loop stops in front of (end of block, a code longer than FAST bits, a copy from closer than
// 16 cells or out of the window) and those near the end of the data or a 1 MiB input piece.
// Output: cells, see MarkerOut; win / wlen: the known history in front of cell 0 (the first chunk
// of a round), nullptr: unknown -> markers.
struct BlockDec {
    const Bits *b = nullptr; MarkerOut *o = nullptr; CrossTracker *ct = nullptr;
    const uint8_t *win = nullptr; int wlen = 0;
    int64_t max_out = 0;
    BlockCodes bc;
    int64_t pos = 0;                        // bit position of the next symbol / behind the block
    bool final = false;
    // the stretch the hot loop may run over, set by prepare()
    uint16_t *c = nullptr; int64_t n = 0, n_stop = 0, cross_bit = 0, fast_limit = 0, limit_bits = 0;
    enum Status { MORE, END, BAD };

    // header at bit position `at` (the BFINAL bit).  A stored block is copied whole: END.
    Status begin(int64_t at)
    {
        pos = at;
        if (pos + 3 > b->nbits) return BAD;
        uint64_t v = b->peek(pos);
        final = v & 1;
        const int type = (int) ((v >> 1) & 3);
        pos += 3;
        if (type == 3) return BAD;
        if (type == 0) {
            pos = (pos + 7) & ~(int64_t) 7;
            if (pos + 32 > b->nbits) return BAD;
            v = b->peek(pos);
            const uint32_t len = (uint32_t) (v & 0xFFFF), nlen = (uint32_t) ((v >> 16) & 0xFFFF);
            if ((len ^ 0xFFFF) != nlen) return BAD;
            pos += 32;
            if (pos + 8 * (int64_t) len > b->nbits || o->n + len > max_out) return BAD;
            const uint8_t *src = b->d + (pos >> 3);
            o->room(len);
            for (uint32_t i = 0; i < len; i++) {
                ct->symbol_end(pos + 8 * (int64_t) (i + 1), o->n);
                o->cells[(size_t) o->n++] = src[i];
            }
            pos += 8 * (int64_t) len;
            return END;
        }
        if (type == 1) fixed_codes(bc);
        else { pos = read_dynamic_header(*b, pos, bc); if (pos < 0) return BAD; }
        bc.pack();
        limit_bits = b->nbits - 48;         // (a symbol needs at most 15 + 5 + 15 + 13 bits)
        return MORE;
    }

    // room for the next stretch of output, and how far the hot loop may go without looking up
    inline void prepare()
    {
        o->room(4096);
        c = o->cells.data();
        n = o->n;
        n_stop = std::min<int64_t>((int64_t) o->cells.size() - 300, max_out);
        cross_bit = ct->next >= NO_BOUNDARY ? INT64_MAX : 8 * ct->next;
        fast_limit = std::min(limit_bits, cross_bit) - 128;
    }
    inline bool hot_ready() const { return pos < fast_limit && n < n_stop && n >= 16; }

    // one symbol, every check.  Needs n < n_stop (room for 258 cells).
    Status careful()
    {
        if (pos > limit_bits) return BAD;
        uint64_t v = b->peek(pos);
        int l;
        int sym = bc.lit.decode(v, l);
        if (sym < 0) return BAD;
        int64_t p = pos + l; v >>= l;
        if (sym < 256) {
            if (p > cross_bit) { ct->symbol_end(p, n); cross_bit = ct->next >= NO_BOUNDARY ? INT64_MAX : 8 * ct->next; }
            c[n++] = (uint16_t) sym;
            pos = p; o->n = n;
            return MORE;
        }
        if (sym == 256) { o->n = n; ct->symbol_end(p, n); pos = p; return END; }
        sym -= 257;
        if (sym >= 29) return BAD;
        const int len = LEN_BASE[sym] + (int) (v & ((1u << LEN_EXTRA[sym]) - 1));
        p += LEN_EXTRA[sym]; v >>= LEN_EXTRA[sym];
        int dl;
        const int ds = bc.dist.decode(v, dl);
        if (ds < 0 || ds >= 30) return BAD;
        p += dl; v >>= dl;
        const int dist = DIST_BASE[ds] + (int) (v & ((1u << DIST_EXTRA[ds]) - 1));
        p += DIST_EXTRA[ds];
        if (p > cross_bit) { ct->symbol_end(p, n); cross_bit = ct->next >= NO_BOUNDARY ? INT64_MAX : 8 * ct->next; }
        pos = p;
        int64_t from = n - dist;
        if (from >= 0) {
            const uint16_t *src = c + from; uint16_t *dst = c + n;
            if (dist >= len) memcpy(dst, src, 2 * (size_t) len);
            else for (int i = 0; i < len; i++) dst[i] = src[i];
            n += len;
        } else {
            if (from < -WSIZE) return BAD;
            if (win) {
                if (wlen + from < 0) return BAD;
                for (int i = 0; i < len; i++, from++) c[n++] = from < 0 ? (uint16_t) win[wlen + from] : c[from];
            } else for (int i = 0; i < len; i++, from++) c[n++] = from < 0 ? (uint16_t) (0x8000 | (WSIZE + from)) : c[from];
        }
        o->n = n;
        return MORE;
    }
};

// The registers of one stream inside the hot loop.
#define KVQ_GZ_HOT_OPEN(S, D)                                                                                   \
    const uint8_t *S##inp = (D).b->d + ((D).pos >> 3);                                                          \
    uint64_t S##bb = (uint64_t) *S##inp++ >> ((D).pos & 7);                                                     \
    int S##left = 8 - (int) ((D).pos & 7);                                                                      \
    { uint64_t w; memcpy(&w, S##inp, 8); S##bb |= w << S##left; S##inp += (63 - S##left) >> 3; S##left |= 56; } \
    const uint8_t *const S##stop = (D).b->d + ((D).fast_limit >> 3);                                            \
    const uint32_t *const S##lt = (D).bc.lt, *const S##dt = (D).bc.dt;                                          \
    uint16_t *const S##c = (D).c;                                                                               \
    int64_t S##n = (D).n;                                                                                       \
    const int64_t S##n_stop = (D).n_stop;                                                                       \
    uint32_t S##e = S##lt[S##bb & M];
// One symbol: a literal is a copy of one cell out of LIT; the distance look-up is made for every symbol and
// masked away for literals; `break` in front of anything unusual, with nothing consumed.  Everything is selected
// with masks (written as ?: the compiler makes a branch of it, and literal-or-copy is a coin toss).
// BZHI(x, n): the low (n & 255) bits of x.
#define KVQ_GZ_HOT_SYMBOL(S, BZHI)                                                                              \
    if (__builtin_expect(S##e & 0x8000, 0)) break;                                                              \
    const uint64_t S##mm = 0ull - ((S##e >> 14) & 1);                                                           \
    const uint32_t S##val = (uint32_t) (BZHI(S##bb, S##e) >> ((S##e >> 8) & 63)) + (S##e >> 16);               \
    const uint64_t S##b1 = S##bb >> (S##e & 63);                                                                \
    const uint32_t S##d = S##dt[S##b1 & M];                                                                     \
    const int64_t S##dist = (int64_t) ((BZHI(S##b1, S##d) >> ((S##d >> 8) & 63)) + (S##d >> 16));              \
    if (__builtin_expect(S##mm & (uint64_t) ((uint64_t) (S##dist - 16) > (uint64_t) (S##n - 16)), 0)) break;    \
    const int S##tot = (int) ((S##e & 0xFF) + ((S##d & 0xFF) & (uint32_t) S##mm));                              \
    const uintptr_t S##s_lit = (uintptr_t) (LIT.cell + S##val), S##s_m = (uintptr_t) (S##c + (S##n - S##dist)); \
    const uint16_t *S##src = (const uint16_t *) (S##s_lit ^ ((S##s_lit ^ S##s_m) & (uintptr_t) S##mm));         \
    const int S##len = 1 + (int) ((S##val - 1) & (uint32_t) S##mm);                                             \
    S##e = S##lt[(S##bb >> S##tot) & M];            /* (>= 16 bits are left: enough for the next code) */       \
    S##bb >>= S##tot; S##left -= S##tot;                                                                        \
    { uint64_t w; memcpy(&w, S##inp, 8); S##bb |= w << S##left; S##inp += (63 - S##left) >> 3; S##left |= 56; } \
    {                                                                                                           \
        uint16_t *dst = S##c + S##n;                                                                            \
        memcpy(dst, S##src, 16); memcpy(dst + 8, S##src + 8, 16);                                               \
        if (__builtin_expect(S##len > 16, 0))                                                                   \
            for (int i = 16; i < S##len; i += 8) memcpy(dst + i, S##src + i, 16);                               \
    }                                                                                                           \
    S##n += S##len;
#define KVQ_GZ_HOT_CLOSE(S, D)                                                                                  \
    (D).pos = 8 * (int64_t) (S##inp - (D).b->d) - S##left;                                                      \
    (D).n = S##n; (D).o->n = S##n;

#define KVQ_GZ_BZHI_PLAIN(x, n) ((x) & ~(~0ull << ((n) & 63)))
#define KVQ_GZ_BZHI_BMI2(x, n) _bzhi_u64((x), (uint32_t) (n))
#define KVQ_GZ_HOT1_BODY(BZHI)                                                                                  \
    {                                                                                                           \
        constexpr uint64_t M = (1u << Huff::FAST) - 1;                                                          \
        KVQ_GZ_HOT_OPEN(a, A)                                                                                   \
        while (an < an_stop && ainp < astop) {                                                                  \
            KVQ_GZ_HOT_SYMBOL(a, BZHI)                                                                          \
        }                                                                                                       \
        KVQ_GZ_HOT_CLOSE(a, A)                                                                                  \
    }
#define KVQ_GZ_HOT2_BODY(BZHI)                                                                                  \
    {                                                                                                           \
        constexpr uint64_t M = (1u << Huff::FAST) - 1;                                                          \
        KVQ_GZ_HOT_OPEN(a, A)                                                                                   \
        KVQ_GZ_HOT_OPEN(b, B)                                                                                   \
        while (an < an_stop && ainp < astop && bn < bn_stop && binp < bstop) {                                  \
            KVQ_GZ_HOT_SYMBOL(a, BZHI)                                                                          \
            KVQ_GZ_HOT_SYMBOL(b, BZHI)                                                                          \
        }                                                                                                       \
        KVQ_GZ_HOT_CLOSE(a, A)                                                                                  \
        KVQ_GZ_HOT_CLOSE(b, B)                                                                                  \
    }
// Built twice: BMI2 (shrx / shlx / bzhi: a third fewer instructions in a loop that is bound by them) and plain.
inline void hot1_plain(BlockDec &A) KVQ_GZ_HOT1_BODY(KVQ_GZ_BZHI_PLAIN)
inline void hot2_plain(BlockDec &A, BlockDec &B) KVQ_GZ_HOT2_BODY(KVQ_GZ_BZHI_PLAIN)
#if defined(__x86_64__)
__attribute__((target("bmi2"))) inline void hot1_bmi2(BlockDec &A) KVQ_GZ_HOT1_BODY(KVQ_GZ_BZHI_BMI2)
__attribute__((target("bmi2"))) inline void hot2_bmi2(BlockDec &A, BlockDec &B) KVQ_GZ_HOT2_BODY(KVQ_GZ_BZHI_BMI2)
inline bool cpu_has_bmi2() { static const bool yes = __builtin_cpu_supports("bmi2") && !getenv("KVQ_GZ_NO_BMI2"); return yes; }
#else
inline void hot1_bmi2(BlockDec &A) { hot1_plain(A); }
inline void hot2_bmi2(BlockDec &A, BlockDec &B) { hot2_plain(A, B); }
inline bool cpu_has_bmi2() { return false; }
#endif
// (both need hot_ready() of their streams)
inline void hot1(BlockDec &A) { if (cpu_has_bmi2()) hot1_bmi2(A); else hot1_plain(A); }
inline void hot2(BlockDec &A, BlockDec &B) { if (cpu_has_bmi2()) hot2_bmi2(A, B); else hot2_plain(A, B); }

// One deflate block at bit position pos (its BFINAL bit) into cells.  Returns the bit position
// behind the block, -1 on invalid data; *final = BFINAL.  max_out bounds the output.
inline int64_t marker_block(const Bits &b, int64_t pos, MarkerOut &o, CrossTracker &ct, bool *final, int64_t max_out,
                            const uint8_t *win = nullptr, int wlen = 0)
{
    BlockDec d;
    d.b = &b; d.o = &o; d.ct = &ct; d.win = win; d.wlen = wlen; d.max_out = max_out;
    BlockDec::Status st = d.begin(pos);
    *final = d.final;
    if (st == BlockDec::BAD) return -1;
    while (st == BlockDec::MORE) {
        d.prepare();
        if (d.n >= max_out) return -1;
        if (d.hot_ready()) hot1(d);
        if (d.n < d.n_stop) st = d.careful();
    }
    return st == BlockDec::END ? d.pos : -1;
}

// Is there a block start at bit position pos?  A non-final dynamic block that decodes to its end
// and is followed by a plausible block header.
inline bool plausible_block_start(const Bits &b, int64_t pos)
{
    if (pos + 17 > b.nbits) return false;
    const uint64_t v = b.peek(pos);
    if ((v & 7) != 4) return false;                     // BFINAL = 0, BTYPE = 10b (dynamic)
    if (((v >> 3) & 31) > 29 || ((v >> 8) & 31) > 29) return false;
    BlockCodes bc;
    if (read_dynamic_header(b, pos + 3, bc) < 0) return false;
    MarkerOut o; CrossTracker ct; ct.next = NO_BOUNDARY; ct.fsize = 0;
    bool fin;
    const int64_t end = marker_block(b, pos, o, ct, &fin, 4 << 20);
    if (end < 0 || o.n < 1024) return false;     // (real blocks of a FastQ hold tens of KB)
    if (end + 3 > b.nbits) return true;
    const uint64_t w = b.peek(end);
    const int type = (int) ((w >> 1) & 3);
    if (type == 3) return false;
    if (type == 2) { BlockCodes nb; return read_dynamic_header(b, end + 3, nb) >= 0; }
    if (type == 0) {
        const int64_t q = (end + 3 + 7) & ~(int64_t) 7;
        if (q + 32 > b.nbits) return false;
        const uint64_t x = b.peek(q);
        return ((x & 0xFFFF) ^ 0xFFFF) == ((x >> 16) & 0xFFFF);
    }
    return true;
}

inline int64_t find_block_start(const Bits &b, int64_t from_bit, int64_t to_bit)
{
    // the cheap tests here, in line (BFINAL = 0, BTYPE = dynamic, HLIT / HDIST in range, code-length code complete):
    // one position in ~1 500 goes on to the full test
    static const uint8_t K[8] = {0, 64, 32, 16, 8, 4, 2, 1};
    to_bit = std::min(to_bit, b.nbits - 17 - 57);
    for (int64_t pos = from_bit; pos < to_bit; pos++) {
        const uint64_t v = b.peek(pos);
        if ((v & 7) != 4 || ((v >> 3) & 31) > 29 || ((v >> 8) & 31) > 29) continue;
        const int ncode = (int) ((v >> 13) & 15) + 4;
        uint64_t w = b.peek(pos + 17) & ~(~0ull << (3 * ncode));        // (a fixed trip count: lengths not sent count as 0)
        int kraft = 0;
        for (int i = 0; i < 19; i++) { kraft += K[w & 7]; w >>= 3; }
        if (kraft != 128) continue;
        if (plausible_block_start(b, pos)) return pos;
    }
    return -1;
}

// ------------------------------------------------------------------------------- one chunk
struct Chunk {
    int64_t start_bit = -1;                 // where decoding starts (chunk 0: known; others: the block start found)
    bool ok = false;                        // decoded without an error up to `end_bit`
    int64_t end_bit = -1;                   // block boundary where it stopped
    int landed = -1;                        // index of the chunk whose start it stopped on, -1: none
    bool at_final = false, at_limit = false; // stopped in front of a final block / at the end of the round
    MarkerOut head;                         // the chunk's output, in cells
    double ms = 0;                          // decoding time, for the trace
    std::vector<Crossing> cross;            // 1 MiB input piece boundaries crossed, `out` counted from the chunk's start
    int64_t out_size() const { return head.n; }
};

struct RoundCtx {
    Bits bits; const uint8_t *file; int64_t fsize;
    int64_t piece_base;                     // file offset where the member's 1 MiB input pieces start
    int64_t stop_bit;                       // chunks stop at the first block boundary at or behind this
    const std::vector<Chunk> *chunks;       // (start_bit of every chunk: landing targets)
};

// stop test at a block boundary; returns true if the chunk ends here
inline bool chunk_stops_here(const RoundCtx &rc, Chunk &c, int self, int64_t pos)
{
    if (pos != c.start_bit)
        for (size_t j = (size_t) self + 1; j < rc.chunks->size(); j++)
            if ((*rc.chunks)[j].start_bit == pos) { c.landed = (int) j; return true; }
    if (pos + 1 > rc.bits.nbits) { c.at_limit = true; return true; }
    if (rc.bits.peek(pos) & 1) { c.at_final = true; return true; }         // the final block belongs to the sequential reader
    if (pos >= rc.stop_bit) { c.at_limit = true; return true; }
    return false;
}

// One chunk being decoded, resumable so that one thread can take two chunks through hot2() together.
// known: the history in front of start_bit is `window` (wlen <= 32 KiB bytes: the round's first chunk, which
// then writes bytes); else the whole chunk goes into marker cells (in a FastQ the markers never die out --
// every header line copies from the one before it -- so there is no point in looking for a clean window).
struct ChunkRun {
    const RoundCtx *rc = nullptr; Chunk *c = nullptr; int self = 0;
    CrossTracker ct; BlockDec d;
    bool in_block = false, finished = true, after_hot = false;

    void start(const RoundCtx &rc_, Chunk &c_, int self_, bool known, const uint8_t *window, int wlen)
    {
        rc = &rc_; c = &c_; self = self_;
        c->ok = false; c->head.n = 0;
        finished = c->start_bit < 0;
        if (finished) return;
        ct.init(rc->piece_base, c->start_bit, rc->fsize);
        d.b = &rc->bits; d.o = &c->head; d.ct = &ct; d.max_out = (int64_t) 1 << 31;
        d.win = known ? (window ? window : (const uint8_t *) "") : nullptr; d.wlen = known ? wlen : 0;
        d.pos = c->start_bit;
        in_block = false; after_hot = false;
    }
    // up to the next stretch for the hot loop (returns true), or to the end of the chunk (false)
    bool advance()
    {
        while (!finished) {
            if (!in_block) {
                if (chunk_stops_here(*rc, *c, self, d.pos)) { c->end_bit = d.pos; c->cross.swap(ct.list); c->ok = true; finished = true; break; }
                const BlockDec::Status st = d.begin(d.pos);
                if (st == BlockDec::BAD) { finished = true; break; }
                in_block = (st == BlockDec::MORE);
                after_hot = false;
                continue;
            }
            d.prepare();
            if (d.n >= d.max_out) { finished = true; break; }
            if (!after_hot && d.hot_ready()) return true;
            after_hot = false;
            if (d.n < d.n_stop) {
                const BlockDec::Status st = d.careful();
                if (st == BlockDec::BAD) { finished = true; break; }
                if (st == BlockDec::END) in_block = false;
            }
        }
        return false;
    }
};

// two chunks on one thread (either may be absent: start_bit < 0)
inline void decode_pair(ChunkRun &x, ChunkRun &y)
{
    bool rx = x.advance(), ry = y.advance();
    while (rx && ry) {
        hot2(x.d, y.d);
        x.after_hot = y.after_hot = true;
        rx = x.advance(); ry = y.advance();
    }
    while (rx) { hot1(x.d); x.after_hot = true; rx = x.advance(); }
    while (ry) { hot1(y.d); y.after_hot = true; ry = y.advance(); }
}

inline void decode_chunk(const RoundCtx &rc, Chunk &c, int self, bool known, const uint8_t *window, int wlen)
{
    ChunkRun x, none;
    x.start(rc, c, self, known, window, wlen);
    decode_pair(x, none);
}

// ------------------------------------------------------------------------------- worker team
// The threads of the parallel rounds, started once per file (three phases per round: starting sixteen threads
// for each costs more than a millisecond of a 30 ms round).  run(n, fn): fn(0 .. n-1), fn(0) on the caller.
class Team {
public:
    explicit Team(int n) { for (int i = 1; i < n; i++) th_.emplace_back([this, i] { work(i); }); }
    ~Team() {
        { std::lock_guard<std::mutex> g(mu_); quit_ = true; gen_++; }
        cv_.notify_all();
        for (auto &t : th_) t.join();
    }
    int size() const { return (int) th_.size() + 1; }
    bool run(int n, const std::function<void(int)> &fn) {
        n = std::min(n, size());
        {
            std::lock_guard<std::mutex> g(mu_);
            fn_ = &fn; n_ = n; pending_ = n - 1; failed_ = 0; gen_++;
        }
        cv_.notify_all();
        try { fn(0); } catch (...) { std::lock_guard<std::mutex> g(mu_); failed_++; }
        std::unique_lock<std::mutex> g(mu_);
        done_.wait(g, [&] { return pending_ == 0; });
        fn_ = nullptr;
        return failed_ == 0;
    }
private:
    void work(int i) {
        uint64_t seen = 0;
        for (;;) {
            const std::function<void(int)> *fn;
            {
                std::unique_lock<std::mutex> g(mu_);
                cv_.wait(g, [&] { return gen_ != seen; });
                seen = gen_;
                if (quit_) return;
                if (i >= n_) continue;
                fn = fn_;
            }
            bool bad = false;
            try { (*fn)(i); } catch (...) { bad = true; }
            std::lock_guard<std::mutex> g(mu_);
            failed_ += bad;
            if (--pending_ == 0) done_.notify_all();
        }
    }
    std::vector<std::thread> th_;
    std::mutex mu_; std::condition_variable cv_, done_;
    const std::function<void(int)> *fn_ = nullptr;
    int n_ = 0, pending_ = 0, failed_ = 0; uint64_t gen_ = 0; bool quit_ = false;
};

// ------------------------------------------------------------------------------- the file reader
// One .gz file in memory -> its inflated bytes and the reference's read positions, in order.
// The sequential part restates the round-1 GzFileSource (engine.cpp) on a memory buffer; parallel
// rounds run inside a member between block boundaries and leave the sequential state exactly where
// the reference's reader would be.
class FileDecoder {
public:
    // `file` must stay valid and be readable 16 bytes past `size`
    FileDecoder(const uint8_t *file, int64_t size, int nthreads)
        : f_(file), size_(size), threads_(std::max(1, nthreads))
    {
        memset(&zs_, 0, sizeof(zs_));
        if (inflateInit2(&zs_, -15) != Z_OK) throw Error(false, "cannot mz_inflateInit()");
        live_ = true;
        remaining_ = size_;
        const char *msg = skip_header(rd_, 0);
        if (msg) throw Error(true, std::string("no valid gzip header found at beginning of file : ") + msg);
        remaining_ -= rd_;
        member_start(rd_);
    }
    ~FileDecoder() { if (live_) inflateEnd(&zs_); }
    FileDecoder(const FileDecoder &) = delete;
    FileDecoder &operator=(const FileDecoder &) = delete;

    int64_t disk_pos() const { return rd_; }
    int64_t produced() const { return produced_; }
    // the next call of next() tries a parallel round (chunks of a member, or members side by side)
    bool parallel_mode() const { return (par_ok_ || ((small_members_ || members_ > 1) && at_member_start_)) && threads_ > 1 && !done_; }
    int64_t parallel_bytes() const { return par_bytes_; }      // inflated bytes that came out of parallel rounds

    // Appends the next stretch of output (one or more buffers) and the read events that fall into it.
    // Returns false when the stream has ended (nothing appended).
    // `spares`: buffers handed out earlier that the caller is done with (their pages are reused:
    // fresh pages cost more than the decoding).
    bool next(std::vector<Bytes> &outs, std::vector<Event> &ev, std::vector<Bytes> *spares = nullptr)
    {
        if (spares) { for (auto &b : *spares) if (spare_.size() < 64) spare_.push_back(std::move(b)); spares->clear(); }
        if (done_) return false;
        if ((small_members_ || members_ > 1) && at_member_start_ && threads_ > 1 && member_round(outs, ev)) return true;
        if (par_ok_ && threads_ > 1) {
            if (parallel_round(outs, ev)) return true;
            par_ok_ = false;
            handover(par_bit_);                     // the sequential reader takes the member from here
        }
        Bytes buf = take_spare((size_t) 4 << 20);
        bool filled = false;
        const int64_t got = fill(buf.data(), (int64_t) buf.size(), &filled, ev);
        buf.resize((size_t) got);
        if (got) outs.push_back(std::move(buf));
        if (!filled) done_ = true;
        return got > 0 || !done_;
    }

private:
    // ---- gzip member header on memory (skip_gz_header, workhorse.c:482-541; fgetc at EOF does not advance)
    const char *skip_header(int64_t &p, int dist) const
    {
        auto get = [&]() -> int { return p < size_ ? (int) f_[p++] : -1; };
        int state = 0, junk = 0, c = get();
        while (state != 2 && junk <= dist && c != -1) {
            if (state == 0 && c == 0x1F) state = 1;
            else if (state == 1 && c == 0x8B) state = 2;
            else { state = 0; junk++; }
            c = get();
        }
        if (state != 2) return "magic bytes not found";
        if (c != 8) return "expected method==DEFLATED";
        const int flags = get();
        if (flags & (0x02 | 0x20 | 0xC0)) return "unsupported flags (CONTINUATION or ENCRYPTED or RESERVED)";
        for (int i = 0; i < 6; i++) (void) get();
        if (flags & 0x04) { int n = get(); n |= get() << 8; while (n-- > 0) (void) get(); }
        if (flags & 0x08) { do c = get(); while (c > 0); }
        if (flags & 0x10) { do c = get(); while (c > 0); }
        return nullptr;
    }

    void member_start(int64_t data_at)
    {
        piece_base_ = data_at;
        member_out0_ = produced_;
        par_ok_ = !small_members_; par_bit_ = 8 * data_at; window_.clear();
        at_member_start_ = true; members_++;
    }

    // ---- the reference's reader (workhorse.c:781-885), zlib instead of miniz
    int64_t fill(uint8_t *dst, int64_t cap, bool *filled, std::vector<Event> &ev)
    {
        int64_t got = 0;
        bool at_member = false;
        zs_.next_out = dst; zs_.avail_out = (uInt) cap;
        do {
            if (zs_.avail_in == 0) {
                at_member_start_ = false;
                const int64_t m = std::min<int64_t>(PIECE, remaining_);
                if (rd_ + m > size_) throw Error(true, "could not read enough bytes from .fastq.gz");
                zs_.next_in = const_cast<Bytef *>(f_ + rd_); zs_.avail_in = (uInt) m;
                rd_ += m; remaining_ -= m;
                ev.push_back(Event{produced_ + got, rd_, false});
            }
            const uInt before = zs_.avail_out;
            int st = inflate(&zs_, Z_SYNC_FLUSH);
            if (st == Z_BUF_ERROR && zs_.avail_out == 0) st = Z_OK;
            if (st != Z_OK && st != Z_STREAM_END)
                throw Error(true, "error while inflating compressed data : status=" + std::to_string(st));
            got += before - zs_.avail_out;
            if (st == Z_STREAM_END && remaining_ + zs_.avail_in > 10) {
                // another member may follow after <= 10 junk bytes (:842-866)
                rd_ -= zs_.avail_in; remaining_ += zs_.avail_in;
                const int64_t pos = rd_;
                if (skip_header(rd_, 10) != nullptr) {
                    remaining_ = 0; zs_.avail_in = 0;
                    ev.push_back(Event{produced_ + got, rd_, true});
                    par_ok_ = false;
                    break;
                }
                remaining_ -= rd_ - pos;
                inflateReset2(&zs_, -15);
                zs_.avail_in = 0;
                ev.push_back(Event{produced_ + got, rd_, true});
                produced_ += got;               // (member_start reads produced_)
                member_start(rd_);
                produced_ -= got;
                at_member = true;               // hand the new member to a parallel round: nothing of it is read yet
                break;
            } else if (st == Z_STREAM_END) {
                break;
            }
        } while (remaining_ + zs_.avail_in > 10 && zs_.avail_out > 0);
        *filled = at_member || (zs_.avail_out == 0);
        produced_ += got;
        return got;
    }

    // sequential reader := "at block boundary `bit` with the history window_, input pieces read up to rd_"
    void handover(int64_t bit)
    {
        inflateReset2(&zs_, -15);
        int64_t nb = bit >> 3;
        if (bit & 7) { inflatePrime(&zs_, 8 - (int) (bit & 7), f_[nb] >> (bit & 7)); nb++; }
        if (!window_.empty()) inflateSetDictionary(&zs_, window_.data(), (uInt) window_.size());
        if (rd_ < nb) rd_ = nb;                 // (cannot happen: every byte taken lies in a piece already read)
        zs_.next_in = const_cast<Bytef *>(f_ + nb);
        zs_.avail_in = (uInt) (rd_ - nb);
        remaining_ = size_ - rd_;
    }

    bool parallel_round(std::vector<Bytes> &outs, std::vector<Event> &ev);
    bool member_round(std::vector<Bytes> &outs, std::vector<Event> &ev);

    Bytes take_spare(size_t n)       // a buffer of n bytes, from the spares when one is large enough
    {
        size_t best = spare_.size();
        for (size_t i = 0; i < spare_.size(); i++)
            if (spare_[i].capacity() >= n && (best == spare_.size() || spare_[i].capacity() < spare_[best].capacity())) best = i;
        Bytes b;
        if (best < spare_.size()) { b.swap(spare_[best]); spare_.erase(spare_.begin() + (long) best); }
        b.resize(n);
        return b;
    }
    std::vector<Bytes> spare_;
    std::vector<Cells> cells_pool_;
    std::unique_ptr<Team> team_;

    const uint8_t *f_; int64_t size_; int threads_;
    z_stream zs_; bool live_ = false;
    int64_t rd_ = 0, remaining_ = 0, produced_ = 0, piece_base_ = 0, member_out0_ = 0, par_bytes_ = 0;
    bool done_ = false, par_ok_ = false, small_members_ = false, at_member_start_ = false; int short_rounds_ = 0, members_ = 0;
    int64_t par_bit_ = 0; std::vector<uint8_t> window_;
};

inline bool FileDecoder::parallel_round(std::vector<Bytes> &outs, std::vector<Event> &ev)
{
    // Two chunks per thread (decode_pair), 2 MiB of compressed bytes each; towards the end of the file the chunks
    // shrink so that every thread still has its pair (down to 256 KiB: below that the sequential reader finishes).
    const int64_t sb = par_bit_ >> 3;
    int64_t CHUNK = 2 << 20;
    if (par_bit_ == 8 * piece_base_) CHUNK = 512 << 10;      // a member's first round is short: the scan behind it starts sooner
    if (size_ - sb < 2 * (int64_t) threads_ * CHUNK) CHUNK = std::max<int64_t>(256 << 10, (size_ - sb) / (2 * (int64_t) threads_));
    if (size_ - sb < 3 * CHUNK) return false;       // not worth a round
    const int T = (int) std::min<int64_t>(2 * (int64_t) threads_, (size_ - sb) / CHUNK);      // chunks
    const int W = (T + 1) / 2;                                                              // threads
    std::vector<Chunk> chunks((size_t) T);
    RoundCtx rc;
    rc.bits = Bits{f_, 8 * size_}; rc.file = f_; rc.fsize = size_; rc.piece_base = piece_base_;
    rc.stop_bit = 8 * std::min<int64_t>(size_, sb + (int64_t) T * CHUNK);
    rc.chunks = &chunks;
    chunks[0].start_bit = par_bit_;
    if ((int) cells_pool_.size() < T) cells_pool_.resize((size_t) T);
    for (int i = 0; i < T; i++) chunks[(size_t) i].head.cells.swap(cells_pool_[(size_t) i]);
    struct GiveBack {                               // the cell buffers go back to the pool whatever happens
        std::vector<Chunk> &ch; std::vector<Cells> &pool; int T;
        ~GiveBack() { for (int i = 0; i < T; i++) if (ch[(size_t) i].head.cells.capacity() > pool[(size_t) i].capacity()) ch[(size_t) i].head.cells.swap(pool[(size_t) i]); }
    } give_back{chunks, cells_pool_, T};
    if (!team_) team_.reset(new Team(threads_));
    auto in_threads = [&](int n, const std::function<void(int)> &fn) { return team_->run(n, fn); };
    const bool trace = getenv("KVQ_GZ_TRACE") != nullptr;
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t0 = now();
    // phase 0: the first block that starts in every chunk (looked for in the chunk's first quarter: deflate
    // blocks of text are tens of KB; data without such blocks -- stored, incompressible -- is not worth a round)
    in_threads(W, [&](int t) {
        for (int i = std::max(1, 2 * t); i < std::min(T, 2 * t + 2); i++)
            chunks[(size_t) i].start_bit = find_block_start(rc.bits, 8 * (sb + (int64_t) i * CHUNK),
                                                            std::min<int64_t>(8 * (sb + (int64_t) i * CHUNK + CHUNK / 4), 8 * size_ - 64));
    });
    int found = 0;
    for (int i = 1; i < T; i++) found += chunks[(size_t) i].start_bit >= 0;
    if (2 * found < T - 1) return false;
    const double t1 = now();
    // phase 1: every chunk from its start to the next start it lands on
    const bool known0 = true;                       // chunk 0 starts at a known boundary with the history in window_ (empty at a member's start)
    in_threads(W, [&](int t) {
        const double ta = now();
        ChunkRun x, y;
        try {
            x.start(rc, chunks[(size_t) (2 * t)], 2 * t, t == 0 && known0, window_.data(), (int) window_.size());
            if (2 * t + 1 < T) y.start(rc, chunks[(size_t) (2 * t + 1)], 2 * t + 1, false, nullptr, 0);
            decode_pair(x, y);
        } catch (...) { chunks[(size_t) (2 * t)].ok = false; if (2 * t + 1 < T) chunks[(size_t) (2 * t + 1)].ok = false; }
        chunks[(size_t) (2 * t)].ms = now() - ta;
        if (2 * t + 1 < T) chunks[(size_t) (2 * t + 1)].ms = chunks[(size_t) (2 * t)].ms;
    });
    const double t2 = now();
    // phase 2a: chain the chunks in stream order (a start nobody lands on was a false positive); the 32 KiB in
    // front of every chunk of the chain from the resolved tail of the one before it
    std::vector<int> order;
    bool broken = false;
    for (int i = 0; i >= 0; i = chunks[(size_t) i].landed) {
        if (!chunks[(size_t) i].ok) { broken = true; break; }
        order.push_back(i);
    }
    std::vector<std::vector<uint8_t>> wins(order.size() + 1);
    wins[0] = window_;
    for (size_t k = 0; k < order.size(); k++) {
        Chunk &c = chunks[(size_t) order[k]];
        const std::vector<uint8_t> &w = wins[k];
        std::vector<uint8_t> &t = wins[k + 1];
        const int64_t need_head = std::min<int64_t>(c.head.n, WSIZE);
        const int64_t need_prev = std::max<int64_t>(0, WSIZE - c.head.n);
        if (need_prev > 0) t.assign(w.end() - std::min<int64_t>(need_prev, (int64_t) w.size()), w.end());
        const int64_t have = (int64_t) w.size();
        const size_t t0n = t.size();
        t.resize(t0n + (size_t) need_head);
        uint8_t *to = t.data() + t0n;
        const uint16_t *from = c.head.cells.data() + (c.head.n - need_head);
        for (int64_t q = 0; q < need_head; q++) {
            const uint16_t x = from[q];
            if (x & 0x8000) {
                const int64_t wi = (int64_t) (x & 0x7FFF) - (WSIZE - have);
                if (wi < 0) { broken = true; break; }
                to[q] = w[(size_t) wi];
            } else to[q] = (uint8_t) x;
        }
        if (broken) { order.resize(k); break; }      // (reaches in front of the member: not a valid stream; zlib will say so)
    }
    // phase 2b: markers -> bytes, every chunk of the chain on its own thread
    std::vector<Bytes> heads(order.size());
    for (size_t k = 0; k < order.size(); k++)
        if (chunks[(size_t) order[k]].head.n) heads[k] = take_spare((size_t) chunks[(size_t) order[k]].head.n);
    std::atomic<int> bad{0};
    auto resolve_one = [&](int k) {
        Chunk &c = chunks[(size_t) order[(size_t) k]];
        if (c.head.n == 0) return;
        const std::vector<uint8_t> &w = wins[(size_t) k];
        const int64_t have = (int64_t) w.size();
        Bytes &hb = heads[(size_t) k];
        const uint16_t *cells = c.head.cells.data();
        if (have == WSIZE || k == 0) {
            // the usual case, branch-free: one table look-up per cell (0..255 -> itself, 0x8000 | k -> window[k]);
            // the round's first chunk holds no markers
            Bytes lut(65536);
            for (int v = 0; v < 256; v++) lut[(size_t) v] = (uint8_t) v;
            if (have) memcpy(lut.data() + 0x8000 + (WSIZE - have), w.data(), (size_t) have);
            const uint8_t *t = lut.data();
            uint8_t *o = hb.data();
            const int64_t n = c.head.n;
            int64_t q = 0;
#if defined(__SSE2__)
            // 16 cells at a time: without a marker among them (4 groups in 5: markers are 5 % of the cells and sit
            // in the header lines) they are narrowed in registers
            for (; q + 16 <= n; q += 16) {
                const __m128i a = _mm_loadu_si128((const __m128i *) (cells + q)), b = _mm_loadu_si128((const __m128i *) (cells + q + 8));
                if ((_mm_movemask_epi8(_mm_or_si128(a, b)) & 0xAAAA) == 0) { _mm_storeu_si128((__m128i *) (o + q), _mm_packus_epi16(a, b)); continue; }
                for (int j = 0; j < 16; j++) o[q + j] = t[cells[q + j]];
            }
#endif
            for (; q < n; q++) o[q] = t[cells[q]];
            return;
        }
        for (int64_t q = 0; q < c.head.n; q++) {
            const uint16_t x = cells[q];
            if (x & 0x8000) {
                const int64_t wi = (int64_t) (x & 0x7FFF) - (WSIZE - have);
                if (wi < 0) { bad++; return; }
                hb[(size_t) q] = w[(size_t) wi];
            } else hb[(size_t) q] = (uint8_t) x;
        }
    };
    in_threads(W, [&](int t) { for (int k = t; k < (int) order.size(); k += W) resolve_one(k); });
    const double t3 = now();
    if (trace) {
        fprintf(stderr, "[gz] t=%.1f round at byte %ld: search %.1f ms, decode %.1f ms, resolve %.1f ms, chain %zu of %d;", fmod(t0, 1e6), (long) sb, t1 - t0, t2 - t1, t3 - t2, order.size(), T);
        for (int i = 0; i < T; i++)
            fprintf(stderr, " [%d: +%ld ok=%d out=%ld land=%d %.1f ms]", i, (long) (chunks[(size_t) i].start_bit / 8 - sb - (int64_t) i * CHUNK),
                    chunks[(size_t) i].ok, (long) chunks[(size_t) i].head.n, chunks[(size_t) i].landed, chunks[(size_t) i].ms);
        fprintf(stderr, "\n");
    }
    if (bad.load() || order.empty()) return false;
    int64_t round_out = 0;
    for (int i : order) round_out += chunks[(size_t) i].out_size();
    if (round_out == 0) return false;               // (a final block right at the start)
    // phase 2c: hand the buffers out in stream order, with the reads of input pieces they stand for
    if (par_bit_ == 8 * piece_base_ && rd_ == piece_base_) {        // the member's first input piece is read before anything comes out
        rd_ = std::min(size_, piece_base_ + PIECE);
        ev.push_back(Event{produced_, rd_, false});
    }
    int64_t at = 0;
    for (size_t k = 0; k < order.size(); k++) {
        Chunk &c = chunks[(size_t) order[k]];
        for (const Crossing &x : c.cross) {
            rd_ = std::min(size_, x.boundary + PIECE);
            ev.push_back(Event{produced_ + at + x.out, rd_, false});
        }
        at += c.out_size();
        if (!heads[k].empty()) outs.push_back(std::move(heads[k]));
    }
    // the sequential state follows: block boundary end_bit, history = tail of the chain
    produced_ += round_out; par_bytes_ += round_out; at_member_start_ = false;
    // A round that ran into the member's final block after a fraction of what its threads decoded: members much
    // shorter than a round (BGZF: 64 KiB each).  More rounds would decode megabytes to keep kilobytes, so the rest of
    // the file goes through the sequential reader.
    if (chunks[(size_t) order.back()].at_final && round_out < (int64_t) T * CHUNK / 2 && ++short_rounds_ >= 2) small_members_ = true;
    par_bit_ = chunks[(size_t) order.back()].end_bit;
    window_.swap(wins[order.size()]);
    if (broken) par_ok_ = false;                    // what the chain has is good; the rest of the member goes sequentially
    if (!par_ok_) handover(par_bit_);
    return true;
}

// Files made of several members (BGZF and the .fastq.gz of Illumina's converters: one member per <= 64 KiB;
// concatenated or block-wise compressed files: members of megabytes).  Members are independent streams, so a
// round decodes the members that start in the next stretch of the file side by side -- candidates are the
// places where a valid member header stands -- and chains them the way the reference walks from member to
// member (end of the deflate data, <= 10 bytes of junk (the trailer), next header: workhorse.c:842-866).  A
// candidate nobody arrives at was not a member.  The read positions follow from the layout: a 1 MiB piece is
// read at every member's data start and whenever a symbol runs over the end of the last piece (CrossTracker).
// Tried at the start of every member but a file's first.
// State on entry and on return true: at a member's data start, nothing of it read (as after fill()'s `at_member`).
inline bool FileDecoder::member_round(std::vector<Bytes> &outs, std::vector<Event> &ev)
{
    const int64_t P = rd_;                                          // data start of the current member
    constexpr int64_t MARGIN = 2 << 20;                             // the end of the file is the sequential reader's
    // at least one candidate per thread (within 512 MiB), beyond that up to 1 MiB of file per thread (short members)
    const int64_t scan_max = std::min(size_ - MARGIN, P + ((int64_t) 512 << 20));
    const int64_t dense_end = P + (int64_t) threads_ * (1 << 20);
    if (scan_max <= P + 4096) return false;
    struct Cand { int64_t data, end = -1; size_t off = 0, len = 0; bool ok = false; std::vector<Crossing> cross; };
    std::vector<Cand> cand;
    cand.push_back(Cand{P});
    for (int64_t q = P; q + 18 < scan_max; q++) {
        if ((int) cand.size() >= threads_ && q >= dense_end) break;
        const void *hit = memchr(f_ + q, 0x1F, (size_t) (scan_max - 18 - q));
        if (!hit) break;
        q = (const uint8_t *) hit - f_;
        if ((int) cand.size() >= threads_ && q >= dense_end) break;
        if (f_[q + 1] != 0x8B || f_[q + 2] != 8 || (f_[q + 3] & (0x02 | 0x20 | 0xC0))) continue;
        int64_t d = q;
        if (skip_header(d, 0) == nullptr && d < scan_max && d > P) cand.push_back(Cand{d});
    }
    if (cand.size() < 4) return false;                              // not that kind of file after all
    // a few long members only (the tail of a file): the chunk rounds keep every thread busy inside each of them
    if ((int) cand.size() < threads_ / 2 && (cand.back().data - P) / (int64_t) cand.size() > (1 << 20)) return false;
    if (!team_) team_.reset(new Team(threads_));
    const int W = (int) std::min<size_t>((size_t) threads_, cand.size());
    if ((int) cells_pool_.size() < W) cells_pool_.resize((size_t) W);
    std::vector<Bytes> bufs((size_t) W);
    for (int t = 0; t < W; t++) bufs[(size_t) t] = take_spare((size_t) 4 << 20);
    auto range = [&](int t) { return std::make_pair(cand.size() * (size_t) t / (size_t) W, cand.size() * (size_t) (t + 1) / (size_t) W); };
    const Bits bits{f_, 8 * size_};
    team_->run(W, [&](int t) {
        // the in-tree decoder (twice zlib's rate), one member at a time: cells with an empty, known history, then bytes
        Bytes &buf = bufs[(size_t) t];
        MarkerOut o;
        o.cells.swap(cells_pool_[(size_t) t]);
        size_t used = 0;
        const auto r = range(t);
        for (size_t i = r.first; i < r.second; i++) {
            Cand &c = cand[i];
            // (a member of more than 32 MiB is left to the chunk rounds: a round holds one member per thread in memory; a chance match of the magic bytes inside a
            //  member's data decodes rubbish for a few symbols and fails)
            const int64_t lim = size_ - MARGIN;
            CrossTracker ct;
            ct.init(c.data, 8 * c.data, size_);
            c.off = used;
            o.n = 0;
            int64_t pos = 8 * c.data;
            bool fin = false, good = true;
            while (!fin && good) {
                pos = marker_block(bits, pos, o, ct, &fin, (int64_t) 32 << 20, (const uint8_t *) "", 0);
                good = pos >= 0 && ((pos + 7) >> 3) <= lim;
            }
            if (!good) continue;
            c.ok = true; c.end = (pos + 7) >> 3; c.len = (size_t) o.n; c.cross.swap(ct.list);
            if (buf.size() - used < c.len + 64) buf.resize(std::max(buf.size() * 2, used + c.len + 64));
            uint8_t *out = buf.data() + used;
            const uint16_t *cells = o.cells.data();
            int64_t q = 0;
#if defined(__SSE2__)
            for (; q + 16 <= o.n; q += 16)
                _mm_storeu_si128((__m128i *) (out + q), _mm_packus_epi16(_mm_loadu_si128((const __m128i *) (cells + q)),
                                                                         _mm_loadu_si128((const __m128i *) (cells + q + 8))));
#endif
            for (; q < o.n; q++) out[q] = (uint8_t) cells[q];
            used += c.len;
        }
        o.cells.swap(cells_pool_[(size_t) t]);
    });
    // chain from the current member on
    std::vector<size_t> chain;
    size_t cur = 0;
    int64_t next_data = -1;
    for (;;) {
        const Cand &c = cand[cur];
        if (!c.ok || size_ - c.end <= MARGIN) break;
        int64_t d = c.end;
        if (skip_header(d, 10) != nullptr) break;                   // (the sequential reader ends the stream there)
        chain.push_back(cur); next_data = d;
        size_t nxt = cur + 1;
        while (nxt < cand.size() && cand[nxt].data < d) nxt++;
        if (nxt >= cand.size() || cand[nxt].data != d) break;
        cur = nxt;
    }
    if (chain.empty()) { for (auto &b : bufs) if (spare_.size() < 64) spare_.push_back(std::move(b)); return false; }
    // events and output in stream order
    int64_t out = produced_;
    ev.push_back(Event{out, std::min(size_, P + std::min<int64_t>(PIECE, size_ - P)), false});    // the piece read at the current member's start
    for (size_t k = 0; k < chain.size(); k++) {
        const Cand &c = cand[chain[k]];
        for (const Crossing &x : c.cross) ev.push_back(Event{out + x.out, std::min(size_, x.boundary + PIECE), false});
        out += (int64_t) c.len;
        const int64_t d = k + 1 < chain.size() ? cand[chain[k + 1]].data : next_data;
        ev.push_back(Event{out, d, true});
        if (k + 1 < chain.size()) ev.push_back(Event{out, std::min(size_, d + std::min<int64_t>(PIECE, size_ - d)), false});
    }
    size_t k = 0;
    for (int t = 0; t < W && k < chain.size(); t++) {
        const auto r = range(t);
        // the thread's members that are in the chain, in order; usually all of them, back to back in its buffer
        size_t k0 = k; bool whole = true; size_t expect = 0;
        while (k < chain.size() && chain[k] < r.second) {
            if (cand[chain[k]].off != expect) whole = false;
            expect = cand[chain[k]].off + cand[chain[k]].len; k++;
        }
        if (k == k0) continue;
        Bytes &buf = bufs[(size_t) t];
        if (whole) { buf.resize(expect); if (expect) outs.push_back(std::move(buf)); }
        else {
            Bytes b2;
            for (size_t j = k0; j < k; j++) b2.insert(b2.end(), buf.begin() + (long) cand[chain[j]].off, buf.begin() + (long) (cand[chain[j]].off + cand[chain[j]].len));
            if (!b2.empty()) outs.push_back(std::move(b2));
        }
    }
    par_bytes_ += out - produced_;
    produced_ = out;
    // the sequential state: at the data start of the member behind the chain, nothing of it read
    inflateReset2(&zs_, -15);
    zs_.avail_in = 0;
    rd_ = next_data; remaining_ = size_ - next_data;
    member_start(next_data);
    return true;
}

} // namespace kvq_gz
