// Host side of libkvarq_b200.so: the C ABI of include/kvarq_b200.h, the FastQ streaming reader
// (plain / gzip / in-memory) that cuts the stream exactly where the reference's fastq_read would
// (csrc/workhorse.c:737-956), the pinned multi-buffered batch pipeline and result marshalling.
// All matching happens in scan_kernel.cu; there is no CPU matcher in this library.
#include <cuda_runtime.h>
#include <zlib.h>
#include <fcntl.h>
#include <unistd.h>
#include <math.h>
#include <sys/stat.h>
#include <sys/mman.h>
#if defined(__x86_64__)
#include <immintrin.h>
#endif
#include <set>
#include <functional>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/kvarq_b200.h"
#include "kvq_device.h"
#include "nccl_shim.h"
#include "gz_parallel.h"

namespace {

constexpr int64_t REF_CHUNK = 1024 * 1024;          // SCANBUFSIZE, workhorse.c:15
constexpr int64_t BATCH_BYTES = 64ll << 20;         // stream bytes per GPU batch
constexpr int64_t BATCH_PAD = 256;                  // zeroed slack behind a batch on the host side
constexpr int64_t DEV_PAD = (KVQ_OWN_ROWS + 3) * KVQ_ROW;   // zeroed slack behind a batch in HBM: every tile loads all its rows unconditionally
constexpr int NSLOT = 4;                // batches: one being cut by the producer, up to three on the GPU

struct KvqError {
    int cls; std::string msg;
    KvqError(int c, const std::string &m) : cls(c), msg(m) {}
};

inline void cuda_check(cudaError_t e, const char *what)
{
    if (e != cudaSuccess)
        throw KvqError(KVQ_ERR_RUNTIME, std::string("CUDA failure in ") + what + ": " + cudaGetErrorString(e));
}
#define CK(call) cuda_check((call), #call)

double now_ms()
{
    using namespace std::chrono;
    return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

// ------------------------------------------------------------------------------- byte sources
// One "file" of the scan.  fill() behaves like the fread / mz_inflate block of fastq_read:
// it delivers up to cap bytes and says whether the window could be filled.
struct Source {
    virtual ~Source() {}
    virtual int64_t fill(char *dst, int64_t cap, bool *window_filled) = 0;
    virtual int64_t disk_size() const = 0;          // on-disk (compressed) size
    virtual int64_t disk_pos() const = 0;           // ftell() of the reference at this point
    virtual bool compressed() const = 0;
    virtual const char *direct() const { return nullptr; }   // next unread byte, if the source is memory
    // Plain files: read up to `max` bytes ahead, with all reader threads, straight to where the
    // following fill() calls will want them (dst = target of the next fill); end_batch() keeps what
    // the batch did not take.
    virtual void preload(char *, int64_t) {}
    virtual void end_batch() {}
};

// A few persistent host threads for the byte-moving work of the reader (page cache -> pinned batch
// buffer): one thread copies at 5-6 GB/s, the H2D link takes 50+.
class WorkPool {
public:
    explicit WorkPool(int n) {
        for (int i = 0; i < n; i++) th_.emplace_back([this] { loop(); });
    }
    ~WorkPool() {
        { std::lock_guard<std::mutex> g(mu_); quit_ = true; }
        cv_.notify_all();
        for (auto &t : th_) t.join();
    }
    int size() const { return (int) th_.size(); }
    // fn(i) for i in [0, n), the caller takes part; returns when all are done
    void run(int n, const std::function<void(int)> &fn) {
        if (n <= 0) return;
        {
            std::lock_guard<std::mutex> g(mu_);
            fn_ = &fn; next_ = 0; total_ = n; pending_ = n; gen_++;
        }
        cv_.notify_all();
        work();
        std::unique_lock<std::mutex> g(mu_);
        done_.wait(g, [&] { return pending_ == 0; });
        fn_ = nullptr;
    }
private:
    void work() {
        for (;;) {
            int i;
            const std::function<void(int)> *fn;
            {
                std::lock_guard<std::mutex> g(mu_);
                if (!fn_ || next_ >= total_) return;
                i = next_++; fn = fn_;
            }
            (*fn)(i);
            std::lock_guard<std::mutex> g(mu_);
            if (--pending_ == 0) done_.notify_all();
        }
    }
    void loop() {
        uint64_t seen = 0;
        for (;;) {
            {
                std::unique_lock<std::mutex> g(mu_);
                cv_.wait(g, [&] { return quit_ || gen_ != seen; });
                if (quit_) return;
                seen = gen_;
            }
            work();
        }
    }
    std::vector<std::thread> th_;
    std::mutex mu_; std::condition_variable cv_, done_;
    const std::function<void(int)> *fn_ = nullptr;
    int next_ = 0, total_ = 0, pending_ = 0; uint64_t gen_ = 0; bool quit_ = false;
};

struct MemSource : Source {
    const char *p; int64_t n, at;
    MemSource(const char *p_, int64_t n_) : p(p_), n(n_), at(0) {}
    int64_t fill(char *dst, int64_t cap, bool *filled) override {
        int64_t m = std::min(cap, n - at);
        if (dst != p + at) memcpy(dst, p + at, m);      // zero-copy batches alias the caller's buffer
        at += m;
        *filled = (m == cap);
        return m;
    }
    int64_t disk_size() const override { return n; }
    int64_t disk_pos() const override { return at; }
    bool compressed() const override { return false; }
    const char *direct() const override { return p + at; }
};

// Copy out of a mapped file into a batch buffer with non-temporal stores: the destination is page-locked memory
// the GPU reads by DMA next, so there is no point in pulling its lines into the cache first (read-for-ownership
// is a third of a plain copy's DRAM traffic).  Measured against pread() on the same threads: 1.4x
// (profiles/microbench/read_modes.cpp).
#if defined(__x86_64__)
__attribute__((target("avx2"))) void stream_copy_avx2(char *dst, const char *src, size_t n)
{
    size_t head = (size_t) (-(intptr_t) dst & 31);
    if (head > n) head = n;
    memcpy(dst, src, head);
    size_t i = head;
    for (; i + 128 <= n; i += 128) {
        const __m256i a = _mm256_loadu_si256((const __m256i *) (src + i)), b = _mm256_loadu_si256((const __m256i *) (src + i + 32));
        const __m256i c = _mm256_loadu_si256((const __m256i *) (src + i + 64)), d = _mm256_loadu_si256((const __m256i *) (src + i + 96));
        _mm256_stream_si256((__m256i *) (dst + i), a); _mm256_stream_si256((__m256i *) (dst + i + 32), b);
        _mm256_stream_si256((__m256i *) (dst + i + 64), c); _mm256_stream_si256((__m256i *) (dst + i + 96), d);
    }
    _mm_sfence();
    memcpy(dst + i, src + i, n - i);
}
void stream_copy(char *dst, const char *src, size_t n)
{
    static const bool avx2 = __builtin_cpu_supports("avx2");
    if (avx2 && n >= 4096) stream_copy_avx2(dst, src, n); else memcpy(dst, src, n);
}
#else
void stream_copy(char *dst, const char *src, size_t n) { memcpy(dst, src, n); }
#endif

struct PlainFileSource : Source {
    int fd; int64_t size, pos = 0; WorkPool *pool;
    const char *map = nullptr;          // the file mapped (regular files; KVQ_READ_MMAP=0: pread instead)
    // file bytes [stg_off, stg_off + stg_len) already in memory at stg_ptr (the batch buffer, or `spill`)
    char *stg_ptr = nullptr; int64_t stg_off = 0, stg_len = 0; std::vector<char> spill;
    PlainFileSource(const char *name, WorkPool *pool_) : pool(pool_) {
        fd = open(name, O_RDONLY);
        struct stat st;
        if (fd < 0 || fstat(fd, &st) != 0) { if (fd >= 0) close(fd); throw KvqError(KVQ_ERR_IO, "cannot open file"); }   // workhorse.c:577
        size = (int64_t) st.st_size;
        static const bool use_map = !(getenv("KVQ_READ_MMAP") && atoi(getenv("KVQ_READ_MMAP")) == 0);
        if (use_map && S_ISREG(st.st_mode) && size > 0) {
            void *m = mmap(nullptr, (size_t) size, PROT_READ, MAP_SHARED, fd, 0);
            if (m != MAP_FAILED) { map = (const char *) m; madvise(m, (size_t) size, MADV_SEQUENTIAL); }
        }
    }
    ~PlainFileSource() override { if (map) munmap((void *) map, (size_t) size); if (fd >= 0) close(fd); }
    void read_range(int fd, char *dst, int64_t off, int64_t n) const {
        if (map) {                          // (the mapping ends at `size`; callers never ask for more)
            stream_copy(dst, map + off, (size_t) n);
            // the pages are not needed again: dropping them from the page table here, on the reader threads, keeps the
            // final munmap from walking a gigabyte of entries on one thread (8 ms per GB after the last batch)
            static const int64_t page = (int64_t) sysconf(_SC_PAGESIZE);
            const int64_t a = (off + page - 1) / page * page, b = (off + n) / page * page;
            if (b - a >= (64 << 10)) madvise((void *) (map + a), (size_t) (b - a), MADV_DONTNEED);
            return;
        }
        while (n > 0) {
            const ssize_t m = pread(fd, dst, (size_t) n, (off_t) off);
            if (m < 0) throw KvqError(KVQ_ERR_IO, "error while reading from file in fastq_read");   // :894
            if (m == 0) { memset(dst, 0, (size_t) n); return; }       // (file shrank under the scan)
            dst += m; off += m; n -= m;
        }
    }
    int64_t fill(char *dst, int64_t cap, bool *filled) override {
        const int64_t m = std::min(cap, size - pos);
        if (m > 0) {
            if (pos >= stg_off && pos + m <= stg_off + stg_len) {
                const char *src = stg_ptr + (pos - stg_off);
                if (src != dst) memcpy(dst, src, (size_t) m);
            } else read_range(fd, dst, pos, m);
        }
        pos += m;
        *filled = (m == cap);       // feof() is set exactly when fread came up short (:901)
        return m;
    }
    void preload(char *dst, int64_t max) override {
        const int64_t m = std::min(max, size - pos);
        if (m <= 0) { stg_len = 0; return; }
        if (map) {      // pages of a mapping behind the file's end raise SIGBUS: a file that shrank since it was opened ends the scan here
            struct stat st;
            if (fstat(fd, &st) != 0 || (int64_t) st.st_size < size)
                throw KvqError(KVQ_ERR_IO, "error while reading from file in fastq_read");          // :894
        }
        int64_t have = 0;
        if (stg_len > 0 && pos >= stg_off && pos < stg_off + stg_len) {     // what the last batch left over
            have = std::min(m, stg_off + stg_len - pos);
            memmove(dst, stg_ptr + (pos - stg_off), (size_t) have);
        }
        const int64_t from = pos + have, n = m - have, part = 2 << 20;
        const int parts = (int) ((n + part - 1) / part);
        std::exception_ptr err; std::mutex emu;
        auto job = [&](int i) {
            try { read_range(fd, dst + have + (int64_t) i * part, from + (int64_t) i * part, std::min<int64_t>(part, n - (int64_t) i * part)); }
            catch (...) { std::lock_guard<std::mutex> g(emu); err = std::current_exception(); }
        };
        if (pool && parts > 1) pool->run(parts, job);
        else for (int i = 0; i < parts; i++) job(i);
        if (err) std::rethrow_exception(err);
        stg_ptr = dst; stg_off = pos; stg_len = m;
    }
    void end_batch() override {
        if (stg_len > 0 && pos < stg_off + stg_len && stg_ptr != spill.data()) {
            std::vector<char> keep(stg_ptr + (pos - stg_off), stg_ptr + stg_len);
            spill.swap(keep);
            stg_ptr = spill.data(); stg_off = pos; stg_len = (int64_t) spill.size();
        } else if (pos >= stg_off + stg_len) stg_len = 0;
    }
    int64_t disk_size() const override { return size; }
    int64_t disk_pos() const override { return pos; }
    bool compressed() const override { return false; }
};

// gzip member header (skip_gz_header, workhorse.c:482-541): magic, method, flags, optional fields
const char *skip_gzip_header(FILE *fd, int dist)
{
    int state = 0, junk = 0, c = fgetc(fd);
    while (state != 2 && junk <= dist && c != EOF) {
        if (state == 0 && c == 0x1F) state = 1;
        else if (state == 1 && c == 0x8B) state = 2;
        else { state = 0; junk++; }
        c = fgetc(fd);
    }
    if (state != 2) return "magic bytes not found";
    if (c != 8) return "expected method==DEFLATED";
    int flags = fgetc(fd);
    if (flags & (0x02 | 0x20 | 0xC0)) return "unsupported flags (CONTINUATION or ENCRYPTED or RESERVED)";
    for (int i = 0; i < 6; i++) (void) fgetc(fd);
    if (flags & 0x04) { int n = fgetc(fd); n |= fgetc(fd) << 8; while (n-- > 0) (void) fgetc(fd); }
    if (flags & 0x08) { do c = fgetc(fd); while (c > 0); }
    if (flags & 0x10) { do c = fgetc(fd); while (c > 0); }
    return nullptr;
}

// .gz input: raw inflate of every member (workhorse.c:781-885), zlib instead of miniz
// File position of the reference's reader as a function of the inflated bytes delivered: `pos`
// holds from output offset `out` on.  A read of compressed input (at_end = false) happens when more
// output is wanted, so a fill that ends exactly at `out` has not issued it yet; the reposition behind
// a member's end (at_end = true) happens in the call that delivers the member's last byte.
struct GzEvent { int64_t out, pos; bool at_end; };

struct GzFileSource : Source {
    FILE *fd; int64_t size, remaining; z_stream zs; std::vector<unsigned char> in;
    int64_t produced = 0; std::vector<GzEvent> *log = nullptr; std::mutex *log_mu = nullptr;   // (GzAhead)
    void note(int64_t got, bool at_end) {
        if (!log) return;
        std::lock_guard<std::mutex> g(*log_mu);
        log->push_back(GzEvent{produced + got, (int64_t) ftell(fd), at_end});
    }
    explicit GzFileSource(const char *name) : in(REF_CHUNK) {
        fd = fopen(name, "rb");
        if (!fd) throw KvqError(KVQ_ERR_IO, "cannot open file");
        memset(&zs, 0, sizeof(zs));
        if (inflateInit2(&zs, -15) != Z_OK) { fclose(fd); throw KvqError(KVQ_ERR_RUNTIME, "cannot mz_inflateInit()"); }
        fseek(fd, 0, SEEK_END); size = remaining = ftell(fd); fseek(fd, 0, SEEK_SET);
        const char *msg = skip_gzip_header(fd, 0);
        if (msg) {
            inflateEnd(&zs); fclose(fd);
            throw KvqError(KVQ_ERR_IO, std::string("no valid gzip header found at beginning of file : ") + msg);
        }
        remaining -= ftell(fd);
    }
    ~GzFileSource() override { inflateEnd(&zs); if (fd) fclose(fd); }
    int64_t fill(char *dst, int64_t cap, bool *filled) override {
        int64_t got = 0;
        zs.next_out = (unsigned char *) dst; zs.avail_out = (unsigned) cap;
        do {
            if (zs.avail_in == 0) {
                int64_t m = std::min<int64_t>(REF_CHUNK, remaining);
                if ((int64_t) fread(in.data(), 1, m, fd) != m)
                    throw KvqError(KVQ_ERR_IO, "could not read enough bytes from .fastq.gz");
                zs.next_in = in.data(); zs.avail_in = (unsigned) m; remaining -= m;
                note(got, false);
            }
            unsigned before = zs.avail_out;
            int st = inflate(&zs, Z_SYNC_FLUSH);
            if (st == Z_BUF_ERROR && zs.avail_out == 0) st = Z_OK;
            if (st != Z_OK && st != Z_STREAM_END)
                throw KvqError(KVQ_ERR_IO, "error while inflating compressed data : status=" + std::to_string(st));
            got += before - zs.avail_out;
            if (st == Z_STREAM_END && remaining + zs.avail_in > 10) {
                // another member may follow after <= 10 junk bytes (:842-866)
                fseek(fd, -((long) zs.avail_in), SEEK_CUR);
                remaining += zs.avail_in;
                long pos = ftell(fd);
                if (skip_gzip_header(fd, 10) != nullptr) { remaining = 0; zs.avail_in = 0; note(got, true); break; }
                remaining -= ftell(fd) - pos;
                inflateReset2(&zs, -15);
                zs.avail_in = 0;
                note(got, true);
            } else if (st == Z_STREAM_END) {
                break;
            }
        } while (remaining + zs.avail_in > 10 && zs.avail_out > 0);
        *filled = (zs.avail_out == 0);
        produced += got;
        return got;
    }
    int64_t disk_size() const override { return size; }
    int64_t disk_pos() const override { return ftell(fd); }
    bool compressed() const override { return true; }
};

// A .gz file inflated ahead of the scan by its own host thread (zlib, one stream per file): the
// files of a scan -- the two mates of a paired-end run, the lanes of a sample -- inflate side by
// side instead of one after the other, and inflate overlaps the cutting of the batches.  What the
// scan sees is byte for byte what GzFileSource delivers, including the reference's file position
// after every fill (GzEvent), which feeds the size estimate of engine.stats().
// With more than one inflate thread per file the stream is decoded chunk-parallel (gz_parallel.h):
// same bytes, same read positions, several times the rate of one zlib stream.
// One chunk-parallel round at a time, the earliest file of the list first: the file being scanned has all the
// inflate threads, and the next one takes them over as soon as that file is done or has its read-ahead full.
// (Side by side with half the threads each, the second mate of a pair is finished long before the scan gets to
// it, and what remains then -- copying it out of the read-ahead -- runs on one thread with the others idle.)
struct GzGate {
    std::mutex mu; std::condition_variable cv;
    std::set<int> wanting; bool busy = false;
    void want(int idx) { std::lock_guard<std::mutex> g(mu); wanting.insert(idx); }          // (kept from round to round)
    void unwant(int idx) {
        { std::lock_guard<std::mutex> g(mu); wanting.erase(idx); }
        cv.notify_all();
    }
    void enter(int idx) {
        std::unique_lock<std::mutex> g(mu);
        cv.wait(g, [&] { return !busy && !wanting.empty() && *wanting.begin() == idx; });
        busy = true;
    }
    void leave() {
        { std::lock_guard<std::mutex> g(mu); busy = false; }
        cv.notify_all();
    }
};

class GzAhead : public Source {
public:
    static constexpr int64_t BLOCK = 4ll << 20, LIMIT = 512ll << 20;    // read-ahead per file
    GzAhead(const std::string &name, int threads, std::shared_ptr<GzGate> gate = nullptr, int index = 0)
        : name_(name), threads_(threads), gate_(std::move(gate)), index_(index) { th_ = std::thread([this] { run(); }); }
    ~GzAhead() override {
        { std::lock_guard<std::mutex> g(mu_); cancel_ = true; }
        cv_.notify_all();
        if (th_.joinable()) th_.join();
    }
    void wait_opened() {                // errors of fastq_open_next surface where the reference raises them
        std::unique_lock<std::mutex> g(mu_);
        cv_.wait(g, [&] { return opened_; });
        if (open_failed_) std::rethrow_exception(err_);
    }
    int64_t fill(char *dst, int64_t cap, bool *filled) override {
        int64_t got = 0;
        std::unique_lock<std::mutex> g(mu_);
        while (got < cap) {
            cv_.wait(g, [&] { return !blocks_.empty() || done_; });
            if (blocks_.empty()) {
                if (err_) std::rethrow_exception(err_);
                break;
            }
            kvq_gz::Bytes &bk = blocks_.front();
            const int64_t m = std::min<int64_t>(cap - got, (int64_t) bk.size() - front_off_);
            g.unlock();
            stream_copy(dst + got, (const char *) bk.data() + front_off_, (size_t) m);
            g.lock();
            got += m; front_off_ += m; buffered_ -= m;
            if (front_off_ == (int64_t) bk.size()) {
                if (spares_.size() < 32) spares_.push_back(std::move(bk));      // its pages are used again by the decoder
                blocks_.pop_front(); front_off_ = 0; cv_.notify_all();
            }
        }
        consumed_ += got;
        while (ev_at_ < events_.size() &&
               (events_[ev_at_].out < consumed_ || (events_[ev_at_].at_end && events_[ev_at_].out == consumed_)))
            pos_ = events_[ev_at_++].pos;
        *filled = (got == cap);
        return got;
    }
    int64_t disk_size() const override { return size_; }
    int64_t disk_pos() const override { return pos_; }
    bool compressed() const override { return true; }
private:
    void run_parallel() {
        // the whole compressed file mapped (readable 64 bytes past its end: the mapping of the file lies on top of
        // a zero-filled reservation one page longer), then round after round
        int fd = open(name_.c_str(), O_RDONLY);
        struct stat sb;
        if (fd < 0 || fstat(fd, &sb) != 0) { if (fd >= 0) close(fd); throw KvqError(KVQ_ERR_IO, "cannot open file"); }
        struct Mapping {
            uint8_t *p = nullptr; size_t n = 0;
            ~Mapping() { if (p) munmap(p, n); }
        } map;
        const size_t page = (size_t) sysconf(_SC_PAGESIZE);
        map.n = (((size_t) sb.st_size + 64 + page - 1) / page + 1) * page;
        void *base = mmap(nullptr, map.n, PROT_READ, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
        if (base == MAP_FAILED) { close(fd); throw KvqError(KVQ_ERR_IO, "cannot map file"); }
        map.p = (uint8_t *) base;
        if (sb.st_size > 0 && mmap(base, (size_t) sb.st_size, PROT_READ, MAP_PRIVATE | MAP_FIXED, fd, 0) == MAP_FAILED) {
            close(fd);
            throw KvqError(KVQ_ERR_IO, "could not read enough bytes from .fastq.gz");
        }
        close(fd);
        madvise(map.p, (size_t) sb.st_size, MADV_WILLNEED);
        struct { const uint8_t *p; const uint8_t *data() const { return p; } } file{map.p};
        try {
            kvq_gz::FileDecoder dec(file.data(), (int64_t) sb.st_size, threads_);
            {
                std::lock_guard<std::mutex> g(mu_);
                size_ = (int64_t) sb.st_size; pos_ = dec.disk_pos(); opened_ = true;
            }
            cv_.notify_all();
            std::vector<kvq_gz::Bytes> outs, spares;
            std::vector<kvq_gz::Event> ev;
            struct Unwant { GzGate *g; int idx; ~Unwant() { if (g) g->unwant(idx); } } unwant{gate_.get(), index_};
            for (;;) {
                {
                    std::unique_lock<std::mutex> g(mu_);
                    if (gate_ && !(buffered_ < LIMIT || cancel_)) gate_->unwant(index_);       // read-ahead full: the next file's turn
                    cv_.wait(g, [&] { return buffered_ < LIMIT || cancel_; });
                    if (cancel_) break;
                    spares.swap(spares_);
                }
                outs.clear(); ev.clear();
                const bool gated = gate_ && dec.parallel_mode();
                if (gated) { gate_->want(index_); gate_->enter(index_); }
                else if (gate_) gate_->unwant(index_);
                struct Leave { GzGate *g; ~Leave() { if (g) g->leave(); } } leave{gated ? gate_.get() : nullptr};
                const bool more = dec.next(outs, ev, &spares);
                if (gated) { gate_->leave(); leave.g = nullptr; }
                {
                    std::lock_guard<std::mutex> g(mu_);
                    for (const kvq_gz::Event &x : ev) events_.push_back(GzEvent{x.out, x.pos, x.at_end});
                    for (auto &o : outs) if (!o.empty()) { buffered_ += (int64_t) o.size(); blocks_.push_back(std::move(o)); }
                    if (!more) done_ = true;
                }
                cv_.notify_all();
                if (!more) break;
            }
        } catch (const kvq_gz::Error &err) {
            throw KvqError(err.io ? KVQ_ERR_IO : KVQ_ERR_RUNTIME, err.what());
        }
    }
    void run() {
        try {
            if (threads_ > 1) { run_parallel(); return; }
            GzFileSource src(name_.c_str());
            {
                std::lock_guard<std::mutex> g(mu_);
                size_ = src.size; pos_ = src.disk_pos(); opened_ = true;
            }
            src.log = &events_; src.log_mu = &mu_;
            cv_.notify_all();
            for (;;) {
                {
                    std::unique_lock<std::mutex> g(mu_);
                    cv_.wait(g, [&] { return buffered_ < LIMIT || cancel_; });
                    if (cancel_) break;
                }
                kvq_gz::Bytes bk((size_t) BLOCK);
                bool filled = false;
                const int64_t got = src.fill((char *) bk.data(), BLOCK, &filled);
                bk.resize((size_t) got);
                {
                    std::lock_guard<std::mutex> g(mu_);
                    if (got) { blocks_.push_back(std::move(bk)); buffered_ += got; }
                    if (!filled) done_ = true;
                }
                cv_.notify_all();
                if (!filled) break;
            }
        } catch (...) {
            std::lock_guard<std::mutex> g(mu_);
            err_ = std::current_exception();
            if (!opened_) { open_failed_ = true; opened_ = true; }
            done_ = true;
            cv_.notify_all();
        }
    }
    std::string name_; int threads_ = 1;
    std::shared_ptr<GzGate> gate_; int index_ = 0;
    std::thread th_;
    std::mutex mu_; std::condition_variable cv_;
    std::deque<kvq_gz::Bytes> blocks_; std::vector<kvq_gz::Bytes> spares_;
    int64_t front_off_ = 0, buffered_ = 0, consumed_ = 0;
    std::vector<GzEvent> events_; size_t ev_at_ = 0;
    int64_t size_ = 0, pos_ = 0;
    bool opened_ = false, open_failed_ = false, done_ = false, cancel_ = false;
    std::exception_ptr err_;
};

bool has_gz_suffix(const std::string &s)
{
    return s.size() >= 3 && s.compare(s.size() - 3, 3, ".gz") == 0;                 // workhorse.c:582
}

} // namespace

namespace {

// ------------------------------------------------------------------------------- stream cutter
// fastq_rewind (workhorse.c:696-718): going backwards, a line that starts with '+' arms the
// search, the next line start holding '@' is where the deferred tail begins.
int64_t deferred_tail(const char *buf, int64_t n)
{
    bool armed = false;
    for (int64_t back = 1; back + 1 < n; back++) {
        const char c = buf[n - back], prev = buf[n - back - 1];
        const bool line_start = (prev == '\n' || prev == '\r');
        if (c == '+' && line_start) armed = true;
        else if (armed && c == '@' && line_start) return back;
    }
    return -1;
}

struct LiveStats {                      // what engine.stats() may read while a scan runs
    std::atomic<int64_t> parsed{0}, total{0}, records{0}, longest{-1}, sigints{0};
    std::atomic<int64_t> rls[KVQ_MAX_READLENGTH];
    std::vector<std::atomic<int64_t>> seqhits, seqbasehits;
    void reset(int nseqs) {
        parsed = 0; total = 0; records = 0; longest = -1; sigints = 0;
        for (auto &r : rls) r = 0;
        seqhits = std::vector<std::atomic<int64_t>>(nseqs);
        seqbasehits = std::vector<std::atomic<int64_t>>(nseqs);
        for (auto &v : seqhits) v = 0;
        for (auto &v : seqbasehits) v = 0;
    }
};

struct HostBatch {
    char *own = nullptr;                // the slot's pinned buffer, BATCH_BYTES + BATCH_PAD
    bool borrowed = false;              // data aliases the caller's memory (in-memory sources): no host copy
    char *data = nullptr;               // own, or a pointer into the caller's buffer
    int64_t nbytes = 0;
    int64_t base_fpos = 0;
    std::vector<int64_t> cuts;          // segment boundaries, cuts[0] == 0, cuts.back() == nbytes
};

// Reproduces the sequence of fastq_read() calls of one reference worker thread and lays the
// returned chunks back to back into batches.
class StreamCutter {
public:
    StreamCutter(std::vector<std::string> names, std::vector<std::pair<const char *, int64_t>> mems,
                 LiveStats *live, WorkPool *pool = nullptr, int inflate_threads = 1)
        : names_(std::move(names)), mems_(std::move(mems)), live_(live), pool_(pool)
    {
        // every .gz file may use all the inflate threads: the parallel rounds take turns (GzGate); with one thread the
        // files of a scan inflate side by side, one zlib stream each (start_ahead)
        gz_threads_ = std::max(1, inflate_threads);
        if (gz_threads_ > 1) gate_ = std::make_shared<GzGate>();
        nfiles_ = names_.empty() ? (int) mems_.size() : (int) names_.size();
        // fastq_open (:659-677): total size first, then the first file
        for (int i = 0; i < nfiles_; i++) {
            if (!names_.empty()) {
                FILE *fd = fopen(names_[i].c_str(), "rb");
                if (!fd) throw KvqError(KVQ_ERR_IO, "cannot open file '" + names_[i] + "' for getting filesize");
                fseek(fd, 0, SEEK_END); size_all_ += ftell(fd); fclose(fd);
            } else size_all_ += mems_[i].second;
        }
        live_->parsed = 0; live_->total = size_all_;
        if (nfiles_ == 0) { finished_ = true; return; }
        ahead_.resize(names_.size());
        start_ahead();
        open_next();
    }

    bool finished() const { return finished_; }

    // Fills one batch; returns false when the stream is exhausted and the batch is empty.
    bool next_batch(HostBatch &b, const std::atomic<int> &stop)
    {
        b.cuts.clear(); b.cuts.push_back(0); b.nbytes = 0;
        int64_t left = (int64_t) carry_.size();
        if (!finished_ && at_eof_ && next_file_ < nfiles_) { open_next(); at_eof_ = false; }
        // in-memory sources are scanned in place: the deferred tail sits right before the next
        // unread byte of the same buffer (it is empty at a file switch)
        const char *direct = src_ ? src_->direct() : nullptr;
        b.borrowed = direct != nullptr;
        b.data = b.borrowed ? const_cast<char *>(direct) - left : b.own;
        if (left && !b.borrowed) memcpy(b.data, carry_.data(), left);
        carry_.clear();
        b.base_fpos = fpos_ - left;
        int64_t win = 0;
        // plain files: everything this batch can take is read now, by all reader threads, to where
        // the fill() calls below expect it (they find it in place and copy nothing)
        static const bool trace = getenv("KVQ_TRACE") != nullptr;
        const double tb0 = trace ? now_ms() : 0;
        // the first batches of a scan are short (8, 24 MiB, then 64): the GPU has its first bytes after 0.3 ms instead of 2.4
        const int64_t room = std::max<int64_t>(2 * REF_CHUNK, batch_no_ == 0 ? (8ll << 20) : batch_no_ == 1 ? (24ll << 20) : BATCH_BYTES);
        batch_no_++;
        if (!finished_ && src_ && !b.borrowed) src_->preload(b.data + left, room - left);
        const double tb1 = trace ? now_ms() : 0;
        while (!finished_) {
            if (stop.load(std::memory_order_relaxed)) { finished_ = true; left = 0; break; }
            if (win + REF_CHUNK > room) break;                              // batch full
            if (at_eof_ && next_file_ < nfiles_) {                          // :749-753
                if (b.borrowed) break;                                      // next buffer = next batch
                open_next();
                src_->preload(b.data + win + left, room - win - left);
            }
            at_eof_ = false;
            bool filled = false;
            const int64_t got = src_->fill(b.data + win + left, REF_CHUNK - left, &filled);
            at_eof_ = !filled;
            if (compressed_) {                                              // :883-884, single precision
                float est = (float) (uint64_t) size_all_ * (float) (uint64_t) (fpos_ + got)
                          / (float) (uint64_t) (done_disk_ + src_->disk_pos());
                live_->total = (int64_t) (uint64_t) est;
            }
            if (got == 0) {                                                 // :905-910
                if (left == 0) { finished_ = true; break; }
                win += left; b.cuts.push_back(win); left = 0;
                continue;
            }
            fpos_ += got; live_->parsed += got;
            if (!at_eof_) {
                const int64_t t = deferred_tail(b.data + win, left + got);
                if (t < 0)
                    throw KvqError(KVQ_ERR_RUNTIME, "could find beginning of record; read " + std::to_string(got)
                                   + " bytes up to " + std::to_string(src_->disk_pos()));
                win += left + got - t; left = t;
            } else {
                win += left + got; left = 0;
            }
            b.cuts.push_back(win);
        }
        if (left) carry_.assign(b.data + win, b.data + win + left);
        if (src_) src_->end_batch();
        b.nbytes = win;
        if (!b.borrowed) memset(b.data + win, 0, BATCH_PAD);
        if (trace) fprintf(stderr, "[kvq trace] producer t=%.3f: preload %.3f ms, cut %.3f ms, %ld bytes\n", fmod(tb0, 1e6), tb1 - tb0, now_ms() - tb1, (long) win);
        return win > 0;
    }

private:
    void open_next()                                                        // fastq_open_next, :559-629
    {
        if (src_) { done_disk_ += src_->disk_pos(); src_.reset(); }
        const int i = next_file_++;
        if (!names_.empty()) {
            if (has_gz_suffix(names_[i])) {
                if (!ahead_[i]) ahead_[i].reset(new GzAhead(names_[i], gz_threads_, gate_, (int) i));
                ahead_[i]->wait_opened();
                src_ = std::move(ahead_[i]);
                start_ahead();
                compressed_ = true;
                live_->total = live_->total * 3;                            // "random guess", :625
            } else {
                if (compressed_)
                    throw KvqError(KVQ_ERR_IO, "error while inflating compressed data : plain file after .gz");
                src_.reset(new PlainFileSource(names_[i].c_str(), pool_));
            }
        } else {
            src_.reset(new MemSource(mems_[i].first, mems_[i].second));
        }
    }

    // up to AHEAD .gz files of the list inflate ahead of the one being scanned
    void start_ahead()
    {
        int running = 0;
        for (size_t i = (size_t) next_file_; i < names_.size() && running < AHEAD; i++) {
            if (!has_gz_suffix(names_[i])) break;                           // (a plain file after .gz is an error anyway)
            if (!ahead_[i]) ahead_[i].reset(new GzAhead(names_[i], gz_threads_, gate_, (int) i));
            running++;
        }
    }
    static constexpr int AHEAD = 4;
    std::vector<std::unique_ptr<GzAhead>> ahead_;
    std::vector<std::string> names_;
    std::vector<std::pair<const char *, int64_t>> mems_;
    LiveStats *live_;
    WorkPool *pool_ = nullptr; int gz_threads_ = 1;
    std::shared_ptr<GzGate> gate_;
    int nfiles_ = 0, next_file_ = 0, batch_no_ = 0;
    std::unique_ptr<Source> src_;
    bool at_eof_ = false, compressed_ = false, finished_ = false;
    int64_t size_all_ = 0, done_disk_ = 0, fpos_ = 0;
    std::vector<char> carry_;
};

} // namespace

namespace {

// ------------------------------------------------------------------------------- device memory
struct DevBuf {
    void *p = nullptr; size_t cap = 0;
    ~DevBuf() { if (p) cudaFree(p); }
    void ensure(size_t n) {
        if (n <= cap && p) return;
        if (p) { cudaFree(p); p = nullptr; cap = 0; }
        size_t want = std::max<size_t>(n, 256);
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) { p = nullptr; throw KvqError(KVQ_ERR_MEMORY, std::string("cudaMalloc: ") + cudaGetErrorString(e)); }
        cap = want;
    }
    template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

struct PinBuf {
    void *p = nullptr; size_t cap = 0;
    ~PinBuf() { if (p) cudaFreeHost(p); }
    void ensure(size_t n) {
        if (n <= cap && p) return;
        if (p) { cudaFreeHost(p); p = nullptr; cap = 0; }
        cudaError_t e = cudaHostAlloc(&p, n, cudaHostAllocDefault);
        if (e != cudaSuccess) { p = nullptr; throw KvqError(KVQ_ERR_MEMORY, std::string("cudaHostAlloc: ") + cudaGetErrorString(e)); }
        cap = n;
    }
    template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

template <class T> void upload(DevBuf &d, const std::vector<T> &v, size_t min_elems = 1)
{
    d.ensure(std::max(v.size(), min_elems) * sizeof(T));
    if (!v.empty()) CK(cudaMemcpy(d.p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
}

// ------------------------------------------------------------------------------- templates
inline uint32_t code_of(unsigned char c) { return KVQ_CODE(c); }

uint32_t kmer_key(const unsigned char *s)            // KVQ_SEED_K bases, base 0 in the low bits
{
    uint32_t k = 0;
    for (int i = 0; i < KVQ_SEED_K; i++) k |= code_of(s[i]) << (2 * i);
    return k;
}

struct SeedTable {
    std::vector<uint32_t> bitmap, off; std::vector<uint4> ent; std::vector<uint8_t> present;
    void build(std::vector<std::pair<uint32_t, uint4>> &items, int pieces, bool flag_short) {
        bitmap.assign(KVQ_BITMAP_WORDS, 0); off.assign(KVQ_NKEYS + 1, 0); ent.resize(items.size());
        present.assign(KVQ_NKEYS, 0);
        for (auto &it : items) { off[it.first + 1]++; bitmap[it.first >> 5] |= 1u << (it.first & 31); present[it.first] = 1; }
        for (int k = 0; k < KVQ_NKEYS; k++) off[k + 1] += off[k];
        std::vector<uint32_t> at(off.begin(), off.end() - 1);
        for (auto &it : items) ent[at[it.first]++] = it.second;
        // direct-mapped first entries + overflow array (see KvqTemplates)
        first.assign(KVQ_NKEYS, make_uint4(0, 0, 0, 0)); ovf.clear();
        for (int k = 0; k < KVQ_NKEYS; k++) {
            const uint32_t b0 = off[k], b1 = off[k + 1];
            if (b0 == b1) continue;
            const uint32_t more = b1 - b0 - 1;
            if (more > 65000u)                  // (a seed word of the scan kernel counts bucket entries in 16 bits)
                throw KvqError(KVQ_ERR_RUNTIME, "template set not supported: more than 65000 templates share one 8-mer at the same seed class");
            if (more >= KVQ_HEAD_MORE) ovf.push_back(make_uint4(more, 0, 0, 0));
            uint4 f = ent[b0];
            f.x |= (more >= KVQ_HEAD_MORE ? KVQ_HEAD_MORE : more) << 24; f.w = (uint32_t) ovf.size();
            for (uint32_t e = b0; e < b1; e++)
                if (flag_short && (int) (ent[e].y >> 16) / pieces < KVQ_T1X_K) f.x |= KVQ_HEAD_SHORT;
            first[k] = f;
            for (uint32_t e = b0 + 1; e < b1; e++) ovf.push_back(ent[e]);
        }
        if (ovf.empty()) ovf.push_back(make_uint4(0, 0, 0, 0));
    }
    std::vector<uint4> first, ovf;
};

// Large sets: one more table (nkeys heads) appended to the shared head / overflow arrays; head
// {seq | more << 24, toff | len << 16, packed_off, index into the shared overflow array}, empty
// heads are all zero (len == 0).
void append_table(std::vector<std::pair<uint32_t, uint4>> &items, uint32_t nkeys,
                  std::vector<uint4> &first_all, std::vector<uint4> &ent_all)
{
    const size_t base = first_all.size();
    first_all.resize(base + nkeys, make_uint4(0, 0, 0, 0));
    std::stable_sort(items.begin(), items.end(),
                     [](const std::pair<uint32_t, uint4> &a, const std::pair<uint32_t, uint4> &b) { return a.first < b.first; });
    for (size_t b0 = 0; b0 < items.size();) {
        size_t b1 = b0;
        while (b1 < items.size() && items[b1].first == items[b0].first) b1++;
        const uint32_t more = (uint32_t) (b1 - b0 - 1);
        if (more >= KVQ_HEAD_MORE) ent_all.push_back(make_uint4(more, 0, 0, 0));
        uint4 f = items[b0].second;
        f.x |= (more >= KVQ_HEAD_MORE ? KVQ_HEAD_MORE : more) << 24; f.w = (uint32_t) ent_all.size();
        first_all[base + items[b0].first] = f;
        for (size_t e = b0 + 1; e < b1; e++) ent_all.push_back(items[e].second);
        b0 = b1;
    }
}

// KVQ_EXACT_PHASE=1: no tile infers its line phase from its own line starts (tests of the exact path)
int exact_phase_knob()
{
    const char *v = getenv("KVQ_EXACT_PHASE");
    return v && atoi(v) > 0 ? 1 : 0;
}

struct TemplateSet {
    int nseqs = 0, nplus = 0; int64_t cov_total = 0;
    std::vector<int32_t> len;
    KvqTemplates dev{}; KvqParams par{};
    DevBuf d_bytes, d_boff, d_packed, d_poff, d_len, d_exh;
    DevBuf d_bm1, d_ent1, d_bm2, d_ent2, d_covoff, d_p2, d_bm2l, d_first1, d_first2, d_t1x, d_t2e, d_t2x;

    static unsigned char comp(unsigned char c) {
        switch (c) { case 'A': return 'T'; case 'T': return 'A'; case 'C': return 'G'; case 'G': return 'C'; default: return c; }
    }

    std::vector<char> cache_key, pending_key;

    void build(const char *const *seqs, const int32_t *lens, int n, int np, const kvq_config_t &cfg)
    {
        if (n < 0 || (n > 0 && (!seqs || !lens))) throw KvqError(KVQ_ERR_TYPE, "seqlist must be sequence of strings");
        {   // same templates and knobs as the previous scan: keep the device tables
            std::vector<char> key;
            auto put = [&](const void *q, size_t m) { key.insert(key.end(), (const char *) q, (const char *) q + m); };
            put(&cfg, sizeof(cfg)); put(&n, sizeof(n)); put(&np, sizeof(np));
            bool ok = true;
            for (int i = 0; i < n && ok; i++) ok = lens[i] >= 0 && seqs[i];
            if (ok) {
                put(lens, sizeof(int32_t) * (size_t) n);
                for (int i = 0; i < n; i++) put(seqs[i], (size_t) lens[i]);
                if (key == cache_key && d_bytes.p) { par.want_hits = 1; par.accumulate = 1; par.exact_phase = exact_phase_knob(); return; }
                pending_key.swap(key);
            } else pending_key.clear();
            cache_key.clear();          // committed as the last statement: a build that throws leaves no key behind
        }
        nseqs = n; nplus = np; len.assign(lens, lens + n);
        for (int i = 0; i < n; i++)
            if (lens[i] < 0 || !seqs[i]) throw KvqError(KVQ_ERR_TYPE, "seqlist must be list of strings");
        if (np > 0) {
            if (n != 2 * np) throw KvqError(KVQ_ERR_TYPE, "nplus must be half the number of sequences");
            for (int i = 0; i < np; i++) {
                if (lens[i] != lens[np + i]) throw KvqError(KVQ_ERR_TYPE, "minus sequence length differs from plus sequence");
                for (int j = 0; j < lens[i]; j++)
                    if ((unsigned char) seqs[np + i][j] != comp((unsigned char) seqs[i][lens[i] - 1 - j]))
                        throw KvqError(KVQ_ERR_TYPE, "seqs[nplus+i] is not the reverse complement of seqs[i]");
            }
        }
        const int E = cfg.maxerrors, O = cfg.minoverlap;
        const int pieces = E + 1;
        const bool seed_ok = E >= 0 && pieces <= KVQ_MAX_PIECES && O >= 1 && O / pieces >= KVQ_SEED_K;

        std::vector<uint8_t> bytes; std::vector<uint32_t> boff(n + 1, 0), packed, poff(n + 1, 0);
        std::vector<uint8_t> seedable(n, 0); std::vector<int32_t> exh;
        std::vector<std::pair<uint32_t, uint4>> it1, it2;
        // short / long template classes for the whole-read T2 probes: the long class holds at most
        // a quarter of the T2 entries (it is everything when no such split exists)
        int short_max = 0;
        if (seed_ok) {
            std::vector<std::pair<int, int64_t>> bylen;             // (length, T2 entries)
            for (int i = 0; i < n; i++)
                if (lens[i] / pieces >= KVQ_SEED_K && lens[i] <= KVQ_MAX_SEEDED_SEQLEN) bylen.push_back({lens[i], lens[i] - KVQ_SEED_K + 1});
            std::sort(bylen.begin(), bylen.end());
            int64_t total = 0, acc = 0;
            for (auto &x : bylen) total += x.second;
            for (size_t k = bylen.size(); k-- > 0;) {
                acc += bylen[k].second;
                if (acc * 4 > total) { short_max = bylen[k].first; break; }     // this length and below are "short"
            }
            // (the longest length ends up short when it alone holds more than a quarter: no long class)
        }
        std::vector<uint32_t> bm2l(KVQ_BITMAP_WORDS, 0), t1x;
        // large sets (kernel variant BIG): 12-mer second level for T1, slot-specific T2 tables
        size_t n_ok = 0;
        for (int i = 0; i < n; i++)
            n_ok += seed_ok && lens[i] / pieces >= KVQ_SEED_K && lens[i] <= KVQ_MAX_SEEDED_SEQLEN && i < (1 << 24);
        const bool t1x_on = n_ok * (size_t) pieces > KVQ_T1X_MIN_PIECES && 5 * pieces <= 32;
        const int OX = KVQ_T1X_K * pieces;                          // overlaps >= OX hold E+1 disjoint 12-mers
        std::vector<std::vector<std::pair<uint32_t, uint4>>> itH(pieces), itT(pieces);
        std::vector<std::pair<uint32_t, uint4>> itX;
        std::vector<uint32_t> t2e, t2x; std::vector<uint8_t> t1mask, t1roles(KVQ_NKEYS, 0);
        if (t1x_on) t1mask.assign(KVQ_NKEYS, 0);
        if (t1x_on) { t2e.assign(KVQ_BITMAP_WORDS, 0); t2x.assign(KVQ_T1X_WORDS, 0); }
        int max_len = 0;
        for (int i = 0; i < n; i++) {
            const unsigned char *s = (const unsigned char *) seqs[i]; const int sl = lens[i];
            boff[i] = (uint32_t) bytes.size(); bytes.insert(bytes.end(), s, s + sl);
            poff[i] = (uint32_t) packed.size();
            const int words = (sl + 15) / 16 + 2;
            packed.resize(packed.size() + words, 0);
            for (int j = 0; j < sl; j++) packed[poff[i] + j / 16] |= code_of(s[j]) << (2 * (j % 16));
            max_len = std::max(max_len, sl);
            const bool ok = seed_ok && sl / pieces >= KVQ_SEED_K && sl <= KVQ_MAX_SEEDED_SEQLEN && i < (1 << 24);
            seedable[i] = ok;
            if (!ok) { exh.push_back(i); continue; }
            // entry = {seq, toff | len << 16, offset of the packed codes, offset of the raw bytes}
            auto entry = [&](int toff) { return make_uint4((unsigned) i, (unsigned) toff | ((unsigned) sl << 16), poff[i], boff[i]); };
            for (int j = 0; j < pieces; j++) {
                const int toff = (int) ((int64_t) j * sl / pieces);
                it1.push_back({kmer_key(s + toff), entry(toff)});
                if (sl / pieces >= KVQ_T1_LONG) {       // small sets: roles of the piece's two 8-mers (scan_kernel.cu, probe_roles)
                    t1roles[kmer_key(s + toff)] |= 2; t1roles[kmer_key(s + toff + KVQ_SEED_K)] |= 4;
                } else t1roles[kmer_key(s + toff)] |= 1;
                if (t1x_on)                             // probe filter: the 3 code bits after the 8-mer (bases 8, 9)
                    t1mask[kmer_key(s + toff)] |= sl / pieces >= KVQ_T1X_K
                        ? (uint8_t) (1u << (code_of(s[toff + 8]) | ((code_of(s[toff + 9]) & 1u) << 2))) : (uint8_t) 0xFF;
                if (sl / pieces >= KVQ_T1X_K) {         // second level: the piece's first 12 bases
                    if (t1x.empty()) t1x.assign(KVQ_T1X_WORDS, 0);
                    uint32_t k = 0;
                    for (int q = 0; q < KVQ_T1X_K; q++) k |= code_of(s[toff + q]) << (2 * q);
                    t1x[k >> 5] |= 1u << (k & 31);
                }
            }
            for (int toff = 0; toff + KVQ_SEED_K <= sl; toff++) {
                const uint32_t key = kmer_key(s + toff);
                it2.push_back({key, entry(toff)});
                if (sl > short_max) bm2l[key >> 5] |= 1u << (key & 31);
                if (!t1x_on) continue;
                // Slot-specific tables.  Head slot j (read offset o_j) answers for read-head-on-template-tail
                // overlaps ov in [O, OX): template offset sl - ov + o_j; tail slot j (read offset rl - O + o_j)
                // for read-tail-on-template-head overlaps ov in [O, OX): template offset ov - O + o_j.
                // Overlaps >= OX are found through the 12-mer slots (any template offset).
                if (sl > O)
                    for (int j = 0; j < pieces; j++) {
                        const int oj = (int) ((int64_t) j * O / pieces);
                        if (toff >= sl - OX + 1 + oj && toff <= sl - O + oj) {
                            itH[j].push_back({key, entry(toff)}); t2e[key >> 5] |= 1u << (key & 31);
                        }
                        if (toff >= oj && toff <= OX - 1 - O + oj && toff <= sl - 1 - O + oj) {
                            itT[j].push_back({key, entry(toff)}); t2e[key >> 5] |= 1u << (key & 31);
                        }
                    }
                if (toff + KVQ_T1X_K <= sl) {
                    uint32_t k = 0;
                    for (int q = 0; q < KVQ_T1X_K; q++) k |= code_of(s[toff + q]) << (2 * q);
                    t2x[k >> 5] |= 1u << (k & 31);
                    itX.push_back({KVQ_X_HASH(k), entry(toff)});
                }
            }
        }
        boff[n] = (uint32_t) bytes.size(); poff[n] = (uint32_t) packed.size();
        bytes.resize(bytes.size() + 64, 0);
        if (t1x_on && t1x.empty()) t1x.assign(KVQ_T1X_WORDS, 0);
        SeedTable t1, t2; t1.build(it1, pieces, t1x_on); t2.build(it2, pieces, false);
        if (t1x_on) {                           // every T2 class back to back (see KvqTemplates)
            std::vector<uint4> first_all, ent_all;
            append_table(it2, KVQ_NKEYS, first_all, ent_all);
            for (int j = 0; j < pieces; j++) append_table(itH[j], KVQ_NKEYS, first_all, ent_all);
            for (int j = 0; j < pieces; j++) append_table(itT[j], KVQ_NKEYS, first_all, ent_all);
            append_table(itX, 1u << KVQ_X_HASH_BITS, first_all, ent_all);
            if (ent_all.empty()) ent_all.push_back(make_uint4(0, 0, 0, 0));
            t2.first.swap(first_all); t2.ovf.swap(ent_all);
        }
        std::vector<uint32_t> covoff(np + 1, 0);
        for (int i = 0; i < np; i++) covoff[i + 1] = covoff[i] + (uint32_t) lens[i];
        cov_total = np > 0 ? covoff[np] : 0;

        upload(d_bytes, bytes); upload(d_boff, boff); upload(d_packed, packed, 4); upload(d_poff, poff);
        upload(d_len, len); upload(d_exh, exh);
        upload(d_bm1, t1x_on ? t1mask : t1roles); upload(d_ent1, t1.ovf); upload(d_first1, t1.first);
        upload(d_bm2, t2.bitmap); upload(d_ent2, t2.ovf); upload(d_first2, t2.first);
        upload(d_bm2l, bm2l); if (t1x_on) { upload(d_t1x, t1x); upload(d_t2e, t2e); upload(d_t2x, t2x); }
        upload(d_covoff, covoff);

        dev.bytes = d_bytes.as<uint8_t>(); dev.byte_off = d_boff.as<uint32_t>();
        dev.packed = d_packed.as<uint32_t>(); dev.packed_off = d_poff.as<uint32_t>(); dev.packed_words = (int32_t) packed.size();
        dev.len = d_len.as<int32_t>();
        dev.exh_list = d_exh.as<int32_t>(); dev.nseqs = n; dev.n_exh = (int) exh.size();
        dev.t1_present = d_bm1.as<uint8_t>(); dev.t1_ent = d_ent1.as<uint4>();
        dev.t2_bitmap = d_bm2.as<uint32_t>(); dev.t2_ent = d_ent2.as<uint4>();
        dev.t1_first = d_first1.as<uint4>(); dev.t2_first = d_first2.as<uint4>();
        dev.t2l_bitmap = d_bm2l.as<uint32_t>(); dev.t1x_bitmap = d_t1x.as<uint32_t>(); dev.t2e_bitmap = d_t2e.as<uint32_t>(); dev.t2x_bitmap = d_t2x.as<uint32_t>();
        dev.cov_off = d_covoff.as<uint32_t>(); dev.nplus = np;

        par.maxerrors = E; par.minoverlap = O; par.minreadlength = cfg.minreadlength;
        par.amin = (int) (signed char) cfg.Amin;
        par.seed_ok = seed_ok; par.npieces = pieces < 1 ? 1 : pieces; par.max_seqlen = max_len;
        par.pieces_inv = (uint32_t) std::min<uint64_t>((1ull << 32) / (uint64_t) par.npieces + 1, 0xFFFFFFFFull);
        par.want_hits = 1; par.accumulate = 1; par.exact_phase = exact_phase_knob();
        par.t1x = t1x_on ? KVQ_T1X_K * pieces : 0;
        par.min_seed_rl = KVQ_SEED_K * par.npieces;
        for (int j = 0; j < KVQ_MAX_PIECES; j++) par.o_piece[j] = (int) ((int64_t) j * O / par.npieces);
        par.n2max = (t1x_on ? 5 : 3) * par.npieces; par.nl_byte = '\n';
        // Fixed read offsets probed against T2, per trimmed read length: E+1 pieces of the whole read
        // (rule d, only if some template is at least as long), of its first minoverlap bases
        // (rule b) and of its last minoverlap bases (rule a); duplicates removed.
        std::vector<int16_t> p2((size_t) (KVQ_MAXRL_FAST + 1) * KVQ_P2W, (int16_t) -1);
        if (seed_ok)
            for (int rl = par.min_seed_rl; rl <= KVQ_MAXRL_FAST; rl++) {
                int16_t *row = p2.data() + (size_t) rl * KVQ_P2W; int nn = 0;
                // a whole-read piece can only lie in a template at least as long as the read: reads longer
                // than the short class probe the long-template bitmap there (KVQ_P2_LONG)
                auto add = [&](int v, bool lng) {
                    for (int i = 0; i < nn; i++)
                        if ((row[i] & KVQ_P2_POS) == v) { if (!lng) row[i] = (int16_t) v; return; }
                    row[nn++] = (int16_t) (v | (lng ? KVQ_P2_LONG : 0));
                };
                if (t1x_on) {                   // fixed layout, slot = class * pieces + j: W, H, T, HX, TX
                    if (max_len >= rl)
                        for (int j = 0; j < pieces; j++)
                            row[j] = (int16_t) ((int) ((int64_t) j * rl / pieces) | (KVQ_T2L && rl > short_max ? KVQ_P2_LONG : 0));
                    if (rl > O) {
                        for (int j = 0; j < pieces; j++) row[pieces + j] = (int16_t) par.o_piece[j];
                        for (int j = 0; j < pieces; j++) row[2 * pieces + j] = (int16_t) (rl - O + par.o_piece[j]);
                        if (rl >= OX) {
                            for (int j = 0; j < pieces; j++) row[3 * pieces + j] = (int16_t) (KVQ_T1X_K * j);
                            for (int j = 0; j < pieces; j++) row[4 * pieces + j] = (int16_t) (rl - OX + KVQ_T1X_K * j);
                        }
                    }
                    continue;
                }
                if (max_len >= rl) for (int j = 0; j < pieces; j++) add((int) ((int64_t) j * rl / pieces), KVQ_T2L && rl > short_max);
                if (rl > O) {
                    for (int j = 0; j < pieces; j++) add(par.o_piece[j], false);
                    for (int j = 0; j < pieces; j++) add(rl - O + par.o_piece[j], false);
                }
            }
        upload(d_p2, p2);
        dev.p2_tab = d_p2.as<int16_t>();
        // the uploads above are synchronous copies from pageable memory on the legacy stream; the scan
        // kernels run on non-blocking streams, so make sure every table has landed before they start
        CK(cudaDeviceSynchronize());
        cache_key.swap(pending_key);
    }
};

} // namespace

namespace {

// tiles of KVQ_TILE bytes inside every segment, in stream order: {start, segment end | first << 62}
void build_tiles(const std::vector<int64_t> &cuts, std::vector<longlong2> &tiles)
{
    tiles.clear();
    for (size_t sgi = 0; sgi + 1 < cuts.size(); sgi++)
        for (int64_t t0 = cuts[sgi]; t0 < cuts[sgi + 1]; t0 += KVQ_TILE) {
            longlong2 ti; ti.x = t0; ti.y = cuts[sgi + 1] | ((int64_t) (t0 == cuts[sgi]) << 62);
            tiles.push_back(ti);
        }
}

struct TileBufs {
    DevBuf d_info, d_status; int ntiles = 0;
    void upload_async(const std::vector<longlong2> &tiles, PinBuf &stage, cudaStream_t st) {
        ntiles = (int) tiles.size();
        const size_t n = std::max<size_t>(1, tiles.size());
        d_info.ensure(n * 16); d_status.ensure(n * 4);
        stage.ensure(n * 16);
        memcpy(stage.p, tiles.data(), tiles.size() * 16);
        CK(cudaMemcpyAsync(d_info.p, stage.p, tiles.size() * 16, cudaMemcpyHostToDevice, st));
    }
    void fill(KvqBatch &b) const {
        b.tile_info = d_info.as<longlong2>(); b.tile_status = d_status.as<uint32_t>(); b.ntiles = ntiles;
    }
};

} // namespace

// ------------------------------------------------------------------------------- engine objects
struct kvq_result {
    std::vector<kvq_hit_t> hits;
    std::vector<char> blob;
    std::vector<int64_t> offsets;
    kvq_stats_t stats{};
    std::vector<int64_t> nseqhits, nseqbasehits, coverage, mutcount;
    kvq_timing_t timing{};
    // HBM-resident scans hand the ordered hits out where the DMA put them: page-locked memory the result owns
    // (v_n >= 0: the accessors return these views instead of the vectors above)
    PinBuf pin; const kvq_hit_t *v_hits = nullptr; const int64_t *v_off = nullptr; const char *v_blob = nullptr; int64_t v_n = -1;
};

namespace {

// A freed result is kept for the next scan: its vectors keep their capacity, so a scan repeated
// on similar input does not page-fault fresh memory for every result array.
std::mutex g_spare_mu;
kvq_result *g_spare = nullptr;

kvq_result *take_result()
{
    std::lock_guard<std::mutex> g(g_spare_mu);
    kvq_result *r = g_spare; g_spare = nullptr;
    return r ? r : new kvq_result();
}

struct Slot {                           // one batch in flight
    HostBatch hb; PinBuf pin;
    DevBuf d_data, d_seg, d_ctr, d_hits, d_blob, d_longs, d_surv; size_t surv_cap = 0;
    PinBuf h_ctr, h_seg, h_tiles, h_live;
    TileBufs tiles;
    cudaStream_t stream = nullptr; cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_copied = nullptr;
    size_t hit_cap = 0, blob_cap = 0, long_cap = 0;
    int64_t data_bytes = 0;
};

struct Accum {                          // device-wide accumulators of one scan, one allocation:
    DevBuf d_all; size_t ns = 0, nc = 0;    // [nseqhits ns][nseqbasehits ns][coverage nc][mutcount 6 nc][tail]
    // tail (filled from the host totals right before a reduce, or by kvq_launch_tail): [readlengths 1024][records][bases]
    // [parsed][longest + 1][ranks whose batch was not complete when the row was reduced]
    static constexpr size_t TAIL = KVQ_HIST + 5;
    size_t row_words() const { return words() + TAIL; }
    unsigned long long *seqhits() const { return d_all.as<unsigned long long>(); }
    unsigned long long *seqbase() const { return seqhits() + ns; }
    unsigned long long *cov() const { return seqhits() + 2 * ns; }
    unsigned long long *mut() const { return seqhits() + 2 * ns + nc; }
    size_t words() const { return 2 * ns + 7 * nc; }
};

struct RawHit { KvqHitRec h; };

struct HitKey { int64_t fpos; uint64_t k2; uint32_t idx; };         // sort key of a hit (finish_result)
struct SortScratch {                    // reused from scan to scan: no fresh pages per result
    std::vector<HitKey> keys, sorted; std::vector<uint32_t> start, bucket, at;
    std::vector<KvqHitRec> hits; std::vector<char> blob;
};

} // namespace

struct kvq_engine {
    int device = 0, sm_count = 148;
    kvq_config_t cfg{0, 20, 10, 1, '!', '!'};      // workhorse.c:71-75
    std::mutex mu;                                  // guards cfg / running / error
    bool running = false;
    std::atomic<int> stop{0};
    int err_cls = 0; std::string err_msg;
    LiveStats live;
    Slot slots[NSLOT]; bool slots_ready = false;
    cudaStream_t copy_stream = nullptr;             // all batch uploads, back to back (concurrent H2D copies share the link badly)
    Accum acc;
    TemplateSet tpl;
    PinBuf pin_out, pin_acc, pin_ord, pin_ctr2;
    std::unique_ptr<WorkPool> pool; int pool_threads = 0;      // reader threads (plain files)
    // multi-GPU: one NCCL communicator per engine (one engine per process and GPU)
    ncclComm_t comm = nullptr; int rank = 0, world = 1;
    DevBuf d_job, d_max; int job_slots = 0; size_t job_row = 0; PinBuf pin_tail;
    // reduce behind the kernels (kvq_scan_staged_reduce): the collective runs on its own stream beside the ordering kernels
    cudaStream_t red_stream = nullptr; cudaEvent_t ev_ra = nullptr, ev_rb = nullptr, ev_r0 = nullptr, ev_r1 = nullptr; DevBuf d_red;               // page-locked staging of the result copies
    SortScratch scratch;

    void set_error(int cls, const std::string &m) { std::lock_guard<std::mutex> g(mu); err_cls = cls; err_msg = m; }
};

struct kvq_staged {
    kvq_engine *e = nullptr;
    DevBuf d_data, d_seg, d_ctr, d_hits, d_blob, d_longs, d_surv; size_t surv_cap = 0;
    TileBufs tiles;
    size_t hit_cap = 0, blob_cap = 0, long_cap = 0;
    // hits put into the reference's order on the device (KvqOrder)
    DevBuf o_keys, o_bkt, o_start, o_cursor, o_members, o_lens, o_src, o_hits, o_offsets, o_blob, o_status;
    int64_t nbytes = 0, base_fpos = 0, parsed = 0, total = 0; int nseg = 0;
};

namespace {

// survivor-queue records to start with (KVQ_SURV_CAP: tests of the overflow re-run)
size_t initial_surv_cap()
{
    const char *v = getenv("KVQ_SURV_CAP");
    const long long n = v ? atoll(v) : 0;
    return n > 0 ? (size_t) n : (size_t) 1 << 20;
}

void init_slots(kvq_engine *e)
{
    if (e->slots_ready) return;
    for (auto &s : e->slots) {
        s.pin.ensure(BATCH_BYTES + BATCH_PAD);
        s.hb.own = s.pin.as<char>(); s.hb.data = s.hb.own;
        s.d_data.ensure(BATCH_BYTES + DEV_PAD);
        const size_t nseg_max = (size_t) (BATCH_BYTES / 4096) + 64;     // cuts are >= 1 byte apart; see launch
        s.d_seg.ensure(nseg_max * sizeof(int64_t)); s.h_seg.ensure(nseg_max * sizeof(int64_t));
        s.d_ctr.ensure(sizeof(KvqCounters)); s.h_ctr.ensure(sizeof(KvqCounters));
        s.long_cap = (size_t) (BATCH_BYTES / 1024) + 64;
        s.d_longs.ensure(s.long_cap * sizeof(KvqLongRec));
        s.surv_cap = initial_surv_cap(); s.d_surv.ensure(s.surv_cap * sizeof(KvqSurvRec));
        s.hit_cap = 1u << 18; s.d_hits.ensure(s.hit_cap * sizeof(KvqHitRec));
        s.blob_cap = 32u << 20; s.d_blob.ensure(s.blob_cap);
        CK(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
        CK(cudaEventCreate(&s.ev0)); CK(cudaEventCreate(&s.ev1));
        CK(cudaEventCreateWithFlags(&s.ev_copied, cudaEventDisableTiming));
    }
    CK(cudaStreamCreateWithFlags(&e->copy_stream, cudaStreamNonBlocking));
    e->slots_ready = true;
}

struct ScanTotals {
    std::vector<KvqHitRec> hits; std::vector<char> blob;
    int64_t rls[KVQ_HIST]; int64_t longest = -1, records = 0, bases = 0, exh = 0, cands = 0;
    unsigned long long err_key = ~0ULL;
    double kernel_ms = 0; int64_t launches = 0, h2d = 0, d2h = 0, stream_bytes = 0;
    bool acc_fetched = false;           // per-template accumulators already in e->pin_acc (copied behind the kernels)
    size_t spec_hits = 0, spec_blob = 0; // hit records / hitseq bytes already in e->pin_out (copied behind the kernels)
    // hits already in the reference's order (ordered on the device): page-locked staging, see kvq_scan_staged
    const KvqHitOut *s_hits = nullptr; const long long *s_off = nullptr; const char *s_blob = nullptr; size_t s_n = 0, s_nb = 0;
    kvq_result *pre = nullptr;          // result taken before the scan: its page-locked buffer is the target of the ordered-hit copies
    ~ScanTotals() { if (pre) kvq_result_free(pre); }
    ScanTotals() { memset(rls, 0, sizeof(rls)); }
};

// device buffers of one batch in flight (a Slot's or a staged stream's)
struct OutBufs {
    DevBuf &ctr, &hits; size_t &hit_cap; DevBuf &blob; size_t &blob_cap; DevBuf &longs; size_t long_cap;
    DevBuf &surv; size_t &surv_cap;
};
template <class S> OutBufs bufs_of(S &s)
{
    return OutBufs{s.d_ctr, s.d_hits, s.hit_cap, s.d_blob, s.blob_cap, s.d_longs, s.long_cap, s.d_surv, s.surv_cap};
}

KvqOut make_out(kvq_engine *e, const OutBufs &ob)
{
    DevBuf &ctr = ob.ctr, &hits = ob.hits, &blob = ob.blob, &longs = ob.longs;
    const size_t hit_cap = ob.hit_cap, blob_cap = ob.blob_cap, long_cap = ob.long_cap;
    KvqOut o{};
    o.ctr = ctr.as<KvqCounters>();
    o.hits = hits.as<KvqHitRec>(); o.hit_cap = hit_cap;
    o.blob = blob.as<uint8_t>(); o.blob_cap = blob_cap;
    o.longs = longs.as<KvqLongRec>(); o.long_cap = long_cap;
    o.surv = ob.surv.as<KvqSurvRec>(); o.surv_cap = ob.surv_cap;
    o.nseqhits = e->acc.seqhits();
    o.nseqbasehits = e->acc.seqbase();
    o.coverage = e->acc.cov();
    o.mutcount = e->acc.mut();
    return o;
}

// Zeroes the accumulators on `st`: the stream the scan kernels run on, or one they wait for
// (the slot streams of a streaming scan wait for the copy stream's events).
void reset_accum(kvq_engine *e, cudaStream_t st)
{
    const TemplateSet &t = e->tpl;
    e->acc.ns = (size_t) std::max(1, t.nseqs); e->acc.nc = (size_t) std::max<int64_t>(1, t.cov_total);
    e->acc.d_all.ensure(e->acc.row_words() * 8);
    CK(cudaMemsetAsync(e->acc.d_all.p, 0, e->acc.row_words() * 8, st));
}

// enqueue memset + scan + emit + long-record kernels for one batch already resident on the device
void enqueue_kernels(kvq_engine *e, const KvqBatch &b, const KvqParams &p, const KvqOut &o,
                     cudaStream_t st, cudaEvent_t ev0, cudaEvent_t ev1)
{
    cuda_check((cudaError_t) kvq_launch_reset(o.ctr, b.tile_status, b.ntiles, st), "reset kernel launch");
    if (ev0) CK(cudaEventRecord(ev0, st));
    cuda_check((cudaError_t) kvq_launch_scan(b, e->tpl.dev, p, o, e->sm_count, st), "scan kernel launch");
    cuda_check((cudaError_t) kvq_launch_emit(b, e->tpl.dev, p, o, e->sm_count, st), "emit kernel launch");
    cuda_check((cudaError_t) kvq_launch_long(b, e->tpl.dev, p, o, e->sm_count, st), "long-record kernel launch");
    if (ev1) CK(cudaEventRecord(ev1, st));
}

void launch_slot(kvq_engine *e, Slot &s, ScanTotals &tot)
{
    const HostBatch &hb = s.hb;
    const int nseg = (int) hb.cuts.size() - 1;
    s.d_seg.ensure(hb.cuts.size() * 8); s.h_seg.ensure(hb.cuts.size() * 8);
    memcpy(s.h_seg.p, hb.cuts.data(), hb.cuts.size() * 8);
    CK(cudaMemcpyAsync(s.d_data.p, hb.data, hb.nbytes, cudaMemcpyHostToDevice, e->copy_stream));
    CK(cudaMemsetAsync(s.d_data.as<char>() + hb.nbytes, 0, DEV_PAD, e->copy_stream));
    CK(cudaEventRecord(s.ev_copied, e->copy_stream));
    CK(cudaStreamWaitEvent(s.stream, s.ev_copied, 0));
    CK(cudaMemcpyAsync(s.d_seg.p, s.h_seg.p, hb.cuts.size() * 8, cudaMemcpyHostToDevice, s.stream));
    std::vector<longlong2> tiles;
    build_tiles(hb.cuts, tiles);
    s.tiles.upload_async(tiles, s.h_tiles, s.stream);
    KvqBatch b{s.d_data.as<uint8_t>(), hb.nbytes, s.d_seg.as<int64_t>(), nseg, hb.base_fpos};
    s.tiles.fill(b);
    KvqOut o = make_out(e, bufs_of(s));
    enqueue_kernels(e, b, e->tpl.par, o, s.stream, s.ev0, s.ev1);
    CK(cudaMemcpyAsync(s.h_ctr.p, s.d_ctr.p, sizeof(KvqCounters), cudaMemcpyDeviceToHost, s.stream));
    s.h_live.ensure(2 * e->acc.ns * 8);
    CK(cudaMemcpyAsync(s.h_live.p, e->acc.d_all.p, 2 * e->acc.ns * 8, cudaMemcpyDeviceToHost, s.stream));
    tot.launches += 4; tot.h2d += hb.nbytes + (int64_t) hb.cuts.size() * 8; tot.d2h += sizeof(KvqCounters);
    tot.stream_bytes += hb.nbytes;
}

// The survivor queue of the scan kernel overflowed, or a tile's inferred line phase was
// contradicted (malformed input): the emit and long-record kernels did nothing.  Run the batch
// again with a queue that fits (its size is known now) / with exact line phases and take the
// counters of that run.
KvqParams rerun_for_survivors(kvq_engine *e, const KvqBatch &b, const OutBufs &ob, KvqCounters &c, PinBuf &h_ctr,
                              cudaStream_t st, ScanTotals &tot)
{
    KvqParams p = e->tpl.par;
    for (int it = 0; it < 4 && (c.surv_count > ob.surv_cap || c.phase_retry); it++) {
        if (c.phase_retry) p.exact_phase = 1;
        if (c.surv_count > ob.surv_cap) {
            ob.surv_cap = (size_t) (c.surv_count + c.surv_count / 8 + 1024);
            ob.surv.ensure(ob.surv_cap * sizeof(KvqSurvRec));
        }
        enqueue_kernels(e, b, p, make_out(e, ob), st, nullptr, nullptr);
        CK(cudaMemcpyAsync(h_ctr.p, ob.ctr.p, sizeof(KvqCounters), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        c = *h_ctr.as<KvqCounters>();
        tot.launches += 4; tot.acc_fetched = false; tot.spec_hits = tot.spec_blob = 0;
    }
    if (c.surv_count > ob.surv_cap || c.phase_retry)
        throw KvqError(KVQ_ERR_RUNTIME, "internal: batch still incomplete after re-running it (survivor queue / line phases)");
    return p;                               // the parameters the counters in `c` were produced with
}

// Fetch hits of a finished batch (re-running it with larger buffers when they overflowed).
void collect_hits(kvq_engine *e, const KvqBatch &b, const OutBufs &ob, const KvqParams &used,
                  const KvqCounters &c, cudaStream_t st, ScanTotals &tot)
{
    DevBuf &d_hits = ob.hits, &d_blob = ob.blob;
    size_t &hit_cap = ob.hit_cap, &blob_cap = ob.blob_cap;
    unsigned long long nh = c.hit_count, nb = c.blob_used;
    if (nh > hit_cap || nb > blob_cap) {
        hit_cap = (size_t) (nh + nh / 4 + 1024); blob_cap = (size_t) (nb + nb / 4 + 4096);
        d_hits.ensure(hit_cap * sizeof(KvqHitRec)); d_blob.ensure(blob_cap);
        KvqParams p = used; p.accumulate = 0;   // (keeps exact_phase when the batch needed it)
        KvqOut o = make_out(e, ob);
        enqueue_kernels(e, b, p, o, st, nullptr, nullptr);
        e->pin_ctr2.ensure(sizeof(KvqCounters));
        CK(cudaMemcpyAsync(e->pin_ctr2.p, ob.ctr.p, sizeof(KvqCounters), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        {
            const KvqCounters &c2 = *e->pin_ctr2.as<KvqCounters>();
            if (c2.phase_retry || c2.surv_count > ob.surv_cap || c2.hit_count != nh)
                throw KvqError(KVQ_ERR_RUNTIME, "internal: re-run with larger hit buffers did not reproduce the batch");
        }
        tot.launches += 4; tot.spec_hits = tot.spec_blob = 0;
    }
    if (nh == 0) return;
    const size_t h0 = tot.hits.size(), b0 = tot.blob.size();
    tot.hits.resize(h0 + nh); tot.blob.resize(b0 + nb);
    // D2H through page-locked staging (DMA at full rate), then a plain copy into the result vectors
    const size_t hbytes = nh * sizeof(KvqHitRec);
    char *ph, *pb;
    if (nh <= tot.spec_hits && nb <= tot.spec_blob) {       // already copied behind the kernels
        ph = e->pin_out.as<char>(); pb = ph + ((tot.spec_hits * sizeof(KvqHitRec) + 255) & ~(size_t) 255);
    } else {
        e->pin_out.ensure(((hbytes + 255) & ~(size_t) 255) + nb + 256);
        ph = e->pin_out.as<char>(); pb = ph + ((hbytes + 255) & ~(size_t) 255);
        CK(cudaMemcpyAsync(ph, d_hits.p, hbytes, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(pb, d_blob.p, nb, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
    }
    memcpy(tot.hits.data() + h0, ph, hbytes); memcpy(tot.blob.data() + b0, pb, nb);
    for (size_t i = h0; i < h0 + nh; i++) tot.hits[i].blob_off += (int64_t) b0;
    tot.d2h += (int64_t) (nh * sizeof(KvqHitRec) + nb);
}

void add_counters(kvq_engine *e, const KvqCounters &c, ScanTotals &tot)
{
    for (int i = 0; i < KVQ_HIST; i++) if (c.rls[i]) { tot.rls[i] += (int64_t) c.rls[i]; e->live.rls[i] += (int64_t) c.rls[i]; }
    tot.longest = std::max<int64_t>(tot.longest, c.longest);
    if (c.longest > e->live.longest) e->live.longest = c.longest;
    tot.records += (int64_t) c.records; e->live.records += (int64_t) c.records;
    tot.bases += (int64_t) c.bases; tot.exh += (int64_t) c.exhaustive_reads; tot.cands += (int64_t) c.candidates;
    tot.err_key = std::min(tot.err_key, c.err_key);
}

void harvest_slot(kvq_engine *e, Slot &s, ScanTotals &tot)
{
    CK(cudaStreamSynchronize(s.stream));
    float ms = 0; CK(cudaEventElapsedTime(&ms, s.ev0, s.ev1)); tot.kernel_ms += ms;
    KvqCounters c = *s.h_ctr.as<KvqCounters>();
    KvqBatch b{s.d_data.as<uint8_t>(), s.hb.nbytes, s.d_seg.as<int64_t>(), (int) s.hb.cuts.size() - 1, s.hb.base_fpos};
    s.tiles.fill(b);
    const KvqParams used = rerun_for_survivors(e, b, bufs_of(s), c, s.h_ctr, s.stream, tot);
    add_counters(e, c, tot);
    if (e->tpl.par.want_hits)
        collect_hits(e, b, bufs_of(s), used, c, s.stream, tot);
    // live per-template counters for engine.stats(): copied behind the batch's kernels (launch_slot)
    if (e->tpl.nseqs > 0) {
        const unsigned long long *a = s.h_live.as<unsigned long long>();
        for (int i = 0; i < e->tpl.nseqs; i++) { e->live.seqhits[i] = (int64_t) a[i]; e->live.seqbasehits[i] = (int64_t) a[e->acc.ns + i]; }
    }
}

} // namespace

namespace {

template <class T> class Chan {        // tiny blocking queue
public:
    void push(T v) { { std::lock_guard<std::mutex> g(mu_); q_.push_back(v); } cv_.notify_one(); }
    T pop() { std::unique_lock<std::mutex> g(mu_); cv_.wait(g, [&] { return !q_.empty(); }); T v = q_.front(); q_.pop_front(); return v; }
private:
    std::mutex mu_; std::condition_variable cv_; std::deque<T> q_;
};

// reference emission order for nthreads=1 (workhorse.c:1107-1175): file order, then template,
// then (a) shifts descending, (b) shifts descending, (c)/(d) shifts ascending
std::string format_error_message(unsigned long long key)
{
    // see report_format_error() in scan_kernel.cu; texts are workhorse.c:1039-1046
    const long long rec = (long long) (key >> 22); const int type = (int) ((key >> 21) & 1);
    const unsigned payload = (unsigned) (key & 0x1FFFFF);
    char buf[256];
    if (type == 0)
        snprintf(buf, sizeof(buf), "record must start with '@' (and not '%c') fpos=%ld", (char) payload, (long) rec);
    else
        snprintf(buf, sizeof(buf), "3rd line of record must start with '+' fpos=%ld", (long) (rec + payload));
    return buf;
}

void fill_stats(kvq_engine *e, kvq_stats_t &st, const int64_t *rls, int64_t longest, int64_t records)
{
    memset(&st, 0, sizeof(st));
    for (int i = 0; i < KVQ_MAX_READLENGTH; i++) st.readlengths[i] = rls[i];
    st.rls_longest = longest; st.records_parsed = records;
    st.parsed = e->live.parsed; st.total = e->live.total; st.sigints = e->live.sigints;
    st.progress = 0;
    if (st.total > 0)                                                       // workhorse.c:1230-1232
        st.progress = ((float) (uint64_t) std::min(st.parsed, st.total)) / (float) (uint64_t) st.total;
    st.nseqs = e->tpl.nseqs;
}

kvq_result *finish_result(kvq_engine *e, ScanTotals &tot, double wall_ms)
{
    const double t_fin = now_ms();
    std::unique_ptr<kvq_result> r(tot.pre ? tot.pre : take_result());
    tot.pre = nullptr;
    // reference emission order: sort 24-byte keys, not the records
    const size_t nh = tot.hits.size();
    typedef HitKey Key;
    SortScratch &sc = e->scratch;
    std::vector<Key> &keys = sc.keys;
    keys.resize(nh);
    for (size_t i = 0; i < nh; i++) {
        const KvqHitRec &h = tot.hits[i];
        const int o = h.rule == 0 ? h.seq_pos : h.rule == 1 ? -h.seq_pos : std::abs(h.seq_pos);
        keys[i] = Key{h.file_pos, ((uint64_t) (uint32_t) h.seq_nr << 34) | ((uint64_t) (uint32_t) h.rule << 32)
                                      | (uint64_t) (uint32_t) (o + (1 << 30)), (uint32_t) i};
    }
    auto before = [](const Key &a, const Key &b) {
        return a.fpos != b.fpos ? a.fpos < b.fpos : a.k2 != b.k2 ? a.k2 < b.k2 : a.idx < b.idx; };
    if (nh > 64) {
        // hits spread evenly over the stream: one counting pass into nh buckets by file position,
        // then a sort inside the (tiny) buckets
        int64_t lo = keys[0].fpos, hi = keys[0].fpos;
        for (const Key &k : keys) { lo = std::min(lo, k.fpos); hi = std::max(hi, k.fpos); }
        const double scale = (double) nh / (double) (hi - lo + 1);       // monotonic is all that is needed
        std::vector<uint32_t> &start = sc.start, &bucket = sc.bucket, &at = sc.at;
        start.assign(nh + 1, 0); bucket.resize(nh);
        for (size_t i = 0; i < nh; i++) {
            bucket[i] = std::min<uint32_t>((uint32_t) nh - 1, (uint32_t) ((double) (keys[i].fpos - lo) * scale));
            start[bucket[i] + 1]++;
        }
        for (size_t i = 0; i < nh; i++) start[i + 1] += start[i];
        std::vector<Key> &sorted = sc.sorted;
        sorted.resize(nh);
        at.assign(start.begin(), start.end() - 1);
        for (size_t i = 0; i < nh; i++) sorted[at[bucket[i]]++] = keys[i];
        for (size_t bk = 0; bk < nh; bk++)
            if (start[bk + 1] - start[bk] > 1) std::sort(sorted.begin() + start[bk], sorted.begin() + start[bk + 1], before);
        keys.swap(sorted);
    } else {
        std::sort(keys.begin(), keys.end(), before);
    }
    const double t_sorted = now_ms();
    static_assert(sizeof(KvqHitOut) == sizeof(kvq_hit_t) && sizeof(kvq_hit_t) == 24, "device hit layout == kvq_hit_t");
    if (tot.s_hits && r->pin.p && (const char *) tot.s_hits >= r->pin.as<char>() && (const char *) tot.s_hits < r->pin.as<char>() + r->pin.cap) {
        // ordered on the device and copied into the result's own page-locked buffer: nothing to move
        r->v_hits = reinterpret_cast<const kvq_hit_t *>(tot.s_hits); r->v_off = reinterpret_cast<const int64_t *>(tot.s_off);
        r->v_blob = tot.s_blob; r->v_n = (int64_t) tot.s_n;
    } else if (tot.s_hits) {            // ordered on the device: three plain copies
        r->hits.resize(tot.s_n); r->offsets.resize(tot.s_n + 1); r->blob.resize(tot.s_nb);
        if (tot.s_n) memcpy(r->hits.data(), tot.s_hits, tot.s_n * sizeof(kvq_hit_t));
        memcpy(r->offsets.data(), tot.s_off, (tot.s_n + 1) * sizeof(int64_t));
        if (tot.s_nb) memcpy(r->blob.data(), tot.s_blob, tot.s_nb);
    } else {
    r->hits.resize(nh); r->offsets.resize(nh + 1); r->blob.resize(tot.blob.size());
    int64_t at = 0;
    for (size_t i = 0; i < nh; i++) {
        const KvqHitRec &h = tot.hits[keys[i].idx];
        r->hits[i] = kvq_hit_t{h.seq_nr, h.seq_pos, h.file_pos, h.length, h.readlength};
        r->offsets[i] = at;
        memcpy(r->blob.data() + at, tot.blob.data() + h.blob_off, h.length);
        at += h.length;
    }
    r->offsets[nh] = at;
    }
    const double t_ordered = now_ms();
    fill_stats(e, r->stats, tot.rls, tot.longest, tot.records);
    const int ns = e->tpl.nseqs; const int64_t nc = e->tpl.cov_total;
    r->nseqhits.assign(ns, 0); r->nseqbasehits.assign(ns, 0);
    {
        const bool want_cov = e->tpl.nplus > 0 && nc > 0;
        const size_t words = want_cov ? e->acc.words() : 2 * e->acc.ns;
        if (!tot.acc_fetched) {
            e->pin_acc.ensure(words * 8);
            CK(cudaMemcpy(e->pin_acc.p, e->acc.d_all.p, words * 8, cudaMemcpyDeviceToHost));
        }
        const int64_t *all = e->pin_acc.as<int64_t>();
        for (int i = 0; i < ns; i++) {
            r->nseqhits[i] = all[i]; r->nseqbasehits[i] = all[e->acc.ns + i];
            e->live.seqhits[i] = all[i]; e->live.seqbasehits[i] = all[e->acc.ns + i];
        }
        if (want_cov) {
            r->coverage.assign(all + 2 * e->acc.ns, all + 2 * e->acc.ns + nc);
            r->mutcount.assign(all + 2 * e->acc.ns + e->acc.nc, all + 2 * e->acc.ns + e->acc.nc + 6 * nc);
            tot.d2h += nc * 56;
        }
    }
    tot.d2h += (int64_t) ns * 16;
    if (getenv("KVQ_TRACE"))
        fprintf(stderr, "[kvq trace] finish: sort %.3f reorder %.3f stats+accumulators %.3f\n",
                t_sorted - t_fin, t_ordered - t_sorted, now_ms() - t_ordered);
    tot.hits.clear(); tot.blob.clear(); sc.hits.swap(tot.hits); sc.blob.swap(tot.blob);     // capacity kept for the next scan
    r->timing.kernel_ms = tot.kernel_ms; r->timing.wall_ms = wall_ms + (now_ms() - t_fin); r->timing.launches = tot.launches;
    r->timing.h2d_bytes = tot.h2d; r->timing.d2h_bytes = tot.d2h; r->timing.bases = tot.bases;
    r->timing.stream_bytes = tot.stream_bytes; r->timing.exhaustive_reads = tot.exh; r->timing.candidates = tot.cands;
    return r.release();
}

// The whole streaming scan: a producer thread cuts the stream into pinned batches while this
// thread keeps up to three batches in flight on the GPU.
kvq_result *run_stream_scan(kvq_engine *e, std::vector<std::string> names,
                            std::vector<std::pair<const char *, int64_t>> mems)
{
    const double t_begin = now_ms();
    const bool trace_stream = getenv("KVQ_TRACE") != nullptr;
    init_slots(e);
    reset_accum(e, e->copy_stream);         // every slot stream waits for an event recorded later on this stream
    {   // reader threads: engine.config(nthreads), at least min(8, cores); KVQ_READ_THREADS overrides
        kvq_config_t cfg; { std::lock_guard<std::mutex> g(e->mu); cfg = e->cfg; }
        int want = std::max<int>(cfg.nthreads, (int) std::min<unsigned>(8u, std::max(1u, std::thread::hardware_concurrency())));
        if (const char *v = getenv("KVQ_READ_THREADS")) want = atoi(v);
        want = std::max(1, std::min(want, 64));
        if (!names.empty() && (!e->pool || e->pool_threads != want)) { e->pool.reset(new WorkPool(want - 1)); e->pool_threads = want; }
    }
    int inflate_threads;
    {   // inflate threads: engine.config(nthreads) when set (> 1), else the machine's hardware threads -- but two of those are
        // left to the thread that copies the inflated bytes into the batches and to the one that drives the GPU (measured,
        // config 3 on a 16-thread host: 16 inflate threads 1.88 Gbp/s, 14: 2.07); KVQ_GZ_THREADS overrides (1: one zlib
        // stream per file)
        kvq_config_t cfg; { std::lock_guard<std::mutex> g(e->mu); cfg = e->cfg; }
        const int hw = (int) std::max(1u, std::thread::hardware_concurrency());
        inflate_threads = cfg.nthreads > 1 ? cfg.nthreads : hw;
        if (hw >= 8) inflate_threads = std::min(inflate_threads, hw - 2);
        if (const char *v = getenv("KVQ_GZ_THREADS")) inflate_threads = atoi(v);
        inflate_threads = std::max(1, std::min(inflate_threads, 128));
    }
    StreamCutter cutter(std::move(names), std::move(mems), &e->live, e->pool.get(), inflate_threads);
    ScanTotals tot;
    Chan<int> free_q, ready_q;
    for (int i = 0; i < NSLOT; i++) free_q.push(i);
    std::unique_ptr<KvqError> producer_err;
    std::atomic<int> abort_flag{0};
    std::thread producer([&] {
        try {
            for (;;) {
                int si = free_q.pop();
                if (si < 0 || abort_flag.load()) break;
                if (!cutter.next_batch(e->slots[si].hb, e->stop)) break;
                ready_q.push(si);
            }
        } catch (const KvqError &err) { producer_err.reset(new KvqError(err)); }
        catch (const std::exception &err) { producer_err.reset(new KvqError(KVQ_ERR_RUNTIME, err.what())); }
        ready_q.push(-1);
    });
    std::unique_ptr<KvqError> consumer_err;
    std::deque<int> inflight;
    try {
        for (;;) {
            const double tq0 = now_ms();
            int si = ready_q.pop();
            if (si < 0) break;
            const double tq1 = now_ms();
            launch_slot(e, e->slots[si], tot);
            const double tq2 = now_ms();
            inflight.push_back(si);
            if ((int) inflight.size() >= NSLOT - 1) {      // keep the copy engine fed: two H2D copies queued behind the running one
                int done = inflight.front(); inflight.pop_front();
                harvest_slot(e, e->slots[done], tot);
                if (trace_stream)
                    fprintf(stderr, "[kvq trace] consumer t=%.3f: wait-ready %.3f launch %.3f harvest %.3f\n",
                            fmod(tq0, 1e6), tq1 - tq0, tq2 - tq1, now_ms() - tq2);
                free_q.push(done);
                if (tot.err_key != ~0ULL) e->stop = 1;          // first format error ends the scan
            }
        }
        while (!inflight.empty()) {
            int done = inflight.front(); inflight.pop_front();
            harvest_slot(e, e->slots[done], tot);
        }
    } catch (const KvqError &err) {
        consumer_err.reset(new KvqError(err));
        abort_flag = 1; e->stop = 1;
        cudaDeviceSynchronize();
        for (int i = 0; i < NSLOT + 1; i++) free_q.push(-1);
        // drain so the producer can finish
        std::thread drain([&] { for (;;) { if (ready_q.pop() < 0) break; } });
        producer.join(); drain.join();
        throw *consumer_err;
    }
    free_q.push(-1);
    producer.join();
    // a format error lies in data already scanned, i.e. before whatever stopped the reader
    if (tot.err_key != ~0ULL) throw KvqError(KVQ_ERR_FASTQ_FORMAT, format_error_message(tot.err_key));
    if (producer_err) throw *producer_err;
    if (trace_stream) fprintf(stderr, "[kvq trace] stream scan: begin t=%.1f, batches done t=%.1f\n", fmod(t_begin, 1e6), fmod(now_ms(), 1e6));
    kvq_result *res = finish_result(e, tot, now_ms() - t_begin);
    if (trace_stream) fprintf(stderr, "[kvq trace] stream scan: result t=%.1f\n", fmod(now_ms(), 1e6));
    return res;
}

} // namespace

// ------------------------------------------------------------------------------- C ABI
namespace {

struct RunGuard {                       // `running` flag of workhorse.c:1258-1265
    kvq_engine *e; bool owns = false;
    explicit RunGuard(kvq_engine *e_) : e(e_) {
        std::lock_guard<std::mutex> g(e->mu);
        if (e->running) return;
        e->running = true; owns = true;
    }
    ~RunGuard() { if (owns) { std::lock_guard<std::mutex> g(e->mu); e->running = false; } }
};

template <class F> int guarded(kvq_engine *e, F &&body)
{
    if (!e) return KVQ_ERR_TYPE;
    try { body(); return KVQ_OK; }
    catch (const KvqError &err) { e->set_error(err.cls, err.msg); return err.cls; }
    catch (const std::bad_alloc &) { e->set_error(KVQ_ERR_MEMORY, "out of memory"); return KVQ_ERR_MEMORY; }
    catch (const std::exception &err) { e->set_error(KVQ_ERR_RUNTIME, err.what()); return KVQ_ERR_RUNTIME; }
}

int scan_entry(kvq_engine *e, std::vector<std::string> names, std::vector<std::pair<const char *, int64_t>> mems,
               const char *const *seqs, const int32_t *seqlens, int nseqs, int nplus, kvq_result **out)
{
    if (out) *out = nullptr;
    return guarded(e, [&] {
        if (!out) throw KvqError(KVQ_ERR_TYPE, "result pointer is NULL");
        RunGuard run(e);
        if (!run.owns) throw KvqError(KVQ_ERR_RUNTIME, "findseqs() already running!");
        e->stop = 0;
        CK(cudaSetDevice(e->device));
        kvq_config_t cfg;
        { std::lock_guard<std::mutex> g(e->mu); cfg = e->cfg; }
        { std::lock_guard<std::mutex> g(e->mu); e->live.reset(nseqs); }
        e->tpl.build(seqs, seqlens, nseqs, nplus, cfg);
        *out = run_stream_scan(e, std::move(names), std::move(mems));
    });
}

} // namespace

extern "C" {

const char *kvq_version(void) { return "kvarq_b200 0.1 (sm_100a)"; }

int kvq_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int kvq_create(kvq_engine **out, int device)
{
    if (!out) return KVQ_ERR_TYPE;
    *out = nullptr;
    int n = kvq_device_count();
    if (n <= 0 || device < 0 || device >= n) return KVQ_ERR_RUNTIME;     // no CPU fallback
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return KVQ_ERR_RUNTIME;
    if (prop.major < 10) return KVQ_ERR_RUNTIME;                         // kernels are sm_100a only
    kvq_engine *e = new kvq_engine();
    e->device = device; e->sm_count = prop.multiProcessorCount;
    *out = e;
    return KVQ_OK;
}

void kvq_destroy(kvq_engine *e)
{
    if (!e) return;
    cudaSetDevice(e->device);
    for (auto &s : e->slots) {
        if (s.stream) cudaStreamDestroy(s.stream);
        if (s.ev0) cudaEventDestroy(s.ev0);
        if (s.ev1) cudaEventDestroy(s.ev1);
        if (s.ev_copied) cudaEventDestroy(s.ev_copied);
    }
    if (e->copy_stream) cudaStreamDestroy(e->copy_stream);
    if (e->red_stream) cudaStreamDestroy(e->red_stream);
    for (cudaEvent_t ev : {e->ev_ra, e->ev_rb, e->ev_r0, e->ev_r1}) if (ev) cudaEventDestroy(ev);
    if (e->comm && kvq_nccl::api().CommDestroy) kvq_nccl::api().CommDestroy(e->comm);
    delete e;
}

int kvq_config(kvq_engine *e, const kvq_config_t *cfg)
{
    if (!e || !cfg) return KVQ_ERR_TYPE;
    std::lock_guard<std::mutex> g(e->mu);
    e->cfg = *cfg;
    return KVQ_OK;
}

int kvq_get_config(kvq_engine *e, kvq_config_t *cfg)
{
    if (!e || !cfg) return KVQ_ERR_TYPE;
    std::lock_guard<std::mutex> g(e->mu);
    *cfg = e->cfg;
    return KVQ_OK;
}

int kvq_findseqs(kvq_engine *e, const char *const *fnames, int nfiles, const char *const *seqs,
                 const int32_t *seqlens, int nseqs, int nplus, kvq_result **out)
{
    std::vector<std::string> names;
    if (nfiles < 0 || (nfiles > 0 && !fnames)) { if (e) e->set_error(KVQ_ERR_TYPE, "fname must be [sequence of] string[s]"); return KVQ_ERR_TYPE; }
    for (int i = 0; i < nfiles; i++) {
        if (!fnames[i]) { e->set_error(KVQ_ERR_TYPE, "fname must be [sequence of] string[s]"); return KVQ_ERR_TYPE; }
        names.push_back(fnames[i]);
    }
    return scan_entry(e, std::move(names), {}, seqs, seqlens, nseqs, nplus, out);
}

int kvq_findseqs_mem(kvq_engine *e, const char *const *bufs, const int64_t *buflens, int nbufs,
                     const char *const *seqs, const int32_t *seqlens, int nseqs, int nplus, kvq_result **out)
{
    std::vector<std::pair<const char *, int64_t>> mems;
    if (nbufs < 0 || (nbufs > 0 && (!bufs || !buflens))) { if (e) e->set_error(KVQ_ERR_TYPE, "bad buffers"); return KVQ_ERR_TYPE; }
    for (int i = 0; i < nbufs; i++) mems.push_back({bufs[i], buflens[i]});
    return scan_entry(e, {}, std::move(mems), seqs, seqlens, nseqs, nplus, out);
}

int64_t kvq_result_nhits(const kvq_result *r) { return !r ? 0 : r->v_n >= 0 ? r->v_n : (int64_t) r->hits.size(); }
const kvq_hit_t *kvq_result_hits(const kvq_result *r) { return !r ? nullptr : r->v_n >= 0 ? r->v_hits : r->hits.data(); }
const char *kvq_result_hitseq_blob(const kvq_result *r) { return !r ? nullptr : r->v_n >= 0 ? r->v_blob : r->blob.data(); }
const int64_t *kvq_result_hitseq_offsets(const kvq_result *r) { return !r ? nullptr : r->v_n >= 0 ? r->v_off : r->offsets.data(); }
const kvq_stats_t *kvq_result_stats(const kvq_result *r) { return r ? &r->stats : nullptr; }
const int64_t *kvq_result_nseqhits(const kvq_result *r) { return r ? r->nseqhits.data() : nullptr; }
const int64_t *kvq_result_nseqbasehits(const kvq_result *r) { return r ? r->nseqbasehits.data() : nullptr; }
const int64_t *kvq_result_coverage(const kvq_result *r, int64_t *n)
{
    if (n) *n = r ? (int64_t) r->coverage.size() : 0;
    return r && !r->coverage.empty() ? r->coverage.data() : nullptr;
}
const int64_t *kvq_result_mutcount(const kvq_result *r, int64_t *n)
{
    if (n) *n = r ? (int64_t) r->mutcount.size() : 0;
    return r && !r->mutcount.empty() ? r->mutcount.data() : nullptr;
}
const kvq_timing_t *kvq_result_timing(const kvq_result *r) { return r ? &r->timing : nullptr; }
void kvq_result_free(kvq_result *r)
{
    if (!r) return;
    {
        std::lock_guard<std::mutex> g(g_spare_mu);
        if (!g_spare) {
            r->hits.clear(); r->blob.clear(); r->offsets.clear();
            r->nseqhits.clear(); r->nseqbasehits.clear(); r->coverage.clear(); r->mutcount.clear();
            r->stats = kvq_stats_t{}; r->timing = kvq_timing_t{};
            r->v_n = -1; r->v_hits = nullptr; r->v_off = nullptr; r->v_blob = nullptr;
            g_spare = r;
            return;
        }
    }
    delete r;
}

int kvq_stats_snapshot(kvq_engine *e, kvq_stats_t *out, int64_t *nseqhits, int64_t *nseqbasehits, int cap)
{
    if (!e || !out) return KVQ_ERR_TYPE;
    std::lock_guard<std::mutex> g(e->mu);
    int64_t rls[KVQ_MAX_READLENGTH];
    for (int i = 0; i < KVQ_MAX_READLENGTH; i++) rls[i] = e->live.rls[i];
    fill_stats(e, *out, rls, e->live.longest, e->live.records);
    const int n = (int) e->live.seqhits.size();
    out->nseqs = n;
    for (int i = 0; i < n && i < cap; i++) {
        if (nseqhits) nseqhits[i] = e->live.seqhits[i];
        if (nseqbasehits) nseqbasehits[i] = e->live.seqbasehits[i];
    }
    return KVQ_OK;
}

int kvq_stop(kvq_engine *e)
{
    if (!e) return KVQ_ERR_TYPE;
    e->stop++;                                                           // workhorse.c:1476
    return KVQ_OK;
}

const char *kvq_last_error(kvq_engine *e, int *cls)
{
    if (!e) { if (cls) *cls = KVQ_ERR_TYPE; return "engine is NULL"; }
    std::lock_guard<std::mutex> g(e->mu);
    if (cls) *cls = e->err_cls;
    return e->err_msg.c_str();
}

} // extern "C"

// ------------------------------------------------------------------------------- HBM-resident stream
extern "C" {

int kvq_stage_mem(kvq_engine *e, const char *const *bufs, const int64_t *buflens, int nbufs, kvq_staged **out)
{
    if (out) *out = nullptr;
    return guarded(e, [&] {
        if (!out || nbufs < 0 || (nbufs > 0 && (!bufs || !buflens))) throw KvqError(KVQ_ERR_TYPE, "bad buffers");
        RunGuard run(e);
        if (!run.owns) throw KvqError(KVQ_ERR_RUNTIME, "findseqs() already running!");
        CK(cudaSetDevice(e->device));
        int64_t total = 0;
        std::vector<std::pair<const char *, int64_t>> mems;
        for (int i = 0; i < nbufs; i++) { mems.push_back({bufs[i], buflens[i]}); total += buflens[i]; }
        std::unique_ptr<kvq_staged> st(new kvq_staged());
        st->e = e;
        st->d_data.ensure((size_t) total + DEV_PAD + 256);
        LiveStats live; live.reset(0);
        StreamCutter cutter(std::vector<std::string>(), std::move(mems), &live);
        std::vector<char> tmp((size_t) (BATCH_BYTES + BATCH_PAD));
        HostBatch hb; hb.own = tmp.data(); hb.data = hb.own;
        std::vector<int64_t> cuts; cuts.push_back(0);
        std::atomic<int> nostop{0};
        int64_t at = 0; bool first = true;
        while (cutter.next_batch(hb, nostop)) {
            if (first) { st->base_fpos = hb.base_fpos; first = false; }
            CK(cudaMemcpy(st->d_data.as<char>() + at, hb.data, (size_t) hb.nbytes, cudaMemcpyHostToDevice));
            for (size_t k = 1; k < hb.cuts.size(); k++) cuts.push_back(at + hb.cuts[k]);
            at += hb.nbytes;
        }
        CK(cudaMemset(st->d_data.as<char>() + at, 0, DEV_PAD));
        st->nbytes = at; st->nseg = (int) cuts.size() - 1;
        st->parsed = live.parsed; st->total = live.total;
        st->d_seg.ensure(cuts.size() * 8);
        CK(cudaMemcpy(st->d_seg.p, cuts.data(), cuts.size() * 8, cudaMemcpyHostToDevice));
        {
            std::vector<longlong2> tiles; PinBuf stage;
            build_tiles(cuts, tiles);
            st->tiles.upload_async(tiles, stage, nullptr);
            CK(cudaDeviceSynchronize());
        }
        st->d_ctr.ensure(sizeof(KvqCounters));
        st->long_cap = (size_t) (at / 1024) + 64; st->d_longs.ensure(st->long_cap * sizeof(KvqLongRec));
        st->surv_cap = initial_surv_cap(); st->d_surv.ensure(st->surv_cap * sizeof(KvqSurvRec));
        st->hit_cap = 1u << 18; st->d_hits.ensure(st->hit_cap * sizeof(KvqHitRec));
        st->blob_cap = 32u << 20; st->d_blob.ensure(st->blob_cap);
        *out = st.release();
    });
}

} // extern "C"

namespace {
void nccl_check(ncclResult_t r, const char *what);
kvq_nccl::Api &need_nccl();
void unpack_reduced(kvq_engine *e, kvq_result *r, const int64_t *all);
}

static int scan_staged_impl(kvq_engine *e, kvq_staged *st, const char *const *seqs, const int32_t *seqlens,
                            int nseqs, int nplus, int want_hits, kvq_result **out, bool reduce, double *reduce_ms)
{
    if (out) *out = nullptr;
    if (reduce_ms) *reduce_ms = 0;
    return guarded(e, [&] {
        if (!out || !st || st->e != e) throw KvqError(KVQ_ERR_TYPE, "bad staged stream");
        RunGuard run(e);
        if (!run.owns) throw KvqError(KVQ_ERR_RUNTIME, "findseqs() already running!");
        const double t_begin = now_ms();
        static const bool trace = getenv("KVQ_TRACE") != nullptr;
        double tp[8]; int ntp = 0;
        auto mark = [&] { if (trace && ntp < 8) tp[ntp++] = now_ms(); };
        e->stop = 0;
        CK(cudaSetDevice(e->device));
        kvq_config_t cfg;
        { std::lock_guard<std::mutex> g(e->mu); cfg = e->cfg; e->live.reset(nseqs); }
        e->tpl.build(seqs, seqlens, nseqs, nplus, cfg);
        e->tpl.par.want_hits = want_hits ? 1 : 0;
        mark();
        init_slots(e);                                  // borrow slot 0's stream and events
        Slot &s0 = e->slots[0];
        reset_accum(e, s0.stream);
        mark();
        e->live.parsed = st->parsed; e->live.total = st->total;
        ScanTotals tot; tot.acc_fetched = true;
        tot.hits.swap(e->scratch.hits); tot.blob.swap(e->scratch.blob); tot.hits.clear(); tot.blob.clear();
        KvqBatch b{st->d_data.as<uint8_t>(), st->nbytes, st->d_seg.as<int64_t>(), st->nseg, st->base_fpos};
        st->tiles.fill(b);
        KvqOut o = make_out(e, bufs_of(*st));
        enqueue_kernels(e, b, e->tpl.par, o, s0.stream, s0.ev0, s0.ev1);
        const bool red = reduce && e->comm != nullptr;
        // the copies of the counters (and the reduce) run on a side stream beside the kernels that order the hits:
        // in stream order they would stand between the scan and the ordering (~20 us of copy latency per step)
        static const bool side_copies = !(getenv("KVQ_SIDE_COPIES") && atoi(getenv("KVQ_SIDE_COPIES")) == 0);
        const bool side = red || (side_copies && want_hits);
        if (side && !e->red_stream) {
            CK(cudaStreamCreateWithFlags(&e->red_stream, cudaStreamNonBlocking));
            CK(cudaEventCreateWithFlags(&e->ev_ra, cudaEventDisableTiming)); CK(cudaEventCreateWithFlags(&e->ev_rb, cudaEventDisableTiming));
            CK(cudaEventCreate(&e->ev_r0)); CK(cudaEventCreate(&e->ev_r1));
        }
        cudaStream_t cs = s0.stream;            // where the counter copies go
        if (side && !red) { CK(cudaEventRecord(e->ev_ra, s0.stream)); CK(cudaStreamWaitEvent(e->red_stream, e->ev_ra, 0)); cs = e->red_stream; }
        CK(cudaMemcpyAsync(s0.h_ctr.p, st->d_ctr.p, sizeof(KvqCounters), cudaMemcpyDeviceToHost, cs));
        const size_t row = e->acc.row_words();
        auto enqueue_reduce = [&](cudaStream_t rs) {
            // tail of the row from the batch's device counters, then the sum over the ranks into d_red (the rank's own
            // row stays intact: a repeated collective needs it) and the sums to the host
            kvq_nccl::Api &a = need_nccl();
            cuda_check((cudaError_t) kvq_launch_tail(st->d_ctr.as<KvqCounters>(), e->acc.seqhits() + e->acc.words(), st->parsed,
                                                     (unsigned long long) st->surv_cap, s0.stream), "tail kernel launch");
            if (rs != s0.stream) { CK(cudaEventRecord(e->ev_ra, s0.stream)); CK(cudaStreamWaitEvent(rs, e->ev_ra, 0)); }
            CK(cudaEventRecord(e->ev_r0, rs));
            nccl_check(a.AllReduce(e->acc.d_all.p, e->d_red.p, row, ncclInt64, ncclSum, e->comm, rs), "ncclAllReduce");
            CK(cudaEventRecord(e->ev_r1, rs));
            CK(cudaMemcpyAsync(e->pin_acc.p, e->d_red.p, row * 8, cudaMemcpyDeviceToHost, rs));
            if (rs != s0.stream) CK(cudaEventRecord(e->ev_rb, rs));
        };
        if (red) {
            e->d_red.ensure(row * 8); e->pin_acc.ensure(row * 8);
            enqueue_reduce(e->red_stream);
        } else {   // the per-template accumulators ride behind the kernels instead of a blocking copy later
            const bool want_cov = e->tpl.nplus > 0 && e->tpl.cov_total > 0;
            const size_t words = want_cov ? e->acc.words() : 2 * e->acc.ns;
            e->pin_acc.ensure(words * 8);
            CK(cudaMemcpyAsync(e->pin_acc.p, e->acc.d_all.p, words * 8, cudaMemcpyDeviceToHost, cs));
            if (side) CK(cudaEventRecord(e->ev_rb, cs));
        }
        // the hits are put into the reference's order on the device (KvqOrder); a last kernel exports the
        // ordered records, the hitseq offsets and the hitseq bytes into the result's page-locked host buffer
        // (mapped into the device's address space under unified addressing) with wide coalesced stores:
        // exactly the bytes of this scan cross PCIe, no staging copy, nothing sized by an earlier scan
        static const bool host_order = getenv("KVQ_HOST_ORDER") != nullptr;
        const bool dev_order = want_hits && !host_order && st->nbytes > 0;
        KvqOrder ord{};
        size_t oh = 0, oo = 0, ob = 0;          // layout of the result's buffer: hits | offsets | blob
        if (dev_order) {
            tot.pre = take_result();
            const size_t cap = std::min<size_t>(st->hit_cap, KVQ_ORDER_MAX);
            oh = 0; oo = (cap * sizeof(KvqHitOut) + 255) & ~(size_t) 255; ob = oo + (((cap + 1) * 8 + 255) & ~(size_t) 255);
            tot.pre->pin.ensure(ob + st->blob_cap + 256);
            char *pin = tot.pre->pin.as<char>();
            st->o_keys.ensure(cap * 16); st->o_bkt.ensure(cap * 4); st->o_start.ensure((cap + 1) * 4); st->o_cursor.ensure((cap + 1) * 4);
            st->o_members.ensure(cap * 4); st->o_lens.ensure((cap + 1) * 4); st->o_src.ensure(cap * 8); st->o_status.ensure(16);
            st->o_hits.ensure(cap * sizeof(KvqHitOut) + 16); st->o_offsets.ensure((cap + 1) * 8 + 16); st->o_blob.ensure(st->blob_cap + 16);
            ord.keys = st->o_keys.as<ulonglong2>(); ord.bkt = st->o_bkt.as<uint32_t>(); ord.start = st->o_start.as<uint32_t>();
            ord.cursor = st->o_cursor.as<uint32_t>(); ord.members = st->o_members.as<uint32_t>(); ord.lens = st->o_lens.as<uint32_t>();
            ord.src = st->o_src.as<unsigned long long>(); ord.status = st->o_status.as<uint32_t>();
            ord.hits = st->o_hits.as<KvqHitOut>(); ord.offsets = st->o_offsets.as<long long>(); ord.blob = st->o_blob.as<uint8_t>();
            ord.out_hits = reinterpret_cast<KvqHitOut *>(pin + oh); ord.out_offsets = reinterpret_cast<long long *>(pin + oo);
            ord.out_blob = reinterpret_cast<uint8_t *>(pin + ob);
            ord.cap = (uint32_t) cap; ord.blob_cap = st->blob_cap; ord.span = st->nbytes;
            cuda_check((cudaError_t) kvq_launch_order(b, o, ord, s0.stream), "order kernels launch");
            e->pin_ord.ensure(16);
            CK(cudaMemcpyAsync(e->pin_ord.p, ord.status, 4, cudaMemcpyDeviceToHost, s0.stream));
        }
        mark();
        if (side) CK(cudaStreamWaitEvent(s0.stream, e->ev_rb, 0));
        CK(cudaStreamSynchronize(s0.stream));
        mark();
        float ms = 0; CK(cudaEventElapsedTime(&ms, s0.ev0, s0.ev1));
        tot.kernel_ms = ms; tot.launches = 4; tot.stream_bytes = st->nbytes; tot.d2h = sizeof(KvqCounters);
        KvqCounters c = *s0.h_ctr.as<KvqCounters>();
        const int64_t launches0 = tot.launches;
        const KvqParams used = rerun_for_survivors(e, b, bufs_of(*st), c, s0.h_ctr, s0.stream, tot);
        add_counters(e, c, tot);
        int red_launches = 0;
        if (red) {
            float rms = 0; CK(cudaEventElapsedTime(&rms, e->ev_r0, e->ev_r1));
            red_launches += 2; tot.d2h += (int64_t) row * 8;
            if (e->pin_acc.as<int64_t>()[e->acc.words() + KVQ_HIST + 4] != 0) {
                // some rank's batch was not complete when its row went into the sum (it has run the batch again by now):
                // every rank sees that in the sum, and every rank repeats the collective with its final row
                enqueue_reduce(s0.stream);
                CK(cudaStreamSynchronize(s0.stream));
                float r2 = 0; CK(cudaEventElapsedTime(&r2, e->ev_r0, e->ev_r1)); rms += r2;
                red_launches += 2; tot.d2h += (int64_t) row * 8;
                if (e->pin_acc.as<int64_t>()[e->acc.words() + KVQ_HIST + 4] != 0)
                    throw KvqError(KVQ_ERR_RUNTIME, "internal: a rank's batch is still incomplete in the repeated reduce");
            }
            tot.acc_fetched = true;
            if (reduce_ms) *reduce_ms = rms;
        }
        if (want_hits) {
            // ordered on the device unless the batch had to be run again or a buffer overflowed
            const bool ordered = dev_order && tot.launches == launches0 && c.hit_count <= st->hit_cap && c.blob_used <= st->blob_cap
                              && *e->pin_ord.as<uint32_t>() == (uint32_t) c.hit_count && c.hit_count <= ord.cap && c.hit_count > 0;
            if (ordered) {
                const size_t nh = (size_t) c.hit_count, nb = (size_t) c.blob_used;
                const char *pin = tot.pre->pin.as<char>();
                tot.s_hits = reinterpret_cast<const KvqHitOut *>(pin + oh); tot.s_off = reinterpret_cast<const long long *>(pin + oo);
                tot.s_blob = pin + ob; tot.s_n = nh; tot.s_nb = nb;
                tot.launches += 7; tot.d2h += (int64_t) (nh * sizeof(KvqHitOut) + (nh + 1) * 8 + nb + 4);
            } else {
                collect_hits(e, b, bufs_of(*st), used, c, s0.stream, tot);
            }
        }
        tot.launches += red_launches;           // (tail kernel + the collective's kernel)
        mark();
        if (tot.err_key != ~0ULL) throw KvqError(KVQ_ERR_FASTQ_FORMAT, format_error_message(tot.err_key));
        *out = finish_result(e, tot, now_ms() - t_begin);
        if (red) unpack_reduced(e, *out, e->pin_acc.as<int64_t>());
        mark();
        if (trace) {
            fprintf(stderr, "[kvq trace] build %.3f reset %.3f enqueue %.3f sync %.3f hits %.3f finish %.3f (kernel %.3f)\n",
                    tp[0] - t_begin, tp[1] - tp[0], tp[2] - tp[1], tp[3] - tp[2], tp[4] - tp[3], tp[5] - tp[4], ms);
        }
    });
}

extern "C" {

int kvq_scan_staged(kvq_engine *e, kvq_staged *st, const char *const *seqs, const int32_t *seqlens,
                    int nseqs, int nplus, int want_hits, kvq_result **out)
{
    return scan_staged_impl(e, st, seqs, seqlens, nseqs, nplus, want_hits, out, false, nullptr);
}

int kvq_scan_staged_reduce(kvq_engine *e, kvq_staged *st, const char *const *seqs, const int32_t *seqlens,
                           int nseqs, int nplus, int want_hits, kvq_result **out, double *reduce_ms)
{
    return scan_staged_impl(e, st, seqs, seqlens, nseqs, nplus, want_hits, out, true, reduce_ms);
}

void kvq_staged_free(kvq_staged *st) { delete st; }

void *kvq_host_alloc(int64_t nbytes)
{
    void *p = nullptr;
    if (nbytes <= 0 || cudaHostAlloc(&p, (size_t) nbytes, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}

void kvq_host_free(void *p) { if (p) cudaFreeHost(p); }

} // extern "C"

// ------------------------------------------------------------------------------- multi-GPU (NCCL)
// One process per GPU, every rank scans its own share of the stream (records are independent, the
// template set is replicated); the per-template counters are combined with ONE ncclAllReduce
// (int64, sum) over NVLink at the end -- SURVEY 8(e).  The counters never leave the device before the
// reduce: the accumulators the kernels wrote are reduced in place on the scan stream.
namespace {

void nccl_check(ncclResult_t r, const char *what)
{
    if (r != ncclSuccess)
        throw KvqError(KVQ_ERR_RUNTIME, std::string("NCCL failure in ") + what + ": " + kvq_nccl::api().GetErrorString(r));
}

kvq_nccl::Api &need_nccl()
{
    kvq_nccl::Api &a = kvq_nccl::api();
    if (!a.handle || !a.error.empty()) throw KvqError(KVQ_ERR_RUNTIME, a.error.empty() ? "NCCL not available" : a.error);
    return a;
}

// the reduced row -> the result's counters
void unpack_reduced(kvq_engine *e, kvq_result *r, const int64_t *all)
{
    const size_t ns = e->acc.ns, nc = e->acc.nc;
    for (int i = 0; i < e->tpl.nseqs; i++) { r->nseqhits[i] = all[i]; r->nseqbasehits[i] = all[ns + i]; }
    if (!r->coverage.empty()) {
        r->coverage.assign(all + 2 * ns, all + 2 * ns + r->coverage.size());
        r->mutcount.assign(all + 2 * ns + nc, all + 2 * ns + nc + r->mutcount.size());
    }
    const int64_t *t = all + e->acc.words();
    // the longest read length is a maximum, not a sum: every rank's histogram is in the sum, so it is
    // the last non-empty bin -- or, beyond the histogram (>= 1024, workhorse.c:397-400), left per rank
    int64_t longest = -1;
    for (int i = 0; i < KVQ_HIST; i++) { r->stats.readlengths[i] = t[i]; if (t[i]) longest = i; }
    r->stats.rls_longest = std::max(r->stats.rls_longest, longest);
    r->stats.records_parsed = t[KVQ_HIST]; r->timing.bases = t[KVQ_HIST + 1]; r->stats.parsed = t[KVQ_HIST + 2];
}

// host totals of the last scan -> tail of the device accumulator row (see Accum)
void upload_tail(kvq_engine *e, const kvq_result *r, cudaStream_t st)
{
    e->pin_tail.ensure(Accum::TAIL * 8);
    int64_t *t = e->pin_tail.as<int64_t>();
    for (int i = 0; i < KVQ_HIST; i++) t[i] = r->stats.readlengths[i];
    t[KVQ_HIST] = r->stats.records_parsed; t[KVQ_HIST + 1] = r->timing.bases;
    t[KVQ_HIST + 2] = r->stats.parsed; t[KVQ_HIST + 3] = r->stats.rls_longest + 1; t[KVQ_HIST + 4] = 0;
    CK(cudaMemcpyAsync(e->acc.seqhits() + e->acc.words(), t, Accum::TAIL * 8, cudaMemcpyHostToDevice, st));
}

void check_result_shape(kvq_engine *e, const kvq_result *r)
{
    if (!r) throw KvqError(KVQ_ERR_TYPE, "result is NULL");
    if (!e->acc.d_all.p || (int) r->nseqhits.size() != e->tpl.nseqs)
        throw KvqError(KVQ_ERR_RUNTIME, "the result is not the engine's last scan");
}

} // namespace

extern "C" {

int kvq_comm_unique_id(char *id)
{
    if (!id) return KVQ_ERR_TYPE;
    kvq_nccl::Api &a = kvq_nccl::api();
    if (!a.handle || !a.error.empty()) return KVQ_ERR_RUNTIME;
    ncclUniqueId u;
    if (a.GetUniqueId(&u) != ncclSuccess) return KVQ_ERR_RUNTIME;
    static_assert(sizeof(u) == KVQ_COMM_ID_BYTES, "ncclUniqueId is 128 bytes");
    memcpy(id, &u, sizeof(u));
    return KVQ_OK;
}

int kvq_comm_init(kvq_engine *e, const char *id, int rank, int world)
{
    return guarded(e, [&] {
        if (!id || world < 1 || rank < 0 || rank >= world) throw KvqError(KVQ_ERR_TYPE, "bad communicator arguments");
        kvq_nccl::Api &a = need_nccl();
        CK(cudaSetDevice(e->device));
        if (e->comm) { a.CommDestroy(e->comm); e->comm = nullptr; }
        ncclUniqueId u; memcpy(&u, id, sizeof(u));
        nccl_check(a.CommInitRank(&e->comm, world, u, rank), "ncclCommInitRank");
        e->rank = rank; e->world = world;
    });
}

int kvq_comm_destroy(kvq_engine *e)
{
    return guarded(e, [&] {
        if (e->comm) { CK(cudaSetDevice(e->device)); need_nccl().CommDestroy(e->comm); e->comm = nullptr; }
        e->rank = 0; e->world = 1;
    });
}

int kvq_comm_info(kvq_engine *e, int *rank, int *world, int *nccl_version)
{
    if (!e) return KVQ_ERR_TYPE;
    if (rank) *rank = e->rank;
    if (world) *world = e->comm ? e->world : 1;
    if (nccl_version) { *nccl_version = 0; kvq_nccl::Api &a = kvq_nccl::api(); if (a.GetVersion) a.GetVersion(nccl_version); }
    return KVQ_OK;
}

// Sums the counters of this engine's last scan over all ranks and writes the sums into `r`
// (coverage, mutcount, nseqhits, nseqbasehits, readlengths, records_parsed, parsed, bases; rls_longest
// is the maximum).  Device accumulators -> ncclAllReduce in place on the scan stream -> one D2H copy.
// Hits stay per rank.  Without a communicator (world 1) the result is left as it is.
int kvq_allreduce(kvq_engine *e, kvq_result *r, double *reduce_ms)
{
    if (reduce_ms) *reduce_ms = 0;
    return guarded(e, [&] {
        check_result_shape(e, r);
        if (!e->comm) return;
        RunGuard run(e);
        if (!run.owns) throw KvqError(KVQ_ERR_RUNTIME, "findseqs() already running!");
        kvq_nccl::Api &a = need_nccl();
        CK(cudaSetDevice(e->device));
        init_slots(e);
        Slot &s0 = e->slots[0];
        cudaStream_t st = s0.stream;
        const size_t n = e->acc.row_words();
        upload_tail(e, r, st);
        CK(cudaEventRecord(s0.ev0, st));
        nccl_check(a.AllReduce(e->acc.d_all.p, e->acc.d_all.p, n, ncclInt64, ncclSum, e->comm, st), "ncclAllReduce");
        CK(cudaEventRecord(s0.ev1, st));
        e->pin_acc.ensure(n * 8);
        CK(cudaMemcpyAsync(e->pin_acc.p, e->acc.d_all.p, n * 8, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        if (reduce_ms) { float ms = 0; CK(cudaEventElapsedTime(&ms, s0.ev0, s0.ev1)); *reduce_ms = ms; }
        unpack_reduced(e, r, e->pin_acc.as<int64_t>());
    });
}

// Batch jobs (BASELINE config 5: many isolates over the ranks): a zeroed [nslots][row] int64 matrix in
// HBM; after each scan the rank stores that isolate's counters into the isolate's row (device to
// device); ONE ncclAllReduce over the whole matrix at the end gives every rank every isolate.
int kvq_job_begin(kvq_engine *e, int nslots, int64_t row_words)
{
    return guarded(e, [&] {
        if (nslots < 1 || row_words < (int64_t) Accum::TAIL) throw KvqError(KVQ_ERR_TYPE, "bad job shape");
        CK(cudaSetDevice(e->device));
        e->job_slots = nslots; e->job_row = (size_t) row_words;
        e->d_job.ensure((size_t) nslots * e->job_row * 8);
        CK(cudaMemset(e->d_job.p, 0, (size_t) nslots * e->job_row * 8));
    });
}

int64_t kvq_job_row_words(kvq_engine *e)     // row of the engine's last scan: 2 nseqs + 7 sum(len plus) + 1028
{
    return e && e->acc.d_all.p ? (int64_t) e->acc.row_words() : 0;
}

int kvq_job_store(kvq_engine *e, int slot, const kvq_result *r)
{
    return guarded(e, [&] {
        check_result_shape(e, r);
        if (slot < 0 || slot >= e->job_slots || e->acc.row_words() != e->job_row) throw KvqError(KVQ_ERR_TYPE, "bad job slot");
        CK(cudaSetDevice(e->device));
        init_slots(e);
        cudaStream_t st = e->slots[0].stream;
        upload_tail(e, r, st);
        CK(cudaMemcpyAsync(e->d_job.as<int64_t>() + (size_t) slot * e->job_row, e->acc.d_all.p, e->job_row * 8,
                           cudaMemcpyDeviceToDevice, st));
        CK(cudaStreamSynchronize(st));
    });
}

int kvq_job_allreduce(kvq_engine *e, int64_t *out, int64_t cap_words, double *reduce_ms)
{
    if (reduce_ms) *reduce_ms = 0;
    return guarded(e, [&] {
        const size_t n = (size_t) e->job_slots * e->job_row;
        if (!e->d_job.p || !out || (size_t) cap_words < n) throw KvqError(KVQ_ERR_TYPE, "bad job buffer");
        CK(cudaSetDevice(e->device));
        init_slots(e);
        Slot &s0 = e->slots[0];
        cudaStream_t st = s0.stream;
        if (e->comm) {
            kvq_nccl::Api &a = need_nccl();
            CK(cudaEventRecord(s0.ev0, st));
            nccl_check(a.AllReduce(e->d_job.p, e->d_job.p, n, ncclInt64, ncclSum, e->comm, st), "ncclAllReduce");
            CK(cudaEventRecord(s0.ev1, st));
        }
        CK(cudaMemcpyAsync(out, e->d_job.p, n * 8, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        if (e->comm && reduce_ms) { float ms = 0; CK(cudaEventElapsedTime(&ms, s0.ev0, s0.ev1)); *reduce_ms = ms; }
    });
}

} // extern "C"
