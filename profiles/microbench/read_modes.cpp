// Development tool: how fast can T host threads move a page-cache file (/dev/shm) into one buffer?
//   pread        what PlainFileSource does today
//   mmap+memcpy  the file mapped, glibc memcpy in 2 MiB pieces
//   mmap+stream  the same with non-temporal 32-byte stores (no read-for-ownership of the destination)
// usage: read_modes [MB]      build: g++ -O2 -mavx2 -pthread read_modes.cpp -o read_modes
#include <fcntl.h>
#include <immintrin.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <atomic>
#include <chrono>
#include <thread>
#include <vector>

static const size_t PART = 2u << 20;

static void stream_copy(uint8_t *dst, const uint8_t *src, size_t n)
{
    size_t i = 0;
    for (; i + 128 <= n; i += 128) {
        __m256i a = _mm256_loadu_si256((const __m256i *) (src + i)), b = _mm256_loadu_si256((const __m256i *) (src + i + 32));
        __m256i c = _mm256_loadu_si256((const __m256i *) (src + i + 64)), d = _mm256_loadu_si256((const __m256i *) (src + i + 96));
        _mm256_stream_si256((__m256i *) (dst + i), a); _mm256_stream_si256((__m256i *) (dst + i + 32), b);
        _mm256_stream_si256((__m256i *) (dst + i + 64), c); _mm256_stream_si256((__m256i *) (dst + i + 96), d);
    }
    for (; i < n; i++) dst[i] = src[i];
    _mm_sfence();
}

int main(int argc, char **argv)
{
    const size_t mb = argc > 1 ? (size_t) atoi(argv[1]) : 1024, n = mb << 20;
    const char *fn = "/dev/shm/kvq_read_modes.bin";
    {
        std::vector<uint8_t> a(n);
        for (size_t i = 0; i < n; i++) a[i] = (uint8_t) (i * 2654435761u >> 24);
        FILE *f = fopen(fn, "wb"); fwrite(a.data(), 1, n, f); fclose(f);
    }
    uint8_t *dst = (uint8_t *) aligned_alloc(4096, n);
    memset(dst, 1, n);
    auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    for (int mode = 0; mode < 3; mode++)
        for (int T : {4, 8, 16, 32}) {
            double best = 1e9;
            for (int rep = 0; rep < 3; rep++) {
                int fd = open(fn, O_RDONLY);
                const double t0 = now();
                const uint8_t *map = nullptr;
                if (mode) map = (const uint8_t *) mmap(nullptr, n, PROT_READ, MAP_SHARED, fd, 0);      // (a fresh mapping per scan: its page faults count)
                std::atomic<size_t> next{0};
                std::vector<std::thread> th;
                for (int t = 0; t < T; t++)
                    th.emplace_back([&] {
                        for (;;) {
                            const size_t at = next.fetch_add(PART);
                            if (at >= n) break;
                            const size_t m = std::min(PART, n - at);
                            if (mode == 0) { if (pread(fd, dst + at, m, (off_t) at) != (ssize_t) m) abort(); }
                            else if (mode == 1) memcpy(dst + at, map + at, m);
                            else stream_copy(dst + at, map + at, m);
                        }
                    });
                for (auto &t : th) t.join();
                best = std::min(best, now() - t0);
                if (map) munmap((void *) map, n);
                close(fd);
            }
            printf("%-12s threads %2d: %.1f GB/s\n", mode == 0 ? "pread" : mode == 1 ? "mmap+memcpy" : "mmap+stream", T, n / best / 1e9);
            fflush(stdout);
        }
    unlink(fn);
}
