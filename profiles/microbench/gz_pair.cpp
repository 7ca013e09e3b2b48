// Development tool: the cell decoder of kvarq_b200/csrc/gz_parallel.h on one thread, four 4 MiB chunks one at a time
// and two at a time (decode_pair).  Build: g++ -O2 -std=c++17 -DGZ_HEADER=\"<path>/gz_parallel.h\" gz_pair.cpp -lz -lpthread
#include GZ_HEADER
#include <stdio.h>
#include <chrono>
using namespace kvq_gz;
int main(int argc, char **argv)
{
    FILE *f = fopen(argv[1], "rb");
    fseek(f, 0, SEEK_END); long n = ftell(f); fseek(f, 0, SEEK_SET);
    std::vector<uint8_t> file((size_t) n + 64, 0);
    if (fread(file.data(), 1, (size_t) n, f) != (size_t) n) return 2;
    Bits b{file.data(), 8 * (int64_t) n};
    const int64_t CH = 4 << 20;
    std::vector<Chunk> ch(5);
    for (int i = 0; i < 5; i++) { ch[i].start_bit = find_block_start(b, 8 * ((i + 1) * CH), 8 * ((i + 1) * CH + CH / 4)); ch[i].head.cells.resize(12 << 20); }
    RoundCtx rc; rc.bits = b; rc.file = file.data(); rc.fsize = n; rc.piece_base = 10; rc.stop_bit = 8 * 6 * CH; rc.chunks = &ch;
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    double best1 = 1e9, best2 = 1e9; long out = 0;
    for (int rep = 0; rep < 8; rep++) {
        double t = now();
        for (int i = 0; i < 4; i++) decode_chunk(rc, ch[i], i, false, nullptr, 0);
        best1 = std::min(best1, now() - t);
        out = 0; for (int i = 0; i < 4; i++) { if (!ch[i].ok) printf("chunk %d failed\n", i); out += ch[i].head.n; }
        t = now();
        for (int i = 0; i < 4; i += 2) { ChunkRun x, y; x.start(rc, ch[i], i, false, nullptr, 0); y.start(rc, ch[i + 1], i + 1, false, nullptr, 0); decode_pair(x, y); }
        best2 = std::min(best2, now() - t);
        long out2 = 0; for (int i = 0; i < 4; i++) out2 += ch[i].head.n;
        if (out2 != out) printf("mismatch %ld %ld\n", out, out2);
    }
    printf("%ld bytes: one at a time %.1f ms (%.0f MB/s), in pairs %.1f ms (%.0f MB/s)\n", out, best1, out / best1 / 1e3, best2, out / best2 / 1e3);
}
