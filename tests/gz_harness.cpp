// TEST INFRASTRUCTURE: drives kvarq_b200/csrc/gz_parallel.h on the CPU.
// usage: gz_harness <threads> <file.gz> <out.bin>   -> writes the inflated bytes, prints the read events
#include "../kvarq_b200/csrc/gz_parallel.h"
#include <stdio.h>
using namespace kvq_gz;
int main(int argc, char **argv)
{
    if (argc < 4) return 2;
    FILE *f = fopen(argv[2], "rb");
    if (!f) return 2;
    fseek(f, 0, SEEK_END); long n = ftell(f); fseek(f, 0, SEEK_SET);
    std::vector<uint8_t> file((size_t) n + 64, 0);
    if (fread(file.data(), 1, (size_t) n, f) != (size_t) n) return 2;
    fclose(f);
    FILE *out = fopen(argv[3], "wb");
    std::vector<Event> ev;
    long long par = 0;
    try {
        FileDecoder d(file.data(), n, atoi(argv[1]));
        std::vector<Bytes> outs, spares;
        bool more = true;
        while (more) {
            more = d.next(outs, ev, &spares);
            for (auto &o : outs) { fwrite(o.data(), 1, o.size(), out); spares.push_back(std::move(o)); }
            outs.clear();
        }
        par = d.parallel_bytes();
    } catch (const Error &e) {
        printf("ERROR %d %s\n", e.io ? 1 : 0, e.what());
    }
    fclose(out);
    for (auto &e : ev) printf("EV %lld %lld %d\n", (long long) e.out, (long long) e.pos, e.at_end ? 1 : 0);
    fprintf(stderr, "parallel_bytes %lld\n", par);
    return 0;
}
