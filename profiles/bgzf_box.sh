#!/bin/bash
# Development tool: member rounds of gz_parallel.h on the GPU box's host cores (a BGZF-like file of 64 KiB members).
g++ -O2 -std=c++17 -o /tmp/gzh tests/gz_harness.cpp -lz -lpthread || exit 1
python - <<'PY'
import gzip, sys
sys.path.insert(0, '.')
from kvarq_b200 import synth
g = synth.genome(2000000, seed=3)
fq = synth.fastq_reads(g, 800000, readlen=250, seed=5, qmean0=37.0, qdrop=20.0).tobytes()
open('/dev/shm/blocks.fastq.gz', 'wb').write(b''.join(gzip.compress(fq[i:i + 65280], 6) for i in range(0, len(fq), 65280)))
print('inflated bytes', len(fq))
PY
for t in 1 4 8 14 16; do
  s=$(date +%s%N); /tmp/gzh $t /dev/shm/blocks.fastq.gz /dev/null > /dev/null; e=$(date +%s%N)
  echo "BGZF threads $t: $(( (e - s) / 1000000 )) ms"
done
