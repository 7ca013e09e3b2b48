#!/bin/bash
# Development tool: the chunk-parallel inflater on the GPU box's host cores (tests/gz_harness.cpp): thread sweep,
# trace of the rounds, and the one-thread decoder with different table sizes (profiles/microbench/gz_pair.cpp).
nproc; grep -m1 "model name" /proc/cpuinfo
g++ -O2 -std=c++17 -o /tmp/gzh tests/gz_harness.cpp -lz -lpthread || exit 1
python - <<'PY'
import gzip, time, sys
sys.path.insert(0, '.')
from kvarq_b200 import synth
g = synth.genome(2000000, seed=3)
fq = synth.fastq_reads(g, 800000, readlen=250, seed=5, qmean0=37.0, qdrop=20.0).tobytes()
t = time.time()
open('/dev/shm/a6.fastq.gz', 'wb').write(gzip.compress(fq, 6))
print('gz bytes', len(fq), 'compress s', time.time() - t)
PY
for F in 9 10 11 12; do
  sed "s/static constexpr int FAST = 11;/static constexpr int FAST = $F;/" kvarq_b200/csrc/gz_parallel.h > /tmp/gz_parallel_f$F.h
  g++ -O2 -std=c++17 -DGZ_HEADER="\"/tmp/gz_parallel_f$F.h\"" -o /tmp/gz_pair_f$F profiles/microbench/gz_pair.cpp -lz -lpthread
done
for r in 1 2; do for F in 9 10 11 12; do echo -n "PAIR FAST $F: "; /tmp/gz_pair_f$F /dev/shm/a6.fastq.gz; done; done
echo -n "PAIR FAST 11 plain: "; KVQ_GZ_NO_BMI2=1 /tmp/gz_pair_f11 /dev/shm/a6.fastq.gz
for t in 1 4 8 16 32; do
  s=$(date +%s%N); /tmp/gzh $t /dev/shm/a6.fastq.gz /dev/null > /dev/null; e=$(date +%s%N)
  echo "GZH threads $t: $(( (e - s) / 1000000 )) ms"
done
KVQ_GZ_TRACE=1 /tmp/gzh 16 /dev/shm/a6.fastq.gz /dev/null 2>&1 > /dev/null | tail -4 | cut -c1-400
