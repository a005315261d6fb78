#!/bin/bash
# End-to-end CLI on a C2-sized FASTA database: load path timings (streaming gz / parallel parser / packed cache).
set -x
mkdir -p /tmp/c2 gpurun_out
python - <<'PY'
import sys, os, time
sys.path.insert(0, "seq-align-gpu_b200")
from swgpu import synth
import numpy as np
t = time.time()
codes, off = synth.config2()
synth.to_fasta("/tmp/c2/db.fa", codes, off)
q = synth.make_query(103, 375)
synth.to_fasta("/tmp/c2/q.fa", q, np.array([0, 375]), prefix="query")
print("fasta written in %.1f s, %d bytes" % (time.time() - t, os.path.getsize("/tmp/c2/db.fa")))
PY
M="--substitution_matrix scoring/BLOSUM62.txt --gapopen -10 --gapextend -1"
export SWGPU_VERBOSE=1
( time SWGPU_PARSE_THREADS=1 bin/smith_waterman $M --files /tmp/c2/q.fa /tmp/c2/db.fa > /tmp/c2/out1.txt ) 2> gpurun_out/cli_c2_t1.log
( time bin/smith_waterman $M --files /tmp/c2/q.fa /tmp/c2/db.fa > /tmp/c2/out2.txt ) 2> gpurun_out/cli_c2_par.log
( time SWGPU_DB_CACHE=1 bin/smith_waterman $M --files /tmp/c2/q.fa /tmp/c2/db.fa > /tmp/c2/out3.txt ) 2> gpurun_out/cli_c2_cache_write.log
( time SWGPU_DB_CACHE=1 bin/smith_waterman $M --files /tmp/c2/q.fa /tmp/c2/db.fa > /tmp/c2/out4.txt ) 2> gpurun_out/cli_c2_cache_read.log
( time SWGPU_DB_CACHE=1 SWGPU_TOPK=10 SWGPU_CIGAR=1 bin/smith_waterman $M --files /tmp/c2/q.fa /tmp/c2/db.fa > gpurun_out/cli_c2_top10.txt ) 2> gpurun_out/cli_c2_top10.log
cmp /tmp/c2/out1.txt /tmp/c2/out2.txt; grep -v "Total Time" /tmp/c2/out2.txt | md5sum; grep -v "Total Time" /tmp/c2/out4.txt | md5sum
tail -n 3 /tmp/c2/out4.txt; nproc
for f in t1 par cache_write cache_read top10; do echo == $f; grep -E "swgpu:|real" gpurun_out/cli_c2_$f.log; done
head -n 6 gpurun_out/cli_c2_top10.txt | cut -c1-200
