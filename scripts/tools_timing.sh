#!/bin/bash
# Wall-clock of the drop-in command-line tools on BASELINE config 2 (what BFQzip_ext.py would run): step 1 + step 3.
set -e
D=${TMPDIR:-/tmp}/bfq_tools_timing; rm -rf $D; mkdir -p $D
python - <<PY
import sys; sys.path.insert(0, ".")
from bfqzip_b200 import synth, fastq
b, q, off = synth.make_reads(${1:-1000000}, ${2:-100}, 1001)
fastq.write_fastq("$D/reads.fastq", b, q, off)
PY
ls -la $D/reads.fastq
for rep in 1 2; do
  T0=$(date +%s.%N); tools/bin/eGap $D/reads.fastq --em --mem 4096 --qs -o $D/out --lcp --lbytes 1 > $D/s1.log 2>&1 || { cat $D/s1.log; exit 1; }
  echo "step1 eGap wall $(echo "$(date +%s.%N) - $T0" | bc) s"; tail -1 $D/s1.log
  T0=$(date +%s.%N); tools/bin/bfq_ext -e $D/out.bwt -q $D/out.bwt.qs -a $D/out.1.lcp -o $D/out -l 250 -s 0 -m 5 > $D/s3.log 2>&1 || { cat $D/s3.log; exit 1; }
  echo "step3 bfq_ext wall $(echo "$(date +%s.%N) - $T0" | bc) s"; tail -1 $D/s3.log
done
grep -i "time\|ms" $D/s1.log $D/s3.log | head -20
md5sum $D/out.fq | cut -c1-32
