#!/bin/bash
# Whole-program wall clock of the drop-in `matFact <file.in>` (bin/matFact) with its own phase
# times (MATFACT_TIMING) -- the figure the reference's reports publish (BASELINE.md section 1:
# whole program, seconds) -- on the samples kept under tests/golden and on the two synthetic
# headline shapes, for every GPU count in $NGPUS (default: 1; e.g. NGPUS="1 2 4 8").
# MATFACT_NGPUS=N is the one-process N-GPU path: users sharded, the R gradient exchanged with the
# low-latency exchange inside CUDA graphs.  Output: one block per input and GPU count.
set -e
cd "$(dirname "$0")/.."
tmp=$(mktemp -d)
gunzip -c tests/golden/instML100k.in.gz > $tmp/instML100k.in
cp tests/golden/inst1000-1e6-1000-1-3.in tests/golden/inst30-40-10-2-10.in $tmp/
python - "$tmp" <<'PY'
import sys
sys.path.insert(0, ".")
from recommender_system_b200 import synth
from recommender_system_b200.io import write_instance
write_instance(sys.argv[1] + "/inst60000-shape.in", synth.inst60000_like())
write_instance(sys.argv[1] + "/instML1M-shape.in", synth.ml_like())
PY
for n in ${NGPUS:-1}; do
for f in inst30-40-10-2-10 instML100k instML1M-shape inst60000-shape inst1000-1e6-1000-1-3; do
	for rep in 1 2; do
		s=$(date +%s%N)
		MATFACT_NGPUS=$n MATFACT_TIMING=1 ./bin/matFact $tmp/$f.in > $tmp/$f.$n.out 2> $tmp/$f.err
		e=$(date +%s%N)
		echo "== $f, $n GPU(s), run $rep: wall $(( (e - s) / 1000000 )) ms, $(wc -l < $tmp/$f.$n.out) lines, md5 $(md5sum < $tmp/$f.$n.out | cut -c1-32)"
		tr '\n' ' ' < $tmp/$f.err; echo
	done
done
done
rm -rf $tmp
