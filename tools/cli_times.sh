#!/bin/bash
# Whole-program wall clock of the drop-in `matFact <file.in>` (bin/matFact) on the samples kept
# under tests/golden, with its own phase times (MATFACT_TIMING) -- the figure the reference's
# reports publish (whole program, seconds).  Output: one block per input.
set -e
cd "$(dirname "$0")/.."
tmp=$(mktemp -d)
gunzip -c tests/golden/instML100k.in.gz > $tmp/instML100k.in
cp tests/golden/inst1000-1e6-1000-1-3.in tests/golden/inst30-40-10-2-10.in $tmp/
for f in inst30-40-10-2-10 instML100k inst1000-1e6-1000-1-3; do
	for rep in 1 2; do
		s=$(date +%s%N)
		MATFACT_TIMING=1 ./bin/matFact $tmp/$f.in > $tmp/$f.out 2> $tmp/$f.err
		e=$(date +%s%N)
		echo "== $f run $rep: wall $(( (e - s) / 1000000 )) ms, $(wc -l < $tmp/$f.out) lines, md5 $(md5sum < $tmp/$f.out | cut -c1-32)"
		tr '\n' ' ' < $tmp/$f.err; echo
	done
done
rm -rf $tmp
