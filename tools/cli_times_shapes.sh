#!/bin/bash
# loop_ms of the drop-in CLI on the two synthetic headline shapes for the GPU counts in $NGPUS
# (one run each; the full table over all inputs is tools/cli_times.sh)
cd "$(dirname "$0")/.."
tmp=$(mktemp -d)
python - "$tmp" <<'PY'
import sys
sys.path.insert(0, ".")
from recommender_system_b200 import synth
from recommender_system_b200.io import write_instance
write_instance(sys.argv[1] + "/inst60000-shape.in", synth.inst60000_like())
write_instance(sys.argv[1] + "/instML1M-shape.in", synth.ml_like())
PY
for n in ${NGPUS:-1}; do for f in instML1M-shape inst60000-shape; do
	MATFACT_NGPUS=$n MATFACT_TIMING=1 ./bin/matFact $tmp/$f.in 2> $tmp/err | md5sum | cut -c1-32 | tr '\n' ' '
	echo "$f $n GPU(s): $(tr '\n' ' ' < $tmp/err)"
done; done
rm -rf $tmp
