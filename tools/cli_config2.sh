#!/bin/bash
# End-to-end CLI timing at BASELINE config 2 (3000 x 100000 filtered SNP columns): FASTA on disk -> host/build/SpydrPick -> listings.
#   bash tools/cli_config2.sh [GPUS] > profiles/r02_cli_config2.txt
# The alignment is written to /dev/shm (the box's page cache would hold it anyway); -v prints the host's stage times.
cd "$(dirname "$0")/.."
GPUS=${1:-1}
W=/dev/shm/spk_cli_c2
rm -rf $W; mkdir -p $W/out
python - <<EOF
import sys, time
sys.path.insert(0, ".")
from spydrpick_b200.synth import generate_alignment, to_fasta
t = time.time()
s, pos, g = generate_alignment(3000, 100000, seed=20260102)
to_fasta(s, "$W/c2.fasta")
print("# synthetic alignment 3000 x 100000 written in %.1f s" % (time.time() - t))
EOF
ls -la $W/c2.fasta
[ -x host/build/SpydrPick ] || python -c "import __graft_entry__ as g; g.build_host()"
for rep in 1 2; do
  echo "# ---- run $rep: SpydrPick c2.fasta -v --gpus $GPUS (default options: filter, reweighting, automatic threshold, ARACNE)"
  rm -rf $W/out/*
  T0=$(date +%s.%N)
  host/build/SpydrPick $W/c2.fasta -v --gpus $GPUS --output-dir $W/out 2> $W/err.txt | grep -v "^$" | tail -60
  echo "exit status ${PIPESTATUS[0]}; wall $(python -c "import time; print('%.2f' % (time.time() - $T0))") s (process start to exit)"
  tail -5 $W/err.txt
  ls -la $W/out
done
rm -rf $W
