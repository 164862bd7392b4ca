#!/bin/bash
# Two-rank run of the PROPhet_b200 command line on one box (one process per GPU, NCCL through the rendezvous file)
# against the single-rank run of the same case: the printed training tables must agree.
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
W=$(mktemp -d)
python - "$W" <<PY
import sys, os
sys.path.insert(0, "$ROOT"); sys.path.insert(0, "$ROOT/tests")
import test_gpu_host as t
from prophet_b200 import synth
cfg = synth.CONFIGS["C2"](8)
for d in ("one", "two"):
    os.makedirs(os.path.join(sys.argv[1], d))
    t._write_case(os.path.join(sys.argv[1], d), cfg, 8, [0, 0, 1, 0, 0, 0, 2, 0], [8, 6], seed=9, niter=6)
PY
(cd $W/one && $ROOT/prophet_b200/PROPhet_b200 -in input_file -train > out.txt 2>&1)
(cd $W/two && RANK=0 WORLD_SIZE=2 LOCAL_RANK=0 $ROOT/prophet_b200/PROPhet_b200 -in input_file -train > out0.txt 2>&1 &
 cd $W/two && RANK=1 WORLD_SIZE=2 LOCAL_RANK=1 $ROOT/prophet_b200/PROPhet_b200 -in input_file -train > out1.txt 2>&1; wait)
echo "--- one rank"; grep -E "^[0-9]+ +[-0-9.e+]+ +[-0-9.e+]+" $W/one/out.txt | tail -6
echo "--- two ranks (rank 0)"; grep -E "^[0-9]+ +[-0-9.e+]+ +[-0-9.e+]+" $W/two/out0.txt | tail -6
tail -3 $W/two/out1.txt
