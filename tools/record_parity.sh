#!/bin/bash
# Parity record of a round: the whole GPU test suite, then the randomised sweep with TMA staging and with the LDGSTS forms.
# outputs: gpurun_out/r02_gpu_tests.log, gpurun_out/r02_fuzz.log
OUT=gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -6 > $OUT/r02_gpu_tests.log
{ echo "== TMA staging (default), seed 21"; python tools/fuzz_parity.py ${1:-150} 21 2>&1 | tail -4
  echo "== ORBX_TMA=0 (LDGSTS staging), seed 22"; ORBX_TMA=0 python tools/fuzz_parity.py ${2:-100} 22 2>&1 | tail -4
  echo "== ORBX_SPLIT=0, seed 23"; ORBX_SPLIT=0 python tools/fuzz_parity.py 40 23 2>&1 | tail -4; } > $OUT/r02_fuzz.log
cat $OUT/r02_gpu_tests.log $OUT/r02_fuzz.log
