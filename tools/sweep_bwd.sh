#!/bin/bash
# tuning sweep (GPU box): rebuild tbe_backward.cu with different ring depths and time the microbench
for S in 4 6 8 12 16; do
  touch torchrec-oldfork_b200/csrc/tbe_backward.cu
  B200_NVCC_EXTRA="-DB200_BWD_STAGES=$S" python torchrec-oldfork_b200/build.py > /dev/null 2>&1 || echo build failed
  echo "== stages $S"; timeout 200 python tools/tbe_microbench.py 2>&1 | grep "backward"
done
