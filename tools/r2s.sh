#!/bin/bash
# Round 2, 1-GPU call: the 2^24-path pipeline rule, the LBVH with per-family stack bounds, the megakernel at 5 / 6 / 8
# resident blocks (96 / 80 / 64 registers) against the 4 of its 128 registers.
#   gpurun --timeout 1500 -- 'bash tools/r2s.sh r2s'
TAG=${1:-r2s}; OUT=gpurun_out; mkdir -p $OUT
L=$PWD/raytracing-in-a-weekend_b200/csrc
timeout 900 python -m pytest tests -m gpu -q -x > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 5 $OUT/${TAG}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 $OUT/${TAG}_smoke.log
for c in C5 C2 C4; do timeout 300 python tools/lbvh_probe.py $c 200 2>&1 | tee -a $OUT/${TAG}_lbvh_vs_sah.txt; done
AB=$OUT/${TAG}_ab_mega_resident_blocks.txt; : > $AB
for rep in 1 2; do
  for lib in b200 minb5 minb6 minb8; do
    echo "== librtw_$lib.so (repeat $rep)" | tee -a $AB
    RTW_B200_LIB=$L/librtw_$lib.so timeout 300 python tools/crossover.py C1,C3 100,1000 3 mega 2>&1 | tee -a $AB
    RTW_B200_LIB=$L/librtw_$lib.so timeout 300 python tools/crossover.py C2,C4,C5 16,64 3 mega 2>&1 | tee -a $AB
  done
done
timeout 300 python tools/crossover.py C2,C4,C5 4,16,64 3 2>&1 | tee $OUT/${TAG}_crossover_default.txt
