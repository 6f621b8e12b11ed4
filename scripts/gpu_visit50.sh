#!/bin/bash
# attention_ws2: share of the exponentials on the FMA pipe (WS_POLY_MASK 0x00 / 0x24 shipped / 0x54 / 0x55)
for m in 0x00 shipped 0x54 0x55; do
  unset DEPTH_B200_LIB
  if [ $m != shipped ]; then export DEPTH_B200_LIB=$PWD/scripts/_bin/libdepthb200_pm$m.so; fi
  echo "== mask $m"; timeout 100 python scripts/attn_bench.py 2>&1 | grep -v Loading
done
