#!/bin/bash
timeout 300 python scripts/proj_probe.py 2>&1 | grep -v Loading
DEPTH_B200_LIB=$PWD/scripts/_bin/libdepthb200_sb2.so timeout 300 python scripts/proj_probe.py 2>&1 | grep -v Loading | sed 's/^/sb2 (older tree): /'
