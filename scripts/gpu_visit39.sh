#!/bin/bash
timeout 200 python scripts/narrow_n_probe.py 2>&1 | grep -v Loading
