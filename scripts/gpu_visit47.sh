#!/bin/bash
timeout 200 python scripts/expand_bn_probe.py 2>&1 | grep -v Loading | tail -12
