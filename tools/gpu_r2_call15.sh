#!/bin/bash
for s in 1 2 3 4 6 8 16; do timeout 300 python tools/rate.py 1 $s $s 2 2>&1 | grep "mb/s"; done
for s in 2 4; do timeout 300 python tools/rate.py 1 $s $s 4 2>&1 | grep "mb/s"; done
