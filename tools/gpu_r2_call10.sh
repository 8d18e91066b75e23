#!/bin/bash
for r in 0 1 2 3; do echo "roles_off=$r"; GSPLIT_DIAG_ROLES=$r timeout 300 python tools/rate.py 1 8 8 2 2>&1 | grep "mb/s"; done
for r in 0 1 2 3; do echo "stop=9 roles_off=$r"; GSPLIT_DIAG_STOP=9 GSPLIT_DIAG_ROLES=$r timeout 300 python tools/rate.py 1 8 8 2 2>&1 | grep "mb/s"; done
