#!/bin/bash
for d in 0 1 16 2 4 8 12 3 15 31; do echo -n "dbg=$d "; B200Q_MXF_DBG=$d timeout 60 python tools/mmq_one.py mxfp4 2880 2880 64 20 2>&1 | tail -1; done
