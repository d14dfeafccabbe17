#!/bin/bash
timeout 600 python tools/dbg_perm.py > gpurun_out/dbg_perm.txt 2>&1
