#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
timeout 300 python tools/trace_e2e.py > $O/n_trace_e2e.txt 2>&1; grep -v "upload lane" $O/n_trace_e2e.txt | tail -45
