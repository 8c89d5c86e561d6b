#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
timeout 300 build/ubench_stage_mt > $O/n_stage_mt.txt 2>&1; cat $O/n_stage_mt.txt
