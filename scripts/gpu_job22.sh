#!/bin/bash
# no-allocation decode entry point: its test alone in a fresh process, then the whole GPU suite (it is collected last)
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -x -m gpu > $O/g22_alone.log 2>&1; tail -15 $O/g22_alone.log
timeout 1500 python -m pytest tests -q -x -m gpu > $O/g22_all.log 2>&1; tail -5 $O/g22_all.log
