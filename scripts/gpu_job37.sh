#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -q -x -m gpu > $O/g37_pytest.log 2>&1; tail -4 $O/g37_pytest.log | cut -c1-300
