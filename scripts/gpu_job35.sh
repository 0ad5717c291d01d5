#!/bin/bash
# near-uniform packed batches persistent (snake order), ragged ones one CTA per item: full GPU suite + the A/B table again
O=gpurun_out; mkdir -p $O/profiles
timeout 1500 python -m pytest tests -q -x -m gpu > $O/g35_pytest.log 2>&1; tail -2 $O/g35_pytest.log
timeout 300 python scripts/varlen_ab.py > $O/g35_varlen.txt 2>&1; cat $O/g35_varlen.txt
