#!/bin/bash
# GPU-box job 35 (round 2): DFT predict with interleaved rotation chains.
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests/test_dft.py -m gpu -q -x > $O/pytest_dft35.log 2>&1; tail -2 $O/pytest_dft35.log
for p in 1 4; do
  timeout 300 python profiles/run_dft.py --pols $p --no-ref > $O/r02_dft35_p$p.json 2> $O/r02_dft35_p$p.err; grep -o '"ours": {[^}]*}' $O/r02_dft35_p$p.json | cut -c1-600; grep -o '"parity": {[^}]*}' $O/r02_dft35_p$p.json
done
