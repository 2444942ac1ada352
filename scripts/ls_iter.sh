#!/bin/bash
# quick loop for the fused linearise+Schur kernel: parity test, C5 stage times, one ncu capture
tag=${1:-ls}
timeout 300 python -m pytest tests/test_global_ba.py -m gpu -x -q -k "fused" 2>&1 | tail -3
timeout 300 python bench.py --no-sub --no-cpu-baseline --steps 10 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['stages_ms'])"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_linschur_tc" -c 1 -f -o gpurun_out/$tag python scripts/run_c5_once.py c5 1 2>&1 | tail -1
