#!/bin/bash
# round 2, call A: cold-box diagnostics of the fp16 mode first, then the tightened parity suite, then the old bench line
mkdir -p gpurun_out
L=gpurun_out/r02a
{ for s in 1 2 3 4 5 6; do SCREAM_POISON=1 timeout 300 python tools/diag_half_cold.py $s; echo "exit $?"; done; } > ${L}_half_cold.log 2>&1
{ for s in 1 2 3; do SCREAM_POISON=1 timeout 300 python __graft_entry__.py --smoke; echo "exit $?"; done; } > ${L}_smoke_poison.log 2>&1
timeout 300 python __graft_entry__.py --smoke > ${L}_smoke.log 2>&1; echo "smoke exit $?" >> ${L}_smoke.log
timeout 1500 python -m pytest tests -m gpu -q -s --maxfail=20 > ${L}_pytest.log 2>&1; tail -15 ${L}_pytest.log
SCREAM_POISON=1 timeout 1500 python -m pytest tests -m gpu -q --maxfail=20 -k "not sixteen" > ${L}_pytest_poison.log 2>&1; tail -5 ${L}_pytest_poison.log
timeout 600 python bench.py > ${L}_bench.log 2>&1; tail -1 ${L}_bench.log | cut -c1-600
grep -h "exit\|MISMATCH\|mismatch" ${L}_half_cold.log ${L}_smoke_poison.log ${L}_smoke.log | sort | uniq -c
