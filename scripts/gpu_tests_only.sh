#!/bin/bash
O=gpurun_out/${1:-tests}; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" $O/pytest.log | tail -30
