#!/bin/bash
OUT=gpurun_out/r02c
mkdir -p "${OUT}"
timeout 1500 python -m pytest tests -m gpu -q -s > "${OUT}/pytest.log" 2>&1; echo "pytest rc=$?" | tee -a "${OUT}/pytest.log"
grep -E "SNR|passed|failed|FAILED|Error" "${OUT}/pytest.log" | head -60
