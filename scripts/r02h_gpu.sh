#!/bin/bash
mkdir -p gpurun_out
python scripts/sanitize_case.py big > gpurun_out/r02h_plain.log 2>&1 || { tail -5 gpurun_out/r02h_plain.log; exit 1; }
tail -3 gpurun_out/r02h_plain.log
timeout 1500 compute-sanitizer --tool memcheck --print-limit 20 python scripts/sanitize_case.py big > gpurun_out/r02h_memcheck.log 2>&1
echo "memcheck rc=$?"; grep -E "ERROR SUMMARY|Invalid|done" gpurun_out/r02h_memcheck.log | head
