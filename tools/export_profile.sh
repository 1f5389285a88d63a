#!/bin/bash
# gpurun_out/<name>.ncu-rep -> profiles/<name>_raw.csv (+ _details.txt): the
# text forms of an ncu capture that are small enough to commit.
set -e
cd "$(dirname "$0")/.."
for rep in "$@"; do
  name=$(basename "$rep" .ncu-rep)
  ncu -i "$rep" --page raw --csv > profiles/${name}_raw.csv 2>/dev/null
  ncu -i "$rep" --page details > profiles/${name}_details.txt 2>/dev/null
  python tools/ncu_ops.py "$rep" ${PIXELS:-1} > profiles/${name}_opcodes.txt 2>/dev/null || true
  ls -la profiles/${name}_*
done
