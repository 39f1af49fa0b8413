#!/bin/bash
# Pattern files of lengths 8/16/32/64 for every <dir>/*.msa: 100 patterns each, written to
# <out>/patterns_<name>.msa_<length>.txt.  This RESTATES the reference's generate_Patterns.sh:1-18 (same two loops, same file
# naming -- the naming is the contract the authors' launch scripts rely on); only the extraction tool it calls is ours.
set -e
DATASET_DIR=$1
OUTPUT_DIR=$2
HERE="$(cd "$(dirname "$0")" && pwd)"
for file in "$DATASET_DIR"/*.msa; do
    name=$(basename "$file" .msa)
    for length in 8 16 32 64; do
        out="${OUTPUT_DIR}/patterns_${name}.msa_${length}.txt"
        python3 "$HERE/extract_patterns_from_msa.py" "$file" "$out" -l "$length" -n 100
        echo "Generated $out"
    done
done
