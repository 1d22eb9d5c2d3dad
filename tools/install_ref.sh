#!/usr/bin/env bash
# Brings the UNMODIFIED Python reference (revollllt/FP8_quantization) next to the checker so that it
# travels to the GPU box with the gpurun snapshot:  /root/reference/{approx,quantization,models,utils,
# debug_params}  ->  oracle/_ref/   (git-ignored: reference sources never enter the history).
# The reference has no setup.py / pyproject.toml, so "installing" it is a copy of its packages; nothing
# is edited (the cupy / timm import stubs live in oracle/ref_shim.py).
# Test infrastructure only: used by tests/test_gpu_dropin.py, bench.py's cpu_baseline /
# gpu_eager_reference legs and `bench.py --impl reference`.  Run by __graft_entry__.build() whenever
# /root/reference is present; a no-op (exit 0) where it is not.
set -euo pipefail
SRC="${FP8_REFERENCE_SRC:-/root/reference}"
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
DST="$HERE/oracle/_ref"
if [ ! -d "$SRC/approx" ]; then
    echo "install_ref: $SRC not present, nothing to do"
    exit 0
fi
rm -rf "$DST"
mkdir -p "$DST"
for d in approx quantization models utils debug_params; do
    cp -r "$SRC/$d" "$DST/$d"
done
find "$DST" -name '__pycache__' -type d -prune -exec rm -rf {} +
( cd "$SRC" && find approx quantization models utils debug_params -type f ! -name '*.pyc' -print0 | sort -z | xargs -0 sha256sum ) > "$DST/SHA256SUMS"
echo "install_ref: copied $(find "$DST" -type f | wc -l) files to $DST"
