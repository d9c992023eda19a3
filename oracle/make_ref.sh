#!/bin/sh
# TEST INFRASTRUCTURE ONLY.  Puts an UNMODIFIED copy of the reference's hot-path modules (efgpnd.py, cg.py, kernels/,
# utils/) under oracle/_ref/ so that the CPU arm of bench.py (`--impl reference`, `cpu_baseline`) times the reference
# ITSELF on the GPU box, where /root/reference does not exist.  oracle/_ref/ is git-ignored (reference sources never
# enter the history) but travels with the gpurun snapshot.  The reference's NUFFT dependency (pytorch_finufft /
# FINUFFT) is not installable here; oracle/fast_shim/ exposes the C/OpenMP ES-kernel NUFFT of oracle/cnufft.c under
# that module name.  Run from anywhere:  sh oracle/make_ref.sh [/root/reference]
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
SRC="${1:-/root/reference}"
[ -f "$SRC/efgpnd.py" ] || { echo "make_ref: $SRC/efgpnd.py not found (nothing to do)"; exit 0; }
DST="$HERE/_ref"
rm -rf "$DST"
mkdir -p "$DST/utils"
cp "$SRC/efgpnd.py" "$SRC/cg.py" "$DST/"
cp -r "$SRC/kernels" "$DST/kernels"
cp "$SRC/utils/kernels.py" "$DST/utils/kernels.py"
[ -f "$SRC/utils/__init__.py" ] && cp "$SRC/utils/__init__.py" "$DST/utils/__init__.py" || : > "$DST/utils/__init__.py"
find "$DST" -name "__pycache__" -type d -prune -exec rm -rf {} +
( cd "$SRC" && sha256sum efgpnd.py cg.py utils/kernels.py kernels/*.py ) > "$DST/SOURCES.sha256"
echo "make_ref: copied $(find "$DST" -name '*.py' | wc -l) files from $SRC into $DST"
