#!/bin/bash
# TEST / BASELINE INFRASTRUCTURE.  Populates oracle/_ref/ (git-ignored, travels to the GPU box with the snapshot) with a
# copy of the reference's Python package where it lies under /root/reference, plus a stub of its un-vendored `learnMSA`
# import (hmm_layer/MsaHMMLayer.py:5).  Nothing under oracle/_ref is product source; bench.py's CPU arm
# (`--impl reference`, cpu_baseline.kind == "reference") drives these objects through oracle/ref_driver.py.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${HMM_REFERENCE_DIR:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -d "$REF/hmm_layer" ]; then
  echo "make_ref: $REF/hmm_layer not found (the GPU box uses the prebuilt oracle/_ref)"; exit 0
fi
rm -rf "$OUT"
mkdir -p "$OUT/learnMSA/msa_hmm"
cp -r "$REF/hmm_layer" "$OUT/hmm_layer"
find "$OUT" -name "__pycache__" -type d -prune -exec rm -rf {} +
touch "$OUT/learnMSA/__init__.py" "$OUT/learnMSA/msa_hmm/__init__.py"
cat > "$OUT/learnMSA/msa_hmm/Utility.py" <<'PY'
def deserialize(x):
    return x
PY
echo "oracle/_ref populated from $REF"
