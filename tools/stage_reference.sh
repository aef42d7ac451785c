#!/usr/bin/env bash
# Stage the reference's Python host code (op wrappers, layers, models, losses, configs) into oracle/_ref/src/
# so the -m gpu tests can run the UNMODIFIED reference modules on the GPU box, where /root/reference does not
# exist.  oracle/_ref/ is git-ignored (never part of the repo history, never part of the product) but not
# gpurun-ignored, so it travels with the snapshot exactly like oracle/_ref/*.so (the reference's kernels).
# Only .py / .yaml files are staged - no binaries, no data.  Test infrastructure only: nothing under
# geot_b200/ reads this directory (tests/test_abi.py::test_product_never_imports_oracle).
set -euo pipefail
REF="${1:-/root/reference}"
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
DST="$HERE/oracle/_ref/src"
if [ ! -d "$REF" ]; then
  echo "stage_reference: $REF not present (GPU box) - keeping whatever is staged in $DST" >&2
  exit 0
fi
rm -rf "$DST"
mkdir -p "$DST"
cd "$REF"
find openpoints pointnet2 pointops utils examples/segmentation cfgs \
     \( -name '*.py' -o -name '*.yaml' \) -not -path '*/build/*' -not -path '*/dist/*' -print0 \
  | while IFS= read -r -d '' f; do
      mkdir -p "$DST/$(dirname "$f")"
      cp "$f" "$DST/$f"
    done
( cd "$DST" && find . -type f | sort | xargs sha256sum ) > "$DST/MANIFEST.sha256"
echo "staged $(find "$DST" -type f | wc -l) files ($(du -sh "$DST" | cut -f1)) into $DST"
