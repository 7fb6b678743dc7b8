#!/bin/sh
# Puts the UNMODIFIED reference modules of the hot path where bench.py's reference arm can import them on the GPU box
# (which has no /root/reference): oracle/_ref/pkg/{segmentation_model,geometry_utils}.py.  oracle/_ref/ is git-ignored
# (the reference's sources stay out of this repository's history) but travels with the working tree.  Run in the
# authoring container; `__graft_entry__.build()` runs it whenever /root/reference is present.
set -e
here="$(cd "$(dirname "$0")" && pwd)"
ref="${1:-/root/reference}"
[ -d "$ref/pkg" ] || { echo "make_ref.sh: $ref/pkg not found" >&2; exit 1; }
mkdir -p "$here/_ref/pkg"
cp -f "$ref/pkg/segmentation_model.py" "$ref/pkg/geometry_utils.py" "$here/_ref/pkg/"
: > "$here/_ref/pkg/__init__.py"
( cd "$ref" && sha256sum pkg/segmentation_model.py pkg/geometry_utils.py ) > "$here/_ref/SHA256SUMS"
echo "oracle/_ref ready: $(wc -l < "$here/_ref/SHA256SUMS") files"
