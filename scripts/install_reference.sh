#!/bin/sh
# Install the UNMODIFIED reference (five loose scripts, no packaging metadata) into baseline/_ref:
# a copy under /tmp gets a pyproject.toml that only names the modules; the installed files are
# byte-identical to /root/reference/src/*.py (checked below). See DESIGN.md section 4.
set -e
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
SRC="${1:-/root/reference}"
rm -rf /tmp/refcopy && cp -r "$SRC" /tmp/refcopy && chmod -R u+w /tmp/refcopy
cat > /tmp/refcopy/pyproject.toml <<'EOF'
[build-system]
requires = ["setuptools"]
build-backend = "setuptools.build_meta"
[project]
name = "csed514-reference"
version = "0.0.0"
[tool.setuptools]
py-modules = ["model", "train", "train_dist", "run1", "run2"]
package-dir = {"" = "src"}
EOF
python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
    --target "$ROOT/baseline/_ref" --upgrade /tmp/refcopy
for f in model train train_dist run1 run2; do cmp "$ROOT/baseline/_ref/$f.py" "$SRC/src/$f.py"; done
echo "reference installed in $ROOT/baseline/_ref (byte-identical)"
