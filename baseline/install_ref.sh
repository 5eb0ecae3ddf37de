#!/usr/bin/env bash
# Populate baseline/_ref with the UNMODIFIED reference sources.
#
# The prescribed `pip install --target baseline/_ref /root/reference` fails because the reference is
# a plain source tree (no setup.py / pyproject.toml):
#   ERROR: Directory '/root/reference' is not installable. Neither 'setup.py' nor 'pyproject.toml' found.
# The reference is run as `python main.py` from its checkout, so "installing" it means copying the
# first-party package (src/), main.py and config/ verbatim.  No file is edited; the three environment
# shims it needs on this image (numpy.Inf, matplotlib, waiting) live in baseline/ref_arm.py.
set -euo pipefail
REF=${1:-/root/reference}
DST="$(cd "$(dirname "$0")" && pwd)/_ref"
python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target "$DST" "$REF" \
  2>&1 | tail -1 || true
mkdir -p "$DST"
rm -rf "$DST/src" "$DST/config" "$DST/main.py"
# first-party code only: the vendored third-party submodule (Mallet, GUI, ...) is not on the benchmark path
rsync -a --exclude 'aux_modules/topicmodeler' --exclude '*.ipynb' --exclude 'checkpoint.pt' "$REF/src" "$DST/" \
  2>/dev/null || { mkdir -p "$DST/src"; (cd "$REF/src" && tar --exclude='aux_modules/topicmodeler' -cf - .) | (cd "$DST/src" && tar -xf -); }
cp "$REF/main.py" "$DST/main.py"
cp -r "$REF/config" "$DST/config"
echo "reference copied to $DST"
