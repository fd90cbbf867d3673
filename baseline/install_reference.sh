#!/usr/bin/env bash
# Places an UNMODIFIED copy of the reference tree under baseline/_ref/ (git-ignored; it still travels
# to the GPU box). The reference is a collection of scripts with no setup.py / pyproject.toml, so
#   python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse \
#          --target baseline/_ref /root/reference
# fails with "Neither 'setup.py' nor 'pyproject.toml' found" (recorded in DESIGN.md); a plain copy is
# the only way to make its train_llm.py scripts importable.
set -euo pipefail
here="$(cd "$(dirname "$0")" && pwd)"
src="${1:-/root/reference}"
mkdir -p "$here/_ref"
if [ -d "$src" ]; then
  rm -rf "$here/_ref/reference_src"
  cp -r "$src" "$here/_ref/reference_src"
  rm -rf "$here/_ref/reference_src/.git"
  echo "reference copied to $here/_ref/reference_src"
else
  echo "reference tree $src not found" >&2
  exit 1
fi
