#!/usr/bin/env bash
# Installs the UNMODIFIED reference into baseline/_ref (git-ignored; travels to the GPU box).
#
# The prescribed command
#   python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse \
#          --target baseline/_ref /root/reference
# fails: "Directory '/root/reference' is not installable. Neither 'setup.py' nor 'pyproject.toml'
# found." -- the reference is a flat script collection with no packaging metadata (SURVEY.md 0.2
# fact 1).  Its "installation" is therefore a verbatim copy of its six files; nothing is edited.
# What the reference needs at run time that this image lacks (collections.Iterable on py3.12,
# matplotlib, the CIFAR-10 files) is provided from OUTSIDE the tree by baseline/shim/sitecustomize.py.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
SRC="${1:-/root/reference}"
DST="$HERE/_ref"
if python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target "$DST" "$SRC" \
     > "$HERE/_pip_install.log" 2>&1; then
  echo "pip install succeeded"
else
  echo "pip install failed (see baseline/_pip_install.log); copying the source tree verbatim"
  rm -rf "$DST"; mkdir -p "$DST/model"
  for f in main.py data.py utils.py plot_curves.py README.md model/resnet.py; do
    cp "$SRC/$f" "$DST/$f"
  done
  chmod -R u+w "$DST"
fi
( cd "$DST" && sha256sum main.py data.py utils.py plot_curves.py model/resnet.py ) > "$HERE/_ref.sha256"
( cd "$SRC" && sha256sum main.py data.py utils.py plot_curves.py model/resnet.py ) | diff - "$HERE/_ref.sha256" \
  && echo "baseline/_ref is byte-identical to $SRC"
