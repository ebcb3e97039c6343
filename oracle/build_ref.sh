#!/bin/sh
# TEST INFRASTRUCTURE ONLY.  Compiles the REFERENCE's own graph_proc.cpp, from the sources where they lie
# under /root/reference (read-only; nothing is copied into the repo), behind oracle/ref_graph_bind.cpp,
# into oracle/_ref/ref_graph.so.  Flags follow csrc/setup.py:39-41 (-O3 -std=c++17 -fopenmp); the Eigen include
# root is NonRigidICP/external (graph_proc.cpp:11 includes <eigen3/Eigen/Dense>).  Skipped when the reference
# tree is absent (GPU box): the prebuilt file travels with the snapshot.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF=${OFX_REFERENCE:-/root/reference}
OUT="$HERE/_ref/ref_graph.so"
[ -d "$REF/csrc/cpu" ] || { echo "build_ref: $REF absent, keeping prebuilt files"; exit 0; }
[ -f "$OUT" ] && [ "$OUT" -nt "$HERE/ref_graph_bind.cpp" ] && exit 0
mkdir -p "$HERE/_ref"
PY=${PYTHON:-python}
INC=$($PY - <<'PYEOF'
import warnings, sysconfig, logging
logging.disable(logging.CRITICAL)
import pybind11
import torch.utils.cpp_extension as E
print(" ".join("-I" + p for p in E.include_paths()), "-I" + pybind11.get_include(), "-I" + sysconfig.get_paths()["include"])
PYEOF
)
CXX=$(test -x /usr/bin/g++ && echo /usr/bin/g++ || echo g++)
$CXX -O3 -std=c++17 -fopenmp -fPIC -shared -w -DTORCH_EXTENSION_NAME=ref_graph \
    -I"$REF/csrc" -I"$REF/NonRigidICP/external" $INC \
    "$HERE/ref_graph_bind.cpp" "$REF/csrc/cpu/graph_proc.cpp" -o "$OUT"
echo "built $OUT"
