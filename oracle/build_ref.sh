#!/usr/bin/env bash
# TEST INFRASTRUCTURE.  Compiles the reference's OWN native sources, where they lie under
# /root/reference, into oracle/_ref/ (git-ignored; travels to the GPU box with gpurun).
# Nothing from the reference is copied into the repository; only built objects are kept.
#
#   MLCodec_CXX   <- DCVC-DC/src/cpp/ops/ops.cpp                (pmf_to_quantized_cdf)
#   MLCodec_rans  <- DCVC-DC/src/cpp/rans/rans.cpp + py_rans/py_rans.cpp
#   MLCodec_rans_hem <- DCVC-HEM/src/cpp/rans/rans_interface.cpp   (module name MLCodec_rans,
#                    kept in its own directory oracle/_ref/hem/)
#
# The only file the reference needs that is absent from its tree is rans64.h (ryg_rans,
# fetched by its cmake at configure time); oracle/third_party/rans64.h restates it.
# The reference's cmake is NOT run (it wants network access for two git fetches).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${DCVC_REFERENCE_ROOT:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -d "$REF/DCVC-DC/src/cpp" ]; then
  echo "build_ref.sh: $REF not present; keeping prebuilt oracle/_ref as is" >&2
  exit 0
fi
mkdir -p "$OUT" "$OUT/hem"
PYINC="$(python -c 'import sysconfig;print(sysconfig.get_paths()["include"])')"
PBINC="$(python -c 'import pybind11;print(pybind11.get_include())')"
SUF="$(python -c 'import sysconfig;print(sysconfig.get_config_var("EXT_SUFFIX"))')"
CXXFLAGS="-O2 -std=c++17 -shared -fPIC -DNDEBUG -I$PYINC -I$PBINC"
DC="$REF/DCVC-DC/src/cpp"
HEM="$REF/DCVC-HEM/src/cpp"

build() { # out, sources...
  local out="$1"; shift
  if [ -f "$out" ]; then
    local newer=0
    for s in "$@" "$HERE/third_party/rans64.h"; do [ "$s" -nt "$out" ] && newer=1; done
    [ "$newer" = 0 ] && return 0
  fi
  echo "  g++ -> ${out#$HERE/}"
  g++ $CXXFLAGS -pthread -I"$HERE/third_party" "${INCS[@]}" "$@" -o "$out"
}

INCS=()
build "$OUT/MLCodec_CXX$SUF" "$DC/ops/ops.cpp"
INCS=(-I"$DC/rans" -I"$DC/py_rans")
build "$OUT/MLCodec_rans$SUF" "$DC/rans/rans.cpp" "$DC/py_rans/py_rans.cpp"
if [ -f "$HEM/rans/rans_interface.cpp" ]; then
  INCS=(-I"$HEM/rans")
  build "$OUT/hem/MLCodec_rans$SUF" "$HEM/rans/rans_interface.cpp"
fi
# A runnable copy of the reference's Python for the boxes that have no /root/reference (the GPU
# box): bench.py's CPU reference arm and its "unmodified reference on the same B200" figure import
# it from here (oracle/ref_loader.py).  oracle/_ref/ is git-ignored build output, never committed.
stage() { # project, subdirs...
  local proj="$1"; shift
  local dst="$OUT/$proj"
  mkdir -p "$dst/src"
  for d in "$@"; do
    rm -rf "$dst/src/$d"
    cp -r "$REF/$proj/src/$d" "$dst/src/$d"
  done
  cp "$REF/$proj/test_video.py" "$dst/test_video.py"
  chmod -R u+w "$dst"
  find "$dst" -name __pycache__ -type d -prune -exec rm -rf {} +
}
stage DCVC-DC models utils transforms
cp "$OUT/MLCodec_CXX$SUF" "$OUT/MLCodec_rans$SUF" "$OUT/DCVC-DC/src/models/"
if [ -d "$REF/DCVC-HEM/src/models" ]; then
  stage DCVC-HEM models layers entropy_models utils
  cp "$OUT/MLCodec_CXX$SUF" "$OUT/DCVC-HEM/src/entropy_models/"
  cp "$OUT/hem/MLCodec_rans$SUF" "$OUT/DCVC-HEM/src/entropy_models/"
fi
echo "oracle/_ref ready"
