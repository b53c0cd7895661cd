#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY -- builds the *unmodified reference* (tomo_TV + its vendored ASTRA fork)
# from the sources where they lie under /root/reference into oracle/_ref/ (git-ignored, but it
# travels to the GPU box).  Nothing from the reference is copied into the repo; only compiled
# artefacts land in oracle/_ref/.  The product (tomo_tv_b200) never links or loads any of this.
#
# Outputs:
#   oracle/_ref/libastra.so                      ASTRA fork (thirdparty/astra-toolbox, commit e44162bb)
#   oracle/_ref/tomoengine<EXT>.so               reference GPU engine   (tomofusion/gpu/utils/tomoengine.cpp)
#   oracle/_ref/multimodal<EXT>.so               reference fusion engine (tomofusion/chemistry/utils/multimodal.cpp)
#   oracle/_ref/ctvlib<EXT>.so                   reference CPU engine   (tomofusion/cpu/utils/ctvlib.cpp)
#
# The reference's own build system (autotools + Makefiles) is NOT run; this is plain g++/nvcc on
# the source files (recipe from SURVEY.md App. C).
set -euo pipefail
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/_ref
OBJ=$OUT/obj
JOBS=${JOBS:-$(nproc)}
A=$REF/thirdparty/astra-toolbox
G=$REF/tomofusion/gpu/utils
EXT=$(python3 -c "import sysconfig;print(sysconfig.get_config_var('EXT_SUFFIX'))")
PYB=$(python3 -m pybind11 --includes)
CUDA=${CUDA_HOME:-/usr/local/cuda}
ARCH="-gencode arch=compute_100,code=sm_100"

if [ ! -d "$REF" ]; then echo "no reference tree at $REF -- keeping prebuilt oracle/_ref"; exit 0; fi
mkdir -p "$OBJ/astra" "$OBJ/tomo"

if [ ! -f "$OUT/libastra.so" ]; then
  echo "[build_ref] ASTRA host sources"
  ls $A/src/*.cpp | xargs -P "$JOBS" -I{} sh -c \
    'o='"$OBJ"'/astra/$(basename {} .cpp).o; [ -f $o ] || g++ -std=c++17 -O2 -fPIC -w -DASTRA_CUDA -DASTRA_BUILDING_CUDA -I'"$A"'/include -I'"$A"'/lib/include -I'"$CUDA"'/include -c {} -o $o'
  echo "[build_ref] ASTRA CUDA sources"
  ls $A/cuda/2d/*.cu $A/cuda/3d/*.cu | xargs -P "$JOBS" -I{} sh -c \
    'o='"$OBJ"'/astra/cu_$(basename $(dirname {}))_$(basename {} .cu).o; [ -f $o ] || nvcc -std=c++17 -O2 -w -DASTRA_CUDA -DASTRA_BUILDING_CUDA -I'"$A"'/include -I'"$A"'/lib/include '"$ARCH"' -Xcompiler -fPIC -c {} -o $o'
  g++ -shared -o "$OUT/libastra.so" $OBJ/astra/*.o -L$CUDA/lib64 -lcudart -lcufft -lpthread
fi

if [ ! -f "$OUT/tomoengine$EXT" ]; then
  echo "[build_ref] reference tomoengine"
  for f in container/matrix_ops regularizers/tv_gd regularizers/tv_fgp; do
    nvcc -w -Xcompiler -fPIC $ARCH -c $G/$f.cu -o $OBJ/tomo/$(basename $f).o
  done
  g++ -fPIC -O3 -std=c++17 -w -I$REF/thirdparty/eigen -c $G/container/Matrix3D.cpp -o $OBJ/tomo/Matrix3D.o
  g++ -fPIC -fopenmp -O3 -w -Wno-div-by-zero -shared -std=c++17 $PYB -I$REF/thirdparty/eigen \
      -DASTRA_CUDA -I$A -I$A/include -I$CUDA/include $G/tomoengine.cpp \
      $OBJ/tomo/Matrix3D.o $OBJ/tomo/matrix_ops.o $OBJ/tomo/tv_gd.o $OBJ/tomo/tv_fgp.o \
      -o "$OUT/tomoengine$EXT" -L"$OUT" -lastra -Wl,-rpath,'$ORIGIN' -L$CUDA/lib64 -lcudart
fi

if [ ! -f "$OUT/multimodal$EXT" ]; then
  echo "[build_ref] reference fusion engine multimodal"
  M=$REF/tomofusion/chemistry/utils
  mkdir -p "$OBJ/chem"
  for f in container/matrix_ops regularizers/tv_gd regularizers/tv_fgp; do
    nvcc -w -Xcompiler -fPIC $ARCH -c $M/$f.cu -o $OBJ/chem/$(basename $f).o
  done
  g++ -fPIC -O3 -std=c++17 -w -I$REF/thirdparty/eigen -I$M -c $M/container/Matrix4D.cpp -o $OBJ/chem/Matrix4D.o
  g++ -fPIC -fopenmp -O3 -w -Wno-div-by-zero -shared -std=c++17 $PYB -I$REF/thirdparty/eigen -I$M \
      -DASTRA_CUDA -I$A -I$A/include -I$CUDA/include $M/multimodal.cpp \
      $OBJ/chem/Matrix4D.o $OBJ/chem/matrix_ops.o $OBJ/chem/tv_gd.o $OBJ/chem/tv_fgp.o \
      -o "$OUT/multimodal$EXT" -L"$OUT" -lastra -Wl,-rpath,'$ORIGIN' -L$CUDA/lib64 -lcudart
fi

if [ ! -f "$OUT/ctvlib$EXT" ]; then
  echo "[build_ref] reference CPU engine ctvlib"
  g++ -fPIC -fopenmp -O3 -ffast-math -w -Wno-div-by-zero -shared -std=c++17 $PYB \
      -I$REF/thirdparty/eigen $REF/tomofusion/cpu/utils/ctvlib.cpp -o "$OUT/ctvlib$EXT"
fi
echo "[build_ref] done:"; ls -la "$OUT"/*.so
