#!/usr/bin/env bash
# Installs the UNMODIFIED reference under baseline/_ref (git-ignored; it travels to the GPU box).
#
# 1. The prescribed `pip install ... --target baseline/_ref /root/reference` cannot work: the reference
#    is a C project with no setup.py / pyproject.toml ("Directory '/root/reference' is not installable").
# 2. So the sources are copied verbatim and built with the reference's OWN Makefile commands:
#       gcc -o cnn cnn.c -lm                      (Makefile:40-41)   -> builds
#       mpicc -o cnnmpi cnnmpi.c -lm              (Makefile:45-46)   -> no MPI in this image; built against
#                                                                       baseline/mpi_shim/mpi.h (our 5-call shm MPI)
#       nvcc -o CUDAcnn CUDAcnn.cu -lm            (Makefile:50-51)   -> 5 compile errors in the reference
#       CUDAMPI.c/.cu                             (no Makefile rule) -> compile errors in the reference
#    The outcome of every build is written to baseline/_ref/BUILD_STATUS.txt.
set -u
HERE="$(cd "$(dirname "$0")" && pwd)"
REF_SRC="${1:-/root/reference}"
DST="$HERE/_ref"
mkdir -p "$DST/src"
if [ -d "$REF_SRC" ]; then cp -f "$REF_SRC"/* "$DST/src/" 2>/dev/null; fi
cd "$DST/src" || exit 1
: > "$DST/BUILD_STATUS.txt"
note() { echo "$1" | tee -a "$DST/BUILD_STATUS.txt"; }

python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target "$DST" "$REF_SRC" \
    > "$DST/pip_install.log" 2>&1 && note "pip install: ok" || note "pip install: FAILED ($(tail -1 "$DST/pip_install.log"))"

gcc -o "$DST/cnn" cnn.c -lm > "$DST/cnn.build.log" 2>&1 && note "cnn (Makefile flags): ok" || note "cnn: FAILED"
gcc -O2 -o "$DST/cnn_O2" cnn.c -lm >> "$DST/cnn.build.log" 2>&1 && note "cnn_O2 (-O2): ok" || note "cnn_O2: FAILED"
if [ -f "$HERE/mpi_shim/mpi.h" ]; then
    gcc -I"$HERE/mpi_shim" -o "$DST/cnnmpi" cnnmpi.c "$HERE/mpi_shim/mpi_shim.c" -lm -lrt > "$DST/cnnmpi.build.log" 2>&1 \
        && note "cnnmpi (unmodified source, in-repo shm mpi.h shim): ok" || note "cnnmpi: FAILED ($(grep -m1 error "$DST/cnnmpi.build.log"))"
fi
if command -v nvcc > /dev/null; then
    nvcc -o "$DST/CUDAcnn" CUDAcnn.cu -lm > "$DST/CUDAcnn.build.log" 2>&1 && note "CUDAcnn: ok" \
        || note "CUDAcnn: FAILED to compile ($(grep -c 'error' "$DST/CUDAcnn.build.log") error lines; first: $(grep -m1 'error' "$DST/CUDAcnn.build.log"))"
    nvcc -I. -c -o "$DST/CUDAMPI.o" CUDAMPI.cu > "$DST/CUDAMPI.build.log" 2>&1 && note "CUDAMPI.cu: ok" \
        || note "CUDAMPI.cu: FAILED to compile ($(grep -c 'error' "$DST/CUDAMPI.build.log") error lines; first: $(grep -m1 'error' "$DST/CUDAMPI.build.log"))"
fi
cat "$DST/BUILD_STATUS.txt" > /dev/null
