#!/bin/bash
# development build of the library with the blocked Cholesky's stage clocks (-DDPVO_CHOL_PROFILE) as
# deep-visual-odometry-slam_b200/libdpvo_b200_prof.so; load it with DPVO_B200_LIB=<path> (tools/chol_stage_profile.py)
set -e
cd "$(dirname "$0")/.."
PKG=deep-visual-odometry-slam_b200
OBJ=/tmp/dpvo_prof_obj; mkdir -p $OBJ
pids=()
for f in $PKG/csrc/*.cu; do
  b=$(basename $f .cu)
  /usr/local/cuda/bin/nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -I include -DDPVO_CHOL_PROFILE -c $f -o $OBJ/$b.o &
  pids+=($!)
done
for p in "${pids[@]}"; do wait $p; done
/usr/local/cuda/bin/nvcc -shared -o $PKG/libdpvo_b200_prof.so $OBJ/*.o -lcudart
echo built $PKG/libdpvo_b200_prof.so
