#!/bin/bash
# tools/build_variant.sh NAME "-DFLAG=.. ..." : rebuilds geo.cu with extra flags and links csrc/variants/libgraphmet_NAME.so
set -e
cd "$(dirname "$0")/../deepmetv2_b200/csrc"
mkdir -p variants build
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -I../../include --expt-relaxed-constexpr"
SRC=${3:-geo}
$NV $2 -c $SRC.cu -o build/${SRC}_$1.o
OBJS=""
for f in radius_graph net_forward net_backward edgeconv_tc loss_opt geo geo_tc knn comm microbench drn_graph drn_net metrics; do
  if [ "$f" == "$SRC" ]; then OBJS="$OBJS build/${SRC}_$1.o"; else OBJS="$OBJS build/$f.o"; fi
done
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o variants/libgraphmet_$1.so $OBJS -lcudart
echo built variants/libgraphmet_$1.so
