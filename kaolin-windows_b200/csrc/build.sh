#!/bin/sh
# Builds libdibr_b200.so (sm_100a only) next to the Python package.
set -e
cd "$(dirname "$0")"
OUT=${DIBR_OUT:-../libdibr_b200.so}
FLAGS="-std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC,-fvisibility=hidden"
pids=""
for f in dibr_api dibr_binning dibr_raster dibr_softmask dibr_softgather dibr_mesh dibr_f64 dibr_epilogue dibr_deftet dibr_peer; do
  nvcc $FLAGS $NVCC_EXTRA -c $f.cu -o /tmp/dibr_$$_$f.o &
  pids="$pids $!"
done
for p in $pids; do wait $p; done
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o $OUT.tmp /tmp/dibr_$$_dibr_api.o /tmp/dibr_$$_dibr_binning.o /tmp/dibr_$$_dibr_raster.o /tmp/dibr_$$_dibr_softmask.o /tmp/dibr_$$_dibr_softgather.o /tmp/dibr_$$_dibr_mesh.o /tmp/dibr_$$_dibr_f64.o /tmp/dibr_$$_dibr_epilogue.o /tmp/dibr_$$_dibr_deftet.o /tmp/dibr_$$_dibr_peer.o
mv $OUT.tmp $OUT
rm -f /tmp/dibr_$$_*.o
echo built $OUT
