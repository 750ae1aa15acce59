# Build a library variant for A/B runs on the GPU box: bash tools/build_variant.sh NAME [extra nvcc flags...]  ->  tools/ab/NAME.so
NAME=$1; shift
cd "$(dirname "$0")/.."
nvcc -std=c++17 -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo --fmad=true --prec-div=true --prec-sqrt=true --ftz=false \
  -shared -Xcompiler -fPIC,-fvisibility=hidden --threads 2 "$@" -o tools/ab/$NAME.so \
  python-ray-tracer_b200/csrc/b200rt.cu python-ray-tracer_b200/csrc/isect_shadow.cu
cuobjdump -sass tools/ab/$NAME.so | awk '/Function :/{name=$3} /^ +\/\*[0-9a-f]{4}\*\//{cnt[name]++} END{for(n in cnt) if (n ~ /k_isect3/ && n ~ /Li8E/) print cnt[n], n}' | sort -rn
