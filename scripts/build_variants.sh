# usage: bash scripts/build_variants.sh tag1 "-DF32_KT=64 ..." tag2 "..."  -- libmaria_atm.so variants differing in the
# compile-time tuning of matm_synth.cu, written to maria_b200/variants/ (git-ignored, travels to the GPU box)
set -e
cd "$(dirname "$0")/../maria_b200"
mkdir -p variants
while [ $# -ge 2 ]; do
  tag=$1; flags=$2; shift 2
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC $flags -c csrc/matm_synth.cu -o variants/matm_synth_$tag.o
  nvcc -shared -o variants/libmaria_atm_$tag.so csrc/matm_core.o csrc/matm_ar.o csrc/matm_smooth.o csrc/matm_setup.o csrc/matm_consumers.o variants/matm_synth_$tag.o -lcudart
  echo "built variants/libmaria_atm_$tag.so ($flags)"
done
