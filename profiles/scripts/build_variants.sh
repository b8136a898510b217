# A/B builds of libdh_b200.so with extra -D flags into profiles/trace_build/ (git-ignored, shipped to the GPU box):
#   bash profiles/scripts/build_variants.sh name1 "-DFOO=1" name2 "-DBAR" ...   -> profiles/trace_build/libdh_<name>.so
# built with -DDH_TUNING: these libraries (and only these) read the DH_G_* / DH_HOST_* environment knobs.
# selected at run time with DH_B200_LIB=<path> (tests/_cabi.py load()); measurement builds only.
set -e
cd "$(dirname "$0")/../.."
mkdir -p profiles/trace_build
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared -Iinclude -DDH_TUNING $flags \
    accv-lab_b200/csrc/dh_kernels.cu accv-lab_b200/csrc/dh_host.cu accv-lab_b200/csrc/dh_pack.cu accv-lab_b200/csrc/dh_gaussians.cu \
    -o profiles/trace_build/libdh_$name.so -lcudart &
done
wait
ls -la profiles/trace_build/
