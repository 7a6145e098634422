# Development sweep on the GPU box: rebuild ncc.cu with different knobs (-DNCC_ROWS band height, -DNCC_MINB CTAs/SM target,
# -DNCC_PF rows of look-ahead) and time it.  Usage: bash tools/ncc_sweep.sh "16 1 1" "16 1 2" ...
set -e
mkdir -p gpurun_out/ncc
[ $# -eq 0 ] && set -- "16 1 1" "16 1 2" "16 24 2" "32 1 2"
for cfg in "$@"; do
  set -- $cfg
  touch cardiac-segmentation-optical-flow_b200/csrc/ncc.cu
  CF_NVCC_EXTRA="-DNCC_ROWS=$1 -DNCC_MINB=$2 -DNCC_PF=$3" python cardiac-segmentation-optical-flow_b200/_build.py > /dev/null
  echo "== NCC_ROWS=$1 NCC_MINB=$2 NCC_PF=$3"
  python tools/kbench.py cfg3 --only ncc --graph 2>&1 | grep "ncc "
done
touch cardiac-segmentation-optical-flow_b200/csrc/ncc.cu
python cardiac-segmentation-optical-flow_b200/_build.py > /dev/null
