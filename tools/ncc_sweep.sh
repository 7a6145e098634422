# Development sweep on the GPU box: rebuild ncc.cu with different band heights / occupancy targets and time it.
set -e
mkdir -p gpurun_out/ncc
for cfg in "16 1" "16 24" "16 32" "24 1" "24 32" "28 1" "28 32" "32 1" "32 32" "56 32"; do
  set -- $cfg
  touch cardiac-segmentation-optical-flow_b200/csrc/ncc.cu
  CF_NVCC_EXTRA="-DNCC_ROWS=$1 -DNCC_MINB=$2" python cardiac-segmentation-optical-flow_b200/_build.py > /dev/null
  echo "== NCC_ROWS=$1 NCC_MINB=$2"
  python tools/kbench.py cfg3 --only ncc --graph 2>&1 | grep "ncc "
done
touch cardiac-segmentation-optical-flow_b200/csrc/ncc.cu
python cardiac-segmentation-optical-flow_b200/_build.py > /dev/null
