# development aid: A/B two builds of the kernels in ONE GPU call.
#   usage (here):  bash scripts/ab.sh "-DMEMB_X=0" "-DMEMB_X=1"   -> builds memb_b200/variants/{a,b}/
#   then on the GPU box: bash scripts/ab_run.sh [bench args]
set -e
i=0
for defs in "$@"; do
  name=$(echo abcdefgh | cut -c$((i+1)))
  MEMB_B200_NVCC_DEFS="$defs" python build_native.py --force > /dev/null 2>&1
  mkdir -p memb_b200/variants/$name
  cp memb_b200/libmemb_b200.so memb_b200/variants/$name/
  echo "$defs" > memb_b200/variants/$name/defs.txt
  cuobjdump -res-usage memb_b200/libmemb_b200.so 2>/dev/null | grep -A1 "trained_decode_kernelILi6ELb1ELb0ELb1ELb0" | tail -1
  i=$((i+1))
done
