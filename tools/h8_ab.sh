# A/B of site_hist8_kernel's register cap on the GPU box: bash tools/h8_ab.sh
for mb in "$@"; do
  bash tools/ab_build.sh "-DBVC_H8_MINBLOCKS=$mb"
  echo "minblocks $mb: $(python tools/prof_run.py C2 1000000 5 | tail -1 | cut -c1-120)"
done
bash tools/ab_build.sh ""
