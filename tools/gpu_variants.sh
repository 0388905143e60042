# usage: bash tools/gpu_variants.sh <tag> <lib suffix>...  -- bench line per library variant under tools/probes/bin
cd $GRAFT_REPO_ROOT
T=$1; shift
mkdir -p gpurun_out
for V in "$@"; do
  VSLAM_LIB=$GRAFT_REPO_ROOT/tools/probes/bin/libvisnav_b200_$V.so timeout 300 python bench.py --no-cpu-baseline --no-extras --e2e-steps 1 > gpurun_out/${T}_$V.json 2> gpurun_out/${T}_$V.err
  python -c "
import json; b=json.load(open('gpurun_out/${T}_$V.json')); print('$V', b['value'], b['stage_ms'])"
done
