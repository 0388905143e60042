# usage: bash tools/gpu_mma_iter.sh <tag>   -- matcher iteration: parity tests, cycle counters, bench line
cd $GRAFT_REPO_ROOT
T=$1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_match_mma.py tests/test_gpu_match.py -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
tail -3 gpurun_out/${T}_pytest.log
VSLAM_LIB=$GRAFT_REPO_ROOT/tools/probes/bin/libvisnav_b200_prof.so timeout 300 python tools/probes/mma_clocks.py 592 > gpurun_out/${T}_mma_clocks.json 2> gpurun_out/${T}_mma_clocks.err; cat gpurun_out/${T}_mma_clocks.json; tail -3 gpurun_out/${T}_mma_clocks.err
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; tail -3 gpurun_out/${T}_bench.err
python -c "
import json; b=json.load(open('gpurun_out/${T}_bench.json')); print(b['value'], b['e2e']['value'], b['stage_ms'], b['roofline_match']['distances_per_s'])"
