cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_match_mma.py tests/test_gpu_detect_describe.py -x -q > gpurun_out/r2d_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2d_pytest.log
tail -4 gpurun_out/r2d_pytest.log
VSLAM_LIB=$GRAFT_REPO_ROOT/tools/probes/bin/libvisnav_b200_prof.so timeout 300 python tools/probes/mma_clocks.py 592 > gpurun_out/r2d_mma_clocks.json 2> gpurun_out/r2d_mma_clocks.err; cat gpurun_out/r2d_mma_clocks.json; tail -3 gpurun_out/r2d_mma_clocks.err
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; tail -3 gpurun_out/r2d_bench.err
python -c "
import json; b=json.load(open('gpurun_out/r2d_bench.json')); print(b['value'], b['e2e']['value'], b['stage_ms'], b['roofline_match']['distances_per_s'])"
