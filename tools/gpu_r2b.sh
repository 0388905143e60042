cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_match_mma.py -x -q > gpurun_out/r2b_pytest_mma.log 2>&1; echo "pytest mma rc=$?" >> gpurun_out/r2b_pytest_mma.log
tail -15 gpurun_out/r2b_pytest_mma.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2b_pytest.log
tail -5 gpurun_out/r2b_pytest.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; tail -3 gpurun_out/r2b_bench.err
python -c "
import json; b=json.load(open('gpurun_out/r2b_bench.json')); print(b['value'], b['e2e']['value'], b['stage_ms'], b['roofline_match']['distances_per_s'])"
