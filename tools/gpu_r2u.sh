cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_detect_describe.py tests/test_gpu_golden.py tests/test_gpu_geometry.py tests/test_cpp_dropin.py -x -q -m gpu > gpurun_out/r2u_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2u_pytest.log
tail -4 gpurun_out/r2u_pytest.log
timeout 200 python tools/probes/single_pair.py > gpurun_out/r2u_single_pair.json 2> gpurun_out/r2u_single_pair.err; cat gpurun_out/r2u_single_pair.json; tail -2 gpurun_out/r2u_single_pair.err
