cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r2v_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2v_pytest.log
tail -5 gpurun_out/r2v_pytest.log
timeout 200 python tools/probes/single_pair.py > gpurun_out/r2v_single_pair.json 2> gpurun_out/r2v_single_pair.err; cat gpurun_out/r2v_single_pair.json; tail -2 gpurun_out/r2v_single_pair.err
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2v_smoke.log 2>&1; tail -2 gpurun_out/r2v_smoke.log
