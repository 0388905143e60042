cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r2a_smi.txt 2>&1
timeout 120 tools/probes/bin/umma_hamming > gpurun_out/r2a_umma.jsonl 2> gpurun_out/r2a_umma.err; echo "umma rc=$?" >> gpurun_out/r2a_umma.err
timeout 200 tools/probes/bin/copy_probe 5 > gpurun_out/r2a_copy.jsonl 2> gpurun_out/r2a_copy.err
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2a_ref.json 2> gpurun_out/r2a_ref.err
timeout 300 python bench.py > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err
tail -3 gpurun_out/r2a_pytest.log; cat gpurun_out/r2a_umma.jsonl gpurun_out/r2a_umma.err gpurun_out/r2a_copy.jsonl
