# usage: bash tools/gpu_ncu_step.sh <tag>  -- ncu --set full of the step's kernels at the bench's launch size (2000 pairs)
# (3 warm-up steps + 1 timed step run as two half-batches on two streams = 10 matching launches each: skipped; captured
# are the 5 whole-batch launches of the first one-stream step bench.py runs for its per-kernel durations)
cd $GRAFT_REPO_ROOT
T=$1
mkdir -p gpurun_out
CMD="python bench.py --no-cpu-baseline --no-extras --steps 1 --warmup 3 --e2e-steps 1"
timeout 300 $CMD > gpurun_out/${T}_plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'response_kernel|response_multi|orient_describe|select_kernel|match_mma' -s 40 -c 5 -o gpurun_out/${T}_step $CMD > gpurun_out/${T}_ncu.log 2>&1
tail -3 gpurun_out/${T}_ncu.log; ls -la gpurun_out/${T}_step.ncu-rep
