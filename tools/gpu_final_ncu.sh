# usage: bash tools/gpu_final_ncu.sh <tag>  -- launch list of a 2-step bench run and ncu --set full of the step's kernels
cd $GRAFT_REPO_ROOT
T=$1
mkdir -p gpurun_out
CMD="python bench.py --no-cpu-baseline --no-extras --steps 2 --warmup 3 --e2e-steps 1"
timeout 300 $CMD > gpurun_out/${T}_plain2.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv $CMD > gpurun_out/${T}_ncu_launches.log 2>&1
tail -2 gpurun_out/${T}_ncu_launches.log
bash tools/gpu_ncu_step.sh $T
