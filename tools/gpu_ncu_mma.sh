cd $GRAFT_REPO_ROOT
T=$1
mkdir -p gpurun_out
timeout 300 python tools/probes/mma_clocks.py 592 > gpurun_out/${T}_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:match_mma -s 1 -c 1 -o gpurun_out/${T}_mma python tools/probes/mma_clocks.py 592 > gpurun_out/${T}_ncu.log 2>&1
tail -5 gpurun_out/${T}_ncu.log; ls -la gpurun_out/${T}_mma.ncu-rep
