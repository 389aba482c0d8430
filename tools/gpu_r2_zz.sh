#!/bin/bash
# round-2 closing evidence on the final tree: other BASELINE configs, launch list of the headline bench, launch list of the 64-utterance solve
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
bash tools/gpu_r2_cfgs.sh > /dev/null 2>&1; cat gpurun_out/r02_z_configs.log
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
echo "launch list rows: $(wc -l < gpurun_out/launches.csv)"
CBX_CFM_GRAPH=0 timeout 300 python tools/cfm_batch_time.py 64 > gpurun_out/b64_plain.log 2>&1 &&
CBX_CFM_GRAPH=0 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 1000 --launch-count 340 --csv --log-file gpurun_out/cfm_b64_launches.csv python tools/cfm_batch_time.py 64 > gpurun_out/ncu_b64.log 2>&1
echo "b64 launch list rows: $(wc -l < gpurun_out/cfm_b64_launches.csv)"
