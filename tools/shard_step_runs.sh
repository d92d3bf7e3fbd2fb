#!/bin/bash
# Reproduces the sharded-step table of profiles/r02_summary.md on an N-GPU box:  bash tools/shard_step_runs.sh N
# (micro-benchmark sections only; every run starts with the parity block and prints one JSON line into gpurun_out/).
N=${1:-2}
B="--no-train --no-config5 --no-e2e --no-cpu-baseline --steps 50 --warmup 10"
mkdir -p gpurun_out
run() {
    tag=$1; shift
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port 29511 \
        bench.py --gpus "$N" $B "$@" > "gpurun_out/shard_n${N}_$tag.log" 2>&1
    python - "gpurun_out/shard_n${N}_$tag.log" <<'PY'
import json, sys
for line in open(sys.argv[1]):
    if line.startswith('{"metric"'):
        d = json.loads(line)
        print(sys.argv[1], "ms/step", round(d.get("ms_per_step", 0), 4), "parity", d.get("parity", {}).get("all_ranks_ok"),
              "mean ulps", d.get("parity", {}).get("mean_max_ulp_vs_oracle"), d.get("error"))
        break
else:
    print(sys.argv[1], "no JSON line")
PY
}
run default                                            # sharded; multicast from 4 ranks on
run sharded_none --step-mode sharded --multicast none  # per-rank loads / stores: bit-exact mean
run sharded_both --step-mode sharded --multicast both
run replicated --step-mode spec                        # the fused replicated-state step
run two_kernels --step-mode checked                    # round 1's reduce-scatter kernel + pulling update
# where the time goes (results are wrong by construction, hence --no-parity):
B200_SHARD_DEBUG=1 run dbg_no_remote_stores --step-mode sharded --multicast none --no-parity
B200_SHARD_DEBUG=2 run dbg_no_remote_loads --step-mode sharded --multicast none --no-parity
B200_SHARD_DEBUG=3 run dbg_all_local --step-mode sharded --multicast none --no-parity
