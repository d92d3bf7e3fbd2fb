N=${1:-4}
B="--no-train --no-config5 --no-e2e --no-cpu-baseline"
run() { tag=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 50 --warmup 10 $B "$@" > gpurun_out/s4_n${N}_$tag.log 2>&1; python - gpurun_out/s4_n${N}_$tag.log <<'PY'
import sys,json
ok=False
for l in open(sys.argv[1]):
    if l.startswith('{"metric"'):
        d=json.loads(l); ok=True; r=d["roofline"]; print(sys.argv[1], "ms/step", round(d["ms_per_step"],4), "kernel", round(r["kernel_ms_avg"],4), "parity", d.get("parity",{}).get("all_ranks_ok"), d.get("parity",{}).get("mean_max_ulp_vs_oracle"))
if not ok: print(sys.argv[1], "NO LINE"); print(open(sys.argv[1]).read()[-1500:])
PY
}
run sharded_both --step-mode sharded --multicast both
run sharded_both_u1 --step-mode sharded --multicast both --unroll 1
