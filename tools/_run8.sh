N=${1:-8}
timeout 400 python -m pytest tests/test_gpu_peer.py -x -q -s -k "sharded_step and $N and (none or both)" 2>&1 | grep -v "^$" | tail -6
B="--no-train --no-config5 --no-e2e --no-cpu-baseline"
run() { tag=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 50 --warmup 10 $B "$@" > gpurun_out/s5_n${N}_$tag.log 2>&1; python - gpurun_out/s5_n${N}_$tag.log <<'PY'
import sys,json
ok=False
for l in open(sys.argv[1]):
    if l.startswith('{"metric"'):
        d=json.loads(l); ok=True; r=d.get("roofline",{}); print(sys.argv[1], "ms/step", round(d.get("ms_per_step",0),4), "kernel", round(r.get("kernel_ms_avg",0),4), "parity", d.get("parity",{}).get("all_ranks_ok"), d.get("parity",{}).get("mean_max_ulp_vs_oracle"), d.get("error"))
if not ok: print(sys.argv[1], "NO LINE"); print(open(sys.argv[1]).read()[-1500:])
PY
}
run auto
run sharded_none --step-mode sharded --multicast none
