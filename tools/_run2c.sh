N=${1:-2}
timeout 500 python -m pytest tests/test_gpu_shard.py tests/test_gpu_peer.py -x -q -s -k "shard" 2>&1 | grep -v "^$" | tail -8
B="--no-train --no-config5 --no-e2e --no-cpu-baseline"
run() { tag=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 50 --warmup 10 $B "$@" > gpurun_out/s4_n${N}_$tag.log 2>&1; python - gpurun_out/s4_n${N}_$tag.log <<'PY'
import sys,json
ok=False
for l in open(sys.argv[1]):
    if l.startswith('{"metric"'):
        d=json.loads(l); ok=True; r=d["roofline"]; print(sys.argv[1], "ms/step", round(d["ms_per_step"],4), "kernel", round(r["kernel_ms_avg"],4), "parity", d.get("parity",{}).get("all_ranks_ok"), r.get("weights_pushed_by"), r.get("multicast_note"))
if not ok: print(sys.argv[1], "NO LINE"); print(open(sys.argv[1]).read()[-1500:])
PY
}
run sharded_both --step-mode sharded --multicast both
run sharded_reduce --step-mode sharded --multicast reduce
run sharded_none --step-mode sharded --multicast none
