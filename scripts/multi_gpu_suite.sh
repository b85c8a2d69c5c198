#!/bin/bash
# Everything that needs N > 1 GPUs in one call (N = $1): parity checks over real peers, the bench line,
# the dense exchange sweep (fused kernel and NCCL arm) and the sparse exchange.  Output: gpurun_out/r2_n$N_*.
# $2: --quick (short sweep) | --no-sweep (no dense exchange sweep);  $3: extra bench.py flags (e.g. --no-resnet)
N=$1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
O=gpurun_out
$TR --master-port 29601 scripts/mgpu_check.py balanced tf32x3 2>&1 | grep "MGPU_CHECK\|step" > $O/r2_n${N}_mgpu_check.log
$TR --master-port 29602 scripts/mgpu_check.py logical tf32x3 2>&1 | grep "MGPU_CHECK\|step" >> $O/r2_n${N}_mgpu_check.log
$TR --master-port 29603 scripts/mgpu_sparse_check.py --bench 2>&1 | grep "^{\|MGPU\|sparse" > $O/r2_n${N}_sparse.log
$TR --master-port 29604 bench.py --gpus $N --steps 30 --warmup 5 $3 2>$O/r2_n${N}_bench.err | grep "^{" > $O/r2_n${N}_bench.json
if [ "$2" != "--no-sweep" ]; then
  $TR --master-port 29606 scripts/ps_exchange_sweep.py $2 --out $O/r2_n${N}_ps_exchange.jsonl > /dev/null 2>&1
fi
cat $O/r2_n${N}_mgpu_check.log; grep MGPU $O/r2_n${N}_sparse.log
python - <<PY
import json
for f in ("$O/r2_n${N}_bench.json",):
    try:
        d = json.load(open(f))
        print(f, "ms/step %.4f value %.0f e2e %.4f parity %s" % (d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"], d.get("parity", {}).get("ok")),
              "resnet %s" % (d.get("resnet50", {}).get("value"),))
    except Exception as e:
        print(f, "failed", e)
for f in ("$O/r2_n${N}_ps_exchange.jsonl",):
    try:
        for l in open(f):
            d = json.loads(l)
            if d["optimizer"] == "momentum" and (d["layout"] == "balanced" or d["shards"] in (1, 16)):
                print(f.split("_")[-1][:8], d["arm"], d["params"], d["shards"], d["layout"], "ms %.3f link %.0f GB/s frac %.2f" % (d["ms"], d["link_gbs_per_dir"], d["link_frac"]), d.get("ranks_bit_identical"), d.get("max_rel_diff_vs_peer"))
    except Exception as e:
        print(f, "failed", e)
PY
