#!/usr/bin/env bash
# One multi-GPU measurement pass on N GPUs of this box: correctness of the peer-memory engines against the NCCL oracle,
# then bench.py for the BASELINE.json configurations (both arms where the reference runs).  Every line of
# gpurun_out/scale_N.jsonl is one bench.py JSON record, prefixed records carry the "case" key.
#   usage: tools/run_scale_suite.sh N [full|short]
set -u
N="${1:-8}"
MODE="${2:-full}"
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
OUT="gpurun_out/scale_${N}.jsonl"
: > "$OUT"
PORT=29600
run() {   # run <case-name> <bench args...>
    local name="$1"; shift
    PORT=$((PORT + 1))
    local line
    if [ "$N" -gt 1 ]; then
        line=$(timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 \
               --master-port "$PORT" bench.py --gpus "$N" "$@" 2>>gpurun_out/scale_${N}.err | grep '^{' | tail -1)
    else
        line=$(timeout 600 python bench.py --gpus 1 "$@" 2>>gpurun_out/scale_${N}.err | grep '^{' | tail -1)
    fi
    if [ -n "$line" ]; then
        python - "$name" "$line" >> "$OUT" <<'PY'
import json, sys
d = json.loads(sys.argv[2]); d["case"] = sys.argv[1]; print(json.dumps(d))
PY
        python - "$name" "$line" <<'PY'
import json, sys
d = json.loads(sys.argv[2])
if "unavailable" in d:
    print(f"{sys.argv[1]:46s} unavailable: {d['unavailable'][:80]}")
else:
    e = d.get("e2e") or {}
    print(f"{sys.argv[1]:46s} {d['value']:10.0f} img/s {d['ms_per_step']:8.3f} ms/step  e2e {e.get('value', 0):10.0f}  cc {d.get('compress_comm_ms_per_step', float('nan')):.3f} ms")
PY
    else
        echo "$name FAILED" | tee -a "$OUT"
    fi
}
if [ "$N" -gt 1 ]; then
    for extra in "" "--nvls"; do
        PORT=$((PORT + 1))
        timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port "$PORT" \
            tools/check_multigpu.py $extra > "gpurun_out/multigpu${N}_check${extra#--}.json" 2>>gpurun_out/scale_${N}.err
        echo "check_multigpu $extra rc=$?"
    done
    PORT=$((PORT + 1))
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port "$PORT" \
        tools/check_graph_multigpu.py > "gpurun_out/multigpu${N}_graph_check.json" 2>>gpurun_out/scale_${N}.err
    echo "check_graph_multigpu rc=$?"
fi
S="--steps 10 --warmup 3"
R="--steps 3 --warmup 3"
run "resnet50 topk1% layerwise EF (headline)"      $S --phase-times
run "resnet50 topk1% layerwise no-EF"               $S --no-ef --skip-e2e
run "resnet50 topk1% layerwise EF nvls"             $S --nvls --skip-e2e
run "REF resnet50 topk1% layerwise"                 --impl reference $R
run "REF resnet50 compress none (per-tensor NCCL)"  --impl reference --compress none $R --skip-e2e
for ratio in 0.001 0.01 0.1; do
    for mode in layerwise entiremodel; do
        run "resnet50 topk ${ratio} ${mode}"        $S --ratio $ratio --compress $mode --skip-e2e
    done
done
if [ "$MODE" = "full" ]; then
    run "REF resnet50 topk 0.1% layerwise"          --impl reference --ratio 0.001 $R --skip-e2e
    run "REF resnet50 topk 10% layerwise"           --impl reference --ratio 0.1 $R --skip-e2e
    for mode in layerwise entiremodel; do
        run "vgg16 randomk1% ${mode}"               --arch vgg16 --method Randomk --compress $mode $S
        run "alexnet QSGD-8bit ${mode}"             --arch alexnet --method RandomDithering --compress $mode $S
        run "alexnet TernGrad ${mode}"              --arch alexnet --method TernGrad --compress $mode $S
    done
    run "resnet9 topk1% layerwise"                  --arch resnet9 $S
    run "REF vgg16 randomk1% layerwise"             --impl reference --arch vgg16 --method Randomk $R
    run "REF vgg16 randomk1% entiremodel"           --impl reference --arch vgg16 --method Randomk --compress entiremodel $R
    run "REF alexnet QSGD-8bit layerwise"           --impl reference --arch alexnet --method RandomDithering $R
    run "REF alexnet TernGrad layerwise"            --impl reference --arch alexnet --method TernGrad $R
    run "REF resnet9 topk1% layerwise"              --impl reference --arch resnet9 $R
fi
echo "suite done: $OUT"
