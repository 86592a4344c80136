#!/bin/bash
# Multi-GPU pass (run under `gpurun --gpus 8`): BASELINE config 4 (8192 images sharded, strong scaling) at N = 1, 2, 4, 8 and
# config 5 (the defense sweep on 4096 synthetic images) on 8 ranks and on 1 rank — the counters must be bit-equal.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
CLI=image-recognition-adversarial-example-attack_b200/cli/defense_experiments.py
SWEEP="--synthetic 4096 --weights random --attacks fgsm pgd --eps_list 0.01568627451 0.03137254902 --output_dir /tmp/dr"
$TR --nproc-per-node 8 --master-port 29611 $CLI $SWEEP > gpurun_out/r2_config5_n8.log 2> gpurun_out/r2_config5_n8.err; echo "config5 n8 rc=$?"
python $CLI $SWEEP > gpurun_out/r2_config5_n1.log 2> gpurun_out/r2_config5_n1.err; echo "config5 n1 rc=$?"
grep counters gpurun_out/r2_config5_n8.log > /tmp/c8; grep counters gpurun_out/r2_config5_n1.log > /tmp/c1
if cmp -s /tmp/c8 /tmp/c1 && [ -s /tmp/c8 ]; then echo "config5 counters: 8 ranks == 1 rank (bit-equal)" | tee gpurun_out/r2_config5_equal.txt; else echo "config5 counters DIFFER" | tee gpurun_out/r2_config5_equal.txt; diff /tmp/c8 /tmp/c1 | tee -a gpurun_out/r2_config5_equal.txt; fi
tail -n 6 gpurun_out/r2_config5_n8.log
for n in 8 4 2; do
  $TR --nproc-per-node $n --master-port $((29620 + n)) bench.py --config shard8192 --gpus $n --steps 5 --warmup 3 > gpurun_out/r2_shard8192_n$n.json 2> gpurun_out/r2_shard8192_n$n.err; echo "shard8192 n$n rc=$?"
done
python bench.py --config shard8192 --gpus 1 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2_shard8192_n1.json 2> gpurun_out/r2_shard8192_n1.err; echo "shard8192 n1 rc=$?"
for n in 8 2; do
  $TR --nproc-per-node $n --master-port $((29640 + n)) bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/r2_bench_n$n.json 2> gpurun_out/r2_bench_n$n.err; echo "bench n$n rc=$?"
done
python - <<'PY'
import json
for n in (1, 2, 4, 8):
    try:
        d = json.loads(open(f"gpurun_out/r2_shard8192_n{n}.json").read().strip().splitlines()[-1])
        print(n, "shard8192", round(d["value"]), "img/s  e2e", round(d["e2e"]["value"]), d["scaling"], d["config"]["attack_success"])
    except Exception as e:
        print(n, "shard8192 failed", e)
PY
