#!/bin/bash
# tools/gpu_scale.sh TAG: the scaling evidence run on one 8-GPU box (gpurun --gpus 8).
#   bench.py at N = 1, 2, 4, 8 (the default workload c4 = the 100k-event tree of configs 4 and 5, the driver's launch line, each
#   with its in-run tally_check: N-way == 1-way, prefix == oracle), config 5 in full, and the GPU tests (run_multi uses every
#   visible device).
tag=${1:-r01}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm --format=csv > $out/${tag}_smi8.txt 2>&1
timeout 600 python bench.py --gpus 1 --steps ${STEPS:-5} --warmup 3 --no-study > $out/${tag}_scale_c4_n1.json 2> $out/${tag}_scale_c4_n1.err; echo "n1 rc=$?"
port=29511
for n in ${SCALE_NS:-2 4 8}; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus $n --steps ${STEPS:-5} --warmup 3 --no-study > $out/${tag}_scale_c4_n$n.json 2> $out/${tag}_scale_c4_n$n.err; echo "n$n rc=$?"
  port=$((port+1))
done
# config 5 in full: 2^36 samples of the 100k-event tree over the 8 GPUs (about 200 s per GPU)
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 \
    tools/run_c5.py --log2-total ${C5_LOG2:-36} > $out/${tag}_c5_full_n8.json 2> $out/${tag}_c5_full_n8.err; echo "c5 rc=$?"; tail -n 1 $out/${tag}_c5_full_n8.json
timeout 600 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest_gpu8.log 2>&1; echo "pytest rc=$?"
tail -n 3 $out/${tag}_pytest_gpu8.log
for f in $out/${tag}_scale_*.json; do echo "== $f"; grep -o '"value": [0-9.e+]*, "unit": "samples/s", "n_gpus": [0-9]*' $f | head -1; done
