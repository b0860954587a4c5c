#!/bin/bash
# tools/gpu_round.sh TAG: the round's evidence run on one B200 (call through gpurun).
#   1. pytest -m gpu      2. smoke()      3. bench.py plain      4. ncu launch list of the same bench command
#   5. ncu --set full of one gate_ring_kernel and one sampler launch on c2 and c4 (tools/ncu_target.py); the sampler
#      capture runs before the bench so that bench.py quotes this build's executed instruction counts
# Everything lands in gpurun_out/TAG_*; nothing printed under ncu is a bench value.
tag=${1:-r01}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/${tag}_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $out/${tag}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $out/${tag}_smoke.log
# sampler capture first: bench.py quotes the executed instruction counts of this build (profiles/sampler_opcounts.json)
if [ -z "$SKIP_NCU" ]; then
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sampler_sliced -s 1 -c 1 -f -o $out/${tag}_sampler_c2 \
    python tools/ncu_target.py c2 > $out/${tag}_ncu_sampler_c2.log 2>&1; echo "ncu sampler rc=$?"
python tools/sampler_opcounts.py $out/${tag}_sampler_c2.ncu-rep 8992 592 $out/${tag}_ncu_source_sampler_c2_opcodes.csv profiles/sampler_opcounts.json 3
cp profiles/sampler_opcounts.json $out/${tag}_sampler_opcounts.json
fi
timeout 900 python bench.py --steps 3 --warmup 3 > $out/${tag}_bench_c2.json 2> $out/${tag}_bench_c2.err; echo "bench rc=$?"
[ -n "$WITH_REFERENCE" ] && { timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $out/${tag}_reference_arm.json 2>&1; echo "reference arm rc=$?"; }
tail -c 3000 $out/${tag}_bench_c2.json
if [ -z "$SKIP_NCU" ]; then
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/${tag}_ncu_launches_bench_c2.csv \
    python bench.py --steps 2 --warmup 3 --no-study > $out/${tag}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
for cfg in c2 c4; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:gate_ring_kernel -s 1 -c 1 -f -o $out/${tag}_gate_$cfg \
      python tools/ncu_target.py $cfg > $out/${tag}_ncu_gate_$cfg.log 2>&1; echo "ncu gate $cfg rc=$?"
done
fi
ls -la $out | tail -20
