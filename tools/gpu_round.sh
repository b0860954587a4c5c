#!/bin/bash
# tools/gpu_round.sh TAG [steps]: one round of evidence on one B200 (call through gpurun).
#   1. pytest -m gpu   2. smoke()   3. bench.py (default workload c4, with tally_check)   4. the CPU arm
#   with NCU=1 also: 5. ncu launch list of the same bench command   6. ncu --set full of one gate-walk and one sampler
#   launch on c2 and c4 (tools/ncu_target.py) -> tools/captures.py writes profiles/captures.json for this build
# Everything lands in gpurun_out/TAG_*; nothing printed under ncu is a bench value.
tag=${1:-r02}
steps=${2:-5}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/${tag}_smi.txt 2>&1
if [ -z "$SKIP_TESTS" ]; then
timeout 1500 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $out/${tag}_pytest_gpu.log
tail -n 5 $out/${tag}_pytest_gpu.log
fi
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $out/${tag}_smoke.log
timeout 900 python bench.py --steps $steps --warmup 3 > $out/${tag}_bench_c4.json 2> $out/${tag}_bench_c4.err; echo "bench rc=$?"
tail -c 6000 $out/${tag}_bench_c4.json; tail -n 5 $out/${tag}_bench_c4.err
[ -n "$WITH_REFERENCE" ] && { timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $out/${tag}_reference_arm.json 2>&1; echo "reference arm rc=$?"; }
if [ -n "$NCU" ]; then
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_ncu_launches_bench_c4.csv \
    python bench.py --steps 1 --warmup 3 --no-study --no-check > $out/${tag}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
for cfg in c4 c2; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:sampler -s 1 -c 1 -f -o $out/${tag}_sampler_$cfg \
      python tools/ncu_target.py $cfg > $out/${tag}_ncu_sampler_$cfg.log 2>&1; echo "ncu sampler $cfg rc=$?"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:gate_ -s 1 -c 1 -f -o $out/${tag}_gate_$cfg \
      python tools/ncu_target.py $cfg > $out/${tag}_ncu_gate_$cfg.log 2>&1; echo "ncu gate $cfg rc=$?"
done
fi
ls -la $out | tail -20
