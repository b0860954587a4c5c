mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/r01c_bench_c2_n2.json 2> gpurun_out/r01c_bench_c2_n2.err; echo "n2 rc=$?"; tail -c 1500 gpurun_out/r01c_bench_c2_n2.json; tail -5 gpurun_out/r01c_bench_c2_n2.err
python -m pytest tests -m gpu -q -k "run_multi" 2>&1 | tail -3
python bench.py --workload c3 --steps 2 --warmup 3 --no-study > gpurun_out/r01c_bench_c3.json 2> gpurun_out/r01c_bench_c3.err; echo "c3 rc=$?"; tail -c 2500 gpurun_out/r01c_bench_c3.json
