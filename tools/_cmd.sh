timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python tools/quick_e2e.py 2>&1 | tee gpurun_out/e2e2.txt
