timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python tools/sort_bench.py 20 2>&1 | tail -4
python tools/bench_variants.py C2 10 2>&1 | grep "mode [01] fwd_ppt 2 bwd_ppt 2"
