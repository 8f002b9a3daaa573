timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 600 python tests/fuzz_stress.py --count 200000 --seed 1 2>&1 | tail -1 | cut -c1-900
