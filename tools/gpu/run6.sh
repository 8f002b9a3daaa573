set -x
timeout 600 compute-sanitizer --tool memcheck --print-limit 5 python bench.py --batch 1024 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e 2>&1 | grep -v "^\s*$" | head -60
