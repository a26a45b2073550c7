DCD_SEARCH_TIMING=1 python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e 2>&1 | grep "dcd search" | tail -1
