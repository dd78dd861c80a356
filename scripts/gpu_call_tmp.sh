python scripts/debug_pool.py 2>&1 | cut -c1-400
