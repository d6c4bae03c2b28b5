timeout 300 python tools/overlap_probe.py 2>&1 | tail -6
