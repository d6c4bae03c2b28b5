timeout 300 python tools/overlap_probe.py 2>&1 | tail -5
OP_B=4 timeout 300 python tools/overlap_probe.py 2>&1 | tail -3
