timeout 200 python tools/chain_stress.py 600 2>&1 | grep -v "mbarrier wait timeout" | tail -5
bash tools/r2_a8.sh
