for f in 0 128 414; do echo "== dbg flags $f (accounting on)"; FINDB200_CHAIN_DBG=$f TRACE_ACCOUNT_ONLY=1 timeout 100 python tools/chain2_trace.py 2>&1 | grep -E "head_fwd|rror"; done > gpurun_out/a6_flags.txt 2>&1; cat gpurun_out/a6_flags.txt
FINDB200_CHAIN_DBG=414 timeout 100 python tools/chain2_trace.py > gpurun_out/a6_trace414.txt 2>&1; cat gpurun_out/a6_trace414.txt
