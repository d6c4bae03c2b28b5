for f in 0 64 96 104; do echo "== dbg flags $f"; FINDB200_CHAIN_DBG=$f TRACE_ACCOUNT_ONLY=1 timeout 100 python tools/chain2_trace.py 2>&1 | grep -E "head_fwd|trunk_fwd|Error|error"; done
