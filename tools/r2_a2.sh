# round 2 call A2: tensor-pipe rate probe, pair-chain timeline + issuer cycle accounting, timing experiments by dbg flag
timeout 120 python tools/umma_rate.py > gpurun_out/a2_umma_rate.json 2> gpurun_out/a2_umma_rate.err; echo rate=$?; cat gpurun_out/a2_umma_rate.json
timeout 100 python tools/chain2_trace.py > gpurun_out/a2_trace.txt 2>&1; cat gpurun_out/a2_trace.txt
for f in 0 2 4 8 16 24 30 32 64 96 104 126; do echo "== dbg flags $f"; FINDB200_CHAIN_DBG=$f TRACE_ACCOUNT_ONLY=1 timeout 100 python tools/chain2_trace.py 2>&1 | grep -E "head_fwd|trunk_fwd|rror"; done > gpurun_out/a2_flags.txt 2>&1; cat gpurun_out/a2_flags.txt
