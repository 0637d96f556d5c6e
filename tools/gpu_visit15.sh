#!/bin/bash
# Visit 15 (last of round 2): smoke() and the default bench line at HEAD.
O=gpurun_out; mkdir -p $O
timeout 40 python -c "import __graft_entry__ as g; g.smoke()" > $O/r02n_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/r02n_smoke.log | cut -c1-200
timeout 80 python bench.py > $O/r02n_bench.json 2> $O/r02n_bench.err; echo "bench rc=$?"; tail -1 $O/r02n_bench.json | cut -c1-1200
