#!/bin/bash
# full GPU suite + smoke + the default bench line (what the driver runs at round end)
O=gpurun_out
mkdir -p $O
timeout 1700 python -m pytest tests -m gpu -x -q > $O/r2_pytest_full.log 2>&1; echo "pytest rc=$?" >> $O/r2_pytest_full.log; tail -4 $O/r2_pytest_full.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r2_smoke.log 2>&1; tail -1 $O/r2_smoke.log
python bench.py > $O/r2_bench_default.json 2> $O/r2_bench_default.err; echo "bench rc=$?"
python bench.py --impl reference --steps 5 --warmup 1 > $O/r2_bench_reference.json 2>/dev/null; echo "ref rc=$?"
