#!/bin/bash
# One GPU session of a development round: GPU tests, smoke, headline bench.  Everything lands in gpurun_out/.
# Usage (on the GPU box, from the repo root):  bash tools/gpu_round.sh [tag]
tag="${1:-r2}"
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x --timeout 900 > "gpurun_out/${tag}_gputest.log" 2>&1
echo "pytest rc=$?" >> "gpurun_out/${tag}_gputest.log"
tail -5 "gpurun_out/${tag}_gputest.log"
python -c "import __graft_entry__ as g; g.smoke()" > "gpurun_out/${tag}_smoke.log" 2>&1
echo "smoke rc=$?" | tee -a "gpurun_out/${tag}_smoke.log"
python bench.py > "gpurun_out/${tag}_bench1.json" 2> "gpurun_out/${tag}_bench1.err"
echo "bench rc=$?"
cat "gpurun_out/${tag}_bench1.json"
