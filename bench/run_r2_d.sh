#!/bin/bash
# round-2 call D (1 GPU): full GPU suite (v1 engine, static-scratch additions), tuned PyTorch arm, in-situ timeline
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2d_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/r2d_pytest.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; echo "bench rc=$?"; cat gpurun_out/r2d_bench.json; tail -3 gpurun_out/r2d_bench.err
python bench.py --impl torch_tuned --steps 20 --warmup 5 --sustain_s 0 > gpurun_out/r2d_tuned.json 2> gpurun_out/r2d_tuned.err; echo "tuned rc=$?"; cat gpurun_out/r2d_tuned.json; tail -5 gpurun_out/r2d_tuned.err
python bench/timeline.py > gpurun_out/r2d_timeline.log 2>&1; echo "timeline rc=$?"; tail -3 gpurun_out/r2d_timeline.log
gzip -f gpurun_out/trace.json
