timeout 900 python scripts/sanitize_paths.py > gpurun_out/sanitize_plain.log 2>&1; tail -2 gpurun_out/sanitize_plain.log
timeout 900 python scripts/sanitize_paths.py > gpurun_out/sanitize_plain2.log 2>&1; tail -1 gpurun_out/sanitize_plain2.log
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
python __graft_entry__.py smoke 2>&1 | tail -1
