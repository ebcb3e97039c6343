#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
timeout 300 python -m pytest tests/test_z_graph_edges_gpu.py tests/test_graph_host.py -x -q -m gpu 2>&1 | tail -15
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r2_gpu_tests.log 2>&1; echo "all gpu tests rc=$?"; tail -4 gpurun_out/r2_gpu_tests.log
timeout 300 python tools/update_graph_probe.py 2>&1 | tail -20
