#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_configs.py -m gpu -q -x -k "solver_builds" 2>&1 | tail -5
