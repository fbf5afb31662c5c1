#!/bin/bash
# Rebuild the in-tree CUDA library from any cwd.
cd "$(dirname "$0")/.." && python -c "from vllm_metal_b200.metal.build import build; print(build(force=True))" 2>&1 | tail -3
