#!/bin/bash
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
echo "== narrow"; python tools/time_configs.py c2 c3 | tail -2
echo "== wide pass 1"; CHAQSDF_ROW8_WIDE=1 python tools/time_configs.py c3 | tail -1
python __graft_entry__.py 2>&1 | tail -2
