#!/bin/bash
python tools/pcie_floor.py
for sub in 1 4; do for pb in 4 8 16; do echo "== C2 sub $sub pb $pb"; CHAQSDF_HULL_SUB=$sub CHAQSDF_HULL_PB=$pb python tools/time_configs.py c2 | tail -1; done; done
echo "== C2 auto"; python tools/time_configs.py c2 c3 | tail -2
for eb in 2 3 4 6 8; do echo "== band8 eb $eb"; CHAQSDF_HULL_EB=$eb python tools/time_bands.py 16384 256 8 | tail -1; done
for pb in 8 16 32; do echo "== band8 sub4 pb $pb"; CHAQSDF_HULL_PB=$pb python tools/time_bands.py 16384 256 8 | tail -1; done
