#!/bin/bash
# Rebuild libvgpu.so and the prebuilt kernels; print their resource usage.
cd "$(dirname "$0")/.." || exit 1
make -C velox-1_b200/csrc 2>&1 | grep -E "error|warning"
python -c "
import importlib,sys
sys.path.insert(0,'.')
importlib.import_module('velox-1_b200.aot_plans').build_all()" 2>&1 | tail -8
for f in velox-1_b200/aot/*.cubin; do echo "$(basename $f) $(cuobjdump -res-usage $f | grep -o 'REG:[0-9]* STACK:[0-9]*') $(grep 'define NT' ${f%.cubin}.cu) $(grep -o '__launch_bounds__(NT, [0-9])' ${f%.cubin}.cu)"; done
