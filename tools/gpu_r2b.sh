#!/bin/bash
# round 2, GPU call B: tier-1 acceptance tuning, then the test files touched since call A
mkdir -p gpurun_out
python tools/tune_accept.py 1e-9 1e-8 1e-7 1e-6 1e-5 1e-4 > gpurun_out/r2b_tune.log 2>&1
cat gpurun_out/r2b_tune.log
timeout 1800 python -m pytest tests/test_gpu_shim.py tests/test_reference_goldens.py tests/test_gpu_transport.py tests/test_gpu_qmie.py tests/test_gpu_qmie_ret.py tests/test_gpu_zarate.py tests/test_gpu_regimes.py -m gpu -q -s -p no:cacheprovider > gpurun_out/r2b_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2b_pytest.log
tail -8 gpurun_out/r2b_pytest.log
python -c "
import __graft_entry__ as ge
from kineticgas_b200 import _lib
print(_lib.lib().kg_build_info().decode(), ge.source_sha())
"
