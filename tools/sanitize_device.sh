#!/bin/sh
# compute-sanitizer over the CUDA path (memcheck + racecheck on the kernels). Needs a GPU.
# ONE tool per gpurun call (B200_PROFILING.md):
#   gpurun --timeout 900 -- 'sh tools/sanitize_device.sh memcheck > gpurun_out/sanitize_memcheck.log 2>&1'
#   gpurun --timeout 900 -- 'sh tools/sanitize_device.sh racecheck > gpurun_out/sanitize_racecheck.log 2>&1'
# Small inputs: the tools slow kernels down 10-100x. Each run must end with "ERROR SUMMARY: 0 errors".
set -x
cd "$(dirname "$0")/.."
tool=${1:-memcheck}
CS=/usr/local/cuda/bin/compute-sanitizer
# search + upload + patch (device-assisted build drives all three), then the smoke search against the oracle
timeout 400 $CS --tool $tool --error-exitcode 9 python tools/check_device_build.py 3000 32 8 40 256 || echo "FAILED: $tool device build"
timeout 400 $CS --tool $tool --error-exitcode 9 python -c "import __graft_entry__ as g; g.smoke()" || echo "FAILED: $tool smoke"
# parity suite at its small shapes (wide rows, odd dims, forced visited-table retries, patch)
timeout 900 $CS --tool $tool --error-exitcode 9 python -m pytest -q -x -m gpu tests/test_gpu_parity.py tests/test_gpu_patch.py -k "not full_size and not large" || echo "FAILED: $tool parity/patch"
if [ "$tool" = memcheck ]; then
    # exact k-NN (tcgen05 filter + re-rank): memcheck only, racecheck does not model TMA/mbarrier traffic
    timeout 600 $CS --tool memcheck --error-exitcode 9 python -m pytest -q -x -m gpu tests/test_gpu_exact_tc.py -k "not full_size" || echo "FAILED: memcheck exact k-NN"
fi
