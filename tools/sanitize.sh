#!/usr/bin/env bash
# compute-sanitizer targets for the native kernels (run on a GPU box, ONE tool per invocation:
#   tools/sanitize.sh memcheck | racecheck | synccheck | initcheck ).
# Uses the smallest tests that cover every kernel family; the peer-flag protocol is exercised by the
# virtual-rank exchange tests (all sends before any receive, so nothing waits on an unlaunched kernel).
set -euo pipefail
tool="${1:-memcheck}"
cd "$(dirname "$0")/.."
exec compute-sanitizer --tool "$tool" --error-exitcode 3 --print-limit 20 \
  python -m pytest -q -x -m gpu \
    "tests/test_fused_exchange_gpu.py::test_deterministic_methods_match_oracle[Topk-kw0-False-2]" \
    "tests/test_fused_exchange_gpu.py::test_quantisers_injected_uniforms" \
    "tests/test_fused_exchange_gpu.py::test_randomk_injected_keys_exact" \
    "tests/test_sgd_gpu.py" "tests/test_pool_gpu.py" "tests/test_loss_gpu.py" \
    "tests/test_bn_gpu.py::test_module_matches_torch_module_in_training" \
    "tests/test_zz_no_alloc.py"
