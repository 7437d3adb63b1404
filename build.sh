#!/usr/bin/env bash
# Build the native libraries in place and, with --wheel, a wheel under dist/.
# (The reference's build.sh runs Maven and bdist_wheel; there is no JVM component here.)
set -euo pipefail
cd "$(dirname "${BASH_SOURCE[0]}")"
python - <<'PY'
import os
from raydp_b200 import _native
print("store  :", _native.build_store(verbose=True))
if os.environ.get("RAYDP_B200_SKIP_CUDA") != "1":
    print("kernels:", _native.build_kernels(verbose=True))
PY
if [[ "${1:-}" == "--wheel" ]]; then
    python setup.py -q bdist_wheel --dist-dir dist
    ls -l dist/
fi
