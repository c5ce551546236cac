#!/bin/bash
# rabit-b200 CI entry (role of the reference's scripts/travis_script.sh): build everything, run the
# CPU test tier, and — when a GPU is visible — the GPU tier.
set -euo pipefail
cd "$(dirname "$0")/.."
python scripts/lint.py
python scripts/gen_api_doc.py --check   # doc/{cpp,c,python}_api.md are generated from the sources
python build.py
./build/bin/unit_tests
python -m pytest tests -x -q -m "not gpu"
# fault scenarios of the host engine under ASan+UBSan and TSan (static sanitizer builds in build/san_*)
python scripts/sanitize.py address
python scripts/sanitize.py thread
if python -c "import torch,sys; sys.exit(0 if torch.cuda.is_available() else 1)"; then
  python -m pytest tests -x -q -m gpu
  python __graft_entry__.py smoke
fi
