"""Stages an UNMODIFIED copy of the reference's `backend/` Python package under baseline/_ref/ (git-ignored, but it
travels to the GPU box with the gpurun snapshot) so that tests/test_gpu_reference_suite.py can run the reference's own
test-suite and service classes over janus_b200/shims/site on a real B200, where /root/reference does not exist.

    python scripts/stage_reference_tests.py          # run HERE, where /root/reference is mounted

Nothing under baseline/_ref/ is tracked; no reference source enters the repository history.
"""
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference/backend"
DST = os.path.join(ROOT, "baseline", "_ref", "backend")


def main() -> int:
    if not os.path.isdir(SRC):
        print(f"{SRC} not present: nothing staged")
        return 1
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    shutil.copytree(SRC, DST, ignore=shutil.ignore_patterns("__pycache__", "*.pyc", "logs", "*.wav", "*.log", "node_modules"))
    n = sum(len(fs) for _, _, fs in os.walk(DST))
    print(f"staged {n} files -> {DST}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
