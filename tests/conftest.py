"""pytest config: registers the `gpu` marker, puts the repo root and the (hyphenated) package
directory on sys.path so `import alegpu`, `import synth`, `from oracle import ...` work."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "ale-libtorch-ppo_b200")
for p in (ROOT, PKG, os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (runs on the B200 box)")
