import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

REFERENCE_DIR = Path(os.environ.get("DLION_REFERENCE_DIR", "/root/reference"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        ngpu = torch.cuda.device_count()
        skip_multi = pytest.mark.skip(reason="needs >= 2 GPUs")
        for item in items:
            if "multigpu" in item.keywords and ngpu < 2:
                item.add_marker(skip_multi)
        return
    skip_gpu = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords or "multigpu" in item.keywords:
            item.add_marker(skip_gpu)


def load_reference_module():
    """The reference's distributed_lion.py, imported by path under another name (ours is
    also called distributed_lion).  Falls back to baseline/_ref on the GPU box."""
    import importlib.util
    for base in (REFERENCE_DIR, ROOT / "baseline" / "_ref"):
        f = base / "distributed_lion.py"
        if f.exists():
            spec = importlib.util.spec_from_file_location("ref_distributed_lion", f)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            return mod
    return None


@pytest.fixture(scope="session")
def ref_mod():
    mod = load_reference_module()
    if mod is None:
        pytest.skip("reference checkout not available")
    return mod
