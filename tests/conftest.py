import importlib.util
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REFERENCE_DIRS = ["/root/reference", os.path.join(ROOT, "baseline", "_ref")]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def load_reference_module(name: str):
    """Import ``<name>.py`` of the upstream project under a private module name (None if absent)."""
    for d in REFERENCE_DIRS:
        path = os.path.join(d, name + ".py")
        if os.path.isfile(path):
            spec = importlib.util.spec_from_file_location(f"_upstream_{name}", path)
            mod = importlib.util.module_from_spec(spec)
            try:
                spec.loader.exec_module(mod)
            except Exception:
                return None
            return mod
    return None


@pytest.fixture(scope="session")
def tiny_bert():
    from fedrec_b200.models.distilbert import DistilBertConfigLite
    return DistilBertConfigLite(n_layers=2, dim=64, hidden_dim=128, n_heads=4, hidden_size=64)


@pytest.fixture(scope="session")
def tiny_data():
    from fedrec_b200.data import make_synthetic_mind
    return make_synthetic_mind(num_news=300, num_train=48, num_valid=12, title_len=12, seed=3)


@pytest.fixture()
def solo_group(monkeypatch):
    """For tests that build a FederatedRunner inside the pytest process: a private rendezvous port, and the process group
    is gone afterwards, so the default port stays free for tests that launch the entrypoints as subprocesses."""
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    monkeypatch.setenv("MASTER_ADDR", "127.0.0.1")
    monkeypatch.setenv("MASTER_PORT", str(port))
    monkeypatch.setenv("RANK", "0")
    monkeypatch.setenv("LOCAL_RANK", "0")
    monkeypatch.setenv("WORLD_SIZE", "1")
    yield port
    import torch.distributed as dist
    if dist.is_initialized():
        dist.destroy_process_group()
