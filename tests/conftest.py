import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with `-m gpu`)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch

        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if not has_gpu:
        skip = pytest.mark.skip(reason="no CUDA device")
        for item in items:
            if "gpu" in item.keywords:
                item.add_marker(skip)


@pytest.fixture(autouse=True)
def _fresh_framework_state():
    """Every test starts from default settings, an empty in-memory registry and empty metric stores."""
    from myfyp_b200.communication.protocols.memory.registry import ServerRegistry
    from myfyp_b200.management.logger import logger
    from myfyp_b200.settings import Settings

    saved = {k: v for k, v in vars(Settings).items() if k.isupper() or k == "_WAIT_HEARTBEATS_CONVERGENCE"}
    yield
    for k, v in saved.items():
        setattr(Settings, k, v) if k != "_WAIT_HEARTBEATS_CONVERGENCE" else type.__setattr__(Settings, k, v)
    ServerRegistry.reset()
    logger.reset_metrics()
    for n in list(logger.get_nodes().keys()):
        try:
            logger.unregister_node(n)
        except Exception:
            pass
