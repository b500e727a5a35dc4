import sys
from pathlib import Path

import pytest

sys.path.insert(0, str(Path(__file__).resolve().parent))
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import common  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle_bin():
    common.build_oracle()
    return common.ORACLE_BIN


@pytest.fixture(scope="session")
def datasets(tmp_path_factory):
    """name -> directory following the run.sh dataset contract (generated once per session)."""
    root = tmp_path_factory.mktemp("datasets")
    dirs = {"bundled": common.bundled_dataset(root / "bundled")}
    for name in common.DATASETS:
        dirs[name] = common.make_dataset(name, root / name)
    for name in common.STAR_DATASETS:
        dirs[name] = common.make_star_dataset(name, root / name)
    return dirs
