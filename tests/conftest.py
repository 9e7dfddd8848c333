import json
import os
import sys

import pytest
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def configs():
    """Parsed reference SS/config/*.yaml (tests/golden/make_configs.py)."""
    with open(os.path.join(ROOT, "human_robot_gym_b200", "data", "shield_configs.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def config_dir(configs, tmp_path_factory):
    """The same configs written back out as YAML files in the reference schema (one file per config)."""
    d = tmp_path_factory.mktemp("shield_config")
    for name, content in configs.items():
        with open(d / (name + ".yaml"), "w") as f:
            yaml.safe_dump(content, f, default_flow_style=None)
    return str(d)
