import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
  if p not in sys.path:
    sys.path.insert(0, p)


def pytest_configure(config):
  config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def g1():
  from asimov_mjlab_b200 import scenes
  return scenes.load_asset("g1_flat")


@pytest.fixture(scope="session")
def go1():
  from asimov_mjlab_b200 import scenes
  return scenes.load_asset("go1_flat")


@pytest.fixture(scope="session")
def asimov_toe():
  from asimov_mjlab_b200 import scenes
  return scenes.load_asset("asimov_toe_flat")


@pytest.fixture(scope="session")
def asimov_toe_full():
  from asimov_mjlab_b200 import scenes
  return scenes.load_asset("asimov_toe_full_flat")


@pytest.fixture(scope="session")
def asimov_toe_rough():
  from asimov_mjlab_b200 import scenes
  return scenes.load_asset("asimov_toe_rough")


@pytest.fixture(scope="session")
def go1_rough_hf():
  from asimov_mjlab_b200 import scenes
  return scenes.load_asset("go1_rough_hf")


@pytest.fixture(scope="session")
def go1_rough():
  from asimov_mjlab_b200 import scenes
  return scenes.load_asset("go1_rough")


@pytest.fixture(scope="session")
def asimov12():
  from asimov_mjlab_b200 import scenes
  return scenes.load_asset("asimov_flat")


@pytest.fixture(scope="session")
def asimov12_full():
  from asimov_mjlab_b200 import scenes
  return scenes.load_asset("asimov_full_flat")
