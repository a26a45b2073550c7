import json
import os
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (HERE, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    with open(os.path.join(HERE, "golden", "indextest.json"), encoding="utf-8") as f:
        return json.load(f)


@pytest.fixture(scope="session")
def data_dir():
    return os.path.join(HERE, "data")


def read_stop_words(path):
    """Index.LoadStopWords, Index.cs:227-230: lines that are non-blank and contain no ';'."""
    out = []
    for s in open(path, encoding="utf-8-sig").read().splitlines():
        if len(s.strip(" ")) > 0 and ";" not in s:
            out.append(s)
    return out


@pytest.fixture(scope="session")
def stop_words(data_dir):
    return read_stop_words(os.path.join(data_dir, "stop.txt"))


@pytest.fixture(scope="session")
def en_voc(data_dir):
    return open(os.path.join(data_dir, "en.voc"), "rb").read()
