import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_cuda():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "fusion_golden.json"), encoding="utf-8") as f:
        return json.load(f)["cases"]


@pytest.fixture(scope="session")
def mmr_golden():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "mmr_golden.json"), encoding="utf-8") as f:
        return json.load(f)["cases"]


@pytest.fixture(scope="session")
def topk_golden():
    """Ranked lists produced by the reference's own brute-force cosine ranking (search_utils.py:56-67, :337-366), run
    in the build container by tests/golden/make_golden.py. Returns (corpus fp32 [n,d], cases) with numpy arrays."""
    import json
    import numpy as np
    with open(os.path.join(ROOT, "tests", "golden", "topk_golden.json"), encoding="utf-8") as f:
        g = json.load(f)
    den = float(g["denominator"])
    corpus = (np.array(g["corpus_num"], dtype=np.float64) / den).astype(np.float32)
    cases = []
    for c in g["cases"]:
        cases.append({"query": (np.array(c["query_num"], dtype=np.float64) / den).astype(np.float32),
                      "min_score": c["min_score"], "scores": np.array(c["scores"], dtype=np.float64),
                      "top_ids": np.array(c["top_ids"], dtype=np.int64), "top_scores": np.array(c["top_scores"], dtype=np.float64)})
    return corpus, cases, int(g["k"])
