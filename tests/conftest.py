import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_pixels():
    with open(os.path.join(GOLDEN, "pixels.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def golden_wire():
    with open(os.path.join(GOLDEN, "videoframe_wire.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def small_frames():
    """[(fmt, w, h, seed, planes, out)] stored raw from the live libswscale."""
    z = np.load(os.path.join(GOLDEN, "small_frames.npz"))
    cases = {}
    for k in z.files:
        n, fmt, wh, seed, what = k.split("_")
        w, h = (int(v) for v in wh.split("x"))
        c = cases.setdefault(int(n), dict(fmt=fmt, w=w, h=h, seed=int(seed[1:]), planes={}))
        if what == "out":
            c["out"] = z[k]
        else:
            c["planes"][int(what[1:])] = z[k]
    out = []
    for n in sorted(cases):
        c = cases[n]
        out.append((c["fmt"], c["w"], c["h"], c["seed"], [c["planes"][i] for i in sorted(c["planes"])], c["out"]))
    return out


def wire_payload(n: int) -> bytes:
    """Payload rule of tests/golden/videoframe_wire.json."""
    return ((np.arange(n, dtype=np.uint64) * 2654435761 % (1 << 32)) >> 24).astype(np.uint8).tobytes()
