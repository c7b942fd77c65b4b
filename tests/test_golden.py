"""Golden vectors: outputs of the REFERENCE's own src/ORBextractor.cc (compiled in place as
oracle/_ref/libreforb_mono.so by tests/golden/make_golden.py) on committed frames.

  not gpu: the oracle restatement reproduces every golden case byte-for-byte
  gpu:     the CUDA path (through the C ABI) reproduces every golden case byte-for-byte
"""
import hashlib
import json
from pathlib import Path

import numpy as np
import pytest

from orb_slam2_test_b200 import synth

G = Path(__file__).resolve().parent / "golden"
META = json.loads((G / "golden.json").read_text())
FRAMES = np.load(G / "frames.npz")
EXPECT = np.load(G / "expected.npz")
KP = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
               ("octave", "<i4"), ("class_id", "<i4")])


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def _check_small(name, kps, desc):
    m = META["small"][name]
    want_k = EXPECT[name + "_kp"].reshape(-1).view(KP)
    want_d = EXPECT[name + "_desc"]
    assert len(kps) == m["n"] == len(want_k)
    for f in KP.names:
        bad = np.nonzero(kps[f] != want_k[f])[0]
        assert bad.size == 0, f"{name}: field {f} differs at {bad[:5]} ({bad.size} of {len(kps)})"
    assert np.array_equal(desc, want_d), f"{name}: {(desc != want_d).any(axis=1).sum()} descriptors differ"
    assert sha(kps.view(np.uint8)) == m["kp_sha256"] and sha(desc) == m["desc_sha256"]


def _params(m):
    return m["nfeatures"], m["scaleFactor"], m["nlevels"], m["iniThFAST"], m["minThFAST"]


@pytest.mark.parametrize("name", list(META["small"]))
def test_oracle_reproduces_reference_golden(name, oracle_built):
    m = META["small"][name]
    img = FRAMES[m["frame"]]
    assert sha(img) == m["input_sha256"]
    kps, desc = oracle_built.Oracle(*_params(m)).extract(img)
    _check_small(name, kps, desc)


def _full_cases():
    return [pytest.param(c, id=f"{c['config']}-s{c['seed']}-f{c['frame']}") for c in META["full"]]


@pytest.mark.parametrize("c", _full_cases())
def test_oracle_full_size_hashes(c, oracle_built):
    img = synth.make_frame((c["rows"], c["cols"]), seed=c["seed"], frame=c["frame"])
    if sha(img) != c["input_sha256"]:
        pytest.skip("this machine's numpy renders different synth bytes; small committed frames cover parity")
    kps, desc = oracle_built.Oracle(c["nfeatures"], 1.2, 8, 20, 7).extract(img)
    assert len(kps) == c["n"]
    assert sha(kps.view(np.uint8)) == c["kp_sha256"] and sha(desc) == c["desc_sha256"]


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(META["small"]))
def test_gpu_reproduces_reference_golden(name):
    from orb_slam2_test_b200 import ORBextractor
    m = META["small"][name]
    ext = ORBextractor(*_params(m))
    kps, desc = ext(FRAMES[m["frame"]])
    _check_small(name, kps, desc)


@pytest.mark.gpu
@pytest.mark.parametrize("c", _full_cases())
def test_gpu_full_size_hashes(c):
    from orb_slam2_test_b200 import ORBextractor
    img = synth.make_frame((c["rows"], c["cols"]), seed=c["seed"], frame=c["frame"])
    if sha(img) != c["input_sha256"]:
        pytest.skip("this machine's numpy renders different synth bytes")
    kps, desc = ORBextractor(c["nfeatures"], 1.2, 8, 20, 7)(img)
    assert len(kps) == c["n"]
    assert sha(kps.view(np.uint8)) == c["kp_sha256"] and sha(desc) == c["desc_sha256"]
