"""pytest configuration. `-m "not gpu"` runs on a CPU-only container (oracle vs goldens, host
logic, ABI symbol checks, gloo world_size=2); `-m gpu` are the parity tests proper: the CUDA path
through the C ABI against the oracle (oracle/_ref + oracle/liboracle.so) and the committed goldens."""
import gzip
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    with gzip.open(os.path.join(GOLDEN, name), "rb") as f:
        return json.loads(f.read().decode())


def expand_rle(s: str) -> bytes:
    out, n = bytearray(), 0
    for ch in s:
        if ch.isdigit():
            n = n * 10 + ord(ch) - 48
        else:
            out += ch.encode() * n
            n = 0
    return bytes(out)


def rle(cig: bytes) -> str:
    if not cig:
        return ""
    a = np.frombuffer(cig, dtype=np.uint8)
    cut = np.flatnonzero(a[1:] != a[:-1]) + 1
    st = np.concatenate(([0], cut))
    en = np.concatenate((cut, [a.size]))
    return "".join("%d%s" % (e - s, chr(a[s])) for s, e in zip(st, en))


def _ensure_built():
    """Build the checker libraries when they are missing (development container); on the GPU box
    the prebuilt files travel with the snapshot."""
    need_oracle = not os.path.exists(os.path.join(ROOT, "oracle", "liboracle.so"))
    if need_oracle:
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "liboracle.so"], check=True)
    import _oracle
    if not os.path.exists(_oracle.ref_path()) and os.path.isdir("/root/reference/WFA2-lib/wavefront"):
        subprocess.run(["make", "-s", "-j4", "-C", os.path.join(ROOT, "oracle"), "all"], check=True)
    if not os.path.exists(os.path.join(ROOT, "tests", "host_emul", "libwfa_emul.so")):
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "tests", "host_emul")], check=True)
    from seqrush_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        _lib.build()


@pytest.fixture(scope="session")
def built():
    _ensure_built()
    return True


@pytest.fixture(scope="session")
def oracle(built):
    import _oracle
    if not os.path.exists(_oracle.ref_path()):
        # never a silent skip: without the checker a parity test proves nothing. oracle/_ref is
        # git-ignored but travels with the gpurun snapshot; build it with `make -C oracle` where
        # /root/reference is mounted.
        pytest.fail("oracle/_ref/%s is missing (build it with `make -C oracle` in the development container; "
                    "it must travel to the GPU box with the snapshot)" % os.path.basename(_oracle.ref_path()))
    return _oracle.Oracle()


@pytest.fixture(scope="session")
def wfa2_golden():
    return load_golden("wfa2_biwfa_affine2p.json.gz")


@pytest.fixture(scope="session")
def seqrush_golden():
    return load_golden("seqrush_pairs.json.gz")


@pytest.fixture(scope="session")
def c1_golden():
    exp = load_golden("c1_expected.json.gz")
    ids, seqs, cur = [], [], None
    with gzip.open(os.path.join(GOLDEN, "DRB1-3123.fa.gz"), "rb") as f:
        for line in f:
            line = line.rstrip(b"\r\n")
            if line.startswith(b">"):
                ids.append(line[1:].split()[0].decode())
                cur = []
                seqs.append(cur)
            elif cur is not None:
                cur.append(line.strip())
    return exp, ids, [b"".join(s) for s in seqs]
