import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, HERE):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    """Vectors produced by the unmodified reference (tests/golden/make_golden.py)."""
    return np.load(os.path.join(HERE, "golden", "reference_vectors.npz"))


@pytest.fixture(scope="session")
def gather_golden():
    """loadFilteredReads / sortedBins vectors from the unmodified reference (tests/golden/make_golden_gather.py)."""
    return np.load(os.path.join(HERE, "golden", "gather_vectors.npz"))


@pytest.fixture(scope="session")
def ingest_golden():
    """FASTA / FASTQ record-reader vectors from the unmodified reference (tests/golden/make_golden_ingest.py)."""
    return np.load(os.path.join(HERE, "golden", "ingest_vectors.npz"))


@pytest.fixture(scope="session")
def oracle():
    from oracle_py import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def dy():
    """The product package with libdyibf.so built."""
    from dream_yara_b200 import build as _b
    _b.build()
    import dream_yara_b200
    return dream_yara_b200
