"""dream_yara_b200 — B200-native (sm_100a CUDA) implementation of DREAM-Yara's IBF read->bin prefilter.

Only what the hot path needs lives here: csrc/ (kernels + the C ABI of include/dyibf.h), the host-side mirror of
the reference's SeqAnBloomFilter class (ibf.py) and of the mapper's filtering stage (mapper.py), and the
multi-GPU shardings (dist.py).  Nothing in this package has a CPU implementation of the filter.
"""
from ._lib import ENC_ASCII, ENC_DNA5, ENC_DNA5Q, DyibfError, lib          # noqa: F401
from .ibf import ClassifyTicket, FilterGroup, SeqAnBloomFilter, SeqAnKDXFilter, as_batch, read_fasta                    # noqa: F401
from .ingest import parseRecords, readRecords                               # noqa: F401
from .mapper import (DisOptions, ReadSet, clasifyLoadedReads, loadFilteredReads, readFilterMetadata,  # noqa: F401
                     sortedBins, loadReads)
