"""Drop-in for the table-comparison entry points of the reference's eSEC_analyser.py: getDissimilarityMatrix (:782-792) and
similarity_manipulations (:843-864).  Same names, arguments and results; the comparison of every pair of tables runs on the
GPU through libe2sec.so (e2sec_dissimilarity, csrc/dissim.cuh).  There is no CPU fallback: without the library or a CUDA
device these raise.  Plotting, clustering and PDF reading of the reference module are out of scope."""
import numpy as np

from .engine import Engine

_engine = None


def _get_engine():
    global _engine
    if _engine is None:
        _engine = Engine(0)
    return _engine


def _as_list(eSEC_matrices):
    """the reference indexes its argument with 0 .. len-1 (a list, or a dict with those keys)"""
    return [eSEC_matrices[k] for k in range(len(eSEC_matrices))]


def getDissimilarityMatrix(eSEC_matrices, engine=None):
    """Dissimilarity matrix of paper [1] over a list/dict of eSEC tables (:782-841).  The normalisation keeps the reference's
    constant 10 (rows per block of the 30-row eSEC table) whatever the row count (:838)."""
    return (engine or _get_engine()).dissimilarity(_as_list(eSEC_matrices), divisor=10.0)


def similarity_manipulations(manipulation_1, manipulation_2, engine=None):
    """Similarity of two e2SEC tables in percent (:843-864; normalised by 8 rows per block, :860)."""
    D = (engine or _get_engine()).dissimilarity([manipulation_1, manipulation_2], divisor=8.0)
    return (1 - D[0, 1]) * 100
