"""B200-native e2SEC extraction hot path of tobi-code/ManipulationLibrary.

Drop-in modules (same names as the reference):
    manipulationlibrary_b200.manipulation_analyser          analyse_maniac_manipulation, _process, esec_to_e2sec
    manipulationlibrary_b200.cython_filter_new.filter_cython_new   region_filter_cython, distance_cython
    manipulationlibrary_b200.eSEC_analyser                  getDissimilarityMatrix, similarity_manipulations
Everything computes through libe2sec.so (hand-written sm_100a CUDA behind the C ABI of include/e2sec.h);
importing those modules without the built library, or using them without a CUDA device, raises.
"""
__all__ = ["build"]
