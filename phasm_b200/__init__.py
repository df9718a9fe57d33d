"""phasm_b200 -- B200-native drop-in for PHASM's ``phasm overlap`` hot path.

Only the overlapper path is here: the CUDA core + C-ABI in ``csrc/``,
the ``ExactOverlapper`` host mirror of the reference's pybind11 class, the GFA2 line writer,
the FASTA reader and the ``overlap`` CLI stage.
"""
__version__ = "0.1.0"
