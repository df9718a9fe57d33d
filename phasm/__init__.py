"""Drop-in ``phasm`` package for the ``phasm overlap`` stage (B200-native).

Only the overlap path of AbeelLab/phasm is provided:
``phasm.overlapper`` (extension module, same API as the reference's pybind11 module),
``phasm.io.gfa`` (``gfa_header`` / ``gfa_line``), ``phasm.io.fasta`` and the ``overlap``
subcommand of ``phasm.cli.assembler``.  layout / chain / phase are out of scope.
"""
__version__ = "0.1.0+b200"
