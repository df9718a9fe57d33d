"""Minimal FASTA reader for the overlap stage.

Stands in for ``dinopy.FastaReader(path).entries()`` and ``dinopy.reverse_complement`` at the
reference's call sites ``phasm/cli/assembler.py:32-40`` (dinopy is not vendored by the
reference).  Record name = header line after ``>``, right-stripped; multi-line records and
blank lines are accepted.  The fast path of the CLI does not use this module (the native
``add_fasta`` does the same parsing in C++); it exists for API parity and for the tests.
"""
from collections import namedtuple

FastaEntry = namedtuple("FastaEntry", ["sequence", "name", "length", "interval"])

_COMP = bytes.maketrans(b"ACGTacgt", b"TGCAtgca")


def reverse_complement(seq):
    if isinstance(seq, str):
        return seq.translate(str.maketrans("ACGTacgt", "TGCAtgca"))[::-1]
    return bytes(seq).translate(_COMP)[::-1]


class FastaReader:
    def __init__(self, path):
        self.path = path

    def entries(self):
        """Yield ``FastaEntry(sequence: bytes, name: bytes, length: int, interval)``."""
        name, parts, start = None, [], 0
        pos = 0
        with open(self.path, "rb") as f:
            for raw in f:
                line = raw.rstrip()
                if line.startswith(b">"):
                    if name is not None:
                        seq = b"".join(parts)
                        yield FastaEntry(seq, name, len(seq), (start, pos))
                    name, parts, start = line[1:], [], pos
                elif line:
                    if name is None:
                        raise ValueError("%s: sequence data before the first '>' line" % self.path)
                    parts.append(line)
                pos += len(raw)
        if name is not None:
            seq = b"".join(parts)
            yield FastaEntry(seq, name, len(seq), (start, pos))
