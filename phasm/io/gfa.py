"""GFA2 line writer of the overlap stage.

Same two functions, same output bytes as the reference's ``phasm/io/gfa.py:230-240``, plus the
edge parser of the consumer side (``gfa2_parse_edge``) and a native whole-file reader; the graph
writers of that file belong to the layout stage and are out of scope.
"""
from typing import Optional


def gfa_line(*fields) -> str:
    """Tab-join the fields (``str()`` of each) and terminate the line."""
    return "\t".join(str(f) for f in fields) + "\n"


def gfa_header(version: str = "2.0", trace_spacing: Optional[int] = None) -> str:
    """``H`` line: ``VN:z:<version>`` and, if given, ``TS:i:<trace_spacing>``."""
    fields = ["H", "VN:z:%s" % version]
    if trace_spacing:
        fields.append("TS:i:%d" % trace_spacing)
    return gfa_line(*fields)


# ---- GFA2 ingest for the consumer of the overlaps (parsing part) ----------
def gfa2_parse_edge(line: str):
    """One ``E`` line -> ``(sid1, sid2, arange, brange, alignment, tags)``; positions may carry the
    GFA2 end marker ``$``.  Same name, result and errors as the reference's ``phasm/io/gfa.py:72-86``."""
    if not line.startswith("E"):
        raise ValueError("Given GFA2 line is not an edge.")
    parts = line.strip().split("\t")

    def pos(field: str) -> int:
        return int(field[:-1]) if field.endswith("$") else int(field)

    arange = (pos(parts[4]), pos(parts[5]))
    brange = (pos(parts[6]), pos(parts[7]))
    return (parts[2].strip(), parts[3].strip(), arange, brange, parts[8].strip(), list(parts[9:]))


def gfa2_read(path):
    """Whole-file form of the reference's two passes (``gfa2_parse_segments`` + ``gfa2_parse_edge`` per
    ``E`` line, ``phasm/cli/assembler.py:55-60,93-95``) done natively (``ovl_gfa_read``): returns
    ``(names, lengths, edges)`` with the edges as a structured array on segment indices."""
    from phasm_b200.gfa_native import read_gfa
    return read_gfa(path)
