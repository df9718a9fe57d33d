"""GFA2 line writer of the overlap stage.

Same two functions, same output bytes as the reference's ``phasm/io/gfa.py:230-240``; the
GFA parsers / graph writers of that file belong to the layout stage and are out of scope.
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
