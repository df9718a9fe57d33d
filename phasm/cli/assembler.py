"""``phasm`` command line: the ``overlap`` subcommand on the B200 overlapper.

Flags, log messages and output bytes follow the reference's
``phasm/cli/assembler.py:29-50`` (stage) and ``:421-452, 663-685`` (argparse, logging):

    phasm [-v...] overlap [-l INT=1000] [-o FILE=stdout] fasta_input

Output = ``H\\tVN:z:2.0``, one ``S\\t<name>\\t<len>\\t*`` per read in FASTA order, one
``E\\t*\\t<a>\\t<b>\\t<astart>\\t<aend>\\t0\\t<bend>\\t*`` per overlap.  The reference's E-line order is
unspecified (suffix-array x unordered_map order); here it is sorted by entry index.

Two equivalent I/O routes: the default one parses the FASTA and formats the GFA2 text in
the native library (``add_fasta`` / ``write_gfa``: no Python object per read or per row);
``--python-io`` runs the reference's per-read / per-row Python loop over the same API
(``add_sequence`` twice per read, ``overlaps``, ``gfa_line``).  Both produce the same bytes.

The other stages of PHASM (layout, chain, phase) are outside this repository's scope.
"""
import argparse
import logging
import sys

from phasm.io import gfa
from phasm.io.fasta import FastaReader, reverse_complement
from phasm.overlapper import ExactOverlapper

logger = logging.getLogger()


def overlap(args):
    out = args.output
    if getattr(args, "gpus", 1) > 1:
        return _overlap_multi(args)
    overlapper = ExactOverlapper()

    if getattr(args, "python_io", False) or not _has_fd(out):
        out.write(gfa.gfa_header())
        logger.info("Building suffix tree and searching for pairwise overlaps...")
        for entry in FastaReader(args.fasta_input).entries():
            name = entry.name.decode('utf-8')
            seq = entry.sequence.decode('utf-8')
            out.write(gfa.gfa_line("S", name, entry.length, "*"))
            overlapper.add_sequence(name + "+", seq)
            overlapper.add_sequence(name + "-", reverse_complement(seq))

        overlaps = overlapper.overlaps(args.min_length)

        logger.info("Writing to GFA2...")
        for aread, bread, astart, aend, bstart, bend in overlaps:
            out.write(gfa.gfa_line(
                "E", "*", aread, bread, astart, aend, bstart, bend, "*"))
    else:
        logger.info("Building suffix tree and searching for pairwise overlaps...")
        overlapper.add_fasta(args.fasta_input, True)
        overlapper.run(args.min_length)
        logger.info("Writing to GFA2...")
        out.flush()
        overlapper.write_gfa(out.fileno(), True, True, True)

    _write_stats(args, overlapper.stats())
    logger.info("Done.")


def _overlap_multi(args):
    """``--gpus N``: one process per GPU, reads sharded by record range, rows gathered to this
    process (rank 0), which writes the same bytes as the single-GPU route (E lines in the same
    canonical order).  No counterpart in the reference (single process)."""
    from phasm_b200 import multi, worker
    if not _has_fd(args.output):
        raise SystemExit("phasm overlap --gpus N needs an output with a file descriptor")
    logger.info("Building suffix tree and searching for pairwise overlaps...")
    port = multi.free_port()
    procs = multi.spawn_workers(args.gpus, args.fasta_input, args.min_length, port)
    try:
        logger.info("Writing to GFA2...")
        st = worker.run_rank(0, args.gpus, port, args.fasta_input, args.min_length, out=args.output)
    finally:
        codes = [p.wait() for p in procs]
    if any(codes):
        raise SystemExit("phasm overlap: a GPU worker failed (exit codes %s)" % codes)
    _write_stats(args, st)
    logger.info("Done.")


def _write_stats(args, st):
    """``--stats json``: counters and device stage times of the run, one JSON object on stderr
    (for roofline accounting; off by default like every flag the reference
    does not have)."""
    if getattr(args, "stats", None) == "json":
        import json
        sys.stderr.write(json.dumps({"phasm_overlap_stats": dict(st)}) + "\n")


def _has_fd(f) -> bool:
    try:
        f.fileno()
        return True
    except (AttributeError, OSError, ValueError):
        return False


def _out_of_scope(args):
    sys.stderr.write("phasm %s: this build provides the `overlap` stage only "
                     "(B200 overlapper); use the reference PHASM for layout/chain/phase.\n"
                     % args.stage)
    return 2


def main(argv=None):
    parser = argparse.ArgumentParser(
        description="PHASM: Haplotype-aware de novo genome assembly.")
    parser.set_defaults(func=None)

    parser.add_argument(
        '-v', '--verbose', action='count', default=0, required=False,
        help="Increase verbosity level, number of levels: 0, 1, 2"
    )

    subparsers = parser.add_subparsers()

    overlap_parser = subparsers.add_parser(
        'overlap', help="Find pairwise exact overlaps between reads with a "
                        "given minimum length."
    )
    overlap_parser.set_defaults(func=overlap)
    overlap_parser.add_argument(
        '-l', '--min-length', type=int, required=False, default=1000,
        help="The minimum overlap length between two reads (default: 1000)."
    )
    overlap_parser.add_argument(
        '-o', '--output', type=argparse.FileType('w'), required=False,
        default=sys.stdout,
        help="The output file. Defaults to standard output."
    )
    overlap_parser.add_argument(
        '--python-io', action='store_true',
        help="Read the FASTA and write the GFA2 lines in Python, row by row, as the "
             "reference does (default: native reader/writer, same bytes)."
    )
    overlap_parser.add_argument(
        '--gpus', type=int, required=False, default=1,
        help="Number of GPUs of this machine to shard the reads over (default: 1)."
    )
    overlap_parser.add_argument(
        '--stats', choices=['json'], required=False, default=None,
        help="Write counters and stage times of the run as one JSON line to stderr."
    )
    overlap_parser.add_argument(
        'fasta_input', help="The fasta file with reads.")

    for stage in ("layout", "chain", "phase"):
        p = subparsers.add_parser(stage, help="(not part of this build)")
        p.set_defaults(func=_out_of_scope, stage=stage)
        p.add_argument('rest', nargs=argparse.REMAINDER)

    args = parser.parse_args(argv)

    # same logging set-up as the reference: root logger at INFO on stderr
    logger.setLevel(logging.INFO)
    phasm_logger = logging.getLogger('phasm')
    phasm_logger.setLevel(logging.WARNING)
    formatter = logging.Formatter("{asctime} - {levelname}:{name}:{message}",
                                  style="{")
    handler = logging.StreamHandler()
    handler.setLevel(logging.DEBUG)
    handler.setFormatter(formatter)
    logger.addHandler(handler)
    if args.verbose > 0:
        phasm_logger.setLevel(logging.INFO)
    if args.verbose > 1:
        logger.setLevel(logging.DEBUG)
        phasm_logger.setLevel(logging.DEBUG)

    if not args.func:
        parser.print_help()
        return 0
    return args.func(args) or 0


if __name__ == "__main__":
    sys.exit(main())
