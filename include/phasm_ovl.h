/*
 * phasm_ovl.h -- C ABI of the B200 (sm_100a) exact overlapper: the drop-in boundary for
 * PHASM's `phasm overlap` hot path.
 *
 * Every entry point names the reference interface it replaces (paths relative to the
 * AbeelLab/phasm checkout).  Plain pointers and sizes only; no C++/torch/pybind11 types.
 * The library is built by `__graft_entry__.build()` into phasm_b200/_lib/libphasm_ovl.so.
 *
 * Conventions
 *   - every call returns 0 on success, a negative OVL_E_* code on failure;
 *     ovl_last_error() gives the message (a NULL ctx returns the last ovl_create error).
 *   - an ovl_ctx is bound to one CUDA device and is NOT thread-safe (the reference object is
 *     not either: src/phasm.cpp:12-15 holds the GIL for the whole call).
 *   - there is NO CPU fallback: without a usable CUDA device ovl_create fails.
 */
#ifndef PHASM_OVL_H
#define PHASM_OVL_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OVL_OK            0
#define OVL_E_CUDA       -1  /* CUDA runtime / driver error (-> RuntimeError)            */
#define OVL_E_NOMEM      -2  /* host or device allocation failed (-> MemoryError)        */
#define OVL_E_ALPHABET   -3  /* a base outside upper-case ACGT (-> ValueError); the       */
                             /* reference compares raw bytes (src/overlapper.h:26), the  */
                             /* 2-bit device layout cannot: documented deviation         */
#define OVL_E_ARG        -4  /* bad argument / bad call order                            */
#define OVL_E_LIMIT      -5  /* a size limit of the device layout was exceeded           */
#define OVL_E_NCCL       -6  /* NCCL error in the multi-GPU path                         */

typedef struct ovl_ctx ovl_ctx;

/* One overlap row.  Replaces OverlapT = tuple<string,string,int,int,int,int>
 * (src/overlapper.h:17): a and b index the entry table (ovl_entry_id gives the strings);
 * bstart is always 0 in the reference (src/overlapper.cpp:81,109) and is not stored. */
typedef struct {
    uint32_t a;      /* entry whose suffix / substring matches        */
    uint32_t b;      /* entry whose prefix / whole sequence matches   */
    uint32_t astart;
    uint32_t aend;
    uint32_t bend;
} ovl_row;

/* Counters and device-side stage times of the last ovl_overlaps* call. */
typedef struct {
    uint64_t entries;          /* M                                                       */
    uint64_t bases;            /* sum of entry lengths (both strands when added as reads) */
    uint64_t ascii_bytes;      /* host->device bytes of sequence data                     */
    uint64_t granules;         /* 64-base packed granules (16 B each)                     */
    uint64_t anchors;          /* entries with |b| >= min_length                          */
    uint64_t candidates;       /* (a,p,b), a != b, emitted by the scan (mirror mode: only  */
                               /* the self-verifying half)                                */
    uint64_t hits;             /* candidates that verified                                */
    uint64_t rows;
    uint64_t compared_bases;   /* sum m_c (only when stats collection is on, else 0)      */
    uint64_t operand_bytes;    /* sum 2*ceil(m_c/4), idem                                 */
    uint64_t scan_positions;   /* A positions with |a|-p >= min_length                    */
    uint32_t kmer;             /* anchor length k = min(max(l,1),32)                      */
    uint32_t anchor_stride;    /* every stride-th A position is probed (1 = dense scan)   */
    uint32_t launches;         /* kernels of this library launched by the call            */
    uint32_t lib_launches;     /* CUB device-wide primitive calls (sort / scan)           */
    uint32_t retries;          /* scan re-runs after a candidate-buffer overflow          */
    uint32_t mirror;           /* 1: strand-mirror halving was on (every entry has its twin) */
    uint64_t verified;         /* candidates actually compared by the verifier            */
    uint64_t hits_sent;        /* multi-GPU mirror mode: hit keys shipped to other ranks  */
    uint64_t hits_received;    /* ... and received from them                              */
    float ms_h2d, ms_pack, ms_anchor, ms_scan, ms_sort, ms_verify, ms_emit, ms_d2h;
    float ms_device;           /* pack..emit: ASCII resident in HBM -> rows resident      */
    float ms_total;            /* h2d..d2h                                                */
    uint32_t big_buckets;      /* A entries with more than 256 hit keys (sorted on the    */
                               /* device by k7_sort_big; there is no fallback path)       */
    uint32_t reserved0;        /* 0 (keeps the 64-bit fields below aligned)               */
    uint64_t scan_survivors;   /* probes that passed the scan's bit filter                */
    uint64_t rows_local;       /* rows this rank emitted (multi-GPU: before the gather)   */
    float ms_gather;           /* multi-GPU: rows of all ranks into rank 0's HBM          */
    float ms_exposed;          /* ovl_overlaps: device time left after the last upload    */
                               /* piece landed (== ms_device when nothing was uploaded)   */
} ovl_stats;

/* ---- lifetime -------------------------------------------------------------------------
 * Replaces ExactOverlapper::ExactOverlapper() (src/overlapper.cpp:19, src/phasm.cpp:13). */
int ovl_create(int device, ovl_ctx **out);
void ovl_destroy(ovl_ctx *ctx);
const char *ovl_last_error(const ovl_ctx *ctx);

/* ---- entry store ----------------------------------------------------------------------
 * ovl_add_sequence replaces ExactOverlapper::addSequence(id, seq) (src/overlapper.cpp:22-26,
 * bound as add_sequence in src/phasm.cpp:14): id and seq are COPIED, the caller keeps
 * ownership; ids are opaque bytes (duplicates allowed). */
int ovl_add_sequence(ovl_ctx *ctx, const char *id, size_t id_len, const char *seq, size_t seq_len);

/* ovl_add_read replaces the pair of calls phasm/cli/assembler.py:39-40 makes per FASTA
 * record: entry 2i = (name+"+", seq), entry 2i+1 = (name+"-", reverse_complement(seq)).
 * The sequence is staged once; the reverse complement is produced on the device. */
int ovl_add_read(ovl_ctx *ctx, const char *name, size_t name_len, const char *seq, size_t seq_len);

/* Bulk form of ovl_add_read / ovl_add_sequence: n records, record i is
 * buf[offsets[i] .. offsets[i+1]); names may be NULL (ids are then empty) or a blob with
 * name_offsets[n+1].  both_strands != 0 adds "+" and "-" entries per record. */
int ovl_add_bulk(ovl_ctx *ctx, const uint8_t *buf, const uint64_t *offsets, uint32_t n,
                 const char *names, const uint64_t *name_offsets, int both_strands);

/* Pre-size the pinned host staging area (optional; avoids regrowth while adding). */
int ovl_reserve(ovl_ctx *ctx, uint64_t seq_bytes, uint32_t n_entries);

/* Multi-GPU only: declare n records whose BASES live on another rank (only their lengths are
 * known here).  Together with the local records they must form the same entry table, in the
 * same order, on every rank; ids of remote entries are "" (+ strand suffix). */
int ovl_add_remote(ovl_ctx *ctx, const uint32_t *lengths, uint32_t n, int both_strands);

int ovl_clear(ovl_ctx *ctx); /* drop all entries, keep device buffers */

int ovl_entry_count(const ovl_ctx *ctx, uint32_t *m);
int ovl_entry_id(const ovl_ctx *ctx, uint32_t idx, const char **id, size_t *id_len);
int ovl_entry_length(const ovl_ctx *ctx, uint32_t idx, uint32_t *len);

/* ---- the hot path ----------------------------------------------------------------------
 * ovl_overlaps replaces ExactOverlapper::overlaps(min_length) (src/overlapper.cpp:28-150,
 * bound in src/phasm.cpp:15): host buffers in, host rows out.  min_length 0 behaves as 1
 * like the reference.  *rows points into memory owned by ctx, valid until the next
 * ovl_overlaps* / ovl_clear / ovl_destroy call.  Rows are sorted by (a,b,astart,aend,bend);
 * the reference's order is unspecified so any comparison sorts.
 * An empty entry table returns 0 rows (the reference segfaults there). */
int ovl_overlaps(ovl_ctx *ctx, uint32_t min_length, const ovl_row **rows, uint64_t *n_rows);

/* The same call split at the PCIe boundary, for device-resident timing:
 *   ovl_upload          host ASCII -> HBM (no-op if nothing was added since the last upload)
 *   ovl_overlaps_device pack .. emit, rows stay in HBM
 *   ovl_fetch_rows      HBM -> host rows                                                  */
int ovl_upload(ovl_ctx *ctx);
int ovl_overlaps_device(ovl_ctx *ctx, uint32_t min_length, uint64_t *n_rows);
int ovl_fetch_rows(ovl_ctx *ctx, const ovl_row **rows, uint64_t *n_rows);

/* Collect sum m_c / operand bytes in the verifier (off by default).  The byte accounting of
 * SURVEY.md 8(d) is defined on the canonical candidates (first k-mer of B at every A position),
 * so this also forces the dense stride-1 scan. */
int ovl_set_collect_stats(ovl_ctx *ctx, int on);
/* Tuning knobs.  "anchor_stride": 0 = auto (a power of two <= min(64, l-k+1) picked by a measured
 * cost model of table build vs scan), else an upper bound (1 forces the dense scan).
 * "pack_blocks_per_sm": K1 grid size per SM (1..32, default 16).  "serial_verify": 1 = chunk-by-chunk early-exit
 * verifier instead of the head/tail two-pass form.  "mirror": 0 = compare every candidate even
 * when every read is present on both strands (default 1: compare one of each strand-mirror
 * pair and emit both rows).  "filter_kb": budget of the scan's bit filter in KiB (default 8192).
 * "upload_chunks": pieces of the host -> HBM upload of ovl_overlaps (1..16, default 8; every piece
 * is packed while the next one is copied).  "cand_per_entry" / "min_cand_cap": first candidate
 * capacity (keys per local entry, floor); an overflow is detected on the device, the buffers grow and
 * the pass repeats (stats.retries).  "peer_timeout_ms": how long a multi-GPU pass waits for a peer. */
int ovl_set_option(ovl_ctx *ctx, const char *name, int64_t value);
int ovl_get_stats(const ovl_ctx *ctx, ovl_stats *out);

/* Device-side stopwatch on the stream every kernel of this ctx is launched on (the CUDA-event
 * bracket bench.py times K steps with): start records an event, stop records a second one,
 * waits for it and returns the elapsed milliseconds between the two. */
int ovl_timer_start(ovl_ctx *ctx);
int ovl_timer_stop(ovl_ctx *ctx, float *ms);

/* ---- text I/O on the host -------------------------------------
 * ovl_add_fasta replaces the dinopy FastaReader loop of phasm/cli/assembler.py:32-40: every
 * record of the FASTA file becomes one read (both_strands != 0: entries name+"+" and name+"-",
 * reverse complement made on the device) or one entry.  Record name = header line after '>',
 * right-stripped; multi-line records and blank lines are accepted (tests/data/test.fasta). */
int ovl_add_fasta(ovl_ctx *ctx, const char *path, int both_strands, uint32_t *n_records);
/* Multi-GPU form: every rank reads the same file; the records are split into nranks contiguous,
 * base-balanced ranges in rank order, only the records [*rec_lo, *rec_hi) of `rank` are staged and
 * the others are declared by name and length, so that every rank holds the same entry table (and
 * rank 0 every id for the GFA2 writer). */
int ovl_add_fasta_shard(ovl_ctx *ctx, const char *path, int both_strands, int rank, int nranks,
                        uint32_t *n_records, uint32_t *rec_lo, uint32_t *rec_hi);

/* ovl_write_gfa replaces the GFA2 writing loop of phasm/cli/assembler.py:30,38,46-48 over
 * gfa_header/gfa_line (phasm/io/gfa.py:230-240), byte for byte: `what` bit0 = header line
 * "H\tVN:z:2.0", bit1 = one "S\t<name>\t<len>\t*" line per record in input order, bit2 = one
 * "E\t*\t<a>\t<b>\t<astart>\t<aend>\t0\t<bend>\t*" line per row of the last ovl_overlaps* call.
 * Writes to the file descriptor fd (the caller flushes its own buffered writer first). */
int ovl_write_gfa(ovl_ctx *ctx, int fd, int what, uint64_t *bytes_written);

/* ---- overlap-consumer filters on the device (OPTIONAL, never applied by
 * ovl_overlaps*) -------------------------------------------------------------------------
 * ovl_filter_rows runs the alignment filters of the reference's `layout` stage over the rows of
 * the last ovl_overlaps* call, in the reference's order and with its streaming semantics
 * (phasm/cli/assembler.py:76-86,96-98; phasm/filter.py:36-122; LocalAlignment.classify,
 * get_overlap_length, get_overhang: phasm/alignments.py:237-258), rows taken in the order this
 * library returns them (= the E-line order ovl_write_gfa writes):
 *   contained_reads     ContainedReads(): a row whose A (B) is contained marks that entry and is
 *                       dropped; a later row touching a marked entry is dropped too;
 *   min_read_length     MinReadLength(n), 0 = off;   min_overlap_length  MinOverlapLength(n), 0 = off;
 *   max_overhang        MaxOverhang(max_overhang_abs, max_overhang_rel): keep iff
 *                       overhang <= min(abs, rel * overlap_length);
 *   collapse_pairs      of the surviving rows keep the LAST one per ordered pair (a, b) -- the row
 *                       that ends up on the edge a->b when the reference adds edges in order
 *                       (assembler.py:72-74, build_assembly_graph).
 * The kept rows come back in the same canonical order in a buffer owned by the ctx (valid until
 * the next ovl_filter_rows / ovl_overlaps* call); the unfiltered rows stay available. */
typedef struct {
    uint32_t contained_reads;
    uint32_t min_read_length;
    uint32_t min_overlap_length;
    uint32_t max_overhang;
    int64_t max_overhang_abs;
    double max_overhang_rel;
    uint32_t collapse_pairs;
    uint32_t reserved0;
} ovl_filter_opts;

typedef struct {
    uint64_t rows_in;
    uint64_t rows_kept;
    uint64_t rows_contained;   /* rows classified A_CONTAINED or B_CONTAINED                  */
    uint64_t entries_removed;  /* entries marked by ContainedReads (its nodes_to_remove)      */
    float ms_device;
    uint32_t reserved0;
} ovl_filter_stats;

int ovl_filter_rows(ovl_ctx *ctx, const ovl_filter_opts *opts, const ovl_row **rows, uint64_t *n_rows,
                    ovl_filter_stats *stats);

/* ---- the same filter for callers that cannot allocate ------------------------------------------
 * ovl_filter_rows_static has NO counterpart in the reference (its filters are Python objects over
 * Python lists); it exists for hosts of this library that must not trigger a cudaMalloc at call
 * time -- a pipeline whose memory pool owns the whole GPU, a latency-bound loop that cannot afford
 * the device-wide synchronisation cudaMalloc / cudaFree imply, a process whose memory budget is
 * fixed at load time.  ovl_filter_rows grows ctx-owned buffers on demand; this variant allocates
 * nothing, ever: no cudaMalloc / cudaFree / host allocation, no ctx, no library primitive that needs
 * temporary storage.  Its working memory is a fixed array inside the library image
 * (OVL_STATIC_SCRATCH_BYTES of HBM on every GPU that loads the library, held until the process
 * exits), which is why the job size is capped:
 *     n_rows <= OVL_STATIC_MAX_ROWS and n_entries <= OVL_STATIC_MAX_ENTRIES, else OVL_E_LIMIT.
 * The caller owns every buffer, all in DEVICE memory of `device`:
 *     d_rows[n_rows]         in   rows in canonical order (as ovl_overlaps* returns them)
 *     d_entry_len[n_entries] in   length of every entry (ovl_entry_length)
 *     d_rows_out[n_rows]     out  kept rows, same order, the first d_counts->rows_kept are valid
 *     d_counts               out  one ovl_static_counts
 * `stream` is a cudaStream_t (NULL = the default stream).  The call only enqueues work and returns;
 * results are ready when the stream reaches that point (synchronise or order a copy behind it).
 * `opts` means what it means for ovl_filter_rows and the kept rows / counters are identical to that
 * call's on the same rows.  A row naming an entry >= n_entries is dropped and counted in rows_bad.
 * `err` (may be NULL) receives the message of a failing call; ovl_last_error is not involved.
 * First call per GPU: looks up the device address of the scratch array (and loads the kernels), so
 * make that call during set-up, before any stream capture.  Sharing: ONE scratch array per GPU for
 * all threads and streams; every call clears what it needs, so calls ordered on one stream are safe
 * back to back, but two calls that may overlap on the same GPU race and must be serialised by the caller. */
#define OVL_STATIC_MAX_ROWS      (1u << 21)   /* 2 097 152 rows    */
#define OVL_STATIC_MAX_ENTRIES   (1u << 19)   /*   524 288 entries */
#define OVL_STATIC_SCRATCH_BYTES 6295808u     /* 6.0 MiB per GPU   */

typedef struct {
    uint64_t rows_kept;
    uint64_t rows_contained;   /* as ovl_filter_stats (0 unless opts->contained_reads)        */
    uint64_t entries_removed;  /* as ovl_filter_stats (0 unless opts->contained_reads)        */
    uint64_t rows_bad;         /* rows dropped because a or b >= n_entries                    */
} ovl_static_counts;

int ovl_filter_rows_static(int device, void *stream, const ovl_row *d_rows, uint64_t n_rows,
                           const uint32_t *d_entry_len, uint32_t n_entries, const ovl_filter_opts *opts,
                           ovl_row *d_rows_out, ovl_static_counts *d_counts, char *err, size_t err_cap);

/* ---- assembly-graph edge list on the device (graph part; OPTIONAL) ------
 * ovl_build_graph replaces build_assembly_graph (phasm/assembly_graph.py:136-179) as the reference's
 * `layout` stage feeds it (phasm/cli/assembler.py:93-100): the rows of the last ovl_overlaps* call
 * (use_filtered == 0) or the rows kept by the last ovl_filter_rows call (use_filtered != 0), taken in
 * order as LocalAlignments with brange = (0, bend).  Every row LocalAlignment.classify
 * (phasm/alignments.py:248-258) calls OVERLAP_AB adds the two directed edges
 *     a -> b        weight = astart,                           overlap_len = max(aend - astart, bend)
 *     b^1 -> a^1    weight = (|b| - bend) - (|a| - aend),      same overlap_len
 * (x^1 = OrientedRead.reverse() = the other strand's entry; needs a strand-paired entry table,
 * else OVL_E_ARG); contained rows add nothing; OVERLAP_BA cannot occur with bstart == 0 but is
 * implemented.  networkx.DiGraph semantics: the LAST row of an ordered node pair decides the
 * attributes, the FIRST its place in the adjacency order, a node exists iff it ends an edge.
 * The edges come back sorted by (u, v) (a CSR adjacency) in a pinned buffer owned by the ctx, valid
 * until the next ovl_build_graph / ovl_overlaps* call.  Nodes are entry indices. */
typedef struct {
    uint32_t u, v;
    int32_t edge_len;       /* the reference's edge attribute `weight` (edge_len)              */
    uint32_t overlap_len;
    uint64_t first_seq;     /* 2 * (row that first added the edge) + (0 | 1: which of its two add_edge calls) */
} ovl_edge;

typedef struct {
    uint64_t rows_in;
    uint64_t rows_overlap;  /* rows of an OVERLAP type (two add_edge calls each)               */
    uint64_t n_edges;       /* networkx.number_of_edges(g)                                     */
    uint64_t n_nodes;       /* networkx.number_of_nodes(g)                                     */
    float ms_device;
    uint32_t reserved0;
} ovl_graph_stats;

int ovl_build_graph(ovl_ctx *ctx, int use_filtered, const ovl_edge **edges, uint64_t *n_edges,
                    ovl_graph_stats *stats);

/* ---- GFA2 ingest for the consumer of the overlaps (parsing part; host only,
 * needs no GPU and no ctx) ----------------------------------------------------------------
 * ovl_gfa_read replaces the two Python passes of the reference's `layout` stage over the GFA2 file
 * (phasm/cli/assembler.py:55-60,93-95): gfa2_parse_segments / gfa2_segment_to_read
 * (phasm/io/gfa.py:34-47,105-107) for the S lines -- a later S line with the same name replaces
 * the earlier one, as the dict does -- and gfa2_parse_edge / gfa2_line_to_la (gfa.py:72-103) for
 * the E lines: fields 2 and 3 are segment names with a trailing orientation character, fields 4..7
 * positions with an optional trailing '$'.  An edge naming an unknown segment is an error
 * (the reference raises KeyError).  Edges come back in file order, on segment indices. */
typedef struct ovl_gfa ovl_gfa;
typedef struct {
    uint32_t a, b;             /* segment indices (first appearance order of the names)      */
    uint32_t astart, aend, bstart, bend;
    uint8_t a_strand, b_strand; /* '+' or '-' (the raw orientation character)               */
    uint8_t reserved[2];
} ovl_gfa_edge;

int ovl_gfa_read(const char *path, ovl_gfa **out);
void ovl_gfa_free(ovl_gfa *g);
int ovl_gfa_counts(const ovl_gfa *g, uint32_t *n_segments, uint64_t *n_edges);
int ovl_gfa_segment(const ovl_gfa *g, uint32_t idx, const char **name, size_t *name_len, uint32_t *length);
int ovl_gfa_edges(const ovl_gfa *g, const ovl_gfa_edge **edges, uint64_t *n_edges);

/* ---- multi-GPU (one process per GPU) ---------------------------------------------------
 * All ranks declare the SAME entry table in the same order (the reads of the other ranks by length
 * only: ovl_add_remote); rank r holds the bases of one contiguous range.  NCCL is used once, as the
 * bootstrap channel (libnccl is dlopen'ed on first use): ovl_comm_unique_id fills a 128-byte id on
 * rank 0, the caller ships it to the other ranks (a socket, torch.distributed, a file, ...) and
 * every rank calls ovl_comm_init.  The first pass then maps every rank's working arena into every
 * other rank (CUDA IPC); from there on all exchanges of a pass happen on the device over NVLink
 * (copy-engine pushes of the packed reads, grid-wide pulls of the item lists:
 * phasm_b200/csrc/k6_exchange.cuh) -- no NCCL call and no host round trip inside a pass.  Every rank calls ovl_overlaps_device / ovl_overlaps_local in step; an error or a
 * capacity overflow on any rank is seen by all of them in the same pass (all raise / all repeat). */
int ovl_comm_unique_id(uint8_t id_out[128]);
int ovl_comm_init(ovl_ctx *ctx, const uint8_t id[128], int rank, int nranks);
/* After ovl_upload on every rank: exchange the local entry ranges and check that they tile the
 * entry table in rank order. */
int ovl_comm_sync_ranges(ovl_ctx *ctx);
/* After a pass the rows live where they were emitted: rank r holds the rows of its A range in its
 * HBM (rank order == canonical order), like the reads.  On rank 0, ovl_overlaps_device reports the
 * row count of the whole job, and ovl_fetch_rows / ovl_comm_gather_rows[_device] PULL the other
 * ranks' rows over NVLink behind rank 0's own (the other ranks do not take part; valid until any
 * rank starts its next pass or changes its entries).  Other ranks get 0 rows from these calls. */
int ovl_comm_gather_rows_device(ovl_ctx *ctx, uint64_t *n_rows);
int ovl_comm_gather_rows(ovl_ctx *ctx, const ovl_row **rows, uint64_t *n_rows);
/* This rank's OWN rows of the last pass (stats.rows_local of them), HBM -> pinned host: every rank
 * of a multi-GPU job can bring its share to its host over its own PCIe link. */
int ovl_fetch_local_rows(ovl_ctx *ctx, const ovl_row **rows, uint64_t *n_rows);
/* ovl_overlaps for a consumer that stays sharded: host buffers in on every rank (chunked upload),
 * the pass, then ovl_fetch_local_rows.  Collective: every rank calls it. */
int ovl_overlaps_local(ovl_ctx *ctx, uint32_t min_length, const ovl_row **rows, uint64_t *n_rows);

/* ---- debug / test hooks (device layout introspection) -----------------------------------
 * Copy the 2-bit packed granules of entry idx (ceil(len/64)*16 bytes) to out; used by the
 * pack / reverse-complement parity tests. */
int ovl_debug_packed(ovl_ctx *ctx, uint32_t idx, uint8_t *out, size_t out_cap, size_t *out_len);
/* The multi-GPU placement rule, evaluated on the host (no device needed): for candidate (a, p, b)
 * with entry lengths la, lb on a job of n_ranks ranks owning the entry ranges rank_e[0..n_ranks],
 * *items = the OVL_HIT-style item bits (2 = suffix-prefix, 1 = containment) the candidate has to
 * compare itself under strand-mirror halving (0: its twins cover it), *verifier = the rank that
 * compares it.  Used by the tests of the "exactly one of every twin pair" property. */
int ovl_debug_mirror_items(int n_ranks, const uint32_t *rank_e, uint32_t a, uint32_t la, uint32_t b,
                           uint32_t lb, uint32_t p, uint32_t *items, int *verifier);

#ifdef __cplusplus
}
#endif
#endif /* PHASM_OVL_H */
