// ovl_io.inl -- host-side text I/O of the C ABI (included at the end of ovl.cu):
//   ovl_add_fasta   FASTA -> entry store.  Replaces the dinopy FastaReader loop of
//                   /root/reference/phasm/cli/assembler.py:32-40 (dinopy is not vendored).
//   ovl_write_gfa   entry store + rows of the last ovl_overlaps* call -> GFA2 text.  Replaces
//                   the per-row Python loop /root/reference/phasm/cli/assembler.py:30,38,46-48
//                   over gfa_header/gfa_line (/root/reference/phasm/io/gfa.py:230-240).
// Pure byte shuffling on the host; no overlap computation happens here.
#include <algorithm>
#include <cmath>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <thread>
#include <unistd.h>

namespace {

// append bytes to the record being staged at the end of the pinned ASCII area
int stage_append(ovl_ctx* c, size_t rec_off, size_t& rec_len, const char* p, size_t n) {
    if (!n) return OVL_OK;
    const size_t need = rec_off + align16(rec_len + n) + SLACK;
    OVL_TRY(pin_ensure(c, c->ascii, need, rec_off + rec_len));
    memcpy(c->ascii.p + rec_off + rec_len, p, n);
    rec_len += n;
    return OVL_OK;
}

inline size_t rstrip(const char* p, size_t n) {
    while (n && (p[n - 1] == '\r' || p[n - 1] == ' ' || p[n - 1] == '\t')) --n;
    return n;
}

inline char* put_u32(char* p, uint32_t v) {
    char tmp[10];
    int n = 0;
    do {
        tmp[n++] = (char)('0' + v % 10);
        v /= 10;
    } while (v);
    while (n) *p++ = tmp[--n];
    return p;
}

int write_all(ovl_ctx* c, int fd, const char* p, size_t n) {
    while (n) {
        const ssize_t w = write(fd, p, n);
        if (w < 0) {
            if (errno == EINTR) continue;
            return fail(c, OVL_E_ARG, "ovl_write_gfa: write failed: %s", strerror(errno));
        }
        p += w;
        n -= (size_t)w;
    }
    return OVL_OK;
}

// One pass over the lines of a FASTA image.  on_header(name, len) opens a record, on_seq(p, n) adds a
// sequence line to it; returns false if sequence data precedes the first header.
template <class H, class Q>
bool fasta_lines(const char* base, const char* end, H on_header, Q on_seq) {
    const char* p = base;
    bool open_rec = false;
    while (p < end) {
        const char* nl = static_cast<const char*>(memchr(p, '\n', (size_t)(end - p)));
        const char* le = nl ? nl : end;
        const size_t n = rstrip(p, (size_t)(le - p));
        if (n && p[0] == '>') {
            if (!on_header(p + 1, n - 1)) return true;  // the callback asked to stop
            open_rec = true;
        } else if (n) {
            if (!open_rec) return false;
            if (!on_seq(p, n)) return true;
        }
        p = nl ? nl + 1 : end;
    }
    return true;
}

// FASTA rules (documented; dinopy's own are unverifiable here): a record starts
// at a line beginning with '>'; its name is the rest of that line, right-stripped; its sequence
// is the concatenation of the following lines (right-stripped, blank lines skipped) up to the
// next '>' line or EOF.  Bytes are taken as they are: anything outside upper-case ACGT is
// rejected later by the pack kernel (OVL_E_ALPHABET).
// nranks > 1: records are split into contiguous, base-balanced ranges in rank order; only the
// records of `rank` are staged, the others are declared by name and length (ovl_add_remote with
// names), so that every rank builds the same entry table and rank 0 can write every id.
int add_fasta_impl(ovl_ctx* c, const char* path, int both_strands, int rank, int nranks, uint32_t* n_records,
                   uint32_t* rec_lo, uint32_t* rec_hi) {
    const int fd = open(path, O_RDONLY);
    if (fd < 0) return fail(c, OVL_E_ARG, "ovl_add_fasta: cannot open '%s': %s", path, strerror(errno));
    struct stat sb;
    if (fstat(fd, &sb) != 0) {
        close(fd);
        return fail(c, OVL_E_ARG, "ovl_add_fasta: cannot stat '%s'", path);
    }
    const size_t fsz = (size_t)sb.st_size;
    if (!fsz) {
        close(fd);
        return OVL_OK;
    }
    void* mp = mmap(nullptr, fsz, PROT_READ, MAP_PRIVATE, fd, 0);
    close(fd);
    if (mp == MAP_FAILED) return fail(c, OVL_E_NOMEM, "ovl_add_fasta: mmap of '%s' failed", path);
    madvise(mp, fsz, MADV_SEQUENTIAL);
    const char* const base = static_cast<const char*>(mp);
    const char* const end = base + fsz;
    int rc = OVL_OK;
    uint32_t nrec = 0, lo = 0, hi = 0xffffffffu;
    try {
        if (nranks > 1) {  // pass 1: record lengths -> this rank's record range
            std::vector<uint64_t> csum{0};
            const bool ok = fasta_lines(base, end,
                [&](const char*, size_t) { csum.push_back(csum.back()); return true; },
                [&](const char*, size_t n) { csum.back() += n; return true; });
            if (!ok) rc = fail(c, OVL_E_ARG, "ovl_add_fasta: '%s' has sequence data before the first '>' line", path);
            const size_t R = csum.size() - 1;
            auto cut = [&](int k) -> uint32_t {  // boundary whose prefix sum is closest to k/nranks of the bases
                if (k <= 0) return 0;
                if (k >= nranks) return (uint32_t)R;
                const double target = (double)csum[R] * k / nranks;
                size_t i = (size_t)(std::lower_bound(csum.begin(), csum.end(), (uint64_t)target) - csum.begin());
                if (i > R) i = R;
                if (i > 0 && fabs((double)csum[i - 1] - target) <= fabs((double)csum[i] - target)) --i;
                return (uint32_t)i;
            };
            lo = cut(rank);
            hi = cut(rank + 1);
            if (hi < lo) hi = lo;
        }
        if (rc == OVL_OK) rc = pin_ensure(c, c->ascii, c->ascii_used + (nranks > 1 ? 0 : fsz) + SLACK, c->ascii_used);
        bool open_rec = false, local = true;
        const char* name = nullptr;
        size_t name_len = 0, rec_off = 0, rec_len = 0;
        auto close_rec = [&]() -> int {
            if (!open_rec) return OVL_OK;
            open_rec = false;
            const uint64_t src = local ? rec_off : NO_SRC;
            if (local) c->ascii_used = rec_off + align16(rec_len);
            ++nrec;
            if (both_strands) {
                OVL_TRY(push_entry(c, name, name_len, "+", src, rec_len, 0));
                return push_entry(c, name, name_len, "-", src, rec_len, 1);
            }
            return push_entry(c, name, name_len, nullptr, src, rec_len, 0);
        };
        if (rc == OVL_OK) {
            const bool ok = fasta_lines(base, end,
                [&](const char* nm, size_t nl) {
                    rc = close_rec();
                    name = nm;
                    name_len = nl;
                    rec_off = c->ascii_used;
                    rec_len = 0;
                    open_rec = true;
                    local = nrec >= lo && nrec < hi;
                    return rc == OVL_OK;
                },
                [&](const char* p, size_t n) {
                    if (local) rc = stage_append(c, rec_off, rec_len, p, n);
                    else rec_len += n;
                    return rc == OVL_OK;
                });
            if (!ok) rc = fail(c, OVL_E_ARG, "ovl_add_fasta: '%s' has sequence data before the first '>' line", path);
        }
        if (rc == OVL_OK) rc = close_rec();
    } catch (const std::bad_alloc&) {
        rc = fail(c, OVL_E_NOMEM, "out of host memory");
    }
    munmap(mp, fsz);
    if (n_records) *n_records = nrec;
    if (rec_lo) *rec_lo = lo;
    if (rec_hi) *rec_hi = hi == 0xffffffffu ? nrec : hi;
    return rc;
}

}  // namespace

extern "C" {

int ovl_add_fasta(ovl_ctx* c, const char* path, int both_strands, uint32_t* n_records) {
    if (!c || !path) return fail(c, OVL_E_ARG, "ovl_add_fasta: NULL argument");
    if (n_records) *n_records = 0;
    OVL_TRY(set_dev(c));
    return add_fasta_impl(c, path, both_strands, 0, 1, n_records, nullptr, nullptr);
}

int ovl_add_fasta_shard(ovl_ctx* c, const char* path, int both_strands, int rank, int nranks, uint32_t* n_records,
                        uint32_t* rec_lo, uint32_t* rec_hi) {
    if (!c || !path || nranks < 1 || rank < 0 || rank >= nranks) return fail(c, OVL_E_ARG, "ovl_add_fasta_shard: bad argument");
    if (n_records) *n_records = 0;
    OVL_TRY(set_dev(c));
    return add_fasta_impl(c, path, both_strands, rank, nranks, n_records, rec_lo, rec_hi);
}

// what: bit0 header line, bit1 segment lines, bit2 edge lines (of the last ovl_overlaps* call).
//   H\tVN:z:2.0                                  assembler.py:30 / gfa.py:234-240
//   S\t<name>\t<len>\t*                          assembler.py:38   one per RECORD, input order
//   E\t*\t<a>\t<b>\t<as>\t<ae>\t0\t<be>\t*       assembler.py:46-48
// A record is an entry that is not a reverse-complement twin (ovl_add_read / _bulk / _fasta
// with both strands: its name is the entry id without the trailing '+'; ovl_add_sequence
// entries are written with their id as it is).
int ovl_write_gfa(ovl_ctx* c, int fd, int what, uint64_t* bytes_written) {
    if (!c || fd < 0) return fail(c, OVL_E_ARG, "ovl_write_gfa: bad argument");
    uint64_t total = 0;
    const size_t M = c->len.size();
    try {
        std::string out;
        if (what & 1) out += "H\tVN:z:2.0\n";
        if (what & 2) {
            out.reserve(out.size() + 24 * M + c->id_blob.size());
            char num[16];
            for (size_t e = 0; e < M; ++e) {
                if (c->flags[e] & 1) continue;
                const bool twin = e + 1 < M && (c->flags[e + 1] & 1);
                const char* id = c->id_blob.data() + c->id_off[e];
                size_t il = (size_t)(c->id_off[e + 1] - c->id_off[e]);
                if (twin && il) --il;  // drop the '+'
                out += "S\t";
                out.append(id, il);
                out += '\t';
                out.append(num, (size_t)(put_u32(num, c->len[e]) - num));
                out += "\t*\n";
            }
        }
        if (!out.empty()) {
            OVL_TRY(write_all(c, fd, out.data(), out.size()));
            total += out.size();
        }
        if (what & 4) {
            const ovl_row* rows = nullptr;
            uint64_t n = 0;
            OVL_TRY(ovl_fetch_rows(c, &rows, &n));
            size_t max_id = 0;
            for (size_t e = 0; e < M; ++e) max_id = std::max(max_id, (size_t)(c->id_off[e + 1] - c->id_off[e]));
            const size_t max_line = 2 * max_id + 64;
            unsigned nt = std::thread::hardware_concurrency();
            if (nt == 0) nt = 1;
            if (nt > 32) nt = 32;
            const uint64_t CHUNK = 1u << 16;  // rows per formatting task
            for (uint64_t r0 = 0; r0 < n; r0 += (uint64_t)nt * CHUNK) {
                const unsigned tasks = (unsigned)std::min<uint64_t>(nt, (n - r0 + CHUNK - 1) / CHUNK);
                std::vector<std::string> bufs(tasks);
                auto work = [&](unsigned t) {
                    const uint64_t lo = r0 + (uint64_t)t * CHUNK, hi = std::min<uint64_t>(lo + CHUNK, n);
                    std::string& b = bufs[t];
                    b.resize((size_t)(hi - lo) * max_line);
                    char* p = &b[0];
                    for (uint64_t i = lo; i < hi; ++i) {
                        const ovl_row& r = rows[i];
                        memcpy(p, "E\t*\t", 4);
                        p += 4;
                        size_t il = (size_t)(c->id_off[r.a + 1] - c->id_off[r.a]);
                        memcpy(p, c->id_blob.data() + c->id_off[r.a], il);
                        p += il;
                        *p++ = '\t';
                        il = (size_t)(c->id_off[r.b + 1] - c->id_off[r.b]);
                        memcpy(p, c->id_blob.data() + c->id_off[r.b], il);
                        p += il;
                        *p++ = '\t';
                        p = put_u32(p, r.astart);
                        *p++ = '\t';
                        p = put_u32(p, r.aend);
                        memcpy(p, "\t0\t", 3);
                        p += 3;
                        p = put_u32(p, r.bend);
                        memcpy(p, "\t*\n", 3);
                        p += 3;
                    }
                    b.resize((size_t)(p - b.data()));
                };
                std::vector<std::thread> th;
                for (unsigned t = 1; t < tasks; ++t) th.emplace_back(work, t);
                work(0);
                for (auto& x : th) x.join();
                for (unsigned t = 0; t < tasks; ++t) {
                    OVL_TRY(write_all(c, fd, bufs[t].data(), bufs[t].size()));
                    total += bufs[t].size();
                }
            }
        }
    } catch (const std::bad_alloc&) {
        return fail(c, OVL_E_NOMEM, "out of host memory");
    } catch (const std::system_error& e) {
        return fail(c, OVL_E_ARG, "ovl_write_gfa: %s", e.what());
    }
    if (bytes_written) *bytes_written = total;
    return OVL_OK;
}

}  // extern "C"

// =========================================================================================
// Optional device post-pass over the rows of the last call: the `layout` stage's alignment
// filters (include/phasm_ovl.h, k8_filter.cuh).
extern "C" int ovl_filter_rows(ovl_ctx* c, const ovl_filter_opts* o, const ovl_row** rows, uint64_t* n_rows,
                               ovl_filter_stats* fs) {
    using namespace ovl;
    if (!c || !o || !rows || !n_rows) return fail(c, OVL_E_ARG, "ovl_filter_rows: NULL argument");
    OVL_TRY(set_dev(c));
    *rows = nullptr;
    *n_rows = 0;
    if (fs) memset(fs, 0, sizeof *fs);
    const uint64_t n = c->n_rows;
    const uint32_t M = (uint32_t)c->len.size();
    if (fs) fs->rows_in = n;
    if (n == 0) {  // an empty pass has an empty, valid filter result (ovl_build_graph(use_filtered) -> empty graph)
        c->n_frows = 0;
        c->frows_valid = true;
        return OVL_OK;
    }
    if (n > 0x7fffffffull) return fail(c, OVL_E_LIMIT, "ovl_filter_rows: more than 2^31 rows (CUB item counts are int)");
    OVL_TRY(gather_rows_pull(c));  // multi-GPU, rank 0: the other ranks' rows are still in their HBM
    cudaStream_t st = c->st;
    const size_t fm_bytes = ((size_t)M * 4 + 7) & ~(size_t)7;
    OVL_TRY(dev_ensure(c, c->d_fmark, fm_bytes + 16));  // firstmark[M] | counters[2]
    OVL_TRY(dev_ensure(c, c->d_fkeep, n * 8));          // keep[n] | keep2[n]
    OVL_TRY(dev_ensure(c, c->d_foff, (n + 1) * 8));
    OVL_TRY(dev_ensure(c, c->d_frows, n * sizeof(Row)));
    uint32_t* fmark = c->d_fmark.as<uint32_t>();
    unsigned long long* fctr = reinterpret_cast<unsigned long long*>(c->d_fmark.as<uint8_t>() + fm_bytes);
    uint32_t* keep = c->d_fkeep.as<uint32_t>();
    uint32_t* keep2 = keep + n;
    CK(cudaEventRecord(c->ev[E_D2H0], st));
    CK(cudaMemsetAsync(fmark, 0xff, (size_t)M * 4, st));
    CK(cudaMemsetAsync(fctr, 0, 16, st));
    FilterParams P;
    P.rows = c->d_rows.as<Row>();
    P.n = n;
    P.ent = c->d_ent.as<uint2>();
    P.firstmark = fmark;
    P.keep = keep;
    P.contained_reads = o->contained_reads;
    P.min_read_length = o->min_read_length;
    P.min_overlap_length = o->min_overlap_length;
    P.max_overhang = o->max_overhang;
    P.collapse_pairs = o->collapse_pairs;
    P.max_overhang_abs = (long long)o->max_overhang_abs;
    P.max_overhang_rel = o->max_overhang_rel;
    P.counters = fctr;
    unsigned blocks = (unsigned)((n + 255) / 256);
    if (blocks > (unsigned)c->n_sm * 16) blocks = (unsigned)c->n_sm * 16;
    if (o->contained_reads) {
        k8_mark<<<blocks, 256, 0, st>>>(P);
        k8_count_marked<<<(M + 255) / 256 < 1024 ? (M + 255) / 256 : 1024, 256, 0, st>>>(fmark, M, fctr + 1, 0xffffffffu);
    }
    k8_keep<<<blocks, 256, 0, st>>>(P);
    const uint32_t* kfinal = keep;
    if (o->collapse_pairs) {
        k8_collapse<<<blocks, 256, 0, st>>>(P, keep2);
        kfinal = keep2;
    }
    CK(cudaGetLastError());
    size_t tb = 0;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, tb, kfinal, c->d_foff.as<uint64_t>(), (int)n, st));
    OVL_TRY(dev_ensure(c, c->d_tmp, tb));
    tb = c->d_tmp.cap;
    CK(cub::DeviceScan::ExclusiveSum(c->d_tmp.p, tb, kfinal, c->d_foff.as<uint64_t>(), (int)n, st));
    k8_compact<<<blocks, 256, 0, st>>>(P.rows, n, kfinal, c->d_foff.as<uint64_t>(), c->d_frows.as<Row>());
    CK(cudaGetLastError());
    uint64_t h[4] = {0, 0, 0, 0};  // last offset, last flag, contained rows, marked entries
    CK(cudaMemcpyAsync(&h[0], c->d_foff.as<uint64_t>() + (n - 1), 8, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(&h[1], kfinal + (n - 1), 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(&h[2], fctr, 16, cudaMemcpyDeviceToHost, st));
    CK(cudaEventRecord(c->ev[E_D2H1], st));
    CK(cudaStreamSynchronize(st));
    const uint64_t kept = h[0] + (h[1] & 0xffffffffull);
    if (kept) {
        OVL_TRY(pin_ensure(c, c->h_frows, kept * sizeof(ovl_row), 0));
        CK(cudaMemcpyAsync(c->h_frows.p, c->d_frows.p, kept * sizeof(ovl_row), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        *rows = reinterpret_cast<const ovl_row*>(c->h_frows.p);
    }
    *n_rows = kept;
    c->n_frows = kept;
    c->frows_valid = true;
    if (fs) {
        fs->rows_kept = kept;
        fs->rows_contained = h[2];
        fs->entries_removed = h[3];
        fs->ms_device = ev_ms(c, E_D2H0, E_D2H1);
    }
    return OVL_OK;
}

// =========================================================================================
// Assembly-graph edge list (k10_graph.cuh; include/phasm_ovl.h has the contract).
extern "C" int ovl_build_graph(ovl_ctx* c, int use_filtered, const ovl_edge** edges, uint64_t* n_edges, ovl_graph_stats* gs) {
    using namespace ovl;
    static_assert(sizeof(ovl_edge) == sizeof(Edge) && sizeof(ovl_edge) == 24, "edge record layout");
    if (!c || !edges || !n_edges) return fail(c, OVL_E_ARG, "ovl_build_graph: NULL argument");
    OVL_TRY(set_dev(c));
    *edges = nullptr;
    *n_edges = 0;
    if (gs) memset(gs, 0, sizeof *gs);
    if (use_filtered && !c->frows_valid) return fail(c, OVL_E_ARG, "ovl_build_graph: no ovl_filter_rows result for the rows of the last pass");
    const uint64_t n = use_filtered ? c->n_frows : c->n_rows;
    const uint32_t M = (uint32_t)c->len.size();
    if (gs) gs->rows_in = n;
    if (n == 0) return OVL_OK;
    if (!c->pairs) return fail(c, OVL_E_ARG, "ovl_build_graph: the entry table is not strand-paired (node reversal is entry ^ 1)");
    if (n >= 0x3fffffffull) return fail(c, OVL_E_LIMIT, "ovl_build_graph: more than 2^30 rows");
    if (!use_filtered) OVL_TRY(gather_rows_pull(c));  // multi-GPU, rank 0: the other ranks' rows are still in their HBM
    cudaStream_t st = c->st;
    OVL_TRY(dev_ensure(c, c->d_gflag, n * 4));
    OVL_TRY(dev_ensure(c, c->d_goff, (n + 1) * 8));
    OVL_TRY(dev_ensure(c, c->d_gmark, (size_t)M * 4 + 16));  // node_mark[M] | counter
    GraphParams P{};
    P.rows = use_filtered ? c->d_frows.as<Row>() : c->d_rows.as<Row>();
    P.n = n;
    P.ent = c->d_ent.as<uint2>();
    P.flag = c->d_gflag.as<uint32_t>();
    P.off = c->d_goff.as<uint64_t>();
    P.node_mark = c->d_gmark.as<uint32_t>();
    unsigned long long* nctr = reinterpret_cast<unsigned long long*>(c->d_gmark.as<uint8_t>() + (((size_t)M * 4 + 7) & ~(size_t)7));
    unsigned blocks = (unsigned)((n + 255) / 256);
    if (blocks > (unsigned)c->n_sm * 16) blocks = (unsigned)c->n_sm * 16;
    CK(cudaEventRecord(c->ev[E_D2H0], st));
    CK(cudaMemsetAsync(P.node_mark, 0, (size_t)M * 4 + 16, st));
    k10_flag<<<blocks, 256, 0, st>>>(P);
    CK(cudaGetLastError());
    size_t tb = 0;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, tb, P.flag, c->d_goff.as<uint64_t>(), (int)n, st));
    OVL_TRY(dev_ensure(c, c->d_tmp, tb));
    tb = c->d_tmp.cap;
    CK(cub::DeviceScan::ExclusiveSum(c->d_tmp.p, tb, P.flag, c->d_goff.as<uint64_t>(), (int)n, st));
    uint64_t h[2] = {0, 0};
    CK(cudaMemcpyAsync(&h[0], c->d_goff.as<uint64_t>() + (n - 1), 8, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(&h[1], P.flag + (n - 1), 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    const uint64_t n_ovl = h[0] + (h[1] & 0xffffffffull), n_items = 2 * n_ovl;
    uint64_t n_e = 0, n_nodes = 0;
    if (n_items) {
        // keys | keys' and seqs | seqs' (radix sort ping-pong), last[] reuses the flag array's owner
        OVL_TRY(dev_ensure(c, c->d_gkeys, 2 * n_items * 8));
        OVL_TRY(dev_ensure(c, c->d_gseqs, 3 * n_items * 4));
        OVL_TRY(dev_ensure(c, c->d_gedges, n_items * sizeof(Edge) + (n_items + 1) * 8));
        uint64_t* k0 = c->d_gkeys.as<uint64_t>();
        uint32_t* s0 = c->d_gseqs.as<uint32_t>();
        P.keys = k0;
        P.seqs = s0;
        P.n_items = n_items;
        k10_items<<<blocks, 256, 0, st>>>(P);
        CK(cudaGetLastError());
        cub::DoubleBuffer<uint64_t> dk(k0, k0 + n_items);
        cub::DoubleBuffer<uint32_t> dv(s0, s0 + n_items);
        const int end_bit = 32 + ceil_log2((uint64_t)M + 1);
        tb = 0;
        CK(cub::DeviceRadixSort::SortPairs(nullptr, tb, dk, dv, (int)n_items, 0, end_bit, st));
        OVL_TRY(dev_ensure(c, c->d_tmp, tb));
        tb = c->d_tmp.cap;
        CK(cub::DeviceRadixSort::SortPairs(c->d_tmp.p, tb, dk, dv, (int)n_items, 0, end_bit, st));
        P.keys = dk.Current();
        P.seqs = dv.Current();
        P.last = s0 + 2 * n_items;
        uint64_t* eoff = reinterpret_cast<uint64_t*>(c->d_gedges.as<uint8_t>() + n_items * sizeof(Edge));
        P.eoff = eoff;
        P.edges = c->d_gedges.as<Edge>();
        unsigned b2 = (unsigned)((n_items + 255) / 256);
        if (b2 > (unsigned)c->n_sm * 16) b2 = (unsigned)c->n_sm * 16;
        k10_runs<<<b2, 256, 0, st>>>(P);
        CK(cudaGetLastError());
        tb = 0;
        CK(cub::DeviceScan::ExclusiveSum(nullptr, tb, P.last, eoff, (int)n_items, st));
        OVL_TRY(dev_ensure(c, c->d_tmp, tb));
        tb = c->d_tmp.cap;
        CK(cub::DeviceScan::ExclusiveSum(c->d_tmp.p, tb, P.last, eoff, (int)n_items, st));
        k10_edges<<<b2, 256, 0, st>>>(P);
        k8_count_marked<<<(M + 255) / 256 < 1024 ? (M + 255) / 256 : 1024, 256, 0, st>>>(P.node_mark, M, nctr, 0u);
        CK(cudaGetLastError());
        uint64_t g[2] = {0, 0};
        CK(cudaMemcpyAsync(&g[0], eoff + (n_items - 1), 8, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(&g[1], nctr, 8, cudaMemcpyDeviceToHost, st));
        CK(cudaEventRecord(c->ev[E_D2H1], st));
        CK(cudaStreamSynchronize(st));
        n_e = g[0] + 1;  // the last item always closes a run
        n_nodes = g[1];
        OVL_TRY(pin_ensure(c, c->h_gedges, n_e * sizeof(ovl_edge), 0));
        CK(cudaMemcpyAsync(c->h_gedges.p, c->d_gedges.p, n_e * sizeof(ovl_edge), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        *edges = reinterpret_cast<const ovl_edge*>(c->h_gedges.p);
    } else {
        CK(cudaEventRecord(c->ev[E_D2H1], st));
        CK(cudaStreamSynchronize(st));
    }
    *n_edges = n_e;
    if (gs) {
        gs->rows_overlap = n_ovl;
        gs->n_edges = n_e;
        gs->n_nodes = n_nodes;
        gs->ms_device = ev_ms(c, E_D2H0, E_D2H1);
    }
    return OVL_OK;
}

// =========================================================================================
// GFA2 ingest (host only): S and E lines -> segment table + index edges.  Restates
// gfa2_segment_to_read / gfa2_parse_segments / gfa2_parse_edge / gfa2_line_to_la of
// /root/reference/phasm/io/gfa.py:34-47,65-107 (see include/phasm_ovl.h).
#include <unordered_map>

struct ovl_gfa {
    std::vector<std::string> names;
    std::vector<uint32_t> lengths;
    std::vector<ovl_gfa_edge> edges;
};

namespace {

inline bool py_space(char ch) { return ch == ' ' || ch == '\t' || ch == '\n' || ch == '\r' || ch == '\v' || ch == '\f'; }

inline void py_strip(const char*& p, size_t& n) {
    while (n && py_space(p[0])) { ++p; --n; }
    while (n && py_space(p[n - 1])) --n;
}

// split [p, p+n) on '\t' into at most `cap` fields; returns the number of fields found (may exceed cap)
inline size_t split_tabs(const char* p, size_t n, const char** f, size_t* fl, size_t cap) {
    size_t k = 0, start = 0;
    for (size_t i = 0; i <= n; ++i) {
        if (i == n || p[i] == '\t') {
            if (k < cap) { f[k] = p + start; fl[k] = i - start; }
            ++k;
            start = i + 1;
        }
    }
    return k;
}

// Python int() of a field (surrounding whitespace allowed, optional sign); `dollar`: drop one trailing '$' first
inline bool parse_pos(const char* p, size_t n, bool dollar, uint32_t* out) {
    if (dollar && n && p[n - 1] == '$') --n;
    py_strip(p, n);
    if (n && p[0] == '+') { ++p; --n; }
    if (!n) return false;
    uint64_t v = 0;
    for (size_t i = 0; i < n; ++i) {
        if (p[i] < '0' || p[i] > '9') return false;
        v = v * 10 + (uint64_t)(p[i] - '0');
        if (v > 0xffffffffull) return false;
    }
    *out = (uint32_t)v;
    return true;
}

int gfa_fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_create_err = buf;
    return code;
}

}  // namespace

extern "C" {

int ovl_gfa_read(const char* path, ovl_gfa** out) {
    if (!path || !out) return gfa_fail(OVL_E_ARG, "ovl_gfa_read: NULL argument");
    *out = nullptr;
    const int fd = open(path, O_RDONLY);
    if (fd < 0) return gfa_fail(OVL_E_ARG, "cannot open '%s': %s", path, strerror(errno));
    struct stat sb;
    if (fstat(fd, &sb) != 0) {
        close(fd);
        return gfa_fail(OVL_E_ARG, "cannot stat '%s': %s", path, strerror(errno));
    }
    const size_t size = (size_t)sb.st_size;
    const char* data = nullptr;
    if (size) {
        void* m = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
        if (m == MAP_FAILED) {
            close(fd);
            return gfa_fail(OVL_E_NOMEM, "cannot map '%s': %s", path, strerror(errno));
        }
        data = static_cast<const char*>(m);
    }
    close(fd);
    ovl_gfa* g = new (std::nothrow) ovl_gfa();
    int rc = g ? OVL_OK : gfa_fail(OVL_E_NOMEM, "out of host memory");
    try {
        std::unordered_map<std::string, uint32_t> index;
        // pass 1: segments (gfa2_parse_segments: a dict, so a repeated name keeps its first slot
        // and takes the later line's length)
        for (size_t pos = 0, ln = 1; rc == OVL_OK && pos < size; ++ln) {
            const char* nl = static_cast<const char*>(memchr(data + pos, '\n', size - pos));
            const size_t len = nl ? (size_t)(nl - (data + pos)) : size - pos;
            const char* p = data + pos;
            pos += len + 1;
            if (!len || p[0] != 'S') continue;
            size_t n = len;
            py_strip(p, n);
            const char* f[4];
            size_t fl[4];
            const size_t k = split_tabs(p, n, f, fl, 4);
            uint32_t L = 0;
            if (k < 4) rc = gfa_fail(OVL_E_ARG, "%s:%zu: S line with %zu fields (4 needed)", path, ln, k);
            else if (!parse_pos(f[2], fl[2], false, &L)) rc = gfa_fail(OVL_E_ARG, "%s:%zu: bad segment length", path, ln);
            if (rc != OVL_OK) break;
            const char* np = f[1];
            size_t nn = fl[1];
            py_strip(np, nn);
            std::string name(np, nn);
            auto it = index.find(name);
            if (it == index.end()) {
                index.emplace(std::move(name), (uint32_t)g->names.size());
                g->names.emplace_back(np, nn);
                g->lengths.push_back(L);
            } else {
                g->lengths[it->second] = L;
            }
        }
        // pass 2: edges
        for (size_t pos = 0, ln = 1; rc == OVL_OK && pos < size; ++ln) {
            const char* nl = static_cast<const char*>(memchr(data + pos, '\n', size - pos));
            const size_t len = nl ? (size_t)(nl - (data + pos)) : size - pos;
            const char* p = data + pos;
            pos += len + 1;
            if (!len || p[0] != 'E') continue;
            size_t n = len;
            py_strip(p, n);
            const char* f[9];
            size_t fl[9];
            const size_t k = split_tabs(p, n, f, fl, 9);
            if (k < 9) {
                rc = gfa_fail(OVL_E_ARG, "%s:%zu: E line with %zu fields (9 needed)", path, ln, k);
                break;
            }
            ovl_gfa_edge e;
            memset(&e, 0, sizeof e);
            uint32_t* seg[2] = {&e.a, &e.b};
            uint8_t* strand[2] = {&e.a_strand, &e.b_strand};
            for (int s = 0; s < 2 && rc == OVL_OK; ++s) {
                const char* sp = f[2 + s];
                size_t sn = fl[2 + s];
                py_strip(sp, sn);
                if (!sn) {
                    rc = gfa_fail(OVL_E_ARG, "%s:%zu: empty segment reference", path, ln);
                    break;
                }
                *strand[s] = (uint8_t)sp[sn - 1];
                auto it = index.find(std::string(sp, sn - 1));
                if (it == index.end()) rc = gfa_fail(OVL_E_ARG, "%s:%zu: edge names unknown segment '%.*s'", path, ln, (int)(sn - 1), sp);
                else *seg[s] = it->second;
            }
            if (rc != OVL_OK) break;
            uint32_t* posf[4] = {&e.astart, &e.aend, &e.bstart, &e.bend};
            for (int q = 0; q < 4; ++q)
                if (!parse_pos(f[4 + q], fl[4 + q], true, posf[q])) {
                    rc = gfa_fail(OVL_E_ARG, "%s:%zu: bad position in field %d", path, ln, 5 + q);
                    break;
                }
            if (rc != OVL_OK) break;
            g->edges.push_back(e);
        }
    } catch (const std::bad_alloc&) {
        rc = gfa_fail(OVL_E_NOMEM, "out of host memory");
    }
    if (data) munmap(const_cast<char*>(data), size);
    if (rc != OVL_OK) {
        delete g;
        return rc;
    }
    *out = g;
    return OVL_OK;
}

void ovl_gfa_free(ovl_gfa* g) { delete g; }

int ovl_gfa_counts(const ovl_gfa* g, uint32_t* n_segments, uint64_t* n_edges) {
    if (!g) return OVL_E_ARG;
    if (n_segments) *n_segments = (uint32_t)g->names.size();
    if (n_edges) *n_edges = g->edges.size();
    return OVL_OK;
}

int ovl_gfa_segment(const ovl_gfa* g, uint32_t idx, const char** name, size_t* name_len, uint32_t* length) {
    if (!g || idx >= g->names.size()) return OVL_E_ARG;
    if (name) *name = g->names[idx].data();
    if (name_len) *name_len = g->names[idx].size();
    if (length) *length = g->lengths[idx];
    return OVL_OK;
}

int ovl_gfa_edges(const ovl_gfa* g, const ovl_gfa_edge** edges, uint64_t* n_edges) {
    if (!g || !edges || !n_edges) return OVL_E_ARG;
    *edges = g->edges.empty() ? nullptr : g->edges.data();
    *n_edges = g->edges.size();
    return OVL_OK;
}

}  // extern "C"
