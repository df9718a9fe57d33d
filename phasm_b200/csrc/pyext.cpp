// pyext.cpp -- the drop-in extension module `phasm.overlapper`.
//
// Mirrors the reference binding /root/reference/src/phasm.cpp:8-18 name for name:
//     ExactOverlapper()                      phasm.cpp:13
//     .add_sequence(id, seq)                 phasm.cpp:14  (positional; str or bytes; copied)
//     .overlaps(min_length) -> list[tuple[str, str, int, int, int, int]]   phasm.cpp:15
// but the work happens in libphasm_ovl.so (sm_100a CUDA) through the C ABI of
// include/phasm_ovl.h.  This file holds no algorithm: argument conversion, error mapping and
// the tuple list.  Extras beyond the reference are additive (add_read, add_fasta,
// overlaps_arrays, write_gfa, stats).
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>

#include <cstdlib>
#include <new>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/phasm_ovl.h"

namespace py = pybind11;

namespace {

[[noreturn]] void raise_for(int code, const std::string& msg) {
    // pybind11 defaults the reference relies on: std::bad_alloc -> MemoryError,
    // std::exception -> RuntimeError
    if (code == OVL_E_NOMEM) throw std::bad_alloc();
    if (code == OVL_E_ALPHABET || code == OVL_E_ARG) throw py::value_error(msg);
    throw std::runtime_error(msg);
}

class Overlapper {
public:
    Overlapper() {
        int dev = 0;
        if (const char* e = std::getenv("PHASM_OVL_DEVICE")) dev = std::atoi(e);
        const int rc = ovl_create(dev, &ctx_);
        if (rc != OVL_OK) {
            ctx_ = nullptr;
            raise_for(rc, std::string("ExactOverlapper: ") + ovl_last_error(nullptr));
        }
    }
    ~Overlapper() { ovl_destroy(ctx_); }
    Overlapper(const Overlapper&) = delete;
    Overlapper& operator=(const Overlapper&) = delete;

    void check(int rc) const {
        if (rc != OVL_OK) raise_for(rc, ovl_last_error(ctx_));
    }

    void add_sequence(const std::string& id, const std::string& seq) {
        check(ovl_add_sequence(ctx_, id.data(), id.size(), seq.data(), seq.size()));
    }

    void add_read(const std::string& name, const std::string& seq) {
        check(ovl_add_read(ctx_, name.data(), name.size(), seq.data(), seq.size()));
    }

    uint32_t add_fasta(const std::string& path, bool both_strands) {
        uint32_t n = 0;
        int rc;
        {
            py::gil_scoped_release nogil;
            rc = ovl_add_fasta(ctx_, path.c_str(), both_strands ? 1 : 0, &n);
        }
        check(rc);
        return n;
    }

    // run the device path; rows stay owned by the ctx
    void run(unsigned min_length, const ovl_row** rows, uint64_t* n) {
        // the reference holds the GIL for the whole call (phasm.cpp:15 has no call guard)
        check(ovl_overlaps(ctx_, min_length, rows, n));
    }

    py::list overlaps(unsigned min_length) {
        const ovl_row* rows = nullptr;
        uint64_t n = 0;
        run(min_length, &rows, &n);
        uint32_t m = 0;
        check(ovl_entry_count(ctx_, &m));
        // one str object per entry, shared by all rows that mention it
        std::vector<py::object> ids(m);
        for (uint32_t e = 0; e < m; ++e) {
            const char* p = nullptr;
            size_t len = 0;
            check(ovl_entry_id(ctx_, e, &p, &len));
            ids[e] = py::reinterpret_steal<py::object>(PyUnicode_DecodeUTF8(p, (Py_ssize_t)len, nullptr));
            if (!ids[e]) throw py::error_already_set();
        }
        py::list out(n);
        py::object zero = py::int_(0);
        for (uint64_t i = 0; i < n; ++i) {
            const ovl_row& r = rows[i];
            PyObject* t = PyTuple_New(6);
            if (!t) throw py::error_already_set();
            PyTuple_SET_ITEM(t, 0, ids[r.a].inc_ref().ptr());
            PyTuple_SET_ITEM(t, 1, ids[r.b].inc_ref().ptr());
            PyTuple_SET_ITEM(t, 2, PyLong_FromUnsignedLong(r.astart));
            PyTuple_SET_ITEM(t, 3, PyLong_FromUnsignedLong(r.aend));
            PyTuple_SET_ITEM(t, 4, zero.inc_ref().ptr());  // bstart is always 0 (overlapper.cpp:81,109)
            PyTuple_SET_ITEM(t, 5, PyLong_FromUnsignedLong(r.bend));
            PyList_SET_ITEM(out.ptr(), (Py_ssize_t)i, t);
        }
        return out;
    }

    py::array_t<uint32_t> overlaps_arrays(unsigned min_length) {
        const ovl_row* rows = nullptr;
        uint64_t n = 0;
        run(min_length, &rows, &n);
        py::array_t<uint32_t> a({(py::ssize_t)n, (py::ssize_t)5});
        if (n) std::memcpy(a.mutable_data(), rows, n * sizeof(ovl_row));
        return a;
    }

    uint64_t write_gfa(int fd, bool header, bool segments, bool edges) {
        uint64_t nb = 0;
        int rc;
        {
            py::gil_scoped_release nogil;
            rc = ovl_write_gfa(ctx_, fd, (header ? 1 : 0) | (segments ? 2 : 0) | (edges ? 4 : 0), &nb);
        }
        check(rc);
        return nb;
    }

    // optional device post-pass: the layout stage's alignment filters over the rows of the last call
    // (ovl_filter_rows; phasm/cli/assembler.py:76-86, phasm/filter.py:36-122)
    py::array_t<uint32_t> filter_rows(bool contained_reads, unsigned min_read_length, unsigned min_overlap_length,
                                      py::object max_overhang_abs, double max_overhang_rel, bool collapse_pairs) {
        ovl_filter_opts o;
        std::memset(&o, 0, sizeof o);
        o.contained_reads = contained_reads ? 1 : 0;
        o.min_read_length = min_read_length;
        o.min_overlap_length = min_overlap_length;
        o.max_overhang = max_overhang_abs.is_none() ? 0 : 1;
        o.max_overhang_abs = max_overhang_abs.is_none() ? 0 : max_overhang_abs.cast<int64_t>();
        o.max_overhang_rel = max_overhang_rel;
        o.collapse_pairs = collapse_pairs ? 1 : 0;
        const ovl_row* rows = nullptr;
        uint64_t n = 0;
        check(ovl_filter_rows(ctx_, &o, &rows, &n, nullptr));
        py::array_t<uint32_t> a({(py::ssize_t)n, (py::ssize_t)5});
        if (n) std::memcpy(a.mutable_data(), rows, n * sizeof(ovl_row));
        return a;
    }

    // optional device post-pass: the edge list of build_assembly_graph (phasm/assembly_graph.py:136-179) over
    // the rows of the last call / the rows kept by the last filter_rows call (ovl_build_graph).  Returns
    // (edges, stats) like phasm_b200.overlapper.ExactOverlapper.build_graph: a structured numpy array
    // (u, v, edge_len, overlap_len, first_seq) sorted by (u, v), and networkx's node / edge counts.
    py::tuple build_graph(bool filtered) {
        const ovl_edge* edges = nullptr;
        uint64_t n = 0;
        ovl_graph_stats gs;
        check(ovl_build_graph(ctx_, filtered ? 1 : 0, &edges, &n, &gs));
        py::module_ np = py::module_::import("numpy");
        py::list fields;
        fields.append(py::make_tuple("u", "<u4"));
        fields.append(py::make_tuple("v", "<u4"));
        fields.append(py::make_tuple("edge_len", "<i4"));
        fields.append(py::make_tuple("overlap_len", "<u4"));
        fields.append(py::make_tuple("first_seq", "<u8"));
        py::object dtype = np.attr("dtype")(fields);
        py::bytes raw(n ? reinterpret_cast<const char*>(edges) : "", (size_t)n * sizeof(ovl_edge));
        py::object arr = np.attr("frombuffer")(raw, dtype).attr("copy")();
        py::dict d;
        d["rows_in"] = gs.rows_in;
        d["rows_overlap"] = gs.rows_overlap;
        d["n_edges"] = gs.n_edges;
        d["n_nodes"] = gs.n_nodes;
        d["ms_device"] = gs.ms_device;
        return py::make_tuple(arr, d);
    }

    // entry ids ("name+" / "name-" for reads) and lengths, in entry order: the node names and read
    // lengths the graph builder needs (phasm/assembly_graph.py of this repo)
    py::list ids() const {
        py::list out;
        const uint32_t m = (uint32_t)size();
        for (uint32_t i = 0; i < m; ++i) {
            const char* p = nullptr;
            size_t n = 0;
            check(ovl_entry_id(ctx_, i, &p, &n));
            out.append(py::str(p ? p : "", n));
        }
        return out;
    }

    py::array_t<uint32_t> lengths() const {
        const uint32_t m = (uint32_t)size();
        py::array_t<uint32_t> a((py::ssize_t)m);
        uint32_t* d = a.mutable_data();
        for (uint32_t i = 0; i < m; ++i) check(ovl_entry_length(ctx_, i, d + i));
        return a;
    }

    uint64_t run_only(unsigned min_length) {
        const ovl_row* rows = nullptr;
        uint64_t n = 0;
        run(min_length, &rows, &n);
        return n;
    }

    py::dict stats() const {
        ovl_stats s;
        check(ovl_get_stats(ctx_, &s));
        py::dict d;
        d["entries"] = s.entries;
        d["bases"] = s.bases;
        d["candidates"] = s.candidates;
        d["hits"] = s.hits;
        d["rows"] = s.rows;
        d["kmer"] = s.kmer;
        d["anchor_stride"] = s.anchor_stride;
        d["launches"] = s.launches;
        d["ms_h2d"] = s.ms_h2d;
        d["ms_device"] = s.ms_device;
        d["ms_d2h"] = s.ms_d2h;
        d["ms_total"] = s.ms_total;
        d["ms_pack"] = s.ms_pack;
        d["ms_anchor"] = s.ms_anchor;
        d["ms_scan"] = s.ms_scan;
        d["ms_verify"] = s.ms_verify;
        d["ms_sort"] = s.ms_sort;
        d["ms_emit"] = s.ms_emit;
        d["ms_exposed"] = s.ms_exposed;
        d["granules"] = s.granules;
        d["anchors"] = s.anchors;
        d["scan_survivors"] = s.scan_survivors;
        d["ascii_bytes"] = s.ascii_bytes;
        d["operand_bytes"] = s.operand_bytes;
        d["mirror"] = s.mirror;
        d["retries"] = s.retries;
        d["big_buckets"] = s.big_buckets;
        return d;
    }

    size_t size() const {
        uint32_t m = 0;
        check(ovl_entry_count(ctx_, &m));
        return m;
    }

private:
    ovl_ctx* ctx_ = nullptr;
};

}  // namespace

PYBIND11_MODULE(overlapper, m) {
    m.doc() = "Optimised C++ all-vs-all exact overlap finder for PHASM";  // phasm.cpp:9-10
    py::class_<Overlapper>(m, "ExactOverlapper")
        .def(py::init<>())
        .def("add_sequence", &Overlapper::add_sequence)
        .def("overlaps", &Overlapper::overlaps)
        // ---- additive extras (not in the reference) ----
        .def("add_read", &Overlapper::add_read)
        .def("add_fasta", &Overlapper::add_fasta, py::arg("path"), py::arg("both_strands") = true)
        .def("overlaps_arrays", &Overlapper::overlaps_arrays)
        .def("run", &Overlapper::run_only)
        .def("write_gfa", &Overlapper::write_gfa, py::arg("fd"), py::arg("header") = true,
             py::arg("segments") = true, py::arg("edges") = true)
        .def("filter_rows", &Overlapper::filter_rows, py::arg("contained_reads") = true,
             py::arg("min_read_length") = 0, py::arg("min_overlap_length") = 0,
             py::arg("max_overhang_abs") = py::int_(1000), py::arg("max_overhang_rel") = 0.8,
             py::arg("collapse_pairs") = false)
        .def("build_graph", &Overlapper::build_graph, py::arg("filtered") = true)
        .def("ids", &Overlapper::ids)
        .def("lengths", &Overlapper::lengths)
        .def("stats", &Overlapper::stats)
        .def("__len__", &Overlapper::size);
}
