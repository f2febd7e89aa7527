// `_ldscore` -- drop-in replacement for the reference's pybind11 extension of the same name.
//
// Same Python surface as reference nldsc/ldscore/_ldscore/ldscore.cpp:17-54 (typed in
// nldsc/ldscore/_ldscore.pyi): LDScoreParams(bfile, *, n_snp, n_org, ld_wind, maf, std_thr,
// rsq_thr, positions), LDScoreResult with seven list attributes, calculate(params).  The body
// forwards to the C ABI in include/nldsc_ld.h; nothing here computes.  Error mapping follows
// the reference: std::invalid_argument -> ValueError (bad PLINK magic, stream.h:97-101),
// std::runtime_error -> RuntimeError.  Unlike the reference the GIL is released for the call.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string_view>
#include <unordered_set>
#include <stdexcept>
#include <algorithm>
#include <string>
#include <vector>

#include "../../include/nldsc_ld.h"

namespace py = pybind11;

namespace {

// Field names and order of reference data.h:33-66.
struct LDScoreParams {
    std::string bedfile;
    int n_snp = 0;
    int n_org = 0;
    double ld_wind = 0;
    std::vector<double> positions;
    double maf = 0;
    double std_thr = 0;
    double rsq_thr = 0;
};

// reference data.h:21-31
struct LDScoreResult {
    std::vector<double> l2, l2d, maf, residuals_std;
    std::vector<int> l2_ws, l2d_ws, l2d_wse;
};

LDScoreResult calculate_impl(LDScoreParams& p, int first_device, int n_devices) {
    if (static_cast<long long>(p.positions.size()) != p.n_snp)
        throw std::invalid_argument("positions must hold n_snp coordinates");
    LDScoreResult r;
    const size_t n = static_cast<size_t>(p.n_snp > 0 ? p.n_snp : 0);
    r.l2.resize(n); r.l2d.resize(n); r.maf.resize(n); r.residuals_std.resize(n);
    r.l2_ws.resize(n); r.l2d_ws.resize(n); r.l2d_wse.resize(n);
    nldsc_ld_params cp{};
    cp.bedfile = p.bedfile.c_str();
    cp.n_snp = p.n_snp;
    cp.n_org = p.n_org;
    cp.ld_wind = p.ld_wind;
    cp.positions = p.positions.data();
    cp.maf = p.maf;
    cp.std_thr = p.std_thr;
    cp.rsq_thr = p.rsq_thr;
    static_assert(sizeof(int) == sizeof(int32_t), "int32 result arrays");
    nldsc_ld_result cr{r.l2.data(), r.l2d.data(), r.maf.data(), r.residuals_std.data(),
                       r.l2_ws.data(), r.l2d_ws.data(), r.l2d_wse.data()};
    char err[512] = {0};
    int rc;
    {
        py::gil_scoped_release nogil;
        rc = n_devices > 0 ? nldsc_ld_calculate_on(&cp, &cr, first_device, n_devices, err, sizeof(err))
                           : nldsc_ld_calculate(&cp, &cr, err, sizeof(err));
    }
    if (rc == NLDSC_ERR_INVALID) throw std::invalid_argument(err);
    if (rc != NLDSC_OK) throw std::runtime_error(err);
    return r;
}

LDScoreResult calculate(LDScoreParams& p) { return calculate_impl(p, -1, 0); }

// ---- .bim table kept on the C++ side ---------------------------------------------------
// The reference parses the .bim with pandas (ldscore/common.py:86-89) and hands its CHR / SNP / BP columns back to
// pandas for the output table (routine.py:32-48).  At 100,000 SNPs that is 0.1-0.15 s of interpreter time per
// chromosome -- more than the GPU needs to score it -- and it serialises on the GIL when one process feeds 8 GPUs.
// read_bim parses the six whitespace-separated columns with the GIL released and keeps the strings here; positions
// come out as NumPy views, the .L2 writer takes the table itself, so no 100,000-element Python list is ever built.
struct BimTable {
    std::vector<std::string> chr, snp;
    std::vector<double> cm;
    std::vector<long long> bp;
    std::vector<double> bp_f;      // BP as written, for --ld-wind-kb windows and non-integer BP columns
    bool bp_integral = true;
};

BimTable read_bim(const std::string& path) {
    BimTable t;
    std::string err;
    {
        py::gil_scoped_release nogil;
        FILE* f = fopen(path.c_str(), "rb");
        if (!f) {
            err = "cannot open " + path;
        } else {
            std::string buf;
            fseek(f, 0, SEEK_END);
            const long sz = ftell(f);
            fseek(f, 0, SEEK_SET);
            buf.resize(sz > 0 ? static_cast<size_t>(sz) : 0);
            if (!buf.empty() && fread(&buf[0], 1, buf.size(), f) != buf.size()) err = "short read on " + path;
            fclose(f);
            const char* p = buf.data();
            const char* end = p + buf.size();
            auto is_sep = [](char c) { return c == '\t' || c == ' '; };
            size_t line_no = 0;
            while (err.empty() && p < end) {
                const char* eol = static_cast<const char*>(memchr(p, '\n', static_cast<size_t>(end - p)));
                const char* le = eol ? eol : end;
                const char* lend = (le > p && le[-1] == '\r') ? le - 1 : le;
                ++line_no;
                if (lend > p) {
                    std::string_view field[6];
                    int nf = 0;
                    const char* q = p;
                    while (q < lend && nf < 6) {
                        while (q < lend && is_sep(*q)) ++q;
                        const char* b = q;
                        while (q < lend && !is_sep(*q)) ++q;
                        if (q > b) field[nf++] = std::string_view(b, static_cast<size_t>(q - b));
                    }
                    if (nf < 4) {
                        err = path + ": line " + std::to_string(line_no) + " has fewer than 4 columns";
                        break;
                    }
                    t.chr.emplace_back(field[0]);
                    t.snp.emplace_back(field[1]);
                    char tmp[64];
                    auto num = [&](std::string_view v) {
                        const size_t n = std::min(v.size(), sizeof(tmp) - 1);
                        memcpy(tmp, v.data(), n);
                        tmp[n] = 0;
                        return tmp;
                    };
                    t.cm.push_back(strtod(num(field[2]), nullptr));
                    char* e2 = nullptr;
                    const char* bs = num(field[3]);
                    const long long bi = strtoll(bs, &e2, 10);
                    if (e2 && *e2 == 0) {
                        t.bp.push_back(bi);
                        t.bp_f.push_back(static_cast<double>(bi));
                    } else {
                        t.bp_integral = false;
                        t.bp.push_back(0);
                        t.bp_f.push_back(strtod(bs, nullptr));
                    }
                }
                p = eol ? eol + 1 : end;
            }
        }
    }
    if (!err.empty()) throw std::runtime_error(err);
    return t;
}

template <class T>
py::array_t<T> table_view(py::object owner, std::vector<T>& v) {
    return py::array_t<T>({static_cast<py::ssize_t>(v.size())}, {static_cast<py::ssize_t>(sizeof(T))}, v.data(), owner);
}

// NumPy view of one result vector; the array keeps the LDScoreResult alive, nothing is copied.
// The list attributes of the reference surface copy the whole vector into a new list on every read
// (ldscore.cpp:43-51) -- seconds at 1.2M SNPs; these do not.
template <class T>
py::array_t<T> view(py::object owner, std::vector<T>& v) {
    return py::array_t<T>({static_cast<py::ssize_t>(v.size())}, {static_cast<py::ssize_t>(sizeof(T))}, v.data(), owner);
}

// The .L2 table exactly as the reference's `out.to_csv(path, sep='\t', index=False, float_format='%.5f')` writes it
// (ldscore/routine.py:32-48,99): header, tab-separated, floats %.5f, NaN -> empty field, integers plain.
// pandas needs seconds for this at 1.2M rows; this is one pass of snprintf with the GIL released.
void write_l2(const std::string& path, const std::vector<std::string>& chr, const std::vector<std::string>& snp,
              py::array_t<long long, py::array::c_style | py::array::forcecast> bp,
              py::array_t<double, py::array::c_style | py::array::forcecast> l2,
              py::array_t<double, py::array::c_style | py::array::forcecast> l2d,
              py::array_t<double, py::array::c_style | py::array::forcecast> maf,
              py::array_t<double, py::array::c_style | py::array::forcecast> rstd,
              py::array_t<int, py::array::c_style | py::array::forcecast> wsa,
              py::array_t<int, py::array::c_style | py::array::forcecast> wsd,
              py::array_t<int, py::array::c_style | py::array::forcecast> wsde, bool extra) {
    const size_t n = chr.size();
    if (snp.size() != n || static_cast<size_t>(bp.size()) != n || static_cast<size_t>(l2.size()) != n ||
        static_cast<size_t>(l2d.size()) != n || static_cast<size_t>(maf.size()) != n || static_cast<size_t>(rstd.size()) != n ||
        static_cast<size_t>(wsa.size()) != n || static_cast<size_t>(wsd.size()) != n || static_cast<size_t>(wsde.size()) != n)
        throw std::invalid_argument("write_l2: columns differ in length");
    const long long* pbp = bp.data();
    const double *pl2 = l2.data(), *pl2d = l2d.data(), *pmaf = maf.data(), *prstd = rstd.data();
    const int *pwsa = wsa.data(), *pwsd = wsd.data(), *pwsde = wsde.data();
    py::gil_scoped_release nogil;
    FILE* f = fopen(path.c_str(), "w");
    if (!f) throw std::runtime_error("cannot open " + path + " for writing");
    std::string buf;
    buf.reserve(size_t(1) << 22);
    buf += extra ? "CHR\tSNP\tBP\tL2\tL2D\tMAF\tWSA\tWSD\tWSDE\tRSTD\n" : "CHR\tSNP\tBP\tL2\tL2D\n";
    char tmp[64];
    auto put_u = [&](unsigned long long u, int min_digits) {       // decimal digits, zero-padded to min_digits
        char d[24];
        int n = 0;
        do { d[n++] = static_cast<char>('0' + u % 10); u /= 10; } while (u);
        while (n < min_digits) d[n++] = '0';
        while (n) buf += d[--n];
    };
    // "%.5f" without printf (which costs ~0.4 us per number, 2.5 s per million rows): scale by 1e5, round, and CHECK the
    // rounding with an exactly rounded residual (one FMA); the rare value that sits within 1e-6 of a rounding boundary,
    // or is too large, goes through snprintf, so every byte is what printf would have written.
    auto put_f = [&](double v, char sep) {
        if (!std::isnan(v)) {
            const double a = std::fabs(v);
            const double r = std::nearbyint(a * 1e5);
            const double res = std::fma(a, 1e5, -r);
            if (a < 1e10 && std::fabs(res) < 0.5 - 1e-6) {
                if (std::signbit(v)) buf += '-';
                const unsigned long long q = static_cast<unsigned long long>(r);
                put_u(q / 100000ull, 1);
                buf += '.';
                put_u(q % 100000ull, 5);
            } else {
                buf.append(tmp, static_cast<size_t>(snprintf(tmp, sizeof(tmp), "%.5f", v)));
            }
        }
        buf += sep;
    };
    auto put_i = [&](long long v, char sep) {
        if (v < 0) { buf += '-'; put_u(0ull - static_cast<unsigned long long>(v), 1); }
        else put_u(static_cast<unsigned long long>(v), 1);
        buf += sep;
    };
    bool ok = true;
    for (size_t i = 0; i < n; ++i) {
        buf += chr[i]; buf += '\t';
        buf += snp[i]; buf += '\t';
        put_i(pbp[i], '\t');
        put_f(pl2[i], '\t');
        put_f(pl2d[i], extra ? '\t' : '\n');
        if (extra) {
            put_f(pmaf[i], '\t');
            put_i(pwsa[i], '\t'); put_i(pwsd[i], '\t'); put_i(pwsde[i], '\t');
            put_f(prstd[i], '\n');
        }
        if (buf.size() > (size_t(1) << 22) - 256) {
            ok = ok && fwrite(buf.data(), 1, buf.size(), f) == buf.size();
            buf.clear();
        }
    }
    ok = ok && fwrite(buf.data(), 1, buf.size(), f) == buf.size();
    ok = (fclose(f) == 0) && ok;
    if (!ok) throw std::runtime_error("short write on " + path);
}

// write_l2 with CHR / SNP / BP taken from a BimTable: no Python list of strings on the way
void write_l2_bim(const std::string& path, BimTable& t,
                  py::array_t<double, py::array::c_style | py::array::forcecast> l2,
                  py::array_t<double, py::array::c_style | py::array::forcecast> l2d,
                  py::array_t<double, py::array::c_style | py::array::forcecast> maf,
                  py::array_t<double, py::array::c_style | py::array::forcecast> rstd,
                  py::array_t<int, py::array::c_style | py::array::forcecast> wsa,
                  py::array_t<int, py::array::c_style | py::array::forcecast> wsd,
                  py::array_t<int, py::array::c_style | py::array::forcecast> wsde, bool extra) {
    if (!t.bp_integral) throw std::invalid_argument("write_l2_bim: the .bim has a non-integer BP column");
    py::array_t<long long> bp({static_cast<py::ssize_t>(t.bp.size())}, {static_cast<py::ssize_t>(sizeof(long long))}, t.bp.data(),
                              py::none());      // a view: write_l2 copies nothing
    write_l2(path, t.chr, t.snp, bp, l2, l2d, maf, rstd, wsa, wsd, wsde, extra);
}

}  // namespace

PYBIND11_MODULE(_ldscore, m) {
    m.doc() = "B200-native LD-score engine behind the reference's _ldscore interface";
    py::class_<LDScoreParams>(m, "LDScoreParams")
        .def(py::init<>())
        .def(py::init([](const std::string& bfile, int n_snp, int n_org, double ld_wind, double maf, double std_thr,
                         double rsq_thr, const std::vector<double>& positions) {
                 LDScoreParams p;
                 p.bedfile = bfile; p.n_snp = n_snp; p.n_org = n_org; p.ld_wind = ld_wind;
                 p.maf = maf; p.std_thr = std_thr; p.rsq_thr = rsq_thr; p.positions = positions;
                 return p;
             }),
             py::arg("bfile"), py::kw_only(), py::arg("n_snp"), py::arg("n_org"), py::arg("ld_wind"), py::arg("maf"),
             py::arg("std_thr"), py::arg("rsq_thr"), py::arg("positions"))
        .def_readwrite("bedfile", &LDScoreParams::bedfile)
        .def_readwrite("n_snp", &LDScoreParams::n_snp)
        .def_readwrite("n_org", &LDScoreParams::n_org)
        .def_readwrite("ld_wind", &LDScoreParams::ld_wind)
        .def_readwrite("positions", &LDScoreParams::positions)
        .def_readwrite("maf", &LDScoreParams::maf)
        .def_readwrite("std_thr", &LDScoreParams::std_thr)
        .def_readwrite("rsq_thr", &LDScoreParams::rsq_thr);
    py::class_<LDScoreResult>(m, "LDScoreResult")
        .def(py::init<>())
        .def_readwrite("l2", &LDScoreResult::l2)
        .def_readwrite("l2d", &LDScoreResult::l2d)
        .def_readwrite("maf", &LDScoreResult::maf)
        .def_readwrite("residuals_std", &LDScoreResult::residuals_std)
        .def_readwrite("l2_ws", &LDScoreResult::l2_ws)
        .def_readwrite("l2d_ws", &LDScoreResult::l2d_ws)
        .def_readwrite("l2d_wse", &LDScoreResult::l2d_wse);
    m.def("calculate", &calculate, py::arg("params"));
    // ---- additions beside the reference surface (names above are unchanged) ----
    m.def("calculate_on", &calculate_impl, py::arg("params"), py::arg("first_device"), py::arg("n_devices") = 1,
          "calculate() on GPUs first_device .. first_device + n_devices - 1 instead of NLDSC_DEVICE / NLDSC_GPUS; "
          "thread-safe, so one process can score several chromosomes at once");
    m.def("device_count", [] {
        int32_t n = 0;
        char err[512] = {0};
        const int rc = nldsc_ld_device_count(&n, err, sizeof(err));
        if (rc != NLDSC_OK) throw std::runtime_error(err);
        return n;
    }, "usable sm_100 devices");
    m.def("warmup", [](int first_device, int n_devices) {
        char err[512] = {0};
        int rc;
        {
            py::gil_scoped_release nogil;
            rc = nldsc_ld_warmup(first_device, n_devices, err, sizeof(err));
        }
        if (rc != NLDSC_OK) throw std::runtime_error(err);
    }, py::arg("first_device") = 0, py::arg("n_devices") = 1,
       "prepare the devices (context, pinned staging, kernel images) in parallel before the first calculate_on");
    m.def("as_arrays", [](py::object r) {
        LDScoreResult& v = r.cast<LDScoreResult&>();
        py::dict d;
        d["l2"] = view(r, v.l2); d["l2d"] = view(r, v.l2d); d["maf"] = view(r, v.maf);
        d["residuals_std"] = view(r, v.residuals_std);
        d["l2_ws"] = view(r, v.l2_ws); d["l2d_ws"] = view(r, v.l2d_ws); d["l2d_wse"] = view(r, v.l2d_wse);
        return d;
    }, py::arg("result"), "zero-copy NumPy views of the seven result vectors");
    py::class_<BimTable>(m, "BimTable", "the columns of a PLINK .bim, parsed and kept on the C++ side")
        .def("__len__", [](const BimTable& t) { return t.snp.size(); })
        .def_property_readonly("bp_integral", [](const BimTable& t) { return t.bp_integral; })
        .def("cm", [](py::object self) { return table_view(self, self.cast<BimTable&>().cm); }, "cM column, float64 view")
        .def("bp", [](py::object self) { return table_view(self, self.cast<BimTable&>().bp); }, "BP column, int64 view")
        .def("bp_float", [](py::object self) { return table_view(self, self.cast<BimTable&>().bp_f); }, "BP column, float64 view")
        .def("chromosomes", [](const BimTable& t) {
            std::vector<std::string> u;
            for (const auto& c : t.chr)
                if (u.empty() || (c != u.back() && std::find(u.begin(), u.end(), c) == u.end())) u.push_back(c);
            return u;
        }, "distinct CHR values, in order of first appearance")
        .def("first_occurrence", [](const BimTable& t) {
            py::array_t<bool> out(static_cast<py::ssize_t>(t.snp.size()));
            bool* o = out.mutable_data();
            std::unordered_set<std::string_view> seen;
            seen.reserve(t.snp.size() * 2);
            for (size_t i = 0; i < t.snp.size(); ++i) o[i] = seen.insert(std::string_view(t.snp[i])).second;
            return out;
        }, "mask of the rows that survive drop_duplicates('SNP')")
        .def("snp", [](const BimTable& t, size_t i) { return t.snp.at(i); })
        .def("snps", [](const BimTable& t) { return t.snp; }, "SNP ids as a Python list (a copy)");
    m.def("read_bim", &read_bim, py::arg("path"), "parse a PLINK .bim (six whitespace-separated columns) with the GIL released");
    m.def("write_l2_bim", &write_l2_bim, py::arg("path"), py::arg("table"), py::arg("l2"), py::arg("l2d"), py::arg("maf"),
          py::arg("residuals_std"), py::arg("l2_ws"), py::arg("l2d_ws"), py::arg("l2d_wse"), py::arg("extra") = false,
          "write_l2 with CHR / SNP / BP taken from a BimTable");
    m.def("write_l2", &write_l2, py::arg("path"), py::arg("chr"), py::arg("snp"), py::arg("bp"), py::arg("l2"), py::arg("l2d"),
          py::arg("maf"), py::arg("residuals_std"), py::arg("l2_ws"), py::arg("l2d_ws"), py::arg("l2d_wse"), py::arg("extra") = false,
          "the .L2 table in the byte layout of the reference's to_csv(sep='\\t', index=False, float_format='%.5f')");
}
