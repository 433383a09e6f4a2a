// csv_writer.hpp -- byte-identical replacement of the reference's CSV output (src/main.cpp:156-174), C++17.
//
// The reference prints every cell with fmt 10.2.1 `{}` on a double: the SHORTEST decimal string that round-trips
// (Dragonbox), laid out in fixed notation when the decimal exponent e10 of the first digit satisfies
// -4 <= e10 < 16 and in exponent notation (d.ddde-05, at least two exponent digits) otherwise; integral values
// print without a decimal point ("0", "1").  std::to_chars(double) yields the same shortest digit string
// (the shortest round-trip digits are unique), so only the layout rule is restated here.
// Rows are formatted in parallel (OpenMP) and written in order.
#pragma once
#include <omp.h>

#include <algorithm>
#include <charconv>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <ctime>
#include <string>
#include <vector>

namespace pfaai_host {

// appends fmt::format("{}", v) to out
inline void append_double(std::string &out, double v) {
    if (std::isnan(v)) { out += std::signbit(v) ? "-nan" : "nan"; return; }
    if (std::isinf(v)) { out += v < 0 ? "-inf" : "inf"; return; }
    if (std::signbit(v)) { out += '-'; v = -v; }
    if (v == 0.0) { out += '0'; return; }
    char buf[40];
    auto res = std::to_chars(buf, buf + sizeof(buf), v, std::chars_format::scientific);   // d[.ddd]e[+-]XX, shortest
    const char *e = static_cast<const char *>(std::memchr(buf, 'e', (size_t)(res.ptr - buf)));
    int e10 = 0;
    std::from_chars(e + 1 + (e[1] == '+' ? 1 : 0), res.ptr, e10);
    char digits[24];
    int nd = 0;
    for (const char *p = buf; p < e; p++)
        if (*p != '.') digits[nd++] = *p;
    if (e10 < -4 || e10 >= 16) {                       // exponent notation
        out += digits[0];
        if (nd > 1) { out += '.'; out.append(digits + 1, (size_t)(nd - 1)); }
        out += 'e';
        out += e10 < 0 ? '-' : '+';
        const int a = e10 < 0 ? -e10 : e10;
        if (a < 10) out += '0';
        out += std::to_string(a);
    } else if (e10 >= nd - 1) {                        // integral: digits followed by zeros, no decimal point
        out.append(digits, (size_t)nd);
        out.append((size_t)(e10 - (nd - 1)), '0');
    } else if (e10 >= 0) {                             // ddd.ddd
        out.append(digits, (size_t)(e10 + 1));
        out += '.';
        out.append(digits + e10 + 1, (size_t)(nd - e10 - 1));
    } else {                                           // 0.000ddd
        out += "0.";
        out.append((size_t)(-e10 - 1), '0');
        out.append(digits, (size_t)nd);
    }
}

inline std::string format_double(double v) {
    std::string s;
    append_double(s, v);
    return s;
}

namespace detail {
inline bool write_header(FILE *fp, const std::string &sep, const std::vector<std::string> &tgtNames) {
    std::string hdr = sep;
    for (size_t j = 0; j < tgtNames.size(); j++) { if (j) hdr += sep; hdr += tgtNames[j]; }
    hdr += '\n';
    return std::fwrite(hdr.data(), 1, hdr.size(), fp) == hdr.size();
}
// rows [r0, r1): formatted in parallel (OpenMP) in batches, written in order
inline bool write_rows(FILE *fp, const std::string &sep, const std::vector<std::string> &qryNames, size_t nT, const double *matrix,
                       size_t r0, size_t r1, std::vector<std::string> &lines) {
    const size_t batch = lines.size();
    bool ok = true;
    for (size_t b0 = r0; b0 < r1 && ok; b0 += batch) {
        const size_t b1 = std::min(r1, b0 + batch);
#pragma omp parallel for schedule(dynamic, 1)
        for (size_t r = b0; r < b1; r++) {
            std::string &s = lines[r - b0];
            s.clear();
            s.reserve(qryNames[r].size() + nT * 20 + 8);
            s += qryNames[r];
            s += sep;
            const double *row = matrix + r * nT;
            for (size_t j = 0; j < nT; j++) { if (j) s += sep; append_double(s, row[j]); }
            s += '\n';
        }
        for (size_t r = b0; r < b1 && ok; r++) ok = std::fwrite(lines[r - b0].data(), 1, lines[r - b0].size(), fp) == lines[r - b0].size();
    }
    return ok;
}
}  // namespace detail

// header: <sep> + target names joined by <sep>; rows: query name, <sep>, values joined by <sep>; '\n' endings.
// Returns false when the file cannot be opened / written.
inline bool write_aji_csv(const std::string &path, const std::string &sep, const std::vector<std::string> &qryNames,
                          const std::vector<std::string> &tgtNames, const double *matrix /*[nQ x nT] row-major*/) {
    FILE *fp = std::fopen(path.c_str(), "w");
    if (!fp) return false;
    std::vector<std::string> lines((size_t)std::max(1, omp_get_max_threads()) * 8);
    bool ok = detail::write_header(fp, sep, tgtNames);
    ok = ok && detail::write_rows(fp, sep, qryNames, tgtNames.size(), matrix, 0, qryNames.size(), lines);
    return (std::fclose(fp) == 0) && ok;
}

// The same file, written WHILE the matrix is being computed: poll(rows_final, finished) reports how many
// leading rows of `matrix` are complete (pfaai_gpu_progress_poll); every batch of finished rows is formatted and written at
// once, so formatting and file I/O overlap the GPU and the device-to-host stores of the later row blocks.
// *overlapped_rows = rows that were written before the compute had finished.
template <class Poll>
inline bool write_aji_csv_streaming(const std::string &path, const std::string &sep, const std::vector<std::string> &qryNames,
                                    const std::vector<std::string> &tgtNames, const double *matrix, Poll poll, size_t *overlapped_rows) {
    FILE *fp = std::fopen(path.c_str(), "w");
    if (!fp) return false;
    const size_t nQ = qryNames.size();
    std::vector<std::string> lines((size_t)std::max(1, omp_get_max_threads()) * 8);
    bool ok = detail::write_header(fp, sep, tgtNames);
    size_t written = 0, overlapped = 0;
    while (ok && written < nQ) {
        int64_t rows_final = 0; int finished = 0;
        if (!poll(rows_final, finished)) { ok = false; break; }
        const size_t upto = std::min(nQ, (size_t)std::max<int64_t>(rows_final, 0));
        if (upto > written) {
            ok = detail::write_rows(fp, sep, qryNames, tgtNames.size(), matrix, written, upto, lines);
            if (!finished) overlapped = upto;
            written = upto;
        } else if (finished) {
            break;                                                  // finished but rows missing: the caller reports the failure
        } else {
            struct timespec ts = {0, 500000};                       // 0.5 ms
            nanosleep(&ts, nullptr);
        }
    }
    if (overlapped_rows) *overlapped_rows = overlapped;
    ok = ok && written == nQ;
    return (std::fclose(fp) == 0) && ok;
}

}  // namespace pfaai_host
