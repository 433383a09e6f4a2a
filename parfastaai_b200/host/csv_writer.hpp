// csv_writer.hpp -- byte-identical replacement of the reference's CSV output (src/main.cpp:156-174), C++17.
//
// The reference prints every cell with fmt 10.2.1 `{}` on a double: the SHORTEST decimal string that round-trips
// (Dragonbox), laid out in fixed notation when the decimal exponent e10 of the first digit satisfies
// -4 <= e10 < 16 and in exponent notation (d.ddde-05, at least two exponent digits) otherwise; integral values
// print without a decimal point ("0", "1").  std::to_chars(double) yields the same shortest digit string
// (the shortest round-trip digits are unique), so only the layout rule is restated here.
// Rows are formatted AND written in parallel (OpenMP, one buffer and one file offset per thread).
#pragma once
#include <fcntl.h>
#include <omp.h>
#include <unistd.h>

#include <algorithm>
#include <cerrno>
#include <charconv>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <ctime>
#include <string>
#include <vector>

namespace pfaai_host {

// writes fmt::format("{}", v) at p (at most 32 characters) and returns the end
inline char *put_double(char *p, double v) {
    if (std::isnan(v)) { if (std::signbit(v)) *p++ = '-'; std::memcpy(p, "nan", 3); return p + 3; }
    if (std::isinf(v)) { if (v < 0) *p++ = '-'; std::memcpy(p, "inf", 3); return p + 3; }
    if (std::signbit(v)) { *p++ = '-'; v = -v; }
    if (v == 0.0) { *p++ = '0'; return p; }
    char buf[40];
    auto res = std::to_chars(buf, buf + sizeof(buf), v, std::chars_format::scientific);   // d[.ddd]e[+-]XX, shortest
    // exponent: the last characters are e[+-]XX[X]
    const char *e = res.ptr - 1;
    while (*e != 'e') e--;
    int e10 = 0;
    for (const char *q = e + 2; q < res.ptr; q++) e10 = e10 * 10 + (*q - '0');
    if (e[1] == '-') e10 = -e10;
    // digits: buf[0], then buf[2 .. e) when there is a fraction
    const int nd = (e - buf > 1) ? (int)(e - buf) - 1 : 1;
    char digits[24];
    digits[0] = buf[0];
    if (nd > 1) std::memcpy(digits + 1, buf + 2, (size_t)(nd - 1));
    if (e10 < -4 || e10 >= 16) {                       // exponent notation
        *p++ = digits[0];
        if (nd > 1) { *p++ = '.'; std::memcpy(p, digits + 1, (size_t)(nd - 1)); p += nd - 1; }
        *p++ = 'e';
        *p++ = e10 < 0 ? '-' : '+';
        int a = e10 < 0 ? -e10 : e10;
        if (a >= 100) { *p++ = (char)('0' + a / 100); a %= 100; *p++ = (char)('0' + a / 10); *p++ = (char)('0' + a % 10); }
        else { *p++ = (char)('0' + a / 10); *p++ = (char)('0' + a % 10); }
    } else if (e10 >= nd - 1) {                        // integral: digits followed by zeros, no decimal point
        std::memcpy(p, digits, (size_t)nd); p += nd;
        for (int k = 0; k < e10 - (nd - 1); k++) *p++ = '0';
    } else if (e10 >= 0) {                             // ddd.ddd
        std::memcpy(p, digits, (size_t)(e10 + 1)); p += e10 + 1;
        *p++ = '.';
        std::memcpy(p, digits + e10 + 1, (size_t)(nd - e10 - 1)); p += nd - e10 - 1;
    } else {                                           // 0.000ddd
        *p++ = '0'; *p++ = '.';
        for (int k = 0; k < -e10 - 1; k++) *p++ = '0';
        std::memcpy(p, digits, (size_t)nd); p += nd;
    }
    return p;
}

// appends fmt::format("{}", v) to out
inline void append_double(std::string &out, double v) {
    char buf[40];
    out.append(buf, (size_t)(put_double(buf, v) - buf));
}

inline std::string format_double(double v) {
    std::string s;
    append_double(s, v);
    return s;
}

// The output file.  header: <sep> + target names joined by <sep>; rows: query name, <sep>, values joined by <sep>; '\n'
// endings -- the bytes of the reference's printOutput (main.cpp:156-174).
// Rows are formatted by the whole OpenMP team, in chunks handed out dynamically, every chunk into its own buffer (raw character
// writes, no string appends).  Buffered writes to ONE file serialise in the kernel whatever the number of writers, so the writing
// is a pipeline stage instead: when batch k is formatted, one thread writes its chunks in order while the others already format
// batch k+1 into the second buffer set (the writer joins them when it is done and simply takes fewer chunks).
class CsvFile {
    int m_fd = -1;
    off_t m_off = 0;
    bool m_ok = false;
    std::vector<std::vector<char>> m_buf[2];   // two buffer sets x chunks per batch
    std::vector<size_t> m_len[2];

  public:
    explicit CsvFile(const std::string &path) {
        m_fd = ::open(path.c_str(), O_WRONLY | O_CREAT | O_TRUNC, 0644);
        m_ok = m_fd >= 0;
    }
    CsvFile(const CsvFile &) = delete;
    CsvFile &operator=(const CsvFile &) = delete;
    ~CsvFile() { if (m_fd >= 0) ::close(m_fd); }
    bool ok() const { return m_ok; }
    bool close() {
        if (m_fd >= 0) { if (::close(m_fd) != 0) m_ok = false; m_fd = -1; }
        return m_ok;
    }
    // all writes of this class are issued in file order by one thread at a time, so a plain write() is equivalent where the
    // target cannot seek (a pipe, /dev/stdout)
    static bool write_all(int fd, const char *p, size_t n, off_t off) {
        while (n > 0) {
            ssize_t w = ::pwrite(fd, p, n, off);
            if (w < 0 && errno == ESPIPE) w = ::write(fd, p, n);
            if (w <= 0) { if (w < 0 && errno == EINTR) continue; return false; }
            p += w; n -= (size_t)w; off += w;
        }
        return true;
    }
    bool header(const std::string &sep, const std::vector<std::string> &tgtNames) {
        std::string hdr = sep;
        for (size_t j = 0; j < tgtNames.size(); j++) { if (j) hdr += sep; hdr += tgtNames[j]; }
        hdr += '\n';
        m_ok = m_ok && write_all(m_fd, hdr.data(), hdr.size(), m_off);
        m_off += (off_t)hdr.size();
        return m_ok;
    }
    // rows [r0, r1) of the row-major matrix [.. x nT]
    bool rows(const std::string &sep, const std::vector<std::string> &qryNames, size_t nT, const double *matrix, size_t r0, size_t r1) {
        if (!m_ok || r1 <= r0) return m_ok;
        const size_t T = (size_t)std::max(1, omp_get_max_threads());
        size_t name_max = 0;
        for (size_t r = r0; r < r1; r++) name_max = std::max(name_max, qryNames[r].size());
        const size_t row_cap = name_max + sep.size() + nT * (32 + sep.size()) + 2;
        const size_t C = 2 * T;                                                    // chunks per batch
        // rows per chunk: the range split evenly, bounded by ~4 MB of buffer per chunk, at least 1
        const size_t per = std::max<size_t>(1, std::min<size_t>((r1 - r0 + C - 1) / C, ((size_t)4 << 20) / row_cap + 1));
        const size_t batch_rows = per * C, n_batches = (r1 - r0 + batch_rows - 1) / batch_rows;
        for (int k = 0; k < 2; k++) { m_buf[k].resize(C); m_len[k].assign(C, 0); }
        bool ok = true;
        off_t off = m_off;
        std::vector<int> next(n_batches, 0);                                       // next chunk of every batch (dynamic hand-out)
#pragma omp parallel num_threads((int)T)
        {
            for (size_t b = 0; b < n_batches; b++) {
                const size_t b0 = r0 + b * batch_rows;
                std::vector<std::vector<char>> &bufs = m_buf[b & 1];
                std::vector<size_t> &lens = m_len[b & 1];
                for (;;) {
                    int c;
#pragma omp atomic capture
                    c = next[b]++;
                    if ((size_t)c >= C) break;
                    const size_t my0 = std::min(r1, b0 + (size_t)c * per), my1 = std::min(r1, my0 + per);
                    std::vector<char> &buf = bufs[(size_t)c];
                    if (buf.size() < per * row_cap) buf.resize(per * row_cap);
                    char *p = buf.data();
                    for (size_t r = my0; r < my1; r++) {
                        std::memcpy(p, qryNames[r].data(), qryNames[r].size()); p += qryNames[r].size();
                        std::memcpy(p, sep.data(), sep.size()); p += sep.size();
                        const double *row = matrix + r * nT;
                        for (size_t j = 0; j < nT; j++) {
                            if (j) { std::memcpy(p, sep.data(), sep.size()); p += sep.size(); }
                            p = put_double(p, row[j]);
                        }
                        *p++ = '\n';
                    }
                    lens[(size_t)c] = (size_t)(p - buf.data());
                }
                // batch b is complete behind this barrier; the thread that wrote batch b-1 has passed it too, so that write is over
                // and its buffer set is free for batch b+1
#pragma omp barrier
#pragma omp single nowait
                {
                    for (size_t c = 0; c < C; c++) {
                        if (lens[c] && !write_all(m_fd, bufs[c].data(), lens[c], off)) ok = false;
                        off += (off_t)lens[c];
                    }
                }
            }
        }
        m_off = off;
        m_ok = m_ok && ok;
        return m_ok;
    }
};

// Returns false when the file cannot be opened / written.
inline bool write_aji_csv(const std::string &path, const std::string &sep, const std::vector<std::string> &qryNames,
                          const std::vector<std::string> &tgtNames, const double *matrix /*[nQ x nT] row-major*/) {
    CsvFile f(path);
    if (!f.ok()) return false;
    f.header(sep, tgtNames);
    f.rows(sep, qryNames, tgtNames.size(), matrix, 0, qryNames.size());
    return f.close();
}

// The same file, written WHILE the matrix is being computed: poll(rows_final, finished) reports how many
// leading rows of `matrix` are complete (pfaai_gpu_progress_poll); every batch of finished rows is formatted and written at
// once, so formatting and file I/O overlap the GPU and the device-to-host stores of the later row blocks.
// *overlapped_rows = rows that were written before the compute had finished.
template <class Poll>
inline bool write_aji_csv_streaming(const std::string &path, const std::string &sep, const std::vector<std::string> &qryNames,
                                    const std::vector<std::string> &tgtNames, const double *matrix, Poll poll, size_t *overlapped_rows) {
    CsvFile f(path);
    if (!f.ok()) return false;
    const size_t nQ = qryNames.size();
    bool ok = f.header(sep, tgtNames);
    size_t written = 0, overlapped = 0;
    while (ok && written < nQ) {
        int64_t rows_final = 0; int finished = 0;
        if (!poll(rows_final, finished)) { ok = false; break; }
        const size_t upto = std::min(nQ, (size_t)std::max<int64_t>(rows_final, 0));
        if (upto > written) {
            ok = f.rows(sep, qryNames, tgtNames.size(), matrix, written, upto);
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
    return f.close() && ok;
}

}  // namespace pfaai_host
