// Host-side reader / writer of the reference's on-disk format: whitespace-separated text, one point per line
// (np.loadtxt(path, dtype=np.float32) at /root/reference/modules/discrete/upsample.py:42 and
// np.savetxt(path, pred, fmt='%.6f') at :57).  With the GPU part of a 2,048-pt cloud at ~1.5 ms, numpy's text
// parser / formatter (5 + 25 ms per file) is the wall-clock bound of the file loop.
// Bit-compatible by construction: values are parsed as double and narrowed to float (what numpy does for float32), and
// written with printf("%.6f") of the widened value (what '%.6f' % numpy.float32 does).  No CUDA here.
#include <errno.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "common.cuh"
#include "puflow_b200.h"

namespace {

bool slurp(const char* path, std::string& buf) {
    FILE* f = fopen(path, "rb");
    if (!f) return false;
    char tmp[1 << 16];
    size_t n;
    while ((n = fread(tmp, 1, sizeof(tmp), f)) > 0) buf.append(tmp, n);
    const bool ok = !ferror(f);
    fclose(f);
    return ok;
}

// ---- exact fast paths (fall back to strtod / snprintf whenever a precondition fails) ----
// Decimal text "[-+]ddd[.ddd]" with at most 19 digits whose digit string is < 2^53 and with at most 22 digits after the
// point: double(mantissa) / 10^k is the correctly rounded double (both operands exact, one IEEE division) -- Clinger's
// fast path -- hence equal to strtod's result.  Returns the end of the number or nullptr.
const double kPow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                           1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
inline const char* fast_decimal(const char* q, const char* stop, double* v) {
    const char* s = q;
    bool neg = false;
    if (s < stop && (*s == '-' || *s == '+')) neg = *s == '-', s++;
    uint64_t mant = 0;
    int digits = 0, frac = 0;
    while (s < stop && *s >= '0' && *s <= '9') mant = mant * 10 + (uint64_t)(*s - '0'), digits++, s++;
    if (s < stop && *s == '.') {
        s++;
        while (s < stop && *s >= '0' && *s <= '9') mant = mant * 10 + (uint64_t)(*s - '0'), digits++, frac++, s++;
    }
    if (digits == 0 || digits > 19 || frac > 22 || mant >= (1ull << 53)) return nullptr;
    if (s < stop && !(*s == ' ' || *s == '\t' || *s == '\r' || *s == '\v' || *s == '\f')) return nullptr;   // exponent, nan, ...
    const double d = (double)mant / kPow10[frac];
    *v = neg ? -d : d;
    return s;
}
// '%.6f' of a float: v * 10^6 = m * 15625 * 2^(e+6) with a 24-bit m, rounded half-to-even on the exact remainder.
inline int fast_fixed6(float f, char* dst) {
    uint32_t bits;
    memcpy(&bits, &f, 4);
    const uint32_t ex = (bits >> 23) & 0xff;
    if (ex == 0xff) return 0;                                            // inf / nan
    const uint64_t m = ex ? ((bits & 0x7fffffu) | 0x800000u) : (bits & 0x7fffffu);
    const int e = (ex ? (int)ex : 1) - 127 - 23 + 6;                     // value * 2^6 = m * 2^e ... times 15625
    uint64_t n = m * 15625ull;
    if (e >= 0) {
        if (e > 24) return 0;                                            // > 2^62: let snprintf do it
        n <<= e;
    } else {
        const int sh = -e;
        if (sh >= 64) {
            n = 0;
        } else {
            const uint64_t q = n >> sh, rem = n & ((1ull << sh) - 1), half = 1ull << (sh - 1);
            n = q + ((rem > half || (rem == half && (q & 1))) ? 1 : 0);
        }
    }
    char tmp[32];
    int len = 0;
    uint64_t ip = n / 1000000ull;
    uint32_t fp = (uint32_t)(n % 1000000ull);
    do tmp[len++] = (char)('0' + ip % 10), ip /= 10; while (ip);
    int o = 0;
    if (bits >> 31) dst[o++] = '-';
    while (len) dst[o++] = tmp[--len];
    dst[o++] = '.';
    for (int i = 5; i >= 0; i--) dst[o + i] = (char)('0' + fp % 10), fp /= 10;
    return o + 6;
}

// Counts data lines and columns.  Lines: '#' starts a comment, blank lines are skipped, every data line must have the
// same number of columns (numpy raises otherwise).
int count(const char* path, const std::string& buf, int64_t* rows_out, int* cols_out) {
    const char* p = buf.c_str();
    const char* end = p + buf.size();
    int64_t rows = 0;
    int cols = -1;
    while (p < end) {
        const char* eol = static_cast<const char*>(memchr(p, '\n', (size_t)(end - p)));
        if (!eol) eol = end;
        const char* stop = static_cast<const char*>(memchr(p, '#', (size_t)(eol - p)));
        if (!stop) stop = eol;
        int c = 0;
        const char* q = p;
        while (q < stop) {
            while (q < stop && (*q == ' ' || *q == '\t' || *q == '\r' || *q == '\v' || *q == '\f')) q++;
            if (q >= stop) break;
            while (q < stop && !(*q == ' ' || *q == '\t' || *q == '\r' || *q == '\v' || *q == '\f')) q++;
            c++;
        }
        if (c > 0) {
            if (cols < 0) cols = c;
            PF_CHECK_ARG(c == cols, "pf_xyz_count: %s: data line %lld has %d columns, expected %d", path,
                         (long long)(rows + 1), c, cols);
            rows++;
        }
        p = eol + 1;
    }
    *rows_out = rows;
    *cols_out = cols < 0 ? 0 : cols;
    return 0;
}

}  // namespace

extern "C" int pf_xyz_count(const char* path, int64_t* rows, int* cols) {
    PF_CHECK_ARG(path && rows && cols, "pf_xyz_count: null pointer");
    std::string buf;
    PF_CHECK_ARG(slurp(path, buf), "pf_xyz_count: cannot read %s: %s", path, strerror(errno));
    return count(path, buf, rows, cols);
}

extern "C" int pf_xyz_read(const char* path, float* out, int64_t rows, int cols) {
    PF_CHECK_ARG(path && out && rows >= 0 && cols > 0, "pf_xyz_read: bad arguments");
    std::string buf;
    PF_CHECK_ARG(slurp(path, buf), "pf_xyz_read: cannot read %s: %s", path, strerror(errno));
    // second pass with the stride known: parse() writes out[row * cols + c] once `cols` is preset
    const char* p = buf.c_str();
    const char* end = p + buf.size();
    int64_t r = 0;
    while (p < end) {
        const char* eol = static_cast<const char*>(memchr(p, '\n', (size_t)(end - p)));
        if (!eol) eol = end;
        const char* stop = static_cast<const char*>(memchr(p, '#', (size_t)(eol - p)));
        if (!stop) stop = eol;
        int c = 0;
        const char* q = p;
        while (q < stop) {
            while (q < stop && (*q == ' ' || *q == '\t' || *q == '\r' || *q == '\v' || *q == '\f')) q++;
            if (q >= stop) break;
            double v;
            const char* e = fast_decimal(q, stop, &v);
            if (!e) {
                char* e2 = nullptr;
                v = strtod(q, &e2);            // buf is NUL-terminated; a number never spans a newline
                e = e2;
            }
            PF_CHECK_ARG(e != q && e <= stop, "pf_xyz_read: %s: cannot parse a number on data line %lld", path,
                         (long long)(r + 1));
            PF_CHECK_ARG(r < rows && c < cols, "pf_xyz_read: %s: more data than the %lld x %d buffer", path,
                         (long long)rows, cols);
            out[r * cols + c] = (float)v;
            c++;
            q = e;
        }
        if (c > 0) {
            PF_CHECK_ARG(c == cols, "pf_xyz_read: %s: data line %lld has %d columns, expected %d", path, (long long)(r + 1),
                         c, cols);
            r++;
        }
        p = eol + 1;
    }
    PF_CHECK_ARG(r == rows, "pf_xyz_read: %s: %lld data lines, expected %lld", path, (long long)r, (long long)rows);
    return 0;
}

extern "C" int pf_xyz_write(const char* path, const float* data, int64_t rows, int cols) {
    PF_CHECK_ARG(path && (data || rows == 0) && rows >= 0 && cols > 0, "pf_xyz_write: bad arguments");
    FILE* f = fopen(path, "wb");
    PF_CHECK_ARG(f, "pf_xyz_write: cannot open %s: %s", path, strerror(errno));
    std::vector<char> buf;
    buf.resize(1 << 20);
    size_t used = 0;
    bool ok = true;
    for (int64_t r = 0; r < rows && ok; r++) {
        if (used + (size_t)cols * 64 > buf.size()) {
            ok = fwrite(buf.data(), 1, used, f) == used;
            used = 0;
        }
        for (int c = 0; c < cols; c++) {
            int w = fast_fixed6(data[r * cols + c], buf.data() + used);
            if (!w) w = snprintf(buf.data() + used, 60, "%.6f", (double)data[r * cols + c]);
            used += (size_t)w;
            buf[used++] = c + 1 < cols ? ' ' : '\n';
        }
    }
    if (ok && used) ok = fwrite(buf.data(), 1, used, f) == used;
    ok = (fclose(f) == 0) && ok;
    PF_CHECK_ARG(ok, "pf_xyz_write: write to %s failed: %s", path, strerror(errno));
    return 0;
}
