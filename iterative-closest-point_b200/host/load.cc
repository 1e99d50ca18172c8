// load.cc — loader / writer with the reference's observable behaviour (src/load.cc), written as a
// single-pass buffer parser instead of the reference's two getline passes + sscanf:
//   * first line is a header and is dropped; n = (#lines) - 1                     (src/load.cc:15-21)
//   * each line yields the first three comma-separated numbers, extra columns are ignored and
//     missing ones stay 0 (what sscanf("%lf,%lf,%lf") does)                         (src/load.cc:26-28)
//   * unopenable file: message on stderr and exit(2)                                (src/load.cc:11-14)
//   * progress lines on stderr, output header "Points_0,Points_1,Points_2", default ostream
//     formatting (6 significant digits)                                             (src/load.cc:9,24,71-79)
#include "load.hh"

#include <cctype>
#include <charconv>
#include <cstdio>
#include <string>

namespace {

// Emulates sscanf(line, "%lf,%lf,%lf", &x, &y, &z) on [s, e): the general (slow) form.
void parse3_general(const char* s, const char* e, double out[3])
{
    out[0] = out[1] = out[2] = 0.0;
    std::string tmp(s, e);               // NUL-terminated copy so strtod cannot run past the line
    const char* c = tmp.c_str();
    for (int k = 0; k < 3; ++k) {
        while (*c && std::isspace((unsigned char)*c)) ++c;
        char* end = nullptr;
        const double v = std::strtod(c, &end);
        if (end == c) return;
        out[k] = v;
        c = end;
        if (k < 2) {
            if (*c != ',') return;
            ++c;
        }
    }
}

// Fast path: three plain decimal fields parsed in place with std::from_chars (bounded by the line, no
// copy, no locale).  Anything from_chars does not take exactly like sscanf's %lf (leading blanks or '+',
// inf/nan, hex floats, a short line) goes through parse3_general, so the observable result is unchanged.
void parse3(const char* s, const char* e, double out[3])
{
    const char* c = s;
    for (int k = 0; k < 3; ++k) {
        if (c >= e || !(std::isdigit((unsigned char)*c) || *c == '-' || *c == '.')) return parse3_general(s, e, out);
        const auto r = std::from_chars(c, e, out[k]);
        if (r.ec != std::errc()) return parse3_general(s, e, out);
        c = r.ptr;
        if (k < 2) {
            if (c >= e || *c != ',') return parse3_general(s, e, out);
            ++c;
        }
    }
}

}  // namespace

GPU::Matrix load_matrix(const char* filename)
{
    std::cerr << "[load] opening " << filename << std::endl;
    std::FILE* f = std::fopen(filename, "rb");
    if (!f) {
        std::cerr << "[load] " << filename << " could not be opened" << std::endl;
        std::exit(2);
    }
    std::string buf;
    char chunk[1 << 16];
    size_t got;
    while ((got = std::fread(chunk, 1, sizeof(chunk), f)) > 0) buf.append(chunk, got);
    std::fclose(f);

    // line starts; a final line without '\n' still counts (getline semantics)
    std::vector<size_t> starts;
    size_t pos = 0;
    while (pos < buf.size()) {
        starts.push_back(pos);
        const size_t nl = buf.find('\n', pos);
        if (nl == std::string::npos) break;
        pos = nl + 1;
    }
    const long n = (long)starts.size() - 1;
    std::cerr << "[load] loading file into matrix" << std::endl;
    if (n < 0) return GPU::Matrix(3, 0);
    GPU::Matrix m(3, n);
    for (long i = 0; i < n; ++i) {
        const size_t b = starts[(size_t)i + 1];
        size_t e = (size_t)i + 2 < starts.size() ? starts[(size_t)i + 2] - 1 : buf.size();
        if (e > b && buf[e - 1] == '\n') --e;
        double v[3];
        parse3(buf.data() + b, buf.data() + e, v);
        m(0, i) = v[0]; m(1, i) = v[1]; m(2, i) = v[2];
    }
    return m;
}

void write_matrix_to(const GPU::Matrix& matrix, const char* filename)
{
    // default ostream formatting of a double == printf("%g") (6 significant digits): src/load.cc:71-79
    std::FILE* f = std::fopen(filename, "wb");
    if (f) {
        std::string buf;
        buf.reserve((size_t)matrix.cols() * 40 + 64);
        buf += "Points_0,Points_1,Points_2\n";
        char line[128];
        for (long i = 0; i < matrix.cols(); i++) {
            const int n = std::snprintf(line, sizeof(line), "%g,%g,%g\n", matrix(0, i), matrix(1, i), matrix(2, i));
            buf.append(line, (size_t)n);
        }
        std::fwrite(buf.data(), 1, buf.size(), f);
        std::fclose(f);
    }
    std::cerr << "[output] output file \"" << filename << "\" was generated." << std::endl;
}

void write_matrix(const GPU::Matrix& matrix) { write_matrix_to(matrix, "output.txt"); }
