// Reader for the calibration file the reference loads with cv::FileStorage
// (ReadObjectYml, FastACVNet_plus/src/FastACVNet_plus_Algorithm.cpp:98-114) and its calibration tool writes
// (Stereo_Calibration/Stereo_Calibration.cpp:165-179): "%YAML:1.0", optional "---", then top-level
//     name: !!opencv-matrix
//        rows: R
//        cols: C
//        dt: d
//        data: [ v, v, ...        <- flow sequence, may span lines, with or without a space after '['
// Other top-level entries (validROIL: [ 0, 0, 640, 480 ], scalars) are skipped.  Keys may come in any order and
// the file may lack a trailing newline.  Only dt d / f (and their "1d"-style spellings) are accepted.
#include <cctype>
#include <cerrno>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>

#include "common.h"

namespace sa {

namespace {

struct Cursor {
    const std::string& s;
    size_t i = 0;
    explicit Cursor(const std::string& str) : s(str) {}
    bool eof() const { return i >= s.size(); }
    void skip_ws() { while (!eof() && (s[i] == ' ' || s[i] == '\t' || s[i] == '\r' || s[i] == '\n')) ++i; }
    void skip_line() { while (!eof() && s[i] != '\n') ++i; if (!eof()) ++i; }
};

bool is_key_char(char c) { return std::isalnum((unsigned char)c) || c == '_'; }

// Reads "key:" at the cursor (leading blanks allowed).  Returns false if the line is not of that form.
bool read_key(Cursor& c, std::string& key)
{
    size_t j = c.i;
    while (j < c.s.size() && (c.s[j] == ' ' || c.s[j] == '\t')) ++j;
    size_t k = j;
    while (k < c.s.size() && is_key_char(c.s[k])) ++k;
    if (k == j || k >= c.s.size() || c.s[k] != ':') return false;
    key.assign(c.s, j, k - j);
    c.i = k + 1;
    return true;
}

std::string rest_of_line(Cursor& c)
{
    size_t j = c.i;
    while (j < c.s.size() && c.s[j] != '\n') ++j;
    std::string r(c.s, c.i, j - c.i);
    c.i = j < c.s.size() ? j + 1 : j;
    size_t a = r.find_first_not_of(" \t\r"), b = r.find_last_not_of(" \t\r");
    return a == std::string::npos ? std::string() : r.substr(a, b - a + 1);
}

}  // namespace

int parse_opencv_yaml(const char* path, std::vector<YamlMat>& out, std::string& err)
{
    std::ifstream f(path, std::ios::binary);
    if (!f.good()) { err = std::string("cannot open calibration file '") + (path ? path : "(null)") + "'"; return SA_ERR_IO; }
    std::stringstream ss;
    ss << f.rdbuf();
    const std::string text = ss.str();
    Cursor c(text);
    out.clear();
    while (!c.eof()) {
        // top-level lines start in column 0
        const char ch = text[c.i];
        if (ch == '%' || ch == '-' || ch == '#' || ch == '\n' || ch == '\r' || ch == ' ' || ch == '\t') { c.skip_line(); continue; }
        std::string key;
        if (!read_key(c, key)) { c.skip_line(); continue; }
        const std::string tag = rest_of_line(c);
        if (tag.find("!!opencv-matrix") == std::string::npos) continue;   // plain scalar / flow list: ignored
        YamlMat m;
        m.name = key;
        std::string dt;
        bool have_data = false;
        // indented members
        while (!c.eof() && (text[c.i] == ' ' || text[c.i] == '\t')) {
            std::string sub;
            if (!read_key(c, sub)) { c.skip_line(); continue; }
            if (sub == "rows" || sub == "cols") {
                const std::string v = rest_of_line(c);
                char* end = nullptr;
                long n = std::strtol(v.c_str(), &end, 10);
                if (end == v.c_str() || n < 0 || n > (1 << 20)) { err = "bad " + sub + " in node '" + key + "'"; return SA_ERR_PARSE; }
                (sub == "rows" ? m.rows : m.cols) = (int)n;
            } else if (sub == "dt") {
                dt = rest_of_line(c);
                if (!dt.empty() && dt.front() == '"') dt = dt.substr(1, dt.size() >= 2 ? dt.size() - 2 : 0);
            } else if (sub == "data") {
                c.skip_ws();
                if (c.eof() || text[c.i] != '[') { err = "node '" + key + "': data is not a flow sequence"; return SA_ERR_PARSE; }
                ++c.i;
                for (;;) {
                    while (!c.eof() && (std::isspace((unsigned char)text[c.i]) || text[c.i] == ',')) ++c.i;
                    if (c.eof()) { err = "node '" + key + "': unterminated data sequence"; return SA_ERR_PARSE; }
                    if (text[c.i] == ']') { ++c.i; break; }
                    // numbers look like 0.  1.  5.17e+02  -6.69e-02  338.82  .5  .Inf  -.Inf  .Nan
                    const char* b = text.c_str() + c.i;
                    char* end = nullptr;
                    double v;
                    if (!std::strncmp(b, ".Inf", 4) || !std::strncmp(b, ".inf", 4)) { v = HUGE_VAL; end = const_cast<char*>(b) + 4; }
                    else if (!std::strncmp(b, "-.Inf", 5) || !std::strncmp(b, "-.inf", 5)) { v = -HUGE_VAL; end = const_cast<char*>(b) + 5; }
                    else if (!std::strncmp(b, ".Nan", 4) || !std::strncmp(b, ".NaN", 4) || !std::strncmp(b, ".nan", 4)) { v = std::strtod("nan", nullptr); end = const_cast<char*>(b) + 4; }
                    else v = std::strtod(b, &end);
                    if (end == b) { err = "node '" + key + "': bad number near offset " + std::to_string(c.i); return SA_ERR_PARSE; }
                    m.data.push_back(v);
                    c.i += (size_t)(end - b);
                }
                have_data = true;
                c.skip_line();
            } else {
                c.skip_line();
            }
        }
        if (dt.empty() || !(dt.back() == 'd' || dt.back() == 'f')) { err = "node '" + key + "': unsupported dt '" + dt + "'"; return SA_ERR_PARSE; }
        if (!have_data || (size_t)m.rows * m.cols != m.data.size()) {
            err = "node '" + key + "': " + std::to_string(m.data.size()) + " values for " + std::to_string(m.rows) + "x" + std::to_string(m.cols);
            return SA_ERR_PARSE;
        }
        out.push_back(std::move(m));
    }
    return SA_OK;
}

}  // namespace sa
