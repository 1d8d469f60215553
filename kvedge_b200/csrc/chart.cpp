// Chart loading, values handling and manifest framing (see chart.hpp).
#include "chart.hpp"

#include <dirent.h>
#include <fnmatch.h>
#include <sys/stat.h>

#include <algorithm>
#include <cerrno>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>

namespace kvedge {
namespace chart {

using tmpl::Error;

// ============================================================================ helpers
static bool is_ws(char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\n' || c == '\v' || c == '\f'; }

std::string trim_space(const std::string &s) {
    size_t a = 0, b = s.size();
    while (a < b && is_ws(s[a])) ++a;
    while (b > a && is_ws(s[b - 1])) --b;
    return s.substr(a, b - a);
}

static std::string read_file(const std::string &path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw Error("open " + path + ": " + strerror(errno));
    std::ostringstream ss;
    ss << f.rdbuf();
    return ss.str();
}

static bool file_exists(const std::string &p) {
    struct stat st;
    return stat(p.c_str(), &st) == 0;
}

// ============================================================================ YAML subset
namespace {

bool in_set(const std::string &s, std::initializer_list<const char *> set) {
    for (const char *x : set)
        if (s == x) return true;
    return false;
}

Value resolve_plain(const std::string &s, bool json_numbers) {
    if (in_set(s, {"", "~", "null", "Null", "NULL"})) return Value::nil();
    if (in_set(s, {"y", "Y", "yes", "Yes", "YES", "on", "On", "ON", "true", "True", "TRUE"})) return Value::boolean(true);
    if (in_set(s, {"n", "N", "no", "No", "NO", "off", "Off", "OFF", "false", "False", "FALSE"})) return Value::boolean(false);
    std::string plain;
    for (char c : s)
        if (c != '_') plain.push_back(c);
    // integers: [-+]? (0x.. | 0o.. | 0b.. | 0[0-7]* | [1-9][0-9]*)
    {
        size_t i = 0;
        bool neg = false;
        if (i < plain.size() && (plain[i] == '-' || plain[i] == '+')) { neg = plain[i] == '-'; ++i; }
        std::string body = plain.substr(i);
        int base = 0;
        std::string digits;
        if (body.size() > 2 && body[0] == '0' && (body[1] == 'x' || body[1] == 'X')) { base = 16; digits = body.substr(2); }
        else if (body.size() > 2 && body[0] == '0' && (body[1] == 'o' || body[1] == 'O')) { base = 8; digits = body.substr(2); }
        else if (body.size() > 2 && body[0] == '0' && (body[1] == 'b' || body[1] == 'B')) { base = 2; digits = body.substr(2); }
        else if (!body.empty() && body[0] == '0') { base = 8; digits = body.size() > 1 ? body.substr(1) : "0"; }
        else if (!body.empty() && body[0] >= '1' && body[0] <= '9') { base = 10; digits = body; }
        bool ok = base != 0 && !digits.empty();
        for (char c : digits) {
            int d = (c >= '0' && c <= '9') ? c - '0' : (c >= 'a' && c <= 'f') ? c - 'a' + 10 : (c >= 'A' && c <= 'F') ? c - 'A' + 10 : 99;
            if (d >= base) ok = false;
        }
        if (ok) {
            errno = 0;
            unsigned long long mag = strtoull(digits.c_str(), nullptr, base);
            if (errno == 0) {
                if (json_numbers) return Value::number(neg && mag ? -(double)mag : (double)mag);  // "-0" is the integer 0
                if (mag <= (unsigned long long)INT64_MAX) return Value::integer(neg ? -(int64_t)mag : (int64_t)mag);
                return Value::number(neg ? -(double)mag : (double)mag);
            }
        }
    }
    // floats: [-+]?(\.[0-9]+|[0-9]+(\.[0-9]*)?)([eE][-+]?[0-9]+)?
    {
        size_t i = 0, n = plain.size();
        if (i < n && (plain[i] == '-' || plain[i] == '+')) ++i;
        size_t d0 = i;
        while (i < n && isdigit((unsigned char)plain[i])) ++i;
        bool int_digits = i > d0, frac_digits = false;
        if (i < n && plain[i] == '.') {
            ++i;
            size_t f0 = i;
            while (i < n && isdigit((unsigned char)plain[i])) ++i;
            frac_digits = i > f0;
        }
        bool ok = int_digits || frac_digits;
        if (ok && i < n && (plain[i] == 'e' || plain[i] == 'E')) {
            ++i;
            if (i < n && (plain[i] == '-' || plain[i] == '+')) ++i;
            size_t e0 = i;
            while (i < n && isdigit((unsigned char)plain[i])) ++i;
            if (i == e0) ok = false;
        }
        if (ok && i == n && n > 0) return Value::number(strtod(plain.c_str(), nullptr));
    }
    if (in_set(plain, {".inf", ".Inf", ".INF", "+.inf", "+.Inf", "+.INF"})) return Value::number(INFINITY);
    if (in_set(plain, {"-.inf", "-.Inf", "-.INF"})) return Value::number(-INFINITY);
    if (in_set(plain, {".nan", ".NaN", ".NAN"})) return Value::number(NAN);
    return Value::str(s);
}

void append_utf8(std::string &o, uint32_t cp) {
    if (cp < 0x80) o.push_back((char)cp);
    else if (cp < 0x800) { o.push_back((char)(0xC0 | (cp >> 6))); o.push_back((char)(0x80 | (cp & 0x3F))); }
    else if (cp < 0x10000) {
        o.push_back((char)(0xE0 | (cp >> 12))); o.push_back((char)(0x80 | ((cp >> 6) & 0x3F))); o.push_back((char)(0x80 | (cp & 0x3F)));
    } else {
        o.push_back((char)(0xF0 | (cp >> 18))); o.push_back((char)(0x80 | ((cp >> 12) & 0x3F)));
        o.push_back((char)(0x80 | ((cp >> 6) & 0x3F))); o.push_back((char)(0x80 | (cp & 0x3F)));
    }
}

struct Yaml {
    std::vector<std::string> lines;
    size_t cur = 0;
    bool json_numbers;

    static int indent_of(const std::string &l) {
        int n = 0;
        while ((size_t)n < l.size() && l[(size_t)n] == ' ') ++n;
        return n;
    }
    static bool blank(const std::string &l) {
        for (char c : l) {
            if (c == '#') return true;
            if (!is_ws(c)) return false;
        }
        return true;
    }
    void skip_blank() {
        while (cur < lines.size() && (blank(lines[cur]) || (lines[cur].compare(0, 3, "---") == 0 && cur == 0))) ++cur;
    }
    // strip a trailing comment (a '#' preceded by white space, outside quotes)
    static std::string strip_comment(const std::string &s) {
        bool sq = false, dq = false;
        for (size_t i = 0; i < s.size(); ++i) {
            char c = s[i];
            if (dq) { if (c == '\\') ++i; else if (c == '"') dq = false; continue; }
            if (sq) {
                if (c == '\'') {
                    if (i + 1 < s.size() && s[i + 1] == '\'') ++i;  // '' is an escaped quote
                    else sq = false;
                }
                continue;
            }
            if (c == '"' && (i == 0 || is_ws(s[i - 1]) || s[i - 1] == ':' || s[i - 1] == '[' || s[i - 1] == ',' || s[i - 1] == '{')) { dq = true; continue; }
            if (c == '\'' && (i == 0 || is_ws(s[i - 1]) || s[i - 1] == '[' || s[i - 1] == ',' || s[i - 1] == '{')) { sq = true; continue; }
            if (c == '#' && (i == 0 || is_ws(s[i - 1]))) return s.substr(0, i);
        }
        return s;
    }
    // position of the ':' that ends a mapping key in `s` (content without indent), or npos
    static size_t key_colon(const std::string &s) {
        if (s.empty()) return std::string::npos;
        size_t i = 0;
        if (s[0] == '"' || s[0] == '\'') {
            char q = s[0];
            ++i;
            while (i < s.size() && s[i] != q) {
                if (q == '"' && s[i] == '\\') ++i;
                ++i;
            }
            ++i;
            while (i < s.size() && s[i] == ' ') ++i;
            if (i < s.size() && s[i] == ':' && (i + 1 == s.size() || is_ws(s[i + 1]))) return i;
            return std::string::npos;
        }
        if (s[0] == '[' || s[0] == '{' || s[0] == '|' || s[0] == '>') return std::string::npos;
        for (; i < s.size(); ++i) {
            if (s[i] == ':' && (i + 1 == s.size() || is_ws(s[i + 1]))) return i;
            if (s[i] == '#' && i > 0 && is_ws(s[i - 1])) return std::string::npos;
        }
        return std::string::npos;
    }
    std::string unquote_double(const std::string &s, size_t *end) {
        std::string o;
        size_t i = 1;
        for (; i < s.size() && s[i] != '"'; ++i) {
            char c = s[i];
            if (c != '\\') { o.push_back(c); continue; }
            ++i;
            if (i >= s.size()) break;
            switch (s[i]) {
            case 'n': o.push_back('\n'); break;
            case 't': o.push_back('\t'); break;
            case 'r': o.push_back('\r'); break;
            case '0': o.push_back('\0'); break;
            case 'a': o.push_back(7); break;
            case 'b': o.push_back(8); break;
            case 'f': o.push_back(12); break;
            case 'v': o.push_back(11); break;
            case 'e': o.push_back(27); break;
            case ' ': o.push_back(' '); break;
            case '"': o.push_back('"'); break;
            case '/': o.push_back('/'); break;
            case '\\': o.push_back('\\'); break;
            case 'N': append_utf8(o, 0x85); break;
            case '_': append_utf8(o, 0xA0); break;
            case 'x': o.push_back((char)strtol(s.substr(i + 1, 2).c_str(), nullptr, 16)); i += 2; break;
            case 'u': append_utf8(o, (uint32_t)strtol(s.substr(i + 1, 4).c_str(), nullptr, 16)); i += 4; break;
            case 'U': append_utf8(o, (uint32_t)strtol(s.substr(i + 1, 8).c_str(), nullptr, 16)); i += 8; break;
            default: throw Error(std::string("yaml: unknown escape \\") + s[i]);
            }
        }
        if (i >= s.size()) throw Error("yaml: unterminated double-quoted scalar");
        *end = i + 1;
        return o;
    }
    std::string unquote_single(const std::string &s, size_t *end) {
        std::string o;
        size_t i = 1;
        for (; i < s.size(); ++i) {
            if (s[i] == '\'') {
                if (i + 1 < s.size() && s[i + 1] == '\'') { o.push_back('\''); ++i; continue; }
                break;
            }
            o.push_back(s[i]);
        }
        if (i >= s.size()) throw Error("yaml: unterminated single-quoted scalar");
        *end = i + 1;
        return o;
    }
    // flow collection or scalar from one piece of text
    Value parse_flow(const std::string &s, size_t &i) {
        while (i < s.size() && is_ws(s[i])) ++i;
        if (i >= s.size()) return Value::nil();
        if (s[i] == '[') {
            Value v = Value::make_list();
            ++i;
            for (;;) {
                while (i < s.size() && is_ws(s[i])) ++i;
                if (i >= s.size()) throw Error("yaml: unterminated flow sequence");
                if (s[i] == ']') { ++i; break; }
                v.list->push_back(parse_flow(s, i));
                while (i < s.size() && is_ws(s[i])) ++i;
                if (i < s.size() && s[i] == ',') ++i;
            }
            return v;
        }
        if (s[i] == '{') {
            Value v = Value::make_map();
            ++i;
            for (;;) {
                while (i < s.size() && is_ws(s[i])) ++i;
                if (i >= s.size()) throw Error("yaml: unterminated flow mapping");
                if (s[i] == '}') { ++i; break; }
                Value k = parse_flow_scalar(s, i, true);
                while (i < s.size() && is_ws(s[i])) ++i;
                Value val;
                if (i < s.size() && s[i] == ':') { ++i; val = parse_flow(s, i); }
                (*v.map)[tmpl::go_print(k)] = val;
                while (i < s.size() && is_ws(s[i])) ++i;
                if (i < s.size() && s[i] == ',') ++i;
            }
            return v;
        }
        return parse_flow_scalar(s, i, false);
    }
    Value parse_flow_scalar(const std::string &s, size_t &i, bool is_key) {
        while (i < s.size() && is_ws(s[i])) ++i;
        if (i < s.size() && s[i] == '"') {
            size_t e;
            std::string sub = s.substr(i);
            std::string o = unquote_double(sub, &e);
            i += e;
            return Value::str(o);
        }
        if (i < s.size() && s[i] == '\'') {
            size_t e;
            std::string sub = s.substr(i);
            std::string o = unquote_single(sub, &e);
            i += e;
            return Value::str(o);
        }
        size_t j = i;
        while (j < s.size() && s[j] != ',' && s[j] != ']' && s[j] != '}' &&
               !(is_key && s[j] == ':' ) && !(s[j] == ':' && j + 1 < s.size() && is_ws(s[j + 1])))
            ++j;
        std::string tok = trim_space(s.substr(i, j - i));
        i = j;
        return resolve_plain(tok, json_numbers);
    }
    Value scalar_from_text(const std::string &text) {
        std::string t = trim_space(text);
        if (t.empty()) return Value::nil();
        if (t[0] == '"') { size_t e; std::string o = unquote_double(t, &e); return Value::str(o); }
        if (t[0] == '\'') { size_t e; std::string o = unquote_single(t, &e); return Value::str(o); }
        if (t[0] == '[' || t[0] == '{') { size_t i = 0; return parse_flow(t, i); }
        return resolve_plain(t, json_numbers);
    }
    Value block_scalar(const std::string &header, int parent_indent) {
        bool literal = header[0] == '|';
        char chomp = 'c';
        int explicit_indent = 0;
        for (size_t i = 1; i < header.size(); ++i) {
            if (header[i] == '-') chomp = 's';
            else if (header[i] == '+') chomp = 'k';
            else if (isdigit((unsigned char)header[i])) explicit_indent = header[i] - '0';
        }
        std::vector<std::string> body;
        int bi = explicit_indent ? parent_indent + explicit_indent : -1;
        while (cur < lines.size()) {
            const std::string &l = lines[cur];
            bool empty = trim_space(l).empty();
            int ind = indent_of(l);
            if (!empty) {
                if (bi < 0) {
                    if (ind <= parent_indent) break;
                    bi = ind;
                }
                if (ind < bi) break;
            }
            body.push_back(empty ? "" : l.substr((size_t)std::min<int>(bi < 0 ? ind : bi, (int)l.size())));
            ++cur;
        }
        size_t trailing = 0;
        while (!body.empty() && body.back().empty()) { body.pop_back(); ++trailing; }
        std::string o;
        if (literal) {
            for (size_t i = 0; i < body.size(); ++i) { o += body[i]; if (i + 1 < body.size()) o += "\n"; }
        } else {
            for (size_t i = 0; i < body.size(); ++i) {
                o += body[i];
                if (i + 1 < body.size()) {
                    if (body[i + 1].empty()) { o += "\n"; while (i + 1 < body.size() && body[i + 1].empty()) { ++i; if (i + 1 < body.size() && body[i + 1].empty()) o += "\n"; } }
                    else if (body[i].empty() || body[i][0] == ' ' || body[i + 1][0] == ' ') o += "\n";
                    else o += " ";
                }
            }
        }
        if (!body.empty()) {
            if (chomp == 'c') o += "\n";
            else if (chomp == 'k') o += std::string(trailing + 1, '\n');
        }
        return Value::str(o);
    }
    Value parse_value_after_key(const std::string &rest_raw, int ind) {
        std::string rest = trim_space(strip_comment(rest_raw));
        if (!rest.empty() && (rest[0] == '|' || rest[0] == '>')) return block_scalar(rest, ind);
        if (!rest.empty()) return scalar_from_text(rest);
        // value on the following lines
        size_t save = cur;
        skip_blank();
        if (cur < lines.size()) {
            int ni = indent_of(lines[cur]);
            std::string c = lines[cur].substr((size_t)ni);
            if (ni > ind) return parse_block(ni);
            if (ni == ind && (c.compare(0, 2, "- ") == 0 || c == "-")) return parse_seq(ni);
        }
        cur = save;
        return Value::nil();
    }
    Value parse_map(int ind) {
        Value v = Value::make_map();
        for (;;) {
            skip_blank();
            if (cur >= lines.size()) break;
            if (indent_of(lines[cur]) != ind) {
                if (indent_of(lines[cur]) > ind) throw Error("yaml: bad indentation at line " + std::to_string(cur + 1));
                break;
            }
            std::string c = lines[cur].substr((size_t)ind);
            size_t kc = key_colon(c);
            if (kc == std::string::npos) break;
            std::string key_txt = trim_space(c.substr(0, kc));
            std::string key;
            if (!key_txt.empty() && key_txt[0] == '"') { size_t e; key = unquote_double(key_txt, &e); }
            else if (!key_txt.empty() && key_txt[0] == '\'') { size_t e; key = unquote_single(key_txt, &e); }
            else {
                // go-yaml types plain keys too, and sigs.k8s.io/yaml turns them back into strings:
                // a key written `n` or `yes` reaches the template as "false" / "true"
                Value kv = resolve_plain(key_txt, false);
                if (kv.kind == Value::Nil) throw Error("yaml: unsupported map key of type <nil> (line " + std::to_string(cur + 1) + ")");
                key = kv.kind == Value::String ? kv.s : tmpl::go_print(kv);
            }
            ++cur;
            (*v.map)[key] = parse_value_after_key(c.substr(kc + 1), ind);
        }
        return v;
    }
    Value parse_seq(int ind) {
        Value v = Value::make_list();
        for (;;) {
            skip_blank();
            if (cur >= lines.size() || indent_of(lines[cur]) != ind) break;
            std::string c = lines[cur].substr((size_t)ind);
            if (!(c.compare(0, 2, "- ") == 0 || c == "-")) break;
            std::string item = c.size() > 1 ? c.substr(2) : "";
            size_t lead = 0;
            while (lead < item.size() && item[lead] == ' ') ++lead;
            std::string body = item.substr(lead);
            if (trim_space(strip_comment(body)).empty()) {
                ++cur;
                size_t save = cur;
                skip_blank();
                if (cur < lines.size() && indent_of(lines[cur]) > ind) v.list->push_back(parse_block(indent_of(lines[cur])));
                else { cur = save; v.list->push_back(Value::nil()); }
                continue;
            }
            if (key_colon(body) != std::string::npos) {
                // "- key: value": a mapping that starts on the dash line
                int vi = ind + 2 + (int)lead;
                lines[cur] = std::string((size_t)vi, ' ') + body;
                v.list->push_back(parse_map(vi));
                continue;
            }
            ++cur;
            std::string t = trim_space(strip_comment(body));
            if (t[0] == '|' || t[0] == '>') v.list->push_back(block_scalar(t, ind));
            else v.list->push_back(scalar_from_text(t));
        }
        return v;
    }
    Value parse_block(int ind) {
        skip_blank();
        if (cur >= lines.size()) return Value::nil();
        std::string c = lines[cur].substr((size_t)indent_of(lines[cur]));
        if (c.compare(0, 2, "- ") == 0 || c == "-") return parse_seq(indent_of(lines[cur]));
        if (key_colon(c) != std::string::npos) return parse_map(indent_of(lines[cur]));
        ++cur;
        (void)ind;
        return scalar_from_text(strip_comment(c));
    }
};

}  // namespace

Value parse_yaml(const std::string &text, bool json_numbers) {
    Yaml y;
    y.json_numbers = json_numbers;
    size_t i = 0;
    while (i <= text.size()) {
        size_t j = text.find('\n', i);
        if (j == std::string::npos) j = text.size();
        std::string l = text.substr(i, j - i);
        if (!l.empty() && l.back() == '\r') l.pop_back();
        for (char c : l) {
            if (c == '\t' ) { /* tabs are not indentation in YAML; keep as content */ break; }
            if (c != ' ') break;
        }
        y.lines.push_back(l);
        i = j + 1;
    }
    y.skip_blank();
    if (y.cur >= y.lines.size()) return Value::nil();
    Value v = y.parse_block(0);
    y.skip_blank();
    if (y.cur < y.lines.size() && y.lines[y.cur].compare(0, 3, "...") != 0 && y.lines[y.cur].compare(0, 3, "---") != 0)
        throw Error("yaml: could not parse line " + std::to_string(y.cur + 1) + ": " + y.lines[y.cur]);
    return v;
}

// ============================================================================ strvals
namespace {

Value set_typed(std::string val, bool force_string) {
    if (force_string) return Value::str(std::move(val));
    if (val.size() <= 5) {   // strings.EqualFold(val, "true" / "false" / "null")
        char low[6] = {0, 0, 0, 0, 0, 0};
        for (size_t k = 0; k < val.size(); ++k) low[k] = (char)tolower((unsigned char)val[k]);
        if (!strcmp(low, "true")) return Value::boolean(true);
        if (!strcmp(low, "false")) return Value::boolean(false);
        if (!strcmp(low, "null")) return Value::nil();
    }
    if (val == "0") return Value::integer(0);
    if (!val.empty() && val[0] != '0' && val.size() <= 20) {
        size_t i = (val[0] == '-' || val[0] == '+') ? 1 : 0;
        bool digits = i < val.size();
        for (size_t k = i; k < val.size() && digits; ++k)
            if (!isdigit((unsigned char)val[k])) digits = false;
        if (digits) {
            errno = 0;
            long long v = strtoll(val.c_str(), nullptr, 10);
            if (errno == 0) return Value::integer(v);
        }
    }
    return Value::str(std::move(val));
}

// key=value[,key=value...]: runs of ordinary characters are appended whole (a value is often a
// 400-character SSH key), only backslashes and the three delimiters are looked at one by one.
std::vector<std::pair<std::string, std::string>> split_set(const std::string &s) {
    std::vector<std::pair<std::string, std::string>> pairs;
    const char *p = s.data();
    size_t i = 0, n = s.size();
    while (i < n) {
        std::string key;
        for (;;) {
            size_t j = i;
            while (j < n && p[j] != '=' && p[j] != '\\' && p[j] != ',') ++j;
            key.append(p + i, j - i);
            i = j;
            if (i >= n || p[i] == '=') break;
            if (p[i] == '\\' && i + 1 < n) { key.append(p + i, 2); i += 2; continue; }
            if (p[i] == ',') throw Error("key \"" + key + "\" has no value");
            key.push_back(p[i++]);   // a lone backslash at the very end
        }
        if (i >= n) throw Error("key \"" + key + "\" has no value");
        ++i;
        std::string val;
        if (i < n && p[i] == '{') {
            size_t j = s.find('}', i);
            if (j == std::string::npos) throw Error("list must terminate with '}'");
            val = s.substr(i, j + 1 - i);
            i = j + 1;
            if (i < n && p[i] == ',') ++i;
        } else {
            for (;;) {
                size_t j = i;
                while (j < n && p[j] != '\\' && p[j] != ',') ++j;
                val.append(p + i, j - i);
                i = j;
                if (i >= n) break;
                if (p[i] == '\\' && i + 1 < n) { val.push_back(p[i + 1]); i += 2; continue; }
                if (p[i] == ',') { ++i; break; }
                val.push_back(p[i++]);
            }
        }
        pairs.emplace_back(std::move(key), std::move(val));
    }
    return pairs;
}

struct PathElem { bool is_index; std::string key; size_t index; };

std::vector<PathElem> split_key(const std::string &key) {
    std::vector<PathElem> parts;
    std::string cur;
    for (size_t i = 0; i < key.size();) {
        char c = key[i];
        if (c == '\\' && i + 1 < key.size()) { cur.push_back(key[i + 1]); i += 2; }
        else if (c == '.') { if (!cur.empty()) parts.push_back({false, cur, 0}); cur.clear(); ++i; }
        else if (c == '[') {
            if (!cur.empty()) parts.push_back({false, cur, 0});
            cur.clear();
            size_t j = key.find(']', i);
            if (j == std::string::npos) throw Error("unterminated index in key " + key);
            parts.push_back({true, "", (size_t)strtoull(key.substr(i + 1, j - i - 1).c_str(), nullptr, 10)});
            i = j + 1;
        } else { cur.push_back(c); ++i; }
    }
    if (!cur.empty()) parts.push_back({false, cur, 0});
    return parts;
}

void assign(Value &root, const std::vector<PathElem> &path, Value value) {
    if (root.kind != Value::MapK) root = Value::make_map();
    Value *cur = &root;
    for (size_t k = 0; k < path.size(); ++k) {
        bool last = k + 1 == path.size();
        bool next_is_index = !last && path[k + 1].is_index;
        Value *slot;
        if (path[k].is_index) {
            if (cur->kind != Value::ListK) throw Error("--set: cannot index into a map with [" + std::to_string(path[k].index) + "]");
            while (cur->list->size() <= path[k].index) cur->list->push_back(Value::nil());
            slot = &(*cur->list)[path[k].index];
        } else {
            if (cur->kind != Value::MapK) throw Error("--set: key \"" + path[k].key + "\" addresses a list, not a map");
            slot = &(*cur->map)[path[k].key];
        }
        if (last) { *slot = std::move(value); return; }
        if (slot->kind != Value::MapK && slot->kind != Value::ListK) *slot = next_is_index ? Value::make_list() : Value::make_map();
        cur = slot;
    }
}

}  // namespace

void apply_set(Value &values, const std::string &expr, bool force_string) {
    for (auto &kv : split_set(expr)) {
        std::string &val = kv.second;
        Value v;
        if (val.size() >= 2 && val.front() == '{' && val.back() == '}') {
            v = Value::make_list();
            std::string inner = val.substr(1, val.size() - 2);
            if (!inner.empty()) {
                size_t i = 0;
                for (;;) {
                    size_t j = inner.find(',', i);
                    std::string item = trim_space(inner.substr(i, j == std::string::npos ? std::string::npos : j - i));
                    v.list->push_back(set_typed(item, force_string));
                    if (j == std::string::npos) break;
                    i = j + 1;
                }
            }
        } else {
            v = set_typed(std::move(val), force_string);
        }
        assign(values, split_key(kv.first), std::move(v));
    }
}

void apply_set_file(Value &values, const std::string &expr) {
    for (const auto &kv : split_set(expr)) assign(values, split_key(kv.first), Value::str(read_file(kv.second)));
}

void apply_set_blob(Value &values, const std::string &key, std::string content) {
    Value v = Value::str(std::string());
    v.s = std::move(content);
    assign(values, split_key(key), std::move(v));
}

// ============================================================================ coalesce
namespace {

// chartutil.coalesceTablesFullKey: dst = the user's table, src = the chart's.
void coalesce_tables(tmpl::Map &dst, const tmpl::Map &src) {
    for (const auto &kv : src) {
        auto it = dst.find(kv.first);
        if (it == dst.end()) {
            dst[kv.first] = kv.second;   // shared, not copied: nothing mutates a value tree after this merge
        } else if (it->second.kind == Value::Nil) {
            dst.erase(it);                                  // "when the YAML value is null, we remove it"
        } else if (kv.second.kind == Value::MapK && it->second.kind == Value::MapK) {
            coalesce_tables(*it->second.map, *kv.second.map);
        }                                                   // otherwise the user's value stands (Helm warns)
    }
}

}  // namespace

Value coalesce_values(const Value &defaults, Value user) {
    if (user.kind != Value::MapK) user = Value::make_map();
    if (defaults.kind == Value::MapK) {
        user.map->reserve(user.map->size() + defaults.map->size());
        coalesce_tables(*user.map, *defaults.map);
    }
    return user;
}

// ============================================================================ chart
namespace {

struct IgnoreRule { std::string pat; bool neg, dir_only, anchored; };

std::vector<IgnoreRule> ignore_rules(const std::string &text) {
    std::vector<IgnoreRule> rules;
    std::istringstream ss(text);
    std::string line;
    while (std::getline(ss, line)) {
        line = trim_space(line);
        if (line.empty() || line[0] == '#') continue;
        IgnoreRule r{line, false, false, false};
        if (r.pat[0] == '!') { r.neg = true; r.pat.erase(0, 1); }
        if (!r.pat.empty() && r.pat.back() == '/') { r.dir_only = true; r.pat.pop_back(); }
        if (!r.pat.empty() && r.pat[0] == '/') { r.anchored = true; r.pat.erase(0, 1); }
        rules.push_back(r);
    }
    return rules;
}

bool ignored(const std::string &rel, bool is_dir, const std::vector<IgnoreRule> &rules) {
    bool ig = false;
    std::string base = rel.substr(rel.find_last_of('/') == std::string::npos ? 0 : rel.find_last_of('/') + 1);
    for (const auto &r : rules) {
        if (r.dir_only && !is_dir) continue;
        bool hit = (r.pat.find('/') != std::string::npos || r.anchored) ? fnmatch(r.pat.c_str(), rel.c_str(), 0) == 0
                                                                        : fnmatch(r.pat.c_str(), base.c_str(), 0) == 0;
        if (hit) ig = !r.neg;
    }
    return ig;
}

void walk(const std::string &root, const std::string &rel, const std::vector<IgnoreRule> &rules,
          std::vector<std::pair<std::string, std::string>> *out) {
    std::string dirpath = root + "/" + rel;
    DIR *d = opendir(dirpath.c_str());
    if (!d) return;
    std::vector<std::string> names;
    while (struct dirent *e = readdir(d)) {
        std::string n = e->d_name;
        if (n != "." && n != "..") names.push_back(n);
    }
    closedir(d);
    std::sort(names.begin(), names.end());
    for (const auto &n : names) {
        std::string r = rel + "/" + n;
        struct stat st;
        if (stat((root + "/" + r).c_str(), &st) != 0) continue;
        bool is_dir = S_ISDIR(st.st_mode);
        if (ignored(r, is_dir, rules)) continue;
        if (is_dir) walk(root, r, rules, out);
        else out->push_back({r, read_file(root + "/" + r)});
    }
}

}  // namespace

std::string Chart::name() const {
    if (metadata.kind == Value::MapK) {
        auto it = metadata.map->find("name");
        if (it != metadata.map->end()) return tmpl::go_print(it->second);
    }
    return "";
}

Chart load_chart(const std::string &dir) {
    Chart c;
    c.dir = dir;
    c.metadata = parse_yaml(read_file(dir + "/Chart.yaml"), false);
    if (c.metadata.kind != Value::MapK) throw Error("Chart.yaml: not a mapping");
    c.values = Value::make_map();
    if (file_exists(dir + "/values.yaml")) {
        Value v = parse_yaml(read_file(dir + "/values.yaml"), true);
        if (v.kind == Value::MapK) c.values = v;
        else if (!v.is_nilish()) throw Error("values.yaml: not a mapping");
    }
    std::vector<IgnoreRule> rules;
    if (file_exists(dir + "/.helmignore")) rules = ignore_rules(read_file(dir + "/.helmignore"));
    walk(dir, "templates", rules, &c.templates);
    return c;
}

// ============================================================================ manifests
static const char *kInstallOrder[] = {
    "Namespace", "NetworkPolicy", "ResourceQuota", "LimitRange", "PodSecurityPolicy", "PodDisruptionBudget",
    "ServiceAccount", "Secret", "SecretList", "ConfigMap", "StorageClass", "PersistentVolume",
    "PersistentVolumeClaim", "CustomResourceDefinition", "ClusterRole", "ClusterRoleList",
    "ClusterRoleBinding", "ClusterRoleBindingList", "Role", "RoleList", "RoleBinding", "RoleBindingList",
    "Service", "DaemonSet", "Pod", "ReplicationController", "ReplicaSet", "Deployment",
    "HorizontalPodAutoscaler", "StatefulSet", "Job", "CronJob", "IngressClass", "Ingress", "APIService"};

static int install_rank(const std::string &kind) {
    for (size_t i = 0; i < sizeof kInstallOrder / sizeof *kInstallOrder; ++i)
        if (kind == kInstallOrder[i]) return (int)i;
    return -1;
}

// releaseutil.SplitManifests: strings.Split on the regexp `(?:^|\s*\n)---\s*` -- "---" at the very
// start of the text or right after a newline; the white space in front of that newline and
// behind the dashes belongs to the separator, and whatever follows the dashes on their line
// starts the next document.  (Go's \s is [\t\n\f\r ].)
static bool re_space(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\f' || c == '\r'; }
namespace {

struct DocRef { const std::string *name; int rank; std::string kind; const char *p; size_t n; };

bool uni_space(char c) { return is_ws(c) || c == '\v' || c == '\f'; }

// Top-level "kind:" of a document head.
std::string doc_kind(const char *d, size_t n) {
    size_t i = 0;
    while (i < n) {
        const char *nl = static_cast<const char *>(memchr(d + i, '\n', n - i));
        const size_t j = nl ? (size_t)(nl - d) : n;
        if (j - i >= 5 && memcmp(d + i, "kind:", 5) == 0) {
            size_t k = i + 5;
            while (k < j && (d[k] == ' ' || d[k] == '\t')) ++k;
            size_t e = k;
            while (e < j && !is_ws(d[e]) && d[e] != '#') ++e;
            std::string kind(d + k, e - k);
            if (kind.size() >= 2 && (kind.front() == '"' || kind.front() == '\'')) kind = kind.substr(1, kind.size() - 2);
            return kind;
        }
        i = j + 1;
    }
    return std::string();
}

// releaseutil.SortManifests over rendered files (sorted by path): documents as views into the
// files' own text -- nothing is copied until the caller assembles its output.
void sorted_docs(const std::vector<std::pair<const std::string *, const std::string *>> &files,
                 std::vector<DocRef> *docs, const std::string **notes) {
    for (const auto &f : files) {
        const std::string &name = *f.first, &text = *f.second;
        const size_t slash = name.find_last_of('/');
        if (name.compare(slash == std::string::npos ? 0 : slash + 1, std::string::npos, "NOTES.txt") == 0) {
            *notes = &text;
            continue;
        }
        size_t a = 0, b = text.size();   // strings.TrimSpace
        while (a < b && uni_space(text[a])) ++a;
        while (b > a && uni_space(text[b - 1])) --b;
        if (a == b) continue;
        // SplitManifests' separator `(?:^|\s*\n)---\s*` over [a, b)
        const char *t = text.data();
        size_t start = a, i = a;
        auto emit = [&](size_t lo, size_t hi) {
            while (lo < hi && uni_space(t[lo])) ++lo;   // TrimLeftFunc(unicode.IsSpace)
            if (lo == hi) return;
            std::string kind = doc_kind(t + lo, hi - lo);
            const int rank = install_rank(kind);
            docs->push_back({&name, rank, std::move(kind), t + lo, hi - lo});
        };
        for (;;) {
            size_t ms, me;
            if (i == a && b - a >= 3 && memcmp(t + a, "---", 3) == 0) {
                ms = a;
                me = a + 3;
            } else {
                // next "\n---": hop from newline to newline (memchr runs at memory speed over the
                // kilobytes of a Secret's payload, which hold no newline at all)
                size_t p = i;
                for (;;) {
                    const char *nl = p + 4 <= b ? static_cast<const char *>(memchr(t + p, '\n', b - p - 3)) : nullptr;
                    if (!nl) { p = std::string::npos; break; }
                    p = (size_t)(nl - t);
                    if (t[p + 1] == '-' && t[p + 2] == '-' && t[p + 3] == '-') break;
                    ++p;
                }
                if (p == std::string::npos) break;
                ms = p;
                while (ms > i && re_space(t[ms - 1])) --ms;
                me = p + 4;
            }
            while (me < b && re_space(t[me])) ++me;
            emit(start, ms);
            start = i = me;
        }
        emit(start, b);
    }
    auto before = [](const DocRef &x, const DocRef &y) {
        if (x.rank < 0 && y.rank < 0) return x.kind < y.kind;  // unknown kinds: alphabetical, stable within a kind
        if (x.rank < 0) return false;
        if (y.rank < 0) return true;
        return x.rank < y.rank;
    };
    if (docs->size() <= 32) {   // a chart's handful of documents: stable insertion sort, no temporary buffer
        for (size_t i = 1; i < docs->size(); ++i) {
            size_t j = i;
            while (j > 0 && before((*docs)[i], (*docs)[j - 1])) --j;
            if (j != i) std::rotate(docs->begin() + (std::ptrdiff_t)j, docs->begin() + (std::ptrdiff_t)i,
                                    docs->begin() + (std::ptrdiff_t)i + 1);
        }
    } else {
        std::stable_sort(docs->begin(), docs->end(), before);
    }
}

}  // namespace

void split_manifests(const std::map<std::string, std::string> &files, std::vector<Manifest> *out, std::string *notes) {
    out->clear();
    notes->clear();
    std::vector<std::pair<const std::string *, const std::string *>> refs;
    for (const auto &kv : files) refs.push_back({&kv.first, &kv.second});  // std::map: sorted by path, like Helm sorts file names
    std::vector<DocRef> docs;
    const std::string *n = nullptr;
    sorted_docs(refs, &docs, &n);
    if (n) *notes = *n;
    for (const auto &d : docs) out->push_back({*d.name, d.kind, std::string(d.p, d.n)});
}

void frame_manifests(const std::vector<std::pair<const std::string *, const std::string *>> &files,
                     std::string *manifest, std::string *notes) {
    static thread_local std::vector<DocRef> docs;   // capacity kept per thread: a fleet frames thousands of releases
    docs.clear();
    const std::string *n = nullptr;
    sorted_docs(files, &docs, &n);
    size_t total = 0;
    for (const auto &d : docs) total += 14 + d.name->size() + 1 + d.n + 1;
    manifest->clear();
    manifest->reserve(total);
    for (const auto &d : docs) {
        manifest->append("---\n# Source: ", 14);
        manifest->append(*d.name);
        manifest->push_back('\n');
        manifest->append(d.p, d.n);
        manifest->push_back('\n');
    }
    notes->clear();
    if (n) {
        size_t a = 0, b = n->size();
        while (a < b && uni_space((*n)[a])) ++a;
        while (b > a && uni_space((*n)[b - 1])) --b;
        notes->assign(*n, a, b - a);
    }
}

}  // namespace chart
}  // namespace kvedge
