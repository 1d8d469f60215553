// Go text/template + Sprig subset (see tmpl.hpp). Host C++: lexer, parser, executor.
#include "tmpl.hpp"

#include <algorithm>
#include <cctype>
#include <charconv>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <deque>
#ifdef __SSE2__
#include <emmintrin.h>
#endif

namespace kvedge {
namespace tmpl {

// ============================================================================ values
bool truth(const Value &v) {
    switch (v.kind) {
    case Value::Nil: case Value::NoValue: return false;
    case Value::Bool: return v.b;
    case Value::Int: return v.i != 0;
    case Value::Float: return v.f != 0;
    case Value::String: return !v.s.empty();
    case Value::ListK: return v.list && !v.list->empty();
    case Value::MapK: return v.map && !v.map->empty();
    case Value::StructK: return true;
    }
    return false;
}

// strconv.FormatFloat(f, 'g', -1, 64) as fmt's %v uses it: shortest digits that round-trip,
// exponent form when exp < -4 || exp >= 21 is JSON's rule; fmt's is exp < -4 || exp >= 6... with
// the shortest flag eprec is 6 -- the reason a values.yaml integer 1234567 prints as
// 1.234567e+06 in Helm.
std::string format_float(double f) {
    if (std::isnan(f)) return "NaN";
    if (std::isinf(f)) return f > 0 ? "+Inf" : "-Inf";
    if (f == 0) return std::signbit(f) ? "-0" : "0";
    // shortest round-trip digits (strconv's 'shortest' mode); printf at growing precision is
    // not the same thing: it breaks exact ties to even and can need one digit more (2^-24)
    char buf[64];
    auto tc = std::to_chars(buf, buf + sizeof buf - 1, f, std::chars_format::scientific);
    *tc.ptr = 0;
    // buf = [-]d.ddddde[+-]XX
    std::string sb(buf);
    bool neg = sb[0] == '-';
    if (neg) sb.erase(0, 1);
    size_t epos = sb.find('e');
    std::string mant = sb.substr(0, epos);
    int x = atoi(sb.c_str() + epos + 1);
    std::string digits;
    for (char c : mant)
        if (c != '.') digits.push_back(c);
    while (digits.size() > 1 && digits.back() == '0') digits.pop_back();
    std::string out = neg ? "-" : "";
    if (x < -4 || x >= 21 || (x >= 6 && true)) {
        if (x < -4 || x >= 6) {
            out += digits.substr(0, 1);
            if (digits.size() > 1) out += "." + digits.substr(1);
            char e[16];
            snprintf(e, sizeof e, "e%c%02d", x < 0 ? '-' : '+', std::abs(x));
            return out + e;
        }
    }
    int dp = x + 1;  // digits before the decimal point
    if (dp <= 0) {
        out += "0." + std::string((size_t)(-dp), '0') + digits;
    } else if ((size_t)dp >= digits.size()) {
        out += digits + std::string((size_t)dp - digits.size(), '0');
    } else {
        out += digits.substr(0, (size_t)dp) + "." + digits.substr((size_t)dp);
    }
    return out;
}

std::string go_print(const Value &v) {
    switch (v.kind) {
    case Value::NoValue: return "<no value>";
    case Value::Nil: return "<nil>";
    case Value::Bool: return v.b ? "true" : "false";
    case Value::Int: return std::to_string(v.i);
    case Value::Float: return format_float(v.f);
    case Value::String: return v.s;
    case Value::ListK: {
        std::string o = "[";
        bool first = true;
        for (const auto &x : *v.list) {
            if (!first) o += " ";
            first = false;
            o += go_print(x);
        }
        return o + "]";
    }
    case Value::MapK: {
        std::string o = "map[";
        bool first = true;
        for (const auto &kv : *v.map) {  // std::map iterates in key order, like fmt sorts map keys
            if (!first) o += " ";
            first = false;
            o += kv.first + ":" + go_print(kv.second);
        }
        return o + "]";
    }
    case Value::StructK: throw Error("cannot print a struct value of type " + v.s);
    }
    return "";
}

static std::string strval(const Value &v) { return v.kind == Value::String ? v.s : go_print(v); }

Value deep_copy(const Value &v) {
    Value o = v;
    if (v.list) {
        o.list = std::make_shared<List>();
        for (const auto &x : *v.list) o.list->push_back(deep_copy(x));
    }
    if (v.map) {
        o.map = std::make_shared<Map>();
        for (const auto &kv : *v.map) (*o.map)[kv.first] = deep_copy(kv.second);
    }
    return o;
}

// Decode one UTF-8 rune at s[i]; returns its width (0 if invalid) and the code point.
static int utf8_rune(const std::string &s, size_t i, uint32_t *cp) {
    unsigned char c = (unsigned char)s[i];
    int w = 0;
    uint32_t v = 0;
    if (c >= 0xC2 && c <= 0xDF) { w = 2; v = c & 0x1F; }
    else if (c >= 0xE0 && c <= 0xEF) { w = 3; v = c & 0x0F; }
    else if (c >= 0xF0 && c <= 0xF4) { w = 4; v = c & 0x07; }
    else return 0;
    if (i + (size_t)w > s.size()) return 0;
    for (int k = 1; k < w; ++k) {
        unsigned char cc = (unsigned char)s[i + k];
        if ((cc & 0xC0) != 0x80) return 0;
        v = (v << 6) | (cc & 0x3F);
    }
    if ((w == 3 && v < 0x800) || (w == 4 && (v < 0x10000 || v > 0x10FFFF)) ||
        (v >= 0xD800 && v <= 0xDFFF))
        return 0;
    *cp = v;
    return w;
}

std::string go_quote(const std::string &s) {
    std::string o = "\"";
    char buf[16];
    for (size_t i = 0; i < s.size();) {
        unsigned char c = (unsigned char)s[i];
        if (c < 0x80) {
            switch (c) {
            case 7: o += "\\a"; break;
            case 8: o += "\\b"; break;
            case 12: o += "\\f"; break;
            case 10: o += "\\n"; break;
            case 13: o += "\\r"; break;
            case 9: o += "\\t"; break;
            case 11: o += "\\v"; break;
            case '\\': o += "\\\\"; break;
            case '"': o += "\\\""; break;
            default:
                if (c < 0x20 || c == 0x7F) {
                    snprintf(buf, sizeof buf, "\\x%02x", c);
                    o += buf;
                } else {
                    o.push_back((char)c);
                }
            }
            ++i;
            continue;
        }
        uint32_t cp = 0;
        int w = utf8_rune(s, i, &cp);
        if (w == 0) {
            snprintf(buf, sizeof buf, "\\x%02x", c);
            o += buf;
            ++i;
            continue;
        }
        // printable approximation of unicode.IsPrint: C1 controls, a few format/separator
        // runes and non-characters are escaped, everything else is kept
        bool printable = !(cp < 0xA0 || cp == 0xAD || (cp >= 0x2028 && cp <= 0x202E) ||
                           (cp >= 0x2060 && cp <= 0x206F) || cp == 0xFEFF ||
                           (cp >= 0xFFF9 && cp <= 0xFFFB) || (cp & 0xFFFE) == 0xFFFE ||
                           (cp >= 0xE000 && cp <= 0xF8FF) || cp == 0x2000 || (cp > 0x2000 && cp <= 0x200F) ||
                           cp == 0x3000 || cp == 0x205F || cp == 0x1680);
        if (printable) {
            o.append(s, i, (size_t)w);
        } else if (cp < 0x10000) {
            snprintf(buf, sizeof buf, "\\u%04x", cp);
            o += buf;
        } else {
            snprintf(buf, sizeof buf, "\\U%08x", cp);
            o += buf;
        }
        i += (size_t)w;
    }
    return o + "\"";
}

// ============================================================================ AST
// Functions, resolved to an id when a template is parsed.
enum Fn : int {
    F_NONE = -1, F_include, F_default, F_trunc, F_trimSuffix, F_trimPrefix, F_replace, F_quote, F_squote, F_indent,
    F_nindent, F_b64enc, F_b64dec, F_eq, F_ne, F_lt, F_le, F_gt, F_ge, F_not, F_printf, F_print, F_println, F_len,
    F_empty, F_required, F_lower, F_upper, F_trim, F_toString, F_hasKey, F_ternary, F_cat, F_contains, F_hasPrefix,
    F_hasSuffix, F_repeat, F_and, F_or
};
static int fn_id(const std::string &name) {
    static const std::map<std::string, int> ids = {
        {"include", F_include}, {"default", F_default}, {"trunc", F_trunc}, {"trimSuffix", F_trimSuffix},
        {"trimPrefix", F_trimPrefix}, {"replace", F_replace}, {"quote", F_quote}, {"squote", F_squote},
        {"indent", F_indent}, {"nindent", F_nindent}, {"b64enc", F_b64enc}, {"b64dec", F_b64dec}, {"eq", F_eq},
        {"ne", F_ne}, {"lt", F_lt}, {"le", F_le}, {"gt", F_gt}, {"ge", F_ge}, {"not", F_not}, {"printf", F_printf},
        {"print", F_print}, {"println", F_println}, {"len", F_len}, {"empty", F_empty}, {"required", F_required},
        {"lower", F_lower}, {"upper", F_upper}, {"trim", F_trim}, {"toString", F_toString}, {"hasKey", F_hasKey},
        {"ternary", F_ternary}, {"cat", F_cat}, {"contains", F_contains}, {"hasPrefix", F_hasPrefix},
        {"hasSuffix", F_hasSuffix}, {"repeat", F_repeat}, {"and", F_and}, {"or", F_or}};
    auto it = ids.find(name);
    return it == ids.end() ? F_NONE : it->second;
}

struct Operand {
    enum Kind { Dot, Field, Var, Str, Num, BoolK, NilK, Ident, Paren } kind = Dot;
    std::string text;                 // Var name / Ident / Str payload
    int fn = F_NONE;                  // Ident: the function's id (F_NONE: unknown, an error when it runs)
    Value lit;                        // Str / Num / Bool / nil literal, built once
    Value num;                        // Num / Bool literal
    std::vector<std::string> fields;  // .A.B chain
    std::shared_ptr<struct Pipeline> paren;
};
struct Command {
    std::vector<Operand> args;
    size_t n_computed = 0;  // arguments that are parenthesised pipelines or bare function names
};
struct Pipeline {
    std::vector<Command> cmds;
    std::string decl;       // "$x" when the pipeline declares / assigns a variable
    bool is_assign = false;
    bool defer_candidate = false;   // `<two or more commands> | b64enc`, no variable: may stay pending
};
struct Var { const std::string *name; Value v; };
struct IncludeMemo { const NodeList *body; const Map *dot; std::string text; };
struct Node {
    enum Kind { Text, Action, If, With, Range, Template } kind = Text;
    std::string text;                 // Text payload / Template name
    std::shared_ptr<Pipeline> pipe;
    NodeList body, else_body;
    bool has_else = false;
    std::vector<std::string> range_vars;
};

// ============================================================================ lexer
namespace {

bool is_space(char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\n'; }

struct Item {
    bool is_action = false;
    std::string text;
};

std::vector<Item> lex(const std::string &src, const std::string &name) {
    std::vector<Item> items;
    size_t i = 0, n = src.size();
    bool trim_next = false;
    while (i < n) {
        size_t j = src.find("{{", i);
        if (j == std::string::npos) j = n;
        std::string text = src.substr(i, j - i);
        if (trim_next) {
            size_t k = 0;
            while (k < text.size() && is_space(text[k])) ++k;
            text.erase(0, k);
            trim_next = false;
        }
        if (j >= n) {
            if (!text.empty()) items.push_back({false, text});
            break;
        }
        size_t k = j + 2;
        bool ltrim = k + 1 < n && src[k] == '-' && is_space(src[k + 1]);
        if (ltrim) {
            while (!text.empty() && is_space(text.back())) text.pop_back();
            ++k;
            while (k < n && is_space(src[k])) ++k;
        }
        if (!text.empty()) items.push_back({false, text});
        if (src.compare(k, 2, "/*") == 0) {
            size_t e = src.find("*/", k + 2);
            if (e == std::string::npos) throw Error(name + ": unclosed comment");
            e += 2;
            if (src.compare(e, 2, "}}") == 0) {
                i = e + 2;
            } else if (e + 3 < n + 1 && e < n && is_space(src[e]) && src.compare(e + 1, 3, "-}}") == 0) {
                i = e + 4;
                trim_next = true;
            } else {
                throw Error(name + ": comment ends before closing delimiter");
            }
            continue;
        }
        size_t p = k;
        for (;;) {
            if (p >= n) throw Error(name + ": unclosed action");
            char c = src[p];
            if (c == '"') {
                ++p;
                while (p < n && src[p] != '"') {
                    if (src[p] == '\\') ++p;
                    if (p < n && src[p] == '\n') throw Error(name + ": unterminated quoted string");
                    ++p;
                }
                ++p;
                continue;
            }
            if (c == '`') {
                size_t q = src.find('`', p + 1);
                if (q == std::string::npos) throw Error(name + ": unterminated raw quoted string");
                p = q + 1;
                continue;
            }
            if (c == '\'') {
                ++p;
                while (p < n && src[p] != '\'') {
                    if (src[p] == '\\') ++p;
                    ++p;
                }
                ++p;
                continue;
            }
            if (c == '}' && p + 1 < n && src[p + 1] == '}') break;
            ++p;
        }
        std::string body = src.substr(k, p - k);
        if (body.size() >= 2 && body.back() == '-' && is_space(body[body.size() - 2])) {
            body.pop_back();
            trim_next = true;
        }
        size_t a = 0, b = body.size();
        while (a < b && is_space(body[a])) ++a;
        while (b > a && is_space(body[b - 1])) --b;
        items.push_back({true, body.substr(a, b - a)});
        i = p + 2;
    }
    return items;
}

// ---- tokens of one action
struct Tok {
    enum Kind { WS, Str, Raw, Char, Decl, Assign, Pipe, LP, RP, Comma, Num, Var, Field, Dot, Ident, End } kind;
    std::string text;
};

bool ident_start(char c) { return std::isalpha((unsigned char)c) || c == '_'; }
bool ident_char(char c) { return std::isalnum((unsigned char)c) || c == '_'; }

std::vector<Tok> tokenize(const std::string &s, const std::string &where) {
    std::vector<Tok> t;
    size_t i = 0, n = s.size();
    auto bad = [&](size_t at) -> Error {
        return Error(where + ": unexpected \"" + s.substr(at, 10) + "\" in action {{" + s + "}}");
    };
    while (i < n) {
        char c = s[i];
        if (is_space(c)) {
            while (i < n && is_space(s[i])) ++i;
            t.push_back({Tok::WS, " "});
        } else if (c == '"') {
            size_t j = i + 1;
            while (j < n && s[j] != '"') {
                if (s[j] == '\\') ++j;
                ++j;
            }
            if (j >= n) throw bad(i);
            t.push_back({Tok::Str, s.substr(i, j + 1 - i)});
            i = j + 1;
        } else if (c == '`') {
            size_t j = s.find('`', i + 1);
            if (j == std::string::npos) throw bad(i);
            t.push_back({Tok::Raw, s.substr(i, j + 1 - i)});
            i = j + 1;
        } else if (c == '\'') {
            size_t j = i + 1;
            while (j < n && s[j] != '\'') {
                if (s[j] == '\\') ++j;
                ++j;
            }
            if (j >= n) throw bad(i);
            t.push_back({Tok::Char, s.substr(i, j + 1 - i)});
            i = j + 1;
        } else if (c == ':' && i + 1 < n && s[i + 1] == '=') {
            t.push_back({Tok::Decl, ":="});
            i += 2;
        } else if (c == '=') {
            t.push_back({Tok::Assign, "="});
            ++i;
        } else if (c == '|') {
            t.push_back({Tok::Pipe, "|"});
            ++i;
        } else if (c == '(') {
            t.push_back({Tok::LP, "("});
            ++i;
        } else if (c == ')') {
            t.push_back({Tok::RP, ")"});
            ++i;
        } else if (c == ',') {
            t.push_back({Tok::Comma, ","});
            ++i;
        } else if (c == '$') {
            size_t j = i + 1;
            while (j < n && ident_char(s[j])) ++j;
            t.push_back({Tok::Var, s.substr(i, j - i)});
            i = j;
        } else if (c == '.' && i + 1 < n && ident_start(s[i + 1])) {
            size_t j = i;
            while (j < n && s[j] == '.' && j + 1 < n && ident_start(s[j + 1])) {
                ++j;
                while (j < n && ident_char(s[j])) ++j;
            }
            t.push_back({Tok::Field, s.substr(i, j - i)});
            i = j;
        } else if (std::isdigit((unsigned char)c) || ((c == '+' || c == '-') && i + 1 < n &&
                   (std::isdigit((unsigned char)s[i + 1]) || s[i + 1] == '.')) ||
                   (c == '.' && i + 1 < n && std::isdigit((unsigned char)s[i + 1]))) {
            size_t j = i + 1;
            while (j < n && (std::isalnum((unsigned char)s[j]) || s[j] == '.' || s[j] == '_' ||
                             ((s[j] == '+' || s[j] == '-') && (s[j - 1] == 'e' || s[j - 1] == 'E'))))
                ++j;
            t.push_back({Tok::Num, s.substr(i, j - i)});
            i = j;
        } else if (c == '.') {
            t.push_back({Tok::Dot, "."});
            ++i;
        } else if (ident_start(c)) {
            size_t j = i;
            while (j < n && ident_char(s[j])) ++j;
            t.push_back({Tok::Ident, s.substr(i, j - i)});
            i = j;
        } else {
            throw bad(i);
        }
    }
    return t;
}

void append_utf8(std::string &o, uint32_t cp) {
    if (cp < 0x80) o.push_back((char)cp);
    else if (cp < 0x800) { o.push_back((char)(0xC0 | (cp >> 6))); o.push_back((char)(0x80 | (cp & 0x3F))); }
    else if (cp < 0x10000) {
        o.push_back((char)(0xE0 | (cp >> 12))); o.push_back((char)(0x80 | ((cp >> 6) & 0x3F)));
        o.push_back((char)(0x80 | (cp & 0x3F)));
    } else {
        o.push_back((char)(0xF0 | (cp >> 18))); o.push_back((char)(0x80 | ((cp >> 12) & 0x3F)));
        o.push_back((char)(0x80 | ((cp >> 6) & 0x3F))); o.push_back((char)(0x80 | (cp & 0x3F)));
    }
}

// Go interpreted string / char literal body -> bytes
std::string unquote(const std::string &tok) {
    std::string body = tok.substr(1, tok.size() - 2), o;
    for (size_t i = 0; i < body.size();) {
        char c = body[i];
        if (c != '\\') { o.push_back(c); ++i; continue; }
        ++i;
        if (i >= body.size()) throw Error("bad escape in string literal");
        c = body[i];
        switch (c) {
        case 'a': o.push_back(7); ++i; break;
        case 'b': o.push_back(8); ++i; break;
        case 'f': o.push_back(12); ++i; break;
        case 'n': o.push_back(10); ++i; break;
        case 'r': o.push_back(13); ++i; break;
        case 't': o.push_back(9); ++i; break;
        case 'v': o.push_back(11); ++i; break;
        case '\\': o.push_back('\\'); ++i; break;
        case '"': o.push_back('"'); ++i; break;
        case '\'': o.push_back('\''); ++i; break;
        case 'x': o.push_back((char)strtol(body.substr(i + 1, 2).c_str(), nullptr, 16)); i += 3; break;
        case 'u': append_utf8(o, (uint32_t)strtol(body.substr(i + 1, 4).c_str(), nullptr, 16)); i += 5; break;
        case 'U': append_utf8(o, (uint32_t)strtol(body.substr(i + 1, 8).c_str(), nullptr, 16)); i += 9; break;
        default:
            if (c >= '0' && c <= '7') {
                o.push_back((char)strtol(body.substr(i, 3).c_str(), nullptr, 8));
                i += 3;
            } else {
                throw Error(std::string("unknown escape \\") + c);
            }
        }
    }
    return o;
}

std::vector<std::string> split_fields(const std::string &f) {  // ".A.B" -> {A, B}
    std::vector<std::string> out;
    size_t i = 1;
    while (i <= f.size()) {
        size_t j = f.find('.', i);
        if (j == std::string::npos) j = f.size();
        out.push_back(f.substr(i, j - i));
        i = j + 1;
    }
    return out;
}

struct PipeParser {
    std::vector<Tok> toks;
    size_t i = 0;
    std::string where;

    Tok peek(bool skip_ws = true) const {
        size_t j = i;
        while (skip_ws && j < toks.size() && toks[j].kind == Tok::WS) ++j;
        return j < toks.size() ? toks[j] : Tok{Tok::End, ""};
    }
    Tok next(bool skip_ws = true) {
        while (skip_ws && i < toks.size() && toks[i].kind == Tok::WS) ++i;
        if (i >= toks.size()) return Tok{Tok::End, ""};
        return toks[i++];
    }
    std::shared_ptr<Pipeline> pipeline(Tok::Kind end) {
        auto p = std::make_shared<Pipeline>();
        size_t save = i;
        Tok t = peek();
        if (t.kind == Tok::Var) {
            next();
            Tok t2 = peek();
            if (t2.kind == Tok::Decl || t2.kind == Tok::Assign) {
                next();
                p->decl = t.text;
                p->is_assign = t2.kind == Tok::Assign;
            } else {
                i = save;
            }
        }
        p->cmds.push_back(command(end));
        while (peek().kind == Tok::Pipe) {
            next();
            p->cmds.push_back(command(end));
        }
        if (peek().kind != end) throw Error(where + ": unexpected \"" + peek().text + "\" in pipeline");
        const Command &last = p->cmds.back();
        p->defer_candidate = p->decl.empty() && p->cmds.size() >= 2 && last.args.size() == 1 &&
                             last.args[0].kind == Operand::Ident && last.args[0].fn == F_b64enc;
        return p;
    }
    Command command(Tok::Kind end) {
        Command c;
        for (;;) {
            Tok t = peek();
            if (t.kind == Tok::Pipe || t.kind == end || t.kind == Tok::End || t.kind == Tok::RP) break;
            c.args.push_back(operand());
        }
        if (c.args.empty()) throw Error(where + ": missing value for command");
        for (size_t k = 1; k < c.args.size(); ++k)
            if (c.args[k].kind == Operand::Paren || c.args[k].kind == Operand::Ident) ++c.n_computed;
        return c;
    }
    Operand operand() {
        Tok t = next();
        Operand op;
        switch (t.kind) {
        case Tok::Dot: op.kind = Operand::Dot; break;
        case Tok::Field: op.kind = Operand::Field; op.fields = split_fields(t.text); break;
        case Tok::Var: {
            op.kind = Operand::Var;
            op.text = t.text;
            Tok nk = peek(false);
            if (nk.kind == Tok::Field) { next(false); op.fields = split_fields(nk.text); }
            break;
        }
        case Tok::Str: op.kind = Operand::Str; op.text = unquote(t.text); break;
        case Tok::Raw: op.kind = Operand::Str; op.text = t.text.substr(1, t.text.size() - 2); break;
        case Tok::Char: {
            std::string u = unquote(t.text);
            uint32_t cp = (unsigned char)u[0];
            if (cp >= 0x80) utf8_rune(u, 0, &cp);
            op.kind = Operand::Num;
            op.num = Value::integer((int64_t)cp);
            break;
        }
        case Tok::Num: {
            std::string txt;
            for (char c : t.text) if (c != '_') txt.push_back(c);
            op.kind = Operand::Num;
            char *endp = nullptr;
            errno = 0;
            long long iv = strtoll(txt.c_str(), &endp, 0);
            if (endp && *endp == 0 && errno == 0) {
                op.num = Value::integer(iv);
            } else {
                double dv = strtod(txt.c_str(), &endp);
                if (!endp || *endp != 0) throw Error(where + ": bad number syntax: \"" + t.text + "\"");
                op.num = Value::number(dv);
            }
            break;
        }
        case Tok::Ident:
            if (t.text == "true" || t.text == "false") { op.kind = Operand::BoolK; op.num = Value::boolean(t.text == "true"); }
            else if (t.text == "nil") op.kind = Operand::NilK;
            else { op.kind = Operand::Ident; op.text = t.text; }
            break;
        case Tok::LP: {
            op.kind = Operand::Paren;
            op.paren = pipeline(Tok::RP);
            next();
            Tok nk = peek(false);
            if (nk.kind == Tok::Field) { next(false); op.fields = split_fields(nk.text); }
            break;
        }
        default:
            throw Error(where + ": unexpected \"" + t.text + "\" in operand");
        }
        if (op.kind == Operand::Str) op.lit = Value::str(op.text);
        else if (op.kind == Operand::Num || op.kind == Operand::BoolK) op.lit = op.num;
        else if (op.kind == Operand::Ident) op.fn = fn_id(op.text);
        return op;
    }
};

std::shared_ptr<Pipeline> parse_pipe(const std::string &text, const std::string &where) {
    PipeParser pp{tokenize(text, where), 0, where};
    return pp.pipeline(Tok::End);
}

bool keyword(const std::string &src, std::string *kw, std::string *rest) {
    static const char *kws[] = {"if", "else", "end", "range", "with", "define", "template", "block", "break", "continue"};
    for (const char *k : kws) {
        size_t L = strlen(k);
        if (src.compare(0, L, k) == 0 && (src.size() == L || !ident_char(src[L]))) {
            *kw = k;
            size_t a = L;
            while (a < src.size() && is_space(src[a])) ++a;
            *rest = src.substr(a);
            return true;
        }
    }
    return false;
}

struct Parser {
    std::vector<Item> items;
    size_t pos = 0;
    std::string name;
    std::map<std::string, NodeList> *defs;

    // terminators: bit0 = end, bit1 = else. Returns which one ended the list ("" at EOF).
    NodeList list(int terminators, std::string *term, std::string *term_rest) {
        NodeList nodes;
        while (pos < items.size()) {
            Item it = items[pos++];
            if (!it.is_action) {
                auto nd = std::make_shared<Node>();
                nd->kind = Node::Text;
                nd->text = it.text;
                nodes.push_back(nd);
                continue;
            }
            std::string kw, rest;
            if (!keyword(it.text, &kw, &rest)) {
                if (it.text.empty()) throw Error(name + ": missing value for command");
                auto nd = std::make_shared<Node>();
                nd->kind = Node::Action;
                nd->pipe = parse_pipe(it.text, name);
                nodes.push_back(nd);
                continue;
            }
            if (kw == "end" || kw == "else") {
                int bit = kw == "end" ? 1 : 2;
                if (!(terminators & bit)) throw Error(name + ": unexpected {{" + kw + "}}");
                *term = kw;
                *term_rest = rest;
                return nodes;
            }
            if (kw == "if" || kw == "with") {
                nodes.push_back(cond(kw, rest));
            } else if (kw == "range") {
                nodes.push_back(range(rest));
            } else if (kw == "define") {
                std::vector<Tok> t = tokenize(rest, name);
                if (t.empty() || (t[0].kind != Tok::Str && t[0].kind != Tok::Raw))
                    throw Error(name + ": define needs a quoted name");
                std::string tname = t[0].kind == Tok::Str ? unquote(t[0].text) : t[0].text.substr(1, t[0].text.size() - 2);
                std::string tm, tr;
                NodeList body = list(1, &tm, &tr);
                if (tm != "end") throw Error(name + ": unexpected EOF in define");
                (*defs)[tname] = body;
            } else if (kw == "template") {
                PipeParser pp{tokenize(rest, name), 0, name};
                Tok t = pp.next();
                if (t.kind != Tok::Str && t.kind != Tok::Raw) throw Error(name + ": template name must be a string");
                auto nd = std::make_shared<Node>();
                nd->kind = Node::Template;
                nd->text = t.kind == Tok::Str ? unquote(t.text) : t.text.substr(1, t.text.size() - 2);
                if (pp.peek().kind != Tok::End) nd->pipe = pp.pipeline(Tok::End);
                nodes.push_back(nd);
            } else {
                throw Error(name + ": {{" + kw + "}} is not supported");
            }
        }
        if (terminators) throw Error(name + ": unexpected EOF");
        term->clear();
        return nodes;
    }

    std::shared_ptr<Node> cond(const std::string &kw, const std::string &rest) {
        auto nd = std::make_shared<Node>();
        nd->kind = kw == "if" ? Node::If : Node::With;
        nd->pipe = parse_pipe(rest, name);
        std::string term, trest;
        nd->body = list(3, &term, &trest);
        if (term == "else") {
            nd->has_else = true;
            std::string kw2, rest2;
            if (!trest.empty() && keyword(trest, &kw2, &rest2) && (kw2 == "if" || kw2 == "with")) {
                nd->else_body.push_back(cond(kw2, rest2));  // shares the single {{end}}
            } else {
                std::string t2, r2;
                nd->else_body = list(1, &t2, &r2);
            }
        }
        return nd;
    }

    std::shared_ptr<Node> range(std::string rest) {
        auto nd = std::make_shared<Node>();
        nd->kind = Node::Range;
        // "$i, $v := pipeline" / "$v := pipeline"
        size_t d = rest.find(":=");
        if (!rest.empty() && rest[0] == '$' && d != std::string::npos) {
            std::string vars = rest.substr(0, d);
            size_t a = 0;
            while (a < vars.size()) {
                while (a < vars.size() && (is_space(vars[a]) || vars[a] == ',')) ++a;
                size_t b = a;
                while (b < vars.size() && !is_space(vars[b]) && vars[b] != ',') ++b;
                if (b > a) nd->range_vars.push_back(vars.substr(a, b - a));
                a = b;
            }
            rest = rest.substr(d + 2);
        }
        nd->pipe = parse_pipe(rest, name);
        std::string term, trest;
        nd->body = list(3, &term, &trest);
        if (term == "else") {
            nd->has_else = true;
            std::string t2, r2;
            nd->else_body = list(1, &t2, &r2);
        }
        return nd;
    }
};

}  // namespace

// ============================================================================ functions
namespace {

int64_t to_int(const Value &v) {
    switch (v.kind) {
    case Value::Bool: return v.b;
    case Value::Int: return v.i;
    case Value::Float: return (int64_t)v.f;
    case Value::String: return strtoll(v.s.c_str(), nullptr, 10);
    default: return 0;
    }
}

std::string replace_all(const std::string &src, const std::string &old_, const std::string &new_) {
    if (old_.empty()) {
        // strings.Replace with empty old: insert `new` before every rune and at the end
        std::string o;
        for (size_t i = 0; i < src.size();) {
            o += new_;
            uint32_t cp;
            int w = (unsigned char)src[i] < 0x80 ? 1 : utf8_rune(src, i, &cp);
            if (w == 0) w = 1;
            o.append(src, i, (size_t)w);
            i += (size_t)w;
        }
        return o + new_;
    }
    std::string o;
    size_t i = 0;
    for (;;) {
        size_t j = src.find(old_, i);
        if (j == std::string::npos) break;
        o.append(src, i, j - i);
        o += new_;
        i = j + old_.size();
    }
    o.append(src, i, std::string::npos);
    return o;
}

const char *basic_kind(const Value &v) {
    switch (v.kind) {
    case Value::Bool: return "bool";
    case Value::Int: return "int";
    case Value::Float: return "float";
    case Value::String: return "string";
    case Value::Nil: case Value::NoValue: return "nil";
    default: return "other";
    }
}

// A function's arguments: pointers at values that live elsewhere (the value tree, literals of
// the AST, temporaries of the caller).
struct ArgView {
    const Value *const *p;
    size_t n;
    const Value &operator[](size_t i) const { return *p[i]; }
    size_t size() const { return n; }
    bool empty() const { return n == 0; }
};

bool eq_fn(const ArgView &a) {
    if (a.size() < 2) throw Error("missing argument for comparison");
    std::string ka = basic_kind(a[0]);
    for (size_t k = 1; k < a.size(); ++k) {
        std::string kb = basic_kind(a[k]);
        if (ka == "nil" || kb == "nil") {
            if (ka == kb) return true;
            continue;
        }
        if (ka != kb) throw Error("incompatible types for comparison");
        if (ka == "other") throw Error("invalid type for comparison");
        bool same = false;
        if (ka == "bool") same = a[0].b == a[k].b;
        else if (ka == "int") same = a[0].i == a[k].i;
        else if (ka == "float") same = a[0].f == a[k].f;
        else same = a[0].s == a[k].s;
        if (same) return true;
    }
    return false;
}

bool cmp_fn(int op, const Value &a, const Value &b) {
    std::string ka = basic_kind(a), kb = basic_kind(b);
    if (ka != kb) throw Error("incompatible types for comparison");
    int c;
    if (ka == "int") c = a.i < b.i ? -1 : a.i > b.i;
    else if (ka == "float") c = a.f < b.f ? -1 : a.f > b.f;
    else if (ka == "string") c = a.s.compare(b.s) < 0 ? -1 : a.s.compare(b.s) > 0;
    else throw Error("invalid type for comparison");
    if (op == F_lt) return c < 0;
    if (op == F_le) return c <= 0;
    if (op == F_gt) return c > 0;
    return c >= 0;
}

std::string go_typename(const Value &v) {
    std::string k = basic_kind(v);
    if (k == "float") return "float64";
    if (k == "nil") return "<nil>";
    if (k == "other") return "interface {}";
    return k;
}

// An argument bound to a Go `string` parameter (Sprig's b64enc, trunc, indent, ... are plain
// func(string) string): text/template refuses anything else instead of printing it --
// "wrong type for value; expected string; got int64", and for nil / a missing key
// "invalid value; expected string" -- so `--set azIotEdgeConfig=12345` aborts the render
// rather than encoding "12345" (aziot-edge-runtime-config-secret.yaml:6).
const std::string &string_arg(const Value &v, const std::string &fn) {
    if (v.kind == Value::String) return v.s;
    if (v.is_nilish()) throw Error("executing at <" + fn + ">: invalid value; expected string");
    throw Error("executing at <" + fn + ">: wrong type for value; expected string; got " + go_typename(v));
}

std::string sprintf_fn(const ArgView &a) {
    if (a.empty()) throw Error("wrong number of args for printf: want at least 1 got 0");
    std::string fmt = strval(a[0]), o;
    size_t ai = 1;
    for (size_t i = 0; i < fmt.size();) {
        char c = fmt[i];
        if (c != '%') { o.push_back(c); ++i; continue; }
        ++i;
        if (i >= fmt.size()) { o += "%!(NOVERB)"; break; }
        char v = fmt[i++];
        if (v == '%') { o.push_back('%'); continue; }
        if (ai >= a.size()) { o += std::string("%!") + v + "(MISSING)"; continue; }
        const Value &x = a[ai++];
        if (v == 's' || v == 'v') o += go_print(x);
        else if (v == 'd') {
            if (x.kind == Value::Int) o += std::to_string(x.i);
            else o += "%!d(" + go_typename(x) + "=" + go_print(x) + ")";
        } else if (v == 'q') o += go_quote(strval(x));
        else if (v == 't') {
            if (x.kind == Value::Bool) o += go_print(x);
            else o += "%!t(" + go_typename(x) + "=" + go_print(x) + ")";
        } else throw Error(std::string("printf verb %") + v + " is not supported");
    }
    if (ai < a.size()) {
        o += "%!(EXTRA ";
        for (size_t k = ai; k < a.size(); ++k) {
            if (k > ai) o += ", ";
            o += go_typename(a[k]) + "=" + go_print(a[k]);
        }
        o += ")";
    }
    return o;
}

std::string ascii_map(const std::string &s, int (*f)(int)) {
    std::string o = s;
    for (auto &c : o)
        if ((unsigned char)c < 0x80) c = (char)f((unsigned char)c);
    return o;
}

}  // namespace

// ============================================================================ executor
// One render = thousands of these calls across a fleet, so the executor is written to touch the
// allocator as little as it can: operands that name a value (.Values.x, $v, literals) are passed
// around as POINTERS into the value tree, never copied; functions are dispatched on an id
// resolved at parse time; arguments sit in a small array on the stack; variables live in one flat
// stack; `include` captures into a buffer whose capacity is reused.
struct Exec {
    const TemplateSet &ts;
    const TemplateSet::ExecOptions &opt;
    int capture_depth = 0;  // > 0 while executing inside include (output is captured as a value)

    // live variables, innermost last.  A deque: references to its elements (a `$x` handed to a
    // function, the dot of a `with $x`) stay valid while later variables come and go.
    std::deque<Var> &vars;
    std::vector<std::string> &capture_pool;   // spare buffers for include
    std::vector<IncludeMemo> *memo;           // this release's memoised includes, or null
    size_t frame_base = 0;      // variables below this belong to the caller of include/template

    static const std::string &dollar() { static const std::string d = "$"; return d; }
    static const Value &no_value() { static const Value v = Value::novalue(); return v; }

    const Value *lookup_var(const std::string &name) const {
        for (size_t k = vars.size(); k-- > frame_base;)
            if (*vars[k].name == name) return &vars[k].v;
        throw Error("undefined variable \"" + name + "\"");
    }

    const Value *field_ref(const Value *v, const std::vector<std::string> &fields) const {
        for (const auto &f : fields) {
            if (v->kind == Value::StructK) {
                auto it = v->map->find(f);
                if (it == v->map->end()) throw Error("can't evaluate field " + f + " in type " + v->s);
                v = &it->second;
            } else if (v->kind == Value::MapK) {
                auto it = v->map->find(f);
                v = it == v->map->end() ? &no_value() : &it->second;
            } else if (v->is_nilish()) {
                throw Error("nil pointer evaluating interface {}." + f);
            } else {
                throw Error("can't evaluate field " + f + " in type " + go_typename(*v));
            }
        }
        return v;
    }

    std::string encode_now(const std::string &s) {
        if (s.empty()) return "";
        if (!opt.encode) throw Error("b64enc: no encoder attached (GPU required)");
        return opt.encode(s);
    }

    // Run a named template with `data` as dot and `$`, appending to out.
    void run_template(const std::string &tname, const Value &data, std::string &out) {
        auto it = ts.templates_.find(tname);
        if (it == ts.templates_.end()) throw Error("template \"" + tname + "\" not defined");
        const size_t saved_base = frame_base, saved_size = vars.size();
        frame_base = saved_size;
        vars.push_back({&dollar(), data});
        try {
            exec_list(it->second, vars[saved_size].v, out);
        } catch (...) {
            vars.resize(saved_size);
            frame_base = saved_base;
            throw;
        }
        vars.resize(saved_size);
        frame_base = saved_base;
    }

    Value call(int fn, const std::string &fname, const ArgView &a) {
        auto need = [&](size_t n) {
            if (a.size() != n)
                throw Error("wrong number of args for " + fname + ": want " + std::to_string(n) + " got " + std::to_string(a.size()));
        };
        auto empty = [](const Value &v) { return !truth(v); };
        switch (fn) {
        case F_include: {
            need(2);
            // same template, same (immutable) dot, no way to see .Template: same text as last time
            const NodeList *body = nullptr;
            if (memo && a[0].kind == Value::String && a[1].kind == Value::MapK) {
                auto it = ts.templates_.find(a[0].s);
                if (it != ts.templates_.end()) {
                    auto ok = ts.memo_ok_.find(&it->second);
                    if (ok != ts.memo_ok_.end() && ok->second) {
                        body = &it->second;
                        for (const auto &m : *memo)
                            if (m.body == body && m.dot == a[1].map.get()) return Value::str(m.text);
                    }
                }
            }
            std::string buf;
            if (!capture_pool.empty()) { buf = std::move(capture_pool.back()); capture_pool.pop_back(); buf.clear(); }
            ++capture_depth;
            try {
                if (a[0].kind == Value::String) run_template(a[0].s, a[1], buf);
                else run_template(strval(a[0]), a[1], buf);
            } catch (...) {
                --capture_depth;
                throw;
            }
            --capture_depth;
            if (body) memo->push_back({body, a[1].map.get(), buf});
            return Value::str(std::move(buf));
        }
        case F_default:
            if (a.empty()) throw Error("wrong number of args for default");
            if (a.size() < 2 || empty(a[1])) return a[0];
            return a[1];
        case F_trunc: {
            need(2);
            int64_t c = to_int(a[0]);
            const std::string &s = string_arg(a[1], fname);
            if (c < 0 && (int64_t)s.size() + c > 0) return Value::str(s.substr((size_t)((int64_t)s.size() + c)));
            if (c >= 0 && (int64_t)s.size() > c) return Value::str(s.substr(0, (size_t)c));
            return Value::str(s);
        }
        case F_trimSuffix: {
            need(2);
            const std::string &suf = string_arg(a[0], fname);
            std::string s = string_arg(a[1], fname);
            if (!suf.empty() && s.size() >= suf.size() && s.compare(s.size() - suf.size(), suf.size(), suf) == 0)
                s.erase(s.size() - suf.size());
            return Value::str(std::move(s));
        }
        case F_trimPrefix: {
            need(2);
            const std::string &pre = string_arg(a[0], fname);
            std::string s = string_arg(a[1], fname);
            if (!pre.empty() && s.compare(0, pre.size(), pre) == 0) s.erase(0, pre.size());
            return Value::str(std::move(s));
        }
        case F_replace:
            need(3);
            return Value::str(replace_all(string_arg(a[2], fname), string_arg(a[0], fname), string_arg(a[1], fname)));
        case F_quote: case F_squote: {
            std::string o;
            bool first = true;
            for (size_t k = 0; k < a.size(); ++k) {
                const Value &x = a[k];
                if (x.is_nilish()) continue;
                if (!first) o += " ";
                first = false;
                o += fn == F_quote ? go_quote(strval(x)) : "'" + go_print(x) + "'";
            }
            return Value::str(std::move(o));
        }
        case F_indent: case F_nindent: {
            need(2);
            const std::string pad((size_t)std::max<int64_t>(0, to_int(a[0])), ' ');
            const std::string &s = string_arg(a[1], fname);
            std::string o;
            o.reserve(s.size() + pad.size() * 8 + 1);
            if (fn == F_nindent) o.push_back('\n');
            o += pad;
            for (char ch : s) {
                o.push_back(ch);
                if (ch == '\n') o += pad;
            }
            return Value::str(std::move(o));
        }
        case F_b64enc: need(1); return Value::str(encode_now(string_arg(a[0], fname)));
        case F_b64dec:
            need(1);
            if (!opt.decode) throw Error("b64dec: no decoder attached (GPU required)");
            return Value::str(opt.decode(string_arg(a[0], fname)));
        case F_eq: return Value::boolean(eq_fn(a));
        case F_ne: need(2); return Value::boolean(!eq_fn(a));
        case F_lt: case F_le: case F_gt: case F_ge: need(2); return Value::boolean(cmp_fn(fn, a[0], a[1]));
        case F_not: need(1); return Value::boolean(!truth(a[0]));
        case F_printf: return Value::str(sprintf_fn(a));
        case F_print: case F_println: {
            std::string o;
            if (fn == F_println) {
                for (size_t k = 0; k < a.size(); ++k) { if (k) o += " "; o += go_print(a[k]); }
                return Value::str(o + "\n");
            }
            bool prev_str = true;
            for (size_t k = 0; k < a.size(); ++k) {
                bool is_s = a[k].kind == Value::String;
                if (k > 0 && !is_s && !prev_str) o += " ";
                o += go_print(a[k]);
                prev_str = is_s;
            }
            return Value::str(std::move(o));
        }
        case F_len:
            need(1);
            if (a[0].kind == Value::String) return Value::integer((int64_t)a[0].s.size());
            if (a[0].kind == Value::ListK) return Value::integer((int64_t)a[0].list->size());
            if (a[0].kind == Value::MapK) return Value::integer((int64_t)a[0].map->size());
            throw Error("len of type " + go_typename(a[0]));
        case F_empty: need(1); return Value::boolean(empty(a[0]));
        case F_required:
            need(2);
            if (a[1].is_nilish() || (a[1].kind == Value::String && a[1].s.empty())) throw Error(strval(a[0]));
            return a[1];
        case F_lower: need(1); return Value::str(ascii_map(string_arg(a[0], fname), ::tolower));
        case F_upper: need(1); return Value::str(ascii_map(string_arg(a[0], fname), ::toupper));
        case F_trim: {
            need(1);
            const std::string &s = string_arg(a[0], fname);
            size_t x = 0, y = s.size();
            while (x < y && (is_space(s[x]) || s[x] == '\v' || s[x] == '\f')) ++x;
            while (y > x && (is_space(s[y - 1]) || s[y - 1] == '\v' || s[y - 1] == '\f')) --y;
            return Value::str(s.substr(x, y - x));
        }
        case F_toString: need(1); return Value::str(strval(a[0]));
        case F_hasKey:
            need(2);
            return Value::boolean(a[0].kind == Value::MapK && a[0].map->count(strval(a[1])) != 0);
        case F_ternary: need(3); return truth(a[2]) ? a[0] : a[1];
        case F_cat: {
            std::string o;
            bool first = true;
            for (size_t k = 0; k < a.size(); ++k) {
                if (a[k].is_nilish()) continue;
                if (!first) o += " ";
                first = false;
                o += strval(a[k]);
            }
            return Value::str(std::move(o));
        }
        case F_contains: need(2); return Value::boolean(string_arg(a[1], fname).find(string_arg(a[0], fname)) != std::string::npos);
        case F_hasPrefix: {
            need(2);
            const std::string &p = string_arg(a[0], fname), &s = string_arg(a[1], fname);
            return Value::boolean(s.compare(0, p.size(), p) == 0);
        }
        case F_hasSuffix: {
            need(2);
            const std::string &p = string_arg(a[0], fname), &s = string_arg(a[1], fname);
            return Value::boolean(s.size() >= p.size() && s.compare(s.size() - p.size(), p.size(), p) == 0);
        }
        case F_repeat: {
            need(2);
            std::string o;
            const std::string &s = string_arg(a[1], fname);
            for (int64_t k = 0; k < to_int(a[0]); ++k) o += s;
            return Value::str(std::move(o));
        }
        default: break;
        }
        throw Error("function \"" + fname + "\" not defined");
    }

    // The value an operand names: a pointer into the value tree or at a literal of the AST, or,
    // for a computed operand, at `tmp`, which then holds the result.
    const Value *eval_operand(const Operand &op, const Value &dot, Value &tmp) {
        switch (op.kind) {
        case Operand::Dot: return &dot;
        case Operand::Field: return field_ref(&dot, op.fields);
        case Operand::Var: return field_ref(lookup_var(op.text), op.fields);
        case Operand::Str: case Operand::Num: case Operand::BoolK: case Operand::NilK: return &op.lit;
        case Operand::Paren:
            tmp = eval_pipe_value(*op.paren, dot);
            return field_ref(&tmp, op.fields);
        case Operand::Ident:
            tmp = call(op.fn, op.text, ArgView{nullptr, 0});
            return &tmp;
        }
        throw Error("bad operand");
    }

    static constexpr size_t kMaxArgs = 16;

    // One command of a pipeline; `piped` is the previous command's value (the last argument).
    // Returns where the result is: in the tree, in the AST, or in `result`.
    const Value *eval_cmd(const Command &cmd, const Value &dot, const Value *piped, Value &result) {
        const Operand &first = cmd.args[0];
        if (first.kind != Operand::Ident) {
            if (cmd.args.size() > 1 || piped) throw Error("can't give argument to non-function");
            return eval_operand(first, dot, result);
        }
        const size_t n_args = cmd.args.size() - 1 + (piped ? 1 : 0);
        if (first.fn == F_and || first.fn == F_or) {
            if (n_args == 0) throw Error("wrong number of args for " + first.text);
            const Value *res = nullptr;
            Value tmp;
            for (size_t k = 0; k < n_args; ++k) {
                res = k + 1 < cmd.args.size() ? eval_operand(cmd.args[k + 1], dot, tmp) : piped;
                if ((first.fn == F_and) != truth(*res)) break;
            }
            if (res == &tmp) { result = std::move(tmp); return &result; }
            return res;
        }
        if (n_args > kMaxArgs) throw Error("too many arguments to " + first.text);
        const Value *argv[kMaxArgs];
        // computed operands (parenthesised pipelines, bare function names) need a home; they are rare
        std::unique_ptr<Value[]> tmps(cmd.n_computed ? new Value[cmd.n_computed] : nullptr);
        size_t t = 0, n = 0;
        for (size_t k = 1; k < cmd.args.size(); ++k) {
            const Operand &op = cmd.args[k];
            const bool computed = op.kind == Operand::Paren || op.kind == Operand::Ident;
            argv[n++] = computed ? eval_operand(op, dot, tmps[t++]) : eval_operand(op, dot, result);
        }
        if (piped) argv[n++] = piped;
        result = call(first.fn, first.text, ArgView{argv, n});
        return &result;
    }

    // The first `upto` commands of a pipeline.  h0/h1 are two holders the commands' results
    // alternate between (a command reads the previous holder while writing the other).
    const Value *eval_pipe_prefix(const Pipeline &p, size_t upto, const Value &dot, Value &h0, Value &h1) {
        const Value *v = nullptr;
        for (size_t k = 0; k < upto; ++k) v = eval_cmd(p.cmds[k], dot, v, (k & 1) ? h1 : h0);
        return v;
    }

    void set_var(const Pipeline &p, const Value &v) {
        if (p.is_assign) {
            for (size_t k = vars.size(); k-- > frame_base;)
                if (*vars[k].name == p.decl) { vars[k].v = v; return; }
            throw Error("undefined variable \"" + p.decl + "\"");
        }
        vars.push_back({&p.decl, v});
    }

    const Value *eval_pipe(const Pipeline &p, const Value &dot, Value &h0, Value &h1) {
        const Value *v = eval_pipe_prefix(p, p.cmds.size(), dot, h0, h1);
        if (!p.decl.empty()) set_var(p, *v);
        return v;
    }

    Value eval_pipe_value(const Pipeline &p, const Value &dot) {
        Value h0, h1;
        const Value *v = eval_pipe(p, dot, h0, h1);
        if (v == &h0) return std::move(h0);
        if (v == &h1) return std::move(h1);
        return *v;
    }

    // `{{ <pipeline> | b64enc }}` printed straight into a top-level file: leave it pending.
    bool try_defer(const Pipeline &p, const Value &dot, std::string &out) {
        if (!p.defer_candidate || !opt.defer_b64 || !opt.pending || capture_depth != 0) return false;
        if (opt.pending->size() >= TemplateSet::kMaxPending) return false;
        Value h0, h1;
        const Value *in = eval_pipe_prefix(p, p.cmds.size() - 1, dot, h0, h1);
        string_arg(*in, "b64enc");               // same refusal as the immediate path
        if (in->s.empty()) return true;          // b64enc "" == ""
        Pending pd;
        pd.len = in->s.size();
        pd.encoded_len = (pd.len + 2) / 3 * 4;
        // a computed input (an include's text) is ours to take; a field of a map of the value
        // tree stays where it is when the caller keeps the tree; anything else is copied
        const Operand &src = p.cmds[0].args[0];
        const bool in_tree = opt.values_outlive_pending && p.cmds.size() == 2 && p.cmds[0].args.size() == 1 &&
                             !src.fields.empty() && (src.kind == Operand::Field || src.kind == Operand::Var);
        if (in == &h0) pd.owned = std::move(h0.s);
        else if (in == &h1) pd.owned = std::move(h1.s);
        else if (in_tree) pd.ext = in->s.data();
        else pd.owned = in->s;
        // the placeholder (TemplateSet::placeholder), written in place
        const size_t id = opt.pending->size();
        for (int k = 0; k < 4; ++k) out.push_back((char)(0x10 + ((id >> (12 - 4 * k)) & 0xF)));
        out.push_back('\x01');
        opt.pending->push_back(std::move(pd));
        return true;
    }

    static void print_value(std::string &out, const Value &v) {
        if (v.kind == Value::String) out += v.s;
        else out += go_print(v);
    }

    void recycle(Value &h) {   // an include's captured text that has been printed: keep its buffer
        if (h.kind == Value::String && h.s.capacity() >= 256 && capture_pool.size() < 8) capture_pool.push_back(std::move(h.s));
    }

    void exec_list(const NodeList &nodes, const Value &dot, std::string &out) {
        for (const auto &nd : nodes) {
            switch (nd->kind) {
            case Node::Text: out += nd->text; break;
            case Node::Action: {
                if (try_defer(*nd->pipe, dot, out)) break;
                Value h0, h1;
                const Value *v = eval_pipe(*nd->pipe, dot, h0, h1);
                if (nd->pipe->decl.empty()) print_value(out, *v);
                recycle(h0);
                recycle(h1);
                break;
            }
            case Node::If: case Node::With: {
                const size_t mark = vars.size();
                Value h0, h1;
                const Value *v = eval_pipe(*nd->pipe, dot, h0, h1);
                if (truth(*v)) {
                    exec_list(nd->body, nd->kind == Node::With ? *v : dot, out);
                } else if (nd->has_else) {
                    exec_list(nd->else_body, dot, out);
                }
                vars.resize(mark);
                break;
            }
            case Node::Range: {
                const size_t mark = vars.size();
                Value h0, h1;
                const Value *vp = eval_pipe(*nd->pipe, dot, h0, h1);
                // the sequence must not move while the body runs (the body may assign variables,
                // which can reallocate `vars`): hold it by value -- a shared_ptr copy, not a deep one
                const Value v = *vp;
                const size_t nv = nd->range_vars.size();
                size_t count = 0;
                auto body = [&](const Value &key, const Value &elem) {
                    const size_t inner = vars.size();
                    if (nv == 1) vars.push_back({&nd->range_vars[0], elem});
                    else if (nv == 2) {
                        vars.push_back({&nd->range_vars[0], key});
                        vars.push_back({&nd->range_vars[1], elem});
                    }
                    exec_list(nd->body, elem, out);
                    vars.resize(inner);
                    ++count;
                };
                if (v.kind == Value::MapK) {
                    for (const auto &kv : *v.map) body(nv == 2 ? Value::str(kv.first) : Value(), kv.second);
                } else if (v.kind == Value::ListK) {
                    int64_t k = 0;
                    for (const auto &x : *v.list) body(Value::integer(k++), x);
                } else if (v.kind == Value::Int) {
                    for (int64_t k = 0; k < v.i; ++k) body(Value::integer(k), Value::integer(k));
                } else if (!v.is_nilish()) {
                    throw Error("range can't iterate over " + go_print(v));
                }
                if (count == 0 && nd->has_else) exec_list(nd->else_body, dot, out);
                vars.resize(mark);
                break;
            }
            case Node::Template: {
                // {{ template }} writes straight into the output, so deferral stays possible
                if (nd->pipe) run_template(nd->text, eval_pipe_value(*nd->pipe, dot), out);
                else run_template(nd->text, Value::nil(), out);
                break;
            }
            }
        }
    }
};

// ============================================================================ TemplateSet
TemplateSet::TemplateSet() {}

void TemplateSet::parse(const std::string &name, const std::string &src) {
    Parser p;
    p.items = lex(src, name);
    p.name = name;
    p.defs = &templates_;
    std::string term, rest;
    NodeList nodes = p.list(0, &term, &rest);
    templates_[name] = nodes;
}

uint64_t TemplateSet::next_session_id() {
    static std::atomic<uint64_t> next{1};
    return next.fetch_add(1);
}

namespace {

// Can executing `nodes` depend on anything but the dot it is given?  The one thing that differs
// between two includes of a template with the same dot inside one release is .Template (it names
// the file being rendered), so a body is memoisable when it never reaches a field chain starting
// at Template, never hands its whole dot to anything but a memoisable template, and only names
// templates by string literals.
struct Purity {
    const std::map<std::string, NodeList> &all;
    std::map<const NodeList *, int> state;   // 1 = in progress (assume pure: recursion), 2 = pure, 3 = impure

    bool operand_ok(const Operand &op, bool whole_dot_allowed) {
        if ((op.kind == Operand::Field) && !op.fields.empty() && op.fields[0] == "Template") return false;
        if (op.kind == Operand::Var && op.text == "$" && !op.fields.empty() && op.fields[0] == "Template") return false;
        if ((op.kind == Operand::Dot || (op.kind == Operand::Var && op.fields.empty())) && !whole_dot_allowed) return false;
        if (op.kind == Operand::Paren) return pipe_ok(*op.paren);
        return true;
    }
    bool cmd_ok(const Command &c) {
        const Operand &f = c.args[0];
        if (f.kind == Operand::Ident && f.fn == F_include) {
            if (c.args.size() < 2 || c.args[1].kind != Operand::Str) return false;   // piped or computed name
            auto it = all.find(c.args[1].text);
            if (it == all.end() || !body_ok(it->second)) return false;
            for (size_t k = 2; k < c.args.size(); ++k)
                if (!operand_ok(c.args[k], true)) return false;
            return true;
        }
        // `default`, `and`, `or`, `ternary`, `required`, `with`-style passing keep a value as it is;
        // anything else given the whole dot could print or walk it
        const bool passes_values = f.kind == Operand::Ident && (f.fn == F_default || f.fn == F_and || f.fn == F_or ||
                                                                 f.fn == F_ternary || f.fn == F_required || f.fn == F_not ||
                                                                 f.fn == F_empty || f.fn == F_hasKey);
        for (size_t k = 0; k < c.args.size(); ++k)
            if (!operand_ok(c.args[k], k == 0 ? c.args.size() == 1 && false : passes_values)) return false;
        return true;
    }
    bool pipe_ok(const Pipeline &p) {
        for (const auto &c : p.cmds)
            if (!cmd_ok(c)) return false;
        return true;
    }
    bool list_ok(const NodeList &nodes) {
        for (const auto &nd : nodes) {
            if (nd->pipe && !pipe_ok(*nd->pipe)) return false;
            if (nd->kind == Node::Template) {
                auto it = all.find(nd->text);
                if (it == all.end() || !body_ok(it->second)) return false;
            }
            if (nd->kind == Node::With || nd->kind == Node::Range) return false;   // dot changes: keep the analysis simple
            if (!list_ok(nd->body) || !list_ok(nd->else_body)) return false;
        }
        return true;
    }
    bool body_ok(const NodeList &nodes) {
        int &st = state[&nodes];
        if (st == 1 || st == 2) return true;
        if (st == 3) return false;
        st = 1;
        const bool ok = list_ok(nodes);
        state[&nodes] = ok ? 2 : 3;
        return ok;
    }
};

}  // namespace

void TemplateSet::finalize() {
    memo_ok_.clear();
    for (const auto &kv : templates_) {
        Purity p{templates_, {}};   // fresh per template: an "in progress" assumption must not leak
        memo_ok_[&kv.second] = p.body_ok(kv.second);
    }
}

const NodeList *TemplateSet::find(const std::string &name) const {
    auto it = templates_.find(name);
    return it == templates_.end() ? nullptr : &it->second;
}

void TemplateSet::execute_into(const NodeList &nodes, const Value &data, const ExecOptions &opt, std::string &out) const {
    // the variable stack and the include buffers are kept per thread: a fleet render executes the
    // same few templates thousands of times on every host thread
    struct Scratch {
        std::deque<Var> vars;
        std::vector<std::string> pool;
        std::vector<IncludeMemo> memo;
        uint64_t session = 0;
        bool busy = false;
    };
    static thread_local Scratch tls;
    std::unique_ptr<Scratch> local;   // only for a nested render from inside a callback
    if (tls.busy) local.reset(new Scratch);
    Scratch &sc = local ? *local : tls;
    struct Guard {
        Scratch &s;
        explicit Guard(Scratch &x) : s(x) { s.busy = true; }
        ~Guard() { s.vars.clear(); s.busy = false; }
    } guard(sc);
    if (sc.session != opt.session_id) {   // another release: its values are other values
        sc.memo.clear();
        sc.session = opt.session_id;
    }
    Exec ex{*this, opt, 0, sc.vars, sc.pool, opt.session_id ? &sc.memo : nullptr};
    sc.vars.push_back({&Exec::dollar(), data});
    ex.exec_list(nodes, data, out);
}

std::string TemplateSet::execute(const std::string &name, const Value &data, const ExecOptions &opt) const {
    const NodeList *nodes = find(name);
    if (!nodes) throw Error("template \"" + name + "\" not defined");
    std::string out;
    execute_into(*nodes, data, opt, out);
    return out;
}

std::string TemplateSet::placeholder(size_t id) {
    std::string s(kPlaceholderLen, '\x01');
    for (int k = 0; k < 4; ++k) s[(size_t)k] = (char)(0x10 + ((id >> (12 - 4 * k)) & 0xF));
    return s;
}

// Does s hold a byte the placeholders use (0x01, 0x10-0x1F)?  Every value string of every
// release goes through here (an 11 KB config.toml each time), eight bytes per step: a byte is
// 0x01 iff (b ^ 0x01) == 0, and in 0x10-0x1F iff ((b ^ 0x10) & 0xF0) == 0; "some byte of v is
// zero" is (v - 0x01..01) & ~v & 0x80..80.
bool TemplateSet::has_sentinel_bytes(const std::string &s) {
    const unsigned char *p = reinterpret_cast<const unsigned char *>(s.data());
    size_t n = s.size(), i = 0;
#ifdef __SSE2__
    // 64 bytes per step: b == 0x01, or (b & 0xF0) == 0x10
    const __m128i one = _mm_set1_epi8(0x01), hi = _mm_set1_epi8((char)0xF0), ten = _mm_set1_epi8(0x10);
    for (; i + 64 <= n; i += 64) {
        __m128i acc = _mm_setzero_si128();
        for (int k = 0; k < 4; ++k) {
            const __m128i v = _mm_loadu_si128(reinterpret_cast<const __m128i *>(p + i + 16 * k));
            acc = _mm_or_si128(acc, _mm_or_si128(_mm_cmpeq_epi8(v, one), _mm_cmpeq_epi8(_mm_and_si128(v, hi), ten)));
        }
        if (_mm_movemask_epi8(acc)) return true;
    }
#endif
    const uint64_t ones = 0x0101010101010101ull, highs = 0x8080808080808080ull;
    for (; i + 8 <= n; i += 8) {
        uint64_t w;
        memcpy(&w, p + i, 8);
        const uint64_t a = w ^ ones;                                   // zero byte <=> 0x01
        const uint64_t b = (w ^ 0x1010101010101010ull) & 0xF0F0F0F0F0F0F0F0ull;   // zero byte <=> 0x10..0x1F
        if ((((a - ones) & ~a) | ((b - ones) & ~b)) & highs) return true;
    }
    for (; i < n; ++i)
        if (p[i] == 0x01 || (p[i] >= 0x10 && p[i] < 0x20)) return true;
    return false;
}

}  // namespace tmpl
}  // namespace kvedge
