// Go text/template + Sprig subset (see tmpl.hpp). Host C++: lexer, parser, executor.
#include "tmpl.hpp"

#include <algorithm>
#include <cctype>
#include <charconv>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

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
struct Operand {
    enum Kind { Dot, Field, Var, Str, Num, BoolK, NilK, Ident, Paren } kind = Dot;
    std::string text;                 // Var name / Ident / Str payload
    Value num;                        // Num / Bool literal
    std::vector<std::string> fields;  // .A.B chain
    std::shared_ptr<struct Pipeline> paren;
};
struct Command { std::vector<Operand> args; };
struct Pipeline {
    std::vector<Command> cmds;
    std::string decl;       // "$x" when the pipeline declares / assigns a variable
    bool is_assign = false;
};
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

bool eq_fn(const std::vector<Value> &a) {
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

bool cmp_fn(const std::string &op, const Value &a, const Value &b) {
    std::string ka = basic_kind(a), kb = basic_kind(b);
    if (ka != kb) throw Error("incompatible types for comparison");
    int c;
    if (ka == "int") c = a.i < b.i ? -1 : a.i > b.i;
    else if (ka == "float") c = a.f < b.f ? -1 : a.f > b.f;
    else if (ka == "string") c = a.s.compare(b.s) < 0 ? -1 : a.s.compare(b.s) > 0;
    else throw Error("invalid type for comparison");
    if (op == "lt") return c < 0;
    if (op == "le") return c <= 0;
    if (op == "gt") return c > 0;
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

std::string sprintf_fn(const std::vector<Value> &a) {
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
struct Exec {
    const TemplateSet &ts;
    const TemplateSet::ExecOptions &opt;
    int capture_depth = 0;  // > 0 while executing inside include (output is captured as a value)

    using Scopes = std::vector<std::map<std::string, Value>>;

    Value lookup_var(const std::string &name, Scopes &sc) {
        for (size_t k = sc.size(); k-- > 0;) {
            auto it = sc[k].find(name);
            if (it != sc[k].end()) return it->second;
        }
        throw Error("undefined variable \"" + name + "\"");
    }

    Value field_chain(Value v, const std::vector<std::string> &fields) {
        for (const auto &f : fields) {
            if (v.kind == Value::StructK) {
                auto it = v.map->find(f);
                if (it == v.map->end()) throw Error("can't evaluate field " + f + " in type " + v.s);
                Value nx = it->second;
                v = nx;
            } else if (v.kind == Value::MapK) {
                auto it = v.map->find(f);
                Value nx = it == v.map->end() ? Value::novalue() : it->second;
                v = nx;
            } else if (v.is_nilish()) {
                throw Error("nil pointer evaluating interface {}." + f);
            } else {
                throw Error("can't evaluate field " + f + " in type " + go_typename(v));
            }
        }
        return v;
    }

    std::string encode_now(const std::string &s) {
        if (s.empty()) return "";
        if (!opt.encode) throw Error("b64enc: no encoder attached (GPU required)");
        return opt.encode(s);
    }

    Value call(const std::string &fn, std::vector<Value> &a, const Value &dot) {
        (void)dot;
        auto need = [&](size_t n) {
            if (a.size() != n)
                throw Error("wrong number of args for " + fn + ": want " + std::to_string(n) + " got " + std::to_string(a.size()));
        };
        auto empty = [](const Value &v) { return !truth(v); };
        if (fn == "include") {
            need(2);
            std::string tname = strval(a[0]);
            auto it = ts.templates_.find(tname);
            if (it == ts.templates_.end()) throw Error("template \"" + tname + "\" not defined");
            std::string out;
            Scopes sc(1);
            sc[0]["$"] = a[1];
            ++capture_depth;
            exec_list(it->second, a[1], sc, out);
            --capture_depth;
            return Value::str(out);
        }
        if (fn == "default") {
            if (a.empty()) throw Error("wrong number of args for default");
            if (a.size() < 2 || empty(a[1])) return a[0];
            return a[1];
        }
        if (fn == "trunc") {
            need(2);
            int64_t c = to_int(a[0]);
            std::string s = string_arg(a[1], fn);
            if (c < 0 && (int64_t)s.size() + c > 0) return Value::str(s.substr((size_t)((int64_t)s.size() + c)));
            if (c >= 0 && (int64_t)s.size() > c) return Value::str(s.substr(0, (size_t)c));
            return Value::str(s);
        }
        if (fn == "trimSuffix") {
            need(2);
            std::string suf = string_arg(a[0], fn), s = string_arg(a[1], fn);
            if (!suf.empty() && s.size() >= suf.size() && s.compare(s.size() - suf.size(), suf.size(), suf) == 0)
                s.erase(s.size() - suf.size());
            return Value::str(s);
        }
        if (fn == "trimPrefix") {
            need(2);
            std::string pre = string_arg(a[0], fn), s = string_arg(a[1], fn);
            if (!pre.empty() && s.compare(0, pre.size(), pre) == 0) s.erase(0, pre.size());
            return Value::str(s);
        }
        if (fn == "replace") { need(3); return Value::str(replace_all(string_arg(a[2], fn), string_arg(a[0], fn), string_arg(a[1], fn))); }
        if (fn == "quote" || fn == "squote") {
            std::string o;
            bool first = true;
            for (const auto &x : a) {
                if (x.is_nilish()) continue;
                if (!first) o += " ";
                first = false;
                o += fn == "quote" ? go_quote(strval(x)) : "'" + go_print(x) + "'";
            }
            return Value::str(o);
        }
        if (fn == "indent" || fn == "nindent") {
            need(2);
            std::string pad((size_t)std::max<int64_t>(0, to_int(a[0])), ' ');
            std::string o = pad + replace_all(string_arg(a[1], fn), "\n", "\n" + pad);
            return Value::str(fn == "nindent" ? "\n" + o : o);
        }
        if (fn == "b64enc") { need(1); return Value::str(encode_now(string_arg(a[0], fn))); }
        if (fn == "b64dec") {
            need(1);
            if (!opt.decode) throw Error("b64dec: no decoder attached (GPU required)");
            return Value::str(opt.decode(string_arg(a[0], fn)));
        }
        if (fn == "eq") return Value::boolean(eq_fn(a));
        if (fn == "ne") { need(2); return Value::boolean(!eq_fn(a)); }
        if (fn == "lt" || fn == "le" || fn == "gt" || fn == "ge") { need(2); return Value::boolean(cmp_fn(fn, a[0], a[1])); }
        if (fn == "not") { need(1); return Value::boolean(!truth(a[0])); }
        if (fn == "printf") return Value::str(sprintf_fn(a));
        if (fn == "print" || fn == "println") {
            std::string o;
            if (fn == "println") {
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
            return Value::str(o);
        }
        if (fn == "len") {
            need(1);
            if (a[0].kind == Value::String) return Value::integer((int64_t)a[0].s.size());
            if (a[0].kind == Value::ListK) return Value::integer((int64_t)a[0].list->size());
            if (a[0].kind == Value::MapK) return Value::integer((int64_t)a[0].map->size());
            throw Error("len of type " + go_typename(a[0]));
        }
        if (fn == "empty") { need(1); return Value::boolean(empty(a[0])); }
        if (fn == "required") {
            need(2);
            if (a[1].is_nilish() || (a[1].kind == Value::String && a[1].s.empty())) throw Error(strval(a[0]));
            return a[1];
        }
        if (fn == "lower") { need(1); return Value::str(ascii_map(string_arg(a[0], fn), ::tolower)); }
        if (fn == "upper") { need(1); return Value::str(ascii_map(string_arg(a[0], fn), ::toupper)); }
        if (fn == "trim") {
            need(1);
            std::string s = string_arg(a[0], fn);
            size_t x = 0, y = s.size();
            while (x < y && (is_space(s[x]) || s[x] == '\v' || s[x] == '\f')) ++x;
            while (y > x && (is_space(s[y - 1]) || s[y - 1] == '\v' || s[y - 1] == '\f')) --y;
            return Value::str(s.substr(x, y - x));
        }
        if (fn == "toString") { need(1); return Value::str(strval(a[0])); }
        if (fn == "hasKey") {
            need(2);
            return Value::boolean(a[0].kind == Value::MapK && a[0].map->count(strval(a[1])) != 0);
        }
        if (fn == "ternary") { need(3); return truth(a[2]) ? a[0] : a[1]; }
        if (fn == "cat") {
            std::string o;
            bool first = true;
            for (const auto &x : a) {
                if (x.is_nilish()) continue;
                if (!first) o += " ";
                first = false;
                o += strval(x);
            }
            return Value::str(o);
        }
        if (fn == "contains") { need(2); return Value::boolean(string_arg(a[1], fn).find(string_arg(a[0], fn)) != std::string::npos); }
        if (fn == "hasPrefix") { need(2); std::string p = string_arg(a[0], fn), s = string_arg(a[1], fn); return Value::boolean(s.compare(0, p.size(), p) == 0); }
        if (fn == "hasSuffix") {
            need(2);
            std::string p = string_arg(a[0], fn), s = string_arg(a[1], fn);
            return Value::boolean(s.size() >= p.size() && s.compare(s.size() - p.size(), p.size(), p) == 0);
        }
        if (fn == "repeat") {
            need(2);
            std::string o, s = string_arg(a[1], fn);
            for (int64_t k = 0; k < to_int(a[0]); ++k) o += s;
            return Value::str(o);
        }
        throw Error("function \"" + fn + "\" not defined");
    }

    Value eval_operand(const Operand &op, const Value &dot, Scopes &sc) {
        switch (op.kind) {
        case Operand::Dot: return dot;
        case Operand::Field: return field_chain(dot, op.fields);
        case Operand::Var: return field_chain(lookup_var(op.text, sc), op.fields);
        case Operand::Str: return Value::str(op.text);
        case Operand::Num: case Operand::BoolK: return op.num;
        case Operand::NilK: return Value::nil();
        case Operand::Paren: return field_chain(eval_pipe(*op.paren, dot, sc), op.fields);
        case Operand::Ident: {
            std::vector<Value> none;
            return call(op.text, none, dot);
        }
        }
        throw Error("bad operand");
    }

    Value eval_cmd(const Command &cmd, const Value &dot, Scopes &sc, const Value *piped) {
        const Operand &first = cmd.args[0];
        if (first.kind == Operand::Ident) {
            const std::string &fn = first.text;
            if (fn == "and" || fn == "or") {
                size_t n_args = cmd.args.size() - 1 + (piped ? 1 : 0);
                if (n_args == 0) throw Error("wrong number of args for " + fn);
                Value res;
                for (size_t k = 0; k < n_args; ++k) {
                    res = k + 1 < cmd.args.size() ? eval_operand(cmd.args[k + 1], dot, sc) : *piped;
                    if ((fn == "and") != truth(res)) return res;
                }
                return res;
            }
            std::vector<Value> args;
            for (size_t k = 1; k < cmd.args.size(); ++k) args.push_back(eval_operand(cmd.args[k], dot, sc));
            if (piped) args.push_back(*piped);
            return call(fn, args, dot);
        }
        if (cmd.args.size() > 1 || piped) throw Error("can't give argument to non-function");
        return eval_operand(first, dot, sc);
    }

    // Evaluate the first `upto` commands of a pipeline.
    Value eval_pipe_prefix(const Pipeline &p, size_t upto, const Value &dot, Scopes &sc) {
        Value v;
        bool has = false;
        for (size_t k = 0; k < upto; ++k) {
            Value nv = eval_cmd(p.cmds[k], dot, sc, has ? &v : nullptr);
            v = nv;
            has = true;
        }
        return v;
    }

    Value eval_pipe(const Pipeline &p, const Value &dot, Scopes &sc) {
        Value v = eval_pipe_prefix(p, p.cmds.size(), dot, sc);
        if (!p.decl.empty()) {
            if (p.is_assign) {
                bool found = false;
                for (size_t k = sc.size(); k-- > 0 && !found;) {
                    auto it = sc[k].find(p.decl);
                    if (it != sc[k].end()) { it->second = v; found = true; }
                }
                if (!found) throw Error("undefined variable \"" + p.decl + "\"");
            } else {
                sc.back()[p.decl] = v;
            }
        }
        return v;
    }

    // `{{ <pipeline> | b64enc }}` printed straight into a top-level file: leave it pending.
    bool try_defer(const Pipeline &p, const Value &dot, Scopes &sc, std::string &out) {
        if (!opt.defer_b64 || !opt.pending || capture_depth != 0 || !p.decl.empty()) return false;
        const Command &last = p.cmds.back();
        if (p.cmds.size() < 2 || last.args.size() != 1 || last.args[0].kind != Operand::Ident ||
            last.args[0].text != "b64enc")
            return false;
        if (opt.pending->size() >= TemplateSet::kMaxPending) return false;
        Value in = eval_pipe_prefix(p, p.cmds.size() - 1, dot, sc);
        string_arg(in, "b64enc");                // same refusal as the immediate path
        std::string s = std::move(in.s);         // a config.toml: do not copy it again
        if (s.empty()) return true;  // b64enc "" == ""
        Pending pd;
        pd.encoded_len = (s.size() + 2) / 3 * 4;
        pd.input = std::move(s);
        // the placeholder (TemplateSet::placeholder) written in place: 4 id bytes, then filler
        const size_t id = opt.pending->size();
        for (int k = 0; k < 4; ++k) out.push_back((char)(0x10 + ((id >> (12 - 4 * k)) & 0xF)));
        out.append(pd.encoded_len - 4, '\x01');
        opt.pending->push_back(std::move(pd));
        return true;
    }

    void exec_list(const NodeList &nodes, const Value &dot, Scopes &sc, std::string &out) {
        for (const auto &nd : nodes) {
            switch (nd->kind) {
            case Node::Text: out += nd->text; break;
            case Node::Action: {
                if (try_defer(*nd->pipe, dot, sc, out)) break;
                Value v = eval_pipe(*nd->pipe, dot, sc);
                if (nd->pipe->decl.empty()) out += go_print(v);
                break;
            }
            case Node::If: case Node::With: {
                sc.emplace_back();
                Value v = eval_pipe(*nd->pipe, dot, sc);
                if (truth(v)) {
                    exec_list(nd->body, nd->kind == Node::With ? v : dot, sc, out);
                } else if (nd->has_else) {
                    exec_list(nd->else_body, dot, sc, out);
                }
                sc.pop_back();
                break;
            }
            case Node::Range: {
                sc.emplace_back();
                Value v = eval_pipe(*nd->pipe, dot, sc);
                std::vector<std::pair<Value, Value>> seq;
                if (v.kind == Value::MapK) {
                    for (const auto &kv : *v.map) seq.push_back({Value::str(kv.first), kv.second});
                } else if (v.kind == Value::ListK) {
                    int64_t k = 0;
                    for (const auto &x : *v.list) seq.push_back({Value::integer(k++), x});
                } else if (v.kind == Value::Int) {
                    for (int64_t k = 0; k < v.i; ++k) seq.push_back({Value::integer(k), Value::integer(k)});
                } else if (!v.is_nilish()) {
                    throw Error("range can't iterate over " + go_print(v));
                }
                if (seq.empty() && nd->has_else) exec_list(nd->else_body, dot, sc, out);
                for (auto &kv : seq) {
                    sc.emplace_back();
                    if (nd->range_vars.size() == 1) sc.back()[nd->range_vars[0]] = kv.second;
                    else if (nd->range_vars.size() == 2) {
                        sc.back()[nd->range_vars[0]] = kv.first;
                        sc.back()[nd->range_vars[1]] = kv.second;
                    }
                    exec_list(nd->body, kv.second, sc, out);
                    sc.pop_back();
                }
                sc.pop_back();
                break;
            }
            case Node::Template: {
                Value d = nd->pipe ? eval_pipe(*nd->pipe, dot, sc) : Value::nil();
                auto it = ts.templates_.find(nd->text);
                if (it == ts.templates_.end()) throw Error("template \"" + nd->text + "\" not defined");
                Scopes inner(1);
                inner[0]["$"] = d;
                // {{ template }} writes straight into the output, so deferral stays possible
                exec_list(it->second, d, inner, out);
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

std::string TemplateSet::execute(const std::string &name, const Value &data, const ExecOptions &opt) const {
    auto it = templates_.find(name);
    if (it == templates_.end()) throw Error("template \"" + name + "\" not defined");
    Exec ex{*this, opt};
    Exec::Scopes sc(1);
    sc[0]["$"] = data;
    std::string out;
    ex.exec_list(it->second, data, sc, out);
    return out;
}

std::string TemplateSet::placeholder(size_t id, size_t encoded_len) {
    std::string s(encoded_len, '\x01');
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
