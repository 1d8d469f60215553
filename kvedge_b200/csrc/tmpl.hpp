// Go text/template + Sprig subset: the host side of the render path.
//
// What Helm does with deployment/helm/templates/* of the reference, restricted to the
// constructs that chart (and neighbours a chart author would reach for next) uses:
//   actions, {{- -}} trimming, comments ......... templates/_helper.tpl:1-75
//   define / include / template ................. _helper.tpl:6,13,20,31
//   if / else / with / range / variables ........ _helper.tpl:22-24, aziot-edge-vm-service.yaml:1,18
//   default trunc trimSuffix printf replace quote indent b64enc eq ...
//                                                 _helper.tpl:7,14,23; aziot-edge-vm.yaml:13;
//                                                 aziot-edge-*-secret.yaml:6
// Strings are byte strings end to end (Go semantics): `trunc` slices bytes and
// --set-file content passes through untouched.
#pragma once
#include <cstdint>
#include <functional>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace kvedge {
namespace tmpl {

struct Error : std::runtime_error {
    using std::runtime_error::runtime_error;
};

struct Value;
using List = std::vector<Value>;

// map[string]interface{}: a vector kept sorted by key (Go prints and ranges over maps in key
// order).  Charts' tables hold a handful of entries and a fleet builds several of them per
// install, so one allocation per table beats one per entry; lookups are a binary search.
class Map {
  public:
    using value_type = std::pair<std::string, Value>;
    using iterator = std::vector<value_type>::iterator;
    using const_iterator = std::vector<value_type>::const_iterator;
    Map();
    Map(const Map &);
    Map &operator=(const Map &);
    ~Map();
    iterator begin();
    iterator end();
    const_iterator begin() const;
    const_iterator end() const;
    iterator find(const std::string &k);
    const_iterator find(const std::string &k) const;
    Value &operator[](const std::string &k);   // inserts a nil Value when the key is new
    size_t count(const std::string &k) const;
    void erase(iterator it);
    size_t size() const;
    bool empty() const;
    void reserve(size_t n);

  private:
    std::vector<value_type> v_;
};

struct Value {
    enum Kind { Nil, NoValue, Bool, Int, Float, String, ListK, MapK, StructK };
    Kind kind = Nil;
    bool b = false;
    int64_t i = 0;
    double f = 0;
    std::string s;                 // String payload; for StructK the Go type name
    std::shared_ptr<List> list;
    std::shared_ptr<Map> map;      // MapK and StructK

    static Value nil() { return Value(); }
    static Value novalue() { Value v; v.kind = NoValue; return v; }
    static Value boolean(bool x) { Value v; v.kind = Bool; v.b = x; return v; }
    static Value integer(int64_t x) { Value v; v.kind = Int; v.i = x; return v; }
    static Value number(double x) { Value v; v.kind = Float; v.f = x; return v; }
    static Value str(std::string x) { Value v; v.kind = String; v.s = std::move(x); return v; }
    static Value make_list() { Value v; v.kind = ListK; v.list = std::make_shared<List>(); return v; }
    static Value make_map() { Value v; v.kind = MapK; v.map = std::make_shared<Map>(); return v; }
    static Value make_struct(std::string type_name) {
        Value v; v.kind = StructK; v.s = std::move(type_name); v.map = std::make_shared<Map>(); return v;
    }
    bool is_nilish() const { return kind == Nil || kind == NoValue; }
};

inline Map::Map() = default;
inline Map::Map(const Map &) = default;
inline Map &Map::operator=(const Map &) = default;
inline Map::~Map() = default;
inline Map::iterator Map::begin() { return v_.begin(); }
inline Map::iterator Map::end() { return v_.end(); }
inline Map::const_iterator Map::begin() const { return v_.begin(); }
inline Map::const_iterator Map::end() const { return v_.end(); }
inline Map::const_iterator Map::find(const std::string &k) const {
    size_t lo = 0, hi = v_.size();
    while (lo < hi) {
        const size_t mid = (lo + hi) / 2;
        const int c = v_[mid].first.compare(k);
        if (c == 0) return v_.begin() + (std::ptrdiff_t)mid;
        if (c < 0) lo = mid + 1; else hi = mid;
    }
    return v_.end();
}
inline Map::iterator Map::find(const std::string &k) {
    const const_iterator it = static_cast<const Map *>(this)->find(k);
    return v_.begin() + (it - v_.cbegin());
}
inline Value &Map::operator[](const std::string &k) {
    size_t lo = 0, hi = v_.size();
    while (lo < hi) {
        const size_t mid = (lo + hi) / 2;
        const int c = v_[mid].first.compare(k);
        if (c == 0) return v_[mid].second;
        if (c < 0) lo = mid + 1; else hi = mid;
    }
    return v_.insert(v_.begin() + (std::ptrdiff_t)lo, value_type(k, Value()))->second;
}
inline size_t Map::count(const std::string &k) const { return find(k) == end() ? 0 : 1; }
inline void Map::erase(iterator it) { v_.erase(it); }
inline size_t Map::size() const { return v_.size(); }
inline bool Map::empty() const { return v_.empty(); }
inline void Map::reserve(size_t n) { v_.reserve(n); }

bool truth(const Value &v);            // text/template truth
std::string go_print(const Value &v);  // fmt.Fprint of one value
std::string format_float(double f);    // Go %v of a float64
std::string go_quote(const std::string &s);  // strconv.Quote
Value deep_copy(const Value &v);

// A b64enc site whose evaluation is deferred to the GPU: the output carries a short placeholder
// (kPlaceholderLen bytes) and the input bytes are remembered here -- as a view when they are a
// value of the tree the caller keeps alive (a config.toml is not copied again), else owned.
struct Pending {
    std::string owned;
    const char *ext = nullptr;
    size_t len = 0;
    size_t encoded_len = 0;
    const char *data() const { return ext ? ext : owned.data(); }
    size_t size() const { return len; }
};

// Called for every `b64enc` the executor cannot defer (its result feeds another function).
using ImmediateEncoder = std::function<std::string(const std::string &)>;

struct Node;
using NodeList = std::vector<std::shared_ptr<Node>>;

class TemplateSet {
  public:
    TemplateSet();
    // Parse one file; `define` blocks inside it land in the shared namespace.
    void parse(const std::string &name, const std::string &src);
    bool has(const std::string &name) const { return templates_.count(name) != 0; }

    struct ExecOptions {
        bool defer_b64 = false;             // leave top-level `... | b64enc` sites pending
        std::vector<Pending> *pending = nullptr;
        bool values_outlive_pending = false;  // the caller keeps `data` alive as long as `pending`
        // Non-zero: all execute_into calls carrying this id render ONE release from the same
        // (immutable) value tree; an `include` of a template that cannot see .Template is then
        // evaluated once per (template, dot) and its text reused -- a chart calls its name helper a
        // dozen times per release.  Ids must never repeat (take them from next_session_id()).
        uint64_t session_id = 0;
        ImmediateEncoder encode;            // required whenever a b64enc must run now
        ImmediateEncoder decode;            // b64dec (optional)
    };
    std::string execute(const std::string &name, const Value &data, const ExecOptions &opt) const;
    // The same for callers that render one template many times: look it up once, append to `out`.
    const NodeList *find(const std::string &name) const;
    void execute_into(const NodeList &nodes, const Value &data, const ExecOptions &opt, std::string &out) const;

    static uint64_t next_session_id();
    // Work out which named templates may be memoised within a release (call once, after the last
    // parse and before templates are executed from several threads).
    void finalize();

    // Placeholder protocol (render_capi.cpp expands it): 4 id bytes in [0x10,0x20), then 0x01.
    static constexpr size_t kPlaceholderLen = 5;
    static std::string placeholder(size_t id);
    static bool has_sentinel_bytes(const std::string &s);
    static constexpr size_t kMaxPending = 65536;

  private:
    friend struct Exec;
    std::map<std::string, NodeList> templates_;
    std::map<const NodeList *, bool> memo_ok_;   // named templates whose text depends on (name, dot) alone
};

}  // namespace tmpl
}  // namespace kvedge
