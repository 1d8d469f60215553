// extern "C" render entry points (declared in include/kvedge_b200.h): load a chart, plan a
// release (render with every top-level `| b64enc` site left pending), encode the pending
// Secret payloads on the GPU in one batch, and frame the result like `helm template`.
//
// Reference path: README.md:60 (`helm install ... --set ... --set-file azIotEdgeConfig=...`)
// -> deployment/helm/templates/* -> the two b64enc sites
// (aziot-edge-runtime-config-secret.yaml:6, aziot-edge-vm-cloud-init-secret.yaml:6).
#include "../../include/kvedge_b200.h"

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "chart.hpp"
#include "tmpl.hpp"

using kvedge::chart::Chart;
using kvedge::chart::Manifest;
using kvedge::tmpl::Pending;
using kvedge::tmpl::TemplateSet;
using kvedge::tmpl::Value;

extern "C" int kv_set_last_error_(int code, const char *msg);  // capi.cu
extern "C" int kv_stage_get_(size_t in_bytes, size_t out_bytes, char **in, char **out);  // capi.cu

struct kv_chart {
    Chart chart;
    TemplateSet ts;
    bool templates_have_sentinels = false;
    // Everything of a render that does not depend on the release, built once at load: a fleet
    // renders ONE chart thousands of times.
    struct File {
        std::string name;                         // "<chart>/templates/<file>"
        const kvedge::tmpl::NodeList *nodes;      // its parsed body
        Value tmpl;                               // .Template of this file
    };
    std::vector<File> files;                      // non-partial templates, sorted by name
    Value chart_meta, caps;                       // .Chart, .Capabilities
    bool defaults_have_sentinels = false;
};

struct kv_release {
    // The `helm template` stream and the NOTES text as planned: every pending b64enc site is a
    // 5-byte placeholder (TemplateSet::kPlaceholderLen), so planning never writes or moves the
    // kilobytes the encodings will take.  The final text is assembled once, when asked for.
    std::string manifest, notes;
    size_t manifest_full_size = 0;      // size of the stream once every placeholder is expanded
    std::vector<Pending> pending;
    std::vector<std::string> filled;    // encoded text per pending site ("" until filled)
    size_t n_filled = 0;
    std::string full_manifest, full_notes;
    bool expanded = false;
    Value keepalive;                    // the render's value tree: pending inputs may point into it
};

namespace {

int fail(int code, const std::string &msg) { return kv_set_last_error_(code, msg.c_str()); }

Value chart_struct(const Value &meta) {
    Value s = Value::make_struct("chart.Metadata");
    auto get = [&](const char *k, Value d) -> Value {
        auto it = meta.map->find(k);
        if (it == meta.map->end() || it->second.is_nilish()) return d;
        return it->second;
    };
    auto str_of = [&](const char *k) { return Value::str(kvedge::tmpl::go_print(get(k, Value::str("")))); };
    (*s.map)["Name"] = get("name", Value::str(""));
    (*s.map)["Home"] = get("home", Value::str(""));
    (*s.map)["Sources"] = get("sources", Value::make_list());
    (*s.map)["Version"] = str_of("version");
    (*s.map)["Description"] = get("description", Value::str(""));
    (*s.map)["Keywords"] = get("keywords", Value::make_list());
    (*s.map)["Maintainers"] = get("maintainers", Value::make_list());
    (*s.map)["Icon"] = get("icon", Value::str(""));
    (*s.map)["APIVersion"] = get("apiVersion", Value::str(""));
    (*s.map)["Condition"] = get("condition", Value::str(""));
    (*s.map)["Tags"] = get("tags", Value::str(""));
    (*s.map)["AppVersion"] = str_of("appVersion");
    (*s.map)["Deprecated"] = Value::boolean(kvedge::tmpl::truth(get("deprecated", Value::boolean(false))));
    (*s.map)["Annotations"] = get("annotations", Value::make_map());
    (*s.map)["KubeVersion"] = get("kubeVersion", Value::str(""));
    (*s.map)["Dependencies"] = get("dependencies", Value::make_list());
    (*s.map)["Type"] = get("type", Value::str(""));
    return s;
}

bool value_has_sentinels(const Value &v) {
    if (v.kind == Value::String) return TemplateSet::has_sentinel_bytes(v.s);
    if (v.list)
        for (const auto &x : *v.list)
            if (value_has_sentinels(x)) return true;
    if (v.map)
        for (const auto &kv : *v.map)
            if (TemplateSet::has_sentinel_bytes(kv.first) || value_has_sentinels(kv.second)) return true;
    return false;
}

void replace_all_inplace(std::string &s, const std::string &from, const std::string &to) {
    size_t i = 0;
    while ((i = s.find(from, i)) != std::string::npos) {
        s.replace(i, from.size(), to);
        i += to.size();
    }
}

std::string gpu_encode_one(const std::string &in) {
    std::string out(kv_b64enc_len(in.size()), '\0');
    int rc = kv_b64enc(reinterpret_cast<const uint8_t *>(in.data()), in.size(), &out[0]);
    if (rc != KV_OK) throw kvedge::tmpl::Error(std::string("b64enc: ") + kv_last_error());
    return out;
}

std::string gpu_decode_one(const std::string &in) {
    std::string out(kv_b64dec_maxlen(in.size()) + 4, '\0');
    size_t n = 0, eo = 0;
    int rc = kv_b64dec(in.data(), in.size(), reinterpret_cast<uint8_t *>(&out[0]), &n, &eo);
    if (rc == KV_ECORRUPT) return std::string("illegal base64 data at input byte ") + std::to_string(eo);
    if (rc != KV_OK) throw kvedge::tmpl::Error(std::string("b64dec: ") + kv_last_error());
    out.resize(n);
    return out;
}

// *user_sentinels: whether any value the caller supplied holds a placeholder byte (the chart's
// own defaults were checked when it was loaded).
Value merged_values(const kv_chart *c, const kv_overrides *ov, bool *user_sentinels = nullptr) {
    if (user_sentinels) *user_sentinels = false;
    if (!ov) return kvedge::tmpl::deep_copy(c->chart.values);
    // Helm collects the flags into a map of their own and then slides the chart's defaults
    // under it (chartutil.CoalesceValues): null deletes a default, lists are not patched.
    Value vals = Value::make_map();
    vals.map->reserve(16);
    for (int i = 0; i < ov->n_set; ++i) kvedge::chart::apply_set(vals, ov->set[i], false);
    for (int i = 0; i < ov->n_set_string; ++i) kvedge::chart::apply_set(vals, ov->set_string[i], true);
    for (int i = 0; i < ov->n_set_file; ++i) kvedge::chart::apply_set_file(vals, ov->set_file[i]);
    for (int i = 0; i < ov->n_blob; ++i)
        kvedge::chart::apply_set_blob(vals, ov->blob_keys[i],
                                      std::string(reinterpret_cast<const char *>(ov->blob_data[i]), ov->blob_len[i]));
    if (user_sentinels) *user_sentinels = value_has_sentinels(vals);
    return kvedge::chart::coalesce_values(c->chart.values, std::move(vals));
}

// What does not depend on the release: .Chart, .Capabilities, and per file its name, parsed body
// and .Template.
void prepare_chart(kv_chart *c) {
    const std::string prefix = c->chart.name() + "/";
    Value caps = Value::make_map();
    Value kube = Value::make_map();
    (*kube.map)["Version"] = Value::str("v1.20.0");
    (*kube.map)["Major"] = Value::str("1");
    (*kube.map)["Minor"] = Value::str("20");
    (*caps.map)["KubeVersion"] = kube;
    Value hv = Value::make_map();
    (*hv.map)["Version"] = Value::str("v3.7.0");
    (*caps.map)["HelmVersion"] = hv;
    c->caps = caps;
    c->chart_meta = chart_struct(c->chart.metadata);
    c->defaults_have_sentinels = value_has_sentinels(c->chart.values);
    for (const auto &t : c->chart.templates) {
        const std::string &rel_path = t.first;
        const std::string base = rel_path.substr(rel_path.find_last_of('/') + 1);
        if (!base.empty() && base[0] == '_') continue;
        kv_chart::File f;
        f.name = prefix + rel_path;
        f.nodes = c->ts.find(f.name);
        f.tmpl = Value::make_map();
        (*f.tmpl.map)["Name"] = Value::str(f.name);
        (*f.tmpl.map)["BasePath"] = Value::str(prefix + "templates");
        c->files.push_back(std::move(f));
    }
    std::sort(c->files.begin(), c->files.end(), [](const kv_chart::File &a, const kv_chart::File &b) { return a.name < b.name; });
}

// Render every non-partial file of the chart, in name order; rendered[k] belongs to c->files[k].
void render_files(const kv_chart *c, const Value &vals, const char *release, const char *ns, bool defer,
                  std::vector<Pending> *pending, std::vector<std::string> *rendered, Value *keepalive) {
    Value rel = Value::make_map();
    rel.map->reserve(6);
    (*rel.map)["Name"] = Value::str(release ? release : "release-name");
    (*rel.map)["Namespace"] = Value::str(ns ? ns : "default");
    (*rel.map)["IsUpgrade"] = Value::boolean(false);
    (*rel.map)["IsInstall"] = Value::boolean(true);
    (*rel.map)["Revision"] = Value::integer(1);
    (*rel.map)["Service"] = Value::str("Helm");
    Value top = Value::make_map();
    top.map->reserve(5);
    (*top.map)["Values"] = vals;
    (*top.map)["Chart"] = c->chart_meta;
    (*top.map)["Release"] = rel;
    (*top.map)["Capabilities"] = c->caps;
    Value &tmpl_slot = (*top.map)["Template"];

    TemplateSet::ExecOptions opt;
    opt.defer_b64 = defer;
    opt.pending = pending;
    opt.values_outlive_pending = keepalive != nullptr;
    opt.session_id = TemplateSet::next_session_id();   // one release, one value tree: includes may be memoised
    if (keepalive) *keepalive = top;
    opt.encode = gpu_encode_one;
    opt.decode = gpu_decode_one;
    rendered->resize(c->files.size());   // the caller's buffers: their capacity is reused
    for (size_t k = 0; k < c->files.size(); ++k) {
        const kv_chart::File &f = c->files[k];
        tmpl_slot = f.tmpl;
        std::string &out = (*rendered)[k];
        out.clear();
        try {
            c->ts.execute_into(*f.nodes, top, opt, out);
        } catch (const kvedge::tmpl::Error &e) {
            throw kvedge::tmpl::Error("template: " + f.name + ": " + e.what());
        }
        if (out.find("<no value>") != std::string::npos)
            replace_all_inplace(out, "<no value>", "");  // missingkey=zero, non-strict
    }
}

void frame(const kv_chart *c, const std::vector<std::string> &rendered, std::string *manifest, std::string *notes) {
    static thread_local std::vector<std::pair<const std::string *, const std::string *>> files;
    files.clear();
    for (size_t k = 0; k < rendered.size(); ++k) files.push_back({&c->files[k].name, &rendered[k]});
    kvedge::chart::frame_manifests(files, manifest, notes);
}

// Walk a planned text: fn_text(ptr, n) for every run of ordinary bytes, fn_site(id) for every
// placeholder (4 id bytes in 0x10-0x1F, then 0x01).
template <typename FT, typename FS>
void walk_planned(const char *text, size_t n, size_t n_pending, FT fn_text, FS fn_site) {
    constexpr size_t kP = TemplateSet::kPlaceholderLen;
    size_t run = 0, i = 0;
    while (i < n) {
        const unsigned char c = (unsigned char)text[i];
        if (c >= 0x10 && c < 0x20 && i + kP <= n && text[i + 4] == '\x01') {
            size_t id = 0;
            bool ok = true;
            for (int k = 0; k < 4; ++k) {
                const unsigned char b = (unsigned char)text[i + (size_t)k];
                ok = ok && b >= 0x10 && b < 0x20;
                id = (id << 4) | (b & 0xF);
            }
            if (ok && id < n_pending) {
                if (i > run) fn_text(text + run, i - run);
                fn_site(id);
                i += kP;
                run = i;
                continue;
            }
        }
        ++i;
    }
    if (n > run) fn_text(text + run, n - run);
}

size_t expanded_size(const std::string &text, const std::vector<Pending> &pending) {
    size_t total = 0;
    walk_planned(text.data(), text.size(), pending.size(), [&](const char *, size_t n) { total += n; },
                 [&](size_t id) { total += pending[id].encoded_len; });
    return total;
}

// The final text: ordinary bytes copied, every placeholder replaced by its encoding, which
// enc_of(id) points at.  Returns the number of bytes written to dst.
template <typename F>
size_t expand_into(char *dst, const std::string &text, const std::vector<Pending> &pending, F enc_of) {
    char *q = dst;
    walk_planned(text.data(), text.size(), pending.size(),
                 [&](const char *p, size_t n) { memcpy(q, p, n); q += n; },
                 [&](size_t id) { memcpy(q, enc_of(id), pending[id].encoded_len); q += pending[id].encoded_len; });
    return (size_t)(q - dst);
}

// The buffer kv_fleet_template hands out.  A fleet's stream is tens of megabytes: fresh from
// malloc it arrives as untouched pages and the first write to each of them is a page fault (more
// time than the encode).  So the last buffer given back through kv_free is kept (one per process)
// and handed out again when it is big enough; kv_shutdown-less by design: it is ordinary memory.
std::mutex g_out_mu;
char *g_out_cached = nullptr;      // owned by the library while g_out_lent is false
size_t g_out_cap = 0;
bool g_out_lent = false;

char *out_buffer_get(size_t bytes) {
    {
        std::lock_guard<std::mutex> lk(g_out_mu);
        if (g_out_cached && !g_out_lent && g_out_cap >= bytes) {
            g_out_lent = true;
            return g_out_cached;
        }
        if (g_out_cached && !g_out_lent) {   // too small: replace it
            free(g_out_cached);
            g_out_cached = nullptr;
            g_out_cap = 0;
        }
    }
    const size_t cap = bytes + bytes / 8 + 4096;
    char *p = static_cast<char *>(malloc(cap));
    if (!p) return nullptr;
    std::lock_guard<std::mutex> lk(g_out_mu);
    if (!g_out_cached) {      // this one becomes the recycled buffer when it comes back
        g_out_cached = p;
        g_out_cap = cap;
        g_out_lent = true;
    }
    return p;
}

// true: p was the recycled buffer and is now parked instead of freed
bool out_buffer_put(void *p) {
    std::lock_guard<std::mutex> lk(g_out_mu);
    if (p && p == g_out_cached && g_out_lent) {
        if (g_out_cap > ((size_t)1 << 30)) {   // do not sit on more than 1 GiB
            g_out_cached = nullptr;
            g_out_cap = 0;
            g_out_lent = false;
            return false;
        }
        g_out_lent = false;
        return true;
    }
    return false;
}

int plan_release(const kv_chart *c, const char *release, const char *ns, const kv_overrides *ov, bool defer,
                 kv_release **out) {
    std::unique_ptr<kv_release> r(new kv_release);
    bool user_sentinels = false;
    Value vals = merged_values(c, ov, &user_sentinels);
    const bool can_defer = defer && !c->templates_have_sentinels && !c->defaults_have_sentinels && !user_sentinels;
    r->pending.reserve(4);
    static thread_local std::vector<std::string> rendered;   // per-file output, capacity kept per thread
    render_files(c, vals, release, ns, can_defer, &r->pending, &rendered, &r->keepalive);
    frame(c, rendered, &r->manifest, &r->notes);
    r->manifest_full_size = expanded_size(r->manifest, r->pending);
    r->filled.assign(r->pending.size(), std::string());
    *out = r.release();
    return KV_OK;
}

}  // namespace

extern "C" {

int kv_chart_load(const char *chart_dir, kv_chart **out) {
    if (!chart_dir || !out) return fail(KV_EINVAL, "null argument");
    try {
        std::unique_ptr<kv_chart> c(new kv_chart);
        c->chart = kvedge::chart::load_chart(chart_dir);
        const std::string prefix = c->chart.name() + "/";
        for (const auto &t : c->chart.templates) {
            c->ts.parse(prefix + t.first, t.second);
            if (TemplateSet::has_sentinel_bytes(t.second)) c->templates_have_sentinels = true;
        }
        c->ts.finalize();
        prepare_chart(c.get());
        *out = c.release();
        return KV_OK;
    } catch (const std::exception &e) {
        return fail(KV_ETEMPLATE, e.what());
    }
}

void kv_chart_free(kv_chart *c) { delete c; }

int kv_release_plan(const kv_chart *chart, const char *release_name, const char *ns,
                    const kv_overrides *overrides, kv_release **out) {
    if (!chart || !out) return fail(KV_EINVAL, "null argument");
    try {
        return plan_release(chart, release_name, ns, overrides, true, out);
    } catch (const std::exception &e) {
        return fail(KV_ETEMPLATE, e.what());
    }
}

int kv_release_pending_count(const kv_release *r) { return r ? (int)r->pending.size() : 0; }

int kv_release_pending_input(const kv_release *r, int i, const uint8_t **data, size_t *len) {
    if (!r || i < 0 || (size_t)i >= r->pending.size() || !data || !len) return fail(KV_EINVAL, "bad pending index");
    *data = reinterpret_cast<const uint8_t *>(r->pending[(size_t)i].data());
    *len = r->pending[(size_t)i].size();
    return KV_OK;
}

int kv_release_pending_fill(kv_release *r, int i, const char *enc, size_t len) {
    if (!r || i < 0 || (size_t)i >= r->pending.size() || !enc) return fail(KV_EINVAL, "bad pending index");
    if (len != r->pending[(size_t)i].encoded_len)
        return fail(KV_EINVAL, "encoded length " + std::to_string(len) + " != " +
                                   std::to_string(r->pending[(size_t)i].encoded_len));
    if (r->filled[(size_t)i].empty()) ++r->n_filled;
    r->filled[(size_t)i].assign(enc, len);
    r->expanded = false;
    return KV_OK;
}

int kv_release_encode_gpu(kv_release *r) {
    if (!r) return fail(KV_EINVAL, "null release");
    const size_t n = r->pending.size();
    if (n == 0) return KV_OK;
    std::vector<int64_t> in_off(n + 1, 0), out_off(n + 1, 0);
    for (size_t i = 0; i < n; ++i) {
        in_off[i + 1] = in_off[i] + (int64_t)r->pending[i].size();
        out_off[i + 1] = out_off[i] + (int64_t)((r->pending[i].encoded_len + 15) / 16 * 16);
    }
    std::string blob;
    blob.reserve((size_t)in_off[n]);
    for (size_t i = 0; i < n; ++i) blob.append(r->pending[i].data(), r->pending[i].size());
    std::string enc((size_t)out_off[n], '\0');
    int rc = kv_b64enc_batch(reinterpret_cast<const uint8_t *>(blob.data()), in_off.data(), (int64_t)n, &enc[0],
                             out_off.data());
    if (rc != KV_OK) return rc;
    for (size_t i = 0; i < n; ++i) r->filled[i].assign(enc, (size_t)out_off[i], r->pending[i].encoded_len);
    r->n_filled = n;
    r->expanded = false;
    return KV_OK;
}

static int finish(kv_release *r) {
    if (r->n_filled < r->pending.size())
        return fail(KV_EINVAL, std::to_string(r->pending.size() - r->n_filled) +
                                   " b64enc site(s) still pending: call kv_release_encode_gpu first");
    if (!r->expanded) {
        auto enc_of = [&](size_t id) { return r->filled[id].data(); };
        r->full_manifest.resize(r->manifest_full_size);
        expand_into(&r->full_manifest[0], r->manifest, r->pending, enc_of);
        r->full_notes.resize(expanded_size(r->notes, r->pending));
        expand_into(&r->full_notes[0], r->notes, r->pending, enc_of);
        r->expanded = true;
    }
    return KV_OK;
}

int kv_release_manifest(kv_release *r, const char **data, size_t *len) {
    if (!r || !data || !len) return fail(KV_EINVAL, "null argument");
    int rc = finish(r);
    if (rc) return rc;
    *data = r->full_manifest.data();
    *len = r->full_manifest.size();
    return KV_OK;
}

int kv_release_notes(kv_release *r, const char **data, size_t *len) {
    if (!r || !data || !len) return fail(KV_EINVAL, "null argument");
    int rc = finish(r);
    if (rc) return rc;
    *data = r->full_notes.data();
    *len = r->full_notes.size();
    return KV_OK;
}

void kv_release_free(kv_release *r) { delete r; }

int kv_helm_template(const char *chart_dir, const char *release_name, const char *ns,
                     const kv_overrides *overrides, char **out, size_t *out_len) {
    if (!out || !out_len) return fail(KV_EINVAL, "null argument");
    kv_chart *c = nullptr;
    int rc = kv_chart_load(chart_dir, &c);
    if (rc) return rc;
    kv_release *r = nullptr;
    rc = kv_release_plan(c, release_name, ns, overrides, &r);
    if (rc == KV_OK) rc = kv_release_encode_gpu(r);
    const char *data = nullptr;
    size_t n = 0;
    if (rc == KV_OK) rc = kv_release_manifest(r, &data, &n);
    if (rc == KV_OK) {
        *out = static_cast<char *>(malloc(n + 1));
        if (!*out) rc = fail(KV_ENOMEM, "malloc");
        else {
            memcpy(*out, data, n);
            (*out)[n] = 0;
            *out_len = n;
        }
    }
    kv_release_free(r);
    kv_chart_free(c);
    return rc;
}

int kv_fleet_template(const kv_chart *chart, int n_releases, const char *const *release_names, const char *ns,
                      const kv_overrides *overrides, char **out, int64_t *out_off) {
    if (!chart || n_releases < 0 || !out || !out_off) return fail(KV_EINVAL, "null argument");
    if (n_releases == 0) {   // nothing to render, nothing to encode
        *out = static_cast<char *>(calloc(1, 1));
        out_off[0] = 0;
        return *out ? KV_OK : fail(KV_ENOMEM, "malloc");
    }
#ifdef KVEDGE_TOOLS
    const bool trace = getenv("KVEDGE_TRACE") && atoi(getenv("KVEDGE_TRACE")) > 0;   // tools build only
#else
    constexpr bool trace = false;
#endif
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t_start = now();
    double t_prev = t_start;
    auto lap = [&](const char *what) {
        if (!trace) return;
        const double t = now();
        fprintf(stderr, "kv_fleet_template: %-28s %8.3f ms\n", what, t - t_prev);
        t_prev = t;
    };
    // Installs are independent (README.md:72).  One team of host threads carries the whole call
    // through its phases -- plan, pack, [one GPU batch], assemble -- and thread t keeps the same
    // installs in every phase, so what a thread allocates while planning it frees itself while
    // assembling (no cross-thread frees in the allocator) and no phase pays for new threads.
    std::vector<std::unique_ptr<kv_release>> rel((size_t)n_releases);
    unsigned n_thr = std::thread::hardware_concurrency();
    if (n_thr == 0) n_thr = 1;
    if (n_thr > 64) n_thr = 64;
    constexpr int kBlock = 8;   // installs per turn of a thread
    const int n_blocks = (n_releases + kBlock - 1) / kBlock;
    if ((int)n_thr > n_blocks) n_thr = (unsigned)(n_blocks > 0 ? n_blocks : 1);

    struct Barrier {
        std::mutex mu;
        std::condition_variable cv;
        unsigned waiting = 0, team;
        unsigned long generation = 0;
        explicit Barrier(unsigned n) : team(n) {}
        void wait() {
            std::unique_lock<std::mutex> lk(mu);
            const unsigned long gen = generation;
            if (++waiting == team) {
                waiting = 0;
                ++generation;
                cv.notify_all();
            } else {
                cv.wait(lk, [&] { return generation != gen; });
            }
        }
    } barrier(n_thr);

    std::atomic<int> first_rc{KV_OK};
    std::mutex err_mu;
    std::string err_msg;
    auto set_error = [&](int code, const std::string &msg) {
        std::lock_guard<std::mutex> lk(err_mu);
        if (first_rc.load() == KV_OK) {
            err_msg = msg;
            first_rc.store(code);
        }
    };
    // shared between the phases; written by thread 0 between barriers
    std::vector<size_t> man_off((size_t)n_releases + 1, 0), unit_base((size_t)n_releases + 1, 0);
    std::vector<int64_t> in_off, enc_off;
    char *stage_in = nullptr, *stage_out = nullptr, *buf = nullptr;
    size_t total_out = 0;
    static std::mutex stage_mu;   // the staging buffers are one per process
    std::unique_lock<std::mutex> stage_lock(stage_mu, std::defer_lock);

    constexpr int kChunkBlocks = 128;   // blocks (of kBlock installs) per GPU batch: ~1000 installs, ~14 MB of payload
    const int n_chunks = (n_blocks + kChunkBlocks - 1) / kChunkBlocks;
    std::vector<std::atomic<unsigned>> packed((size_t)(n_chunks > 0 ? n_chunks : 1));
    std::vector<std::atomic<int>> encoded((size_t)(n_chunks > 0 ? n_chunks : 1));
    for (auto &x : packed) x.store(0);
    for (auto &x : encoded) x.store(0);
    std::thread gpu_thread;
    auto gpu_driver = [&]() {
        for (int c = 0; c < n_chunks; ++c) {
            while (packed[(size_t)c].load(std::memory_order_acquire) < n_thr) std::this_thread::yield();
            const int i_lo = c * kChunkBlocks * kBlock, i_hi = std::min(n_releases, (c + 1) * kChunkBlocks * kBlock);
            const size_t u_lo = unit_base[(size_t)i_lo], u_hi = unit_base[(size_t)i_hi];
            if (u_hi > u_lo && first_rc.load() == KV_OK) {
                int rc = kv_b64enc_batch(reinterpret_cast<const uint8_t *>(stage_in), &in_off[u_lo], (int64_t)(u_hi - u_lo), stage_out,
                                         &enc_off[u_lo]);
                if (rc) set_error(rc, kv_last_error());
            }
            encoded[(size_t)c].store(1, std::memory_order_release);
        }
    };

    auto for_my_installs = [&](unsigned t, auto fn) {
        for (int b = (int)t; b < n_blocks; b += (int)n_thr)
            for (int i = b * kBlock; i < n_releases && i < (b + 1) * kBlock; ++i) fn(i);
    };
    auto team_member = [&](unsigned t) {
        // ---- plan
        for_my_installs(t, [&](int i) {
            if (first_rc.load() != KV_OK) return;
            try {
                kv_release *r = nullptr;
                int rc = plan_release(chart, release_names ? release_names[i] : nullptr, ns,
                                      overrides ? &overrides[i] : nullptr, true, &r);
                if (rc) throw kvedge::tmpl::Error(kv_last_error());
                rel[(size_t)i].reset(r);
            } catch (const std::exception &e) {
                set_error(KV_ETEMPLATE, e.what());
            }
        });
        barrier.wait();
        if (t == 0 && first_rc.load() == KV_OK) {
            lap("plan (all host threads)");
            // One packed batch over every pending Secret payload of the fleet.  The planned
            // manifests hold 5-byte placeholders and know their final size, so the output is laid
            // out before anything is encoded; afterwards every output byte is written exactly
            // once: ordinary text from the planned manifest, encodings from the staging blob.
            for (int i = 0; i < n_releases; ++i) {
                man_off[(size_t)i + 1] = man_off[(size_t)i] + rel[(size_t)i]->manifest_full_size;
                unit_base[(size_t)i + 1] = unit_base[(size_t)i] + rel[(size_t)i]->pending.size();
            }
            const size_t n_units = unit_base[(size_t)n_releases];
            in_off.assign(n_units + 1, 0);
            enc_off.assign(n_units + 1, 0);
            size_t u = 0;
            for (auto &r : rel)
                for (auto &p : r->pending) {
                    in_off[u + 1] = in_off[u] + (int64_t)p.size();
                    enc_off[u + 1] = enc_off[u] + (int64_t)((p.encoded_len + 15) / 16 * 16);
                    ++u;
                }
            total_out = man_off[(size_t)n_releases];
            stage_lock.lock();
            int rc = kv_stage_get_((size_t)in_off[n_units], (size_t)enc_off[n_units], &stage_in, &stage_out);
            if (rc) set_error(rc, kv_last_error());
            else if (!(buf = out_buffer_get(total_out + 1))) set_error(KV_ENOMEM, "malloc");
        }
        barrier.wait();
        // ---- pack | encode | assemble, pipelined over chunks of installs: while the GPU (driven by
        // one extra thread, which mostly sleeps in the driver) encodes chunk c, the team packs
        // chunk c + 1 and assembles the chunks that are back; every thread still touches only the
        // installs it planned, and frees them when it has assembled them
        if (first_rc.load() != KV_OK) {
            for_my_installs(t, [&](int i) { rel[(size_t)i].reset(); });
            return;
        }
        if (t == 0) gpu_thread = std::thread(gpu_driver);
        auto my_installs_of_chunk = [&](int c, auto fn) {
            const int lo = c * kChunkBlocks, hi = std::min(n_blocks, (c + 1) * kChunkBlocks);
            for (int b = lo + (int)((t + n_thr - (unsigned)(lo % (int)n_thr)) % n_thr); b < hi; b += (int)n_thr)
                for (int i = b * kBlock; i < n_releases && i < (b + 1) * kBlock; ++i) fn(i);
        };
        auto assemble = [&](int c) {
            const bool ok = first_rc.load() == KV_OK;
            my_installs_of_chunk(c, [&](int i) {
                if (ok) {
                    kv_release &r = *rel[(size_t)i];
                    const int64_t *eo = &enc_off[unit_base[(size_t)i]];
                    expand_into(buf + man_off[(size_t)i], r.manifest, r.pending, [&](size_t id) { return stage_out + eo[id]; });
                }
                rel[(size_t)i].reset();
            });
        };
        int next_asm = 0;
        for (int c = 0; c < n_chunks; ++c) {
            my_installs_of_chunk(c, [&](int i) {
                size_t u = unit_base[(size_t)i];
                for (auto &p : rel[(size_t)i]->pending) memcpy(stage_in + in_off[u++], p.data(), p.size());
            });
            packed[(size_t)c].fetch_add(1, std::memory_order_release);
            while (next_asm < c && encoded[(size_t)next_asm].load(std::memory_order_acquire)) assemble(next_asm++);
        }
        for (; next_asm < n_chunks; ++next_asm) {
            while (!encoded[(size_t)next_asm].load(std::memory_order_acquire)) std::this_thread::yield();
            assemble(next_asm);
        }
    };
    {
        std::vector<std::thread> team;
        for (unsigned t = 1; t < n_thr; ++t) team.emplace_back(team_member, t);
        team_member(0);
        for (auto &th : team) th.join();
        if (gpu_thread.joinable()) gpu_thread.join();
    }
    lap("pack | encode | assemble (pipelined)");
    if (first_rc.load() != KV_OK) {
        if (buf) kv_free(buf);
        return fail(first_rc.load(), err_msg);
    }
    for (int i = 0; i <= n_releases; ++i) out_off[i] = (int64_t)man_off[(size_t)i];
    buf[total_out] = 0;
    *out = buf;
    return KV_OK;
}

// ---- lint (BASELINE.json config 1) ------------------------------------------------------
static bool dns1123(const std::string &s) {
    if (s.empty() || s.size() > 253) return false;
    auto alnum = [](char c) { return (c >= 'a' && c <= 'z') || (c >= '0' && c <= '9'); };
    size_t i = 0;
    while (i < s.size()) {
        size_t j = s.find('.', i);
        if (j == std::string::npos) j = s.size();
        if (j == i || !alnum(s[i]) || !alnum(s[j - 1])) return false;
        for (size_t k = i; k < j; ++k)
            if (!alnum(s[k]) && s[k] != '-') return false;
        i = j + 1;
        if (j == s.size()) break;
        if (i == s.size()) return false;
    }
    return true;
}

static bool semver(const std::string &v) {
    size_t i = 0;
    if (i < v.size() && v[i] == 'v') ++i;
    for (int part = 0; part < 3; ++part) {
        size_t d0 = i;
        while (i < v.size() && isdigit((unsigned char)v[i])) ++i;
        if (i == d0) return false;
        if (part < 2) {
            if (i >= v.size() || v[i] != '.') return false;
            ++i;
        }
    }
    if (i == v.size()) return true;
    if (v[i] != '-' && v[i] != '+') return false;
    for (size_t k = i + 1; k < v.size(); ++k)
        if (!isalnum((unsigned char)v[k]) && v[k] != '.' && v[k] != '-' && v[k] != '+') return false;
    return i + 1 < v.size();
}

int kv_helm_lint(const char *chart_dir, const kv_overrides *overrides, int *ok, char **messages, size_t *len) {
    if (!chart_dir || !ok || !messages || !len) return fail(KV_EINVAL, "null argument");
    std::vector<std::string> msgs;
    kv_chart *c = nullptr;
    if (kv_chart_load(chart_dir, &c) != KV_OK) {
        msgs.push_back(std::string("[ERROR] Chart.yaml: ") + kv_last_error());
    } else {
        const Value &meta = c->chart.metadata;
        auto get = [&](const char *k) -> std::string {
            auto it = meta.map->find(k);
            if (it == meta.map->end() || it->second.is_nilish()) return "";
            return kvedge::tmpl::go_print(it->second);
        };
        if (get("name").empty()) msgs.push_back("[ERROR] Chart.yaml: name is required");
        std::string api = get("apiVersion");
        if (api.empty()) msgs.push_back("[ERROR] Chart.yaml: apiVersion is required. The value must be either \"v1\" or \"v2\"");
        else if (api != "v1" && api != "v2")
            msgs.push_back("[ERROR] Chart.yaml: apiVersion '" + api + "' is not valid. The value must be either \"v1\" or \"v2\"");
        std::string ver = get("version");
        if (ver.empty()) msgs.push_back("[ERROR] Chart.yaml: version is required");
        else if (!semver(ver)) msgs.push_back("[ERROR] Chart.yaml: version '" + ver + "' is not a valid SemVer");
        std::string type = get("type");
        if (!type.empty() && type != "application" && type != "library")
            msgs.push_back("[ERROR] Chart.yaml: chart type is not valid in apiVersion 'v2'. It must be 'application' or 'library'");
        if (get("icon").empty()) msgs.push_back("[INFO] Chart.yaml: icon is recommended");
        try {
            bool user_sentinels = false;
            Value vals = merged_values(c, overrides, &user_sentinels);
            std::vector<Pending> pending;
            const bool defer = !c->templates_have_sentinels && !c->defaults_have_sentinels && !user_sentinels;
            std::vector<std::string> rendered;
            Value keep;
            render_files(c, vals, "test-release", "default", defer, &pending, &rendered, &keep);
            std::map<std::string, std::string> files;
            for (size_t k = 0; k < rendered.size(); ++k) files[c->files[k].name] = rendered[k];
            std::vector<Manifest> ms;
            std::string notes;
            kvedge::chart::split_manifests(files, &ms, &notes);
            std::stable_sort(ms.begin(), ms.end(), [](const Manifest &a, const Manifest &b) { return a.name < b.name; });
            for (const auto &m : ms) {
                std::string shortn = m.name.substr(m.name.find('/') + 1);
                Value doc;
                try {
                    doc = kvedge::chart::parse_yaml(m.content, false);
                } catch (const std::exception &e) {
                    msgs.push_back("[ERROR] " + shortn + ": unable to parse YAML: " + e.what());
                    continue;
                }
                if (doc.kind != Value::MapK) {
                    msgs.push_back("[ERROR] " + shortn + ": document is not a mapping");
                    continue;
                }
                for (const char *req : {"apiVersion", "kind"})
                    if (!doc.map->count(req)) msgs.push_back("[ERROR] " + shortn + ": " + req + " is required");
                std::string nm;
                auto md = doc.map->find("metadata");
                if (md != doc.map->end() && md->second.kind == Value::MapK) {
                    auto ni = md->second.map->find("name");
                    if (ni != md->second.map->end() && !ni->second.is_nilish()) nm = kvedge::tmpl::go_print(ni->second);
                }
                if (nm.empty()) msgs.push_back("[ERROR] " + shortn + ": object name not present in metadata");
                else if (!dns1123(nm))
                    msgs.push_back("[ERROR] " + shortn + ": object name does not conform to Kubernetes naming requirements: \"" + nm + "\"");
            }
        } catch (const std::exception &e) {
            msgs.push_back(std::string("[ERROR] templates/: ") + e.what());
        }
        kv_chart_free(c);
    }
    *ok = 1;
    std::string joined;
    for (const auto &m : msgs) {
        if (m.compare(0, 7, "[ERROR]") == 0) *ok = 0;
        joined += m + "\n";
    }
    *messages = static_cast<char *>(malloc(joined.size() + 1));
    if (!*messages) return fail(KV_ENOMEM, "malloc");
    memcpy(*messages, joined.data(), joined.size());
    (*messages)[joined.size()] = 0;
    *len = joined.size();
    return KV_OK;
}

void kv_free(void *p) {
    if (!out_buffer_put(p)) free(p);
}

}  // extern "C"
