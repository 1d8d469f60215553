// Helm chart handling around the template engine: what Helm v3 does with
// deployment/helm/ of the reference before and after rendering (README.md:60;
// deployment/helm/Chart.yaml, values.yaml, .helmignore, templates/*).
#pragma once
#include <map>
#include <string>
#include <vector>

#include "tmpl.hpp"

namespace kvedge {
namespace chart {

using tmpl::Value;

// YAML subset sufficient for Chart.yaml / values.yaml: block mappings and sequences,
// flow [] {} collections, plain / quoted / block scalars, comments.  Plain scalars are
// typed like gopkg.in/yaml.v2 does; with json_numbers every number becomes a float64
// (sigs.k8s.io/yaml goes through JSON, which is how Helm reads values.yaml).
Value parse_yaml(const std::string &text, bool json_numbers);

// pkg/strvals: --set / --set-string expressions (README.md:60) and --set-file.
void apply_set(Value &values, const std::string &expr, bool force_string);
void apply_set_file(Value &values, const std::string &expr);                 // key=path
void apply_set_blob(Value &values, const std::string &key, std::string content);

// pkg/chartutil CoalesceValues: the chart's defaults slide UNDER the user's values (--set and
// friends, collected into a map of their own first).  A user value of null deletes the key
// when the chart has a default for it; tables merge key by key; lists and scalars of the user
// win whole (`--set ports[0].port=1` replaces the chart's list, it does not patch it).
Value coalesce_values(const Value &defaults, Value user);

struct Chart {
    std::string dir;
    Value metadata;                                // Chart.yaml as a map
    Value values;                                  // values.yaml as a map
    std::vector<std::pair<std::string, std::string>> templates;  // "templates/x.yaml" -> source
    std::string name() const;
};
Chart load_chart(const std::string &dir);          // honours .helmignore

struct Manifest {
    std::string name;      // "<chart>/templates/<file>"
    std::string kind;
    std::string content;
};

// pkg/releaseutil SortManifests + InstallOrder; NOTES.txt is returned apart.
void split_manifests(const std::map<std::string, std::string> &files, std::vector<Manifest> *out,
                     std::string *notes);

// The same, straight to the `helm template` stream ("---\n# Source: <name>\n<doc>\n" per document
// in install order) and the trimmed NOTES text.  `files`: (name, rendered text), sorted by name.
void frame_manifests(const std::vector<std::pair<const std::string *, const std::string *>> &files,
                     std::string *manifest, std::string *notes);

std::string trim_space(const std::string &s);

}  // namespace chart
}  // namespace kvedge
