// extender_codec.h — the wire formats either side of the dealer path: the JSON the
// kube-scheduler posts to /scheduler/filter and /scheduler/priorities, the JSON the extender
// answers with, and the name -> row resolution in between, sized for 100k+ candidate names.
//
// What it replaces in the reference (paths relative to /root/reference/pkg/):
//   routes/routes.go:47-53    json.NewDecoder(body).Decode(&extenderArgs)      -> DecodeExtenderArgs
//   scheduler/predicate.go:21-38  split names into canSchedule / canNotSchedule -> ExtenderRoutes::Predicate
//   routes/routes.go:75-87    json.Marshal(extenderFilterResult)               -> EncodeFilterResult
//   scheduler/priority.go:30-37 + routes.go:113-119  HostPriorityList + Marshal -> ExtenderRoutes::Prioritize
//   routes/routes.go:125-169 + scheduler/bind.go:25-84  bind verb: args, pod fetch, Dealer.Bind, result -> ExtenderRoutes::Bind
//   routes/routes.go:212-240  json.Marshal(nodeMaps) of the status route        -> Dealer::StatusJSON
//
// The structs are those of k8s.io/kube-scheduler v0.18.0 (go.mod:19), extender/v1/types.go, which
// is NOT vendored under /root/reference: ExtenderArgs{Pod, Nodes, NodeNames}, ExtenderFilterResult
// {Nodes, NodeNames, FailedNodes, Error}, HostPriority{Host, Score}, ExtenderBindingArgs{PodName,
// PodNamespace, PodUID, Node}, ExtenderBindingResult{Error} — no json tags, so the keys are the
// field names.  v1.Pod keys are those of k8s.io/api v0.18.0 (metadata.name/namespace/uid/labels/
// annotations, spec.nodeName, spec.containers[].name/.resources.limits).  Decoding and encoding
// restate Go 1.16 encoding/json: case-insensitive key match, unknown keys skipped, null leaves the
// zero value, strings unescaped with U+FFFD for bad input; on output map keys sorted, HTML-unsafe
// characters as < > &, control characters as \u00xx, U+2028/9 escaped.
// Parity status: UNPINNED against Go itself (no Go toolchain in this image) — pinned against a
// Python restatement of the same published rules (tests/test_extender_codec.py).
#pragma once
#include <deque>
#include <string_view>

#include "dealer_host.h"

namespace ngs_host {

struct ExtenderArgs {
    bool has_pod = false;  // "Pod" present and not null
    Pod pod;
    bool has_nodes = false;       // "Nodes" not null (never read: the extender is nodeCacheCapable)
    bool has_node_names = false;  // NodeNames != nil (routes.go:62)
    std::vector<std::string_view> node_names;  // views into the body, or into `unescaped`
    std::deque<std::string> unescaped;         // names that needed unescaping (stable addresses)
    std::vector<uint8_t> name_verbatim;        // 1: json.Marshal writes this name exactly as it stood in the body
    std::string_view node_names_raw;           // the array's bytes in the body: the candidate-set cache key
    bool node_names_cached = false;            // the array equalled `cached_node_names`: node_names was NOT filled
};

// json.NewDecoder(body).Decode(&args).  Returns "" or the decoder's error text ("EOF", "unexpected EOF",
// "invalid character 'x' looking for beginning of value", "json: cannot unmarshal number into Go struct
// field ExtenderArgs.NodeNames of type string", ...).  `out` may hold views into `body`.
// cached_node_names: the bytes of a NodeNames array tokenised before; when the body's array is the same
// bytes it is stepped over (one memcmp) and only out->node_names_cached is set.
Error DecodeExtenderArgs(std::string_view body, ExtenderArgs* out, const std::string* cached_node_names = nullptr);

struct BindingArgs {
    std::string PodName, PodNamespace, PodUID, Node;
};
Error DecodeBindingArgs(std::string_view body, BindingArgs* out);
std::string EncodeBindingResult(const std::string& error);  // {"Error":"..."}

// resource.Quantity: ParseQuantity + Value() (round up to an integer), apimachinery v0.18.0
bool ParseQuantityValue(std::string_view s, int64_t* out);

// encoding/json's string encoder with escapeHTML on (what json.Marshal uses)
void AppendJSONString(std::string* out, std::string_view s);

// json.Marshal(&ExtenderFilterResult{NodeNames: &can, FailedNodes: cannot, Error: ""}).  `failed` must be
// in ascending key order without duplicates (Go sorts map keys).
std::string EncodeFilterResult(const std::vector<std::string_view>& can,
                               const std::vector<std::pair<std::string_view, std::string>>& failed);
// the error shape of routes.go:54-69: {"Nodes":null,"NodeNames":null,"FailedNodes":null,"Error":"..."}
std::string EncodeFilterError(const std::string& error);
std::string EncodeHostPriorityList(const std::vector<std::string_view>& hosts, const std::vector<int>& scores);

// PredicateRoute + Predicate.Handler and PrioritizeRoute + Prioritize.Handler, body in -> body out.
// A candidate list whose bytes equal the previous request's is not tokenised or resolved again
// (kube-scheduler resends the same filtered list for the priorities call, and between pods when
// the cluster does not change), and a list that names every row in row order is handed to the C
// ABI as "all rows" (idx == NULL), so no index vector crosses PCIe.
class ExtenderRoutes {
  public:
    ExtenderRoutes(Dealer* d, PolicySpec spec, bool isLoadSchedule) : d_(d), spec_(std::move(spec)), load_(isLoadSchedule) {}
    std::string Predicate(std::string_view body);                     // always a 200 body (routes.go:75-87)
    std::string Prioritize(std::string_view body, bool* panicked);    // *panicked: the reference panics (routes.go:103,110)
    // BindRoute + Bind.Handler + NewNanoGPUBind (routes.go:125-169, scheduler/bind.go:25-59): the pod is fetched from
    // the ClusterSource, checked (uid, not completed) and bound through Dealer.Bind; *http_status = 200 or 500
    std::string Bind(std::string_view body, int* http_status);
    struct Counters {
        uint64_t requests = 0, cache_hits = 0, all_rows = 0;
        uint64_t ns_decode = 0, ns_resolve = 0, ns_dealer = 0, ns_encode = 0;  // where the wall time of the routes went
    };
    const Counters& counters() const { return c_; }

  private:
    std::mutex mu_;  // the candidate-set cache is shared state: requests are served one at a time
    Dealer* d_;
    PolicySpec spec_;
    bool load_;
    Counters c_;
    // candidate-set cache
    std::string raw_;                        // copy of the NodeNames array bytes
    std::string names_store_, tok_store_;    // names that needed unescaping / their re-escaped tokens
    std::vector<std::string_view> names_;    // views into raw_ (or names_store_)
    std::vector<std::string_view> tok_;      // each name as json.Marshal writes it, quotes included: views into raw_ (or tok_store_)
    size_t tok_bytes_ = 0;
    std::vector<uint32_t> order_;            // indices of names_ in ascending name order
    CandidateSet set_;
    bool valid_ = false;
    bool prepare(const ExtenderArgs& a);     // fills / reuses the cache; false when it had to rebuild
};

}  // namespace ngs_host
